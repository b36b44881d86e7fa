"""CUDA-event timing of every GEMM of one c3 layer (fwd + bwd), each timed alone (not a bench line).
    python tests/bench_gemms.py            # BB=64 by default
"""
import os
import sys

import torch

sys.path.insert(0, __file__.rsplit("/tools/", 1)[0])
sys.path.insert(0, __file__.rsplit("/tools/", 1)[0] + "/tests")
from mocov2_whisper_flamingo_b200 import _lib  # noqa: E402
from gpu_util import ptr, stream  # noqa: E402

L = _lib.lib()
if os.environ.get("SM_LIMIT"):        # experiment: persistent grids sized for fewer SMs (is a GEMM chip-bound or SM-bound?)
    _lib.set_sm_limit(int(os.environ["SM_LIMIT"]))
B, Lq, Lk, D = int(os.environ.get("BB", 64)), 448, 750, int(os.environ.get("DD", 1024))
M, Mk = B * Lq, B * Lk
E = _lib


def timeit(fn, n=10, warm=3):
    for _ in range(warm):
        fn()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    torch.cuda.synchronize()
    e0.record()
    for _ in range(n):
        fn()
    e1.record()
    torch.cuda.synchronize()
    return e0.elapsed_time(e1) / n * 1e3


def bf(*s):
    return torch.randn(*s, device="cuda").to(torch.bfloat16)


# name, (M, N, K), a_mn, b_mn, epi
CASES = [
    ("fwd  Q      ", (M, D, D), 0, 0, E.EPI_BF16),
    ("fwd  KV     ", (Mk, 2 * D, D), 0, 0, E.EPI_BF16),
    ("fwd  out+gate", (M, D, D), 0, 0, E.EPI_GATE_RES),
    ("fwd  FFN1+gelu", (M, 4 * D, D), 0, 0, E.EPI_BIAS_GELU),
    ("fwd  FFN2+gate", (M, D, 4 * D), 0, 0, E.EPI_GATE_RES),
    ("dgrad FFN2+gelu'", (M, 4 * D, D), 0, 1, E.EPI_GELU_BWD),
    ("wgrad FFN2  ", (D, 4 * D, M), 1, 1, E.EPI_F32),
    ("dgrad FFN1  ", (M, D, 4 * D), 0, 1, E.EPI_BF16),
    ("wgrad FFN1  ", (4 * D, D, M), 1, 1, E.EPI_F32),
    ("dgrad out   ", (M, D, D), 0, 1, E.EPI_BF16),
    ("wgrad out   ", (D, D, M), 1, 1, E.EPI_F32),
    ("wgrad Q     ", (D, D, M), 1, 1, E.EPI_F32),
    ("wgrad KV    ", (2 * D, D, Mk), 1, 1, E.EPI_F32),
    ("dgrad Q     ", (M, D, D), 0, 1, E.EPI_BF16),
    ("dgrad KV->dxa", (Mk, D, 2 * D), 0, 1, E.EPI_F32),
]
tot_t, tot_f = 0.0, 0.0
for name, (m, n, k), a_mn, b_mn, epi in CASES:
    A = bf(k, m) if a_mn else bf(m, k)
    Bm = bf(k, n) if b_mn else bf(n, k)
    out_f32 = epi in (E.EPI_F32, E.EPI_GATE_RES)
    Cm = torch.zeros(m, n, device="cuda", dtype=torch.float32 if out_f32 else torch.bfloat16)
    C2 = torch.empty(m, n, device="cuda", dtype=torch.bfloat16) if epi in (E.EPI_GATE_RES, E.EPI_BIAS_GELU) else None
    bias = torch.randn(n, device="cuda") if epi in (E.EPI_BF16, E.EPI_GATE_RES, E.EPI_BIAS_GELU) and not b_mn else None
    aux = None
    if epi == E.EPI_GATE_RES:
        aux = torch.randn(m, n, device="cuda")
    elif epi == E.EPI_GELU_BWD:
        aux = bf(m, n)
    gate = torch.full((1,), 0.7, device="cuda") if epi == E.EPI_GATE_RES else None
    acc = 1 if epi == E.EPI_F32 else 0

    def run():
        _lib.check(L.gxa_gemm_bf16(ptr(A), A.shape[1], a_mn, ptr(Bm), Bm.shape[1], b_mn, m, n, k, epi, ptr(Cm), n, ptr(C2), n,
                                   ptr(bias), ptr(aux), n if aux is not None else 0, ptr(gate), acc, 0, 0, stream()), name)
    t = timeit(run)
    fl = 2.0 * m * n * k
    tot_t += t
    tot_f += fl
    print(f"{name:18s} M={m:6d} N={n:5d} K={k:6d} a_mn={a_mn} b_mn={b_mn}  {t:8.1f} us  {fl / t / 1e6:7.1f} TFLOP/s")
    del A, Bm, Cm, C2, aux
print(f"all GEMMs of one layer: {tot_t:8.1f} us  {tot_f / tot_t / 1e6:7.1f} TFLOP/s")
