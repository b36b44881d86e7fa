"""A few launches of the attention forward / backward kernels at the BASELINE config-3 shape WITH training dropout (the
bench's configuration), for `ncu --set full` captures and CUDA-event timing.  Not a bench line.
    python tools/prof_attn.py [B] [p_drop]
"""
import sys

import torch

sys.path.insert(0, __file__.rsplit("/tools/", 1)[0])
sys.path.insert(0, __file__.rsplit("/tools/", 1)[0] + "/tests")
from mocov2_whisper_flamingo_b200 import _lib  # noqa: E402
from gpu_util import ptr, stream  # noqa: E402

L = _lib.lib()
B = int(sys.argv[1]) if len(sys.argv) > 1 else 64
p = float(sys.argv[2]) if len(sys.argv) > 2 else 0.1
H, Lq, Lk = 16, 448, 750
D = 64 * H
g = torch.Generator(device="cuda").manual_seed(0)
q = torch.randn(B, Lq, D, device="cuda", generator=g).to(torch.bfloat16)
kv = torch.randn(B, Lk, 2 * D, device="cuda", generator=g).to(torch.bfloat16)
dctx = torch.randn(B, Lq, D, device="cuda", generator=g).to(torch.bfloat16)
ctx = torch.empty_like(q)
lse = torch.empty(B, H, Lq, device="cuda")
delta = torch.empty_like(lse)
wsb = torch.empty(L.gxa_attn_bwd_workspace_bytes(B, H, Lq), dtype=torch.uint8, device="cuda")
dq, dkv = torch.empty_like(q), torch.empty_like(kv)
mask = torch.zeros(B, Lk, dtype=torch.uint8, device="cuda")
kvend = torch.full((B,), Lk, dtype=torch.int32, device="cuda")
k, v = kv[..., :D], kv[..., D:]
seed = 0x1234ABCD5678EF01


def fwd():
    _lib.check(L.gxa_attn_fwd_dropout(ptr(q), D, ptr(k), ptr(v), 2 * D, ptr(mask), ptr(kvend), ptr(ctx), D, ptr(lse), B, H,
                                      Lq, Lk, p, seed, stream()), "fwd")


def bwd():
    _lib.check(L.gxa_attn_bwd_dropout(ptr(q), D, ptr(k), ptr(v), 2 * D, ptr(mask), ptr(kvend), ptr(ctx), ptr(dctx), D,
                                      ptr(lse), ptr(delta), ptr(dq), D, ptr(dkv), ptr(dkv[..., D:]), 2 * D, ptr(wsb), B, H,
                                      Lq, Lk, p, seed, stream()), "bwd")


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


fl = 4.0 * B * Lq * Lk * D
t = timeit(fwd)
print(f"attn fwd  B={B} p={p}: {t:8.1f} us  {fl / t / 1e6:7.1f} TFLOP/s (algorithmic 4*B*Lq*Lk*D)")
t = timeit(bwd)
print(f"attn bwd  B={B} p={p}: {t:8.1f} us  {2.5 * fl / t / 1e6:7.1f} TFLOP/s (algorithmic 10*B*Lq*Lk*D; includes delta + dq finalize + memset)")
