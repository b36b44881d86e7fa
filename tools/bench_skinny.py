"""Few-row linear layers of the decode step timed alone (CUDA events, back-to-back launches of one kernel; PDL off so
that launches do not overlap): gxa_linear_few_rows_bf16 vs the 128-row-tile GEMM on the same operands."""
import os
import sys

os.environ.setdefault("GXA_PDL", "0")
import torch

sys.path.insert(0, __file__.rsplit("/tools/", 1)[0])
sys.path.insert(0, __file__.rsplit("/tools/", 1)[0] + "/tests")
from mocov2_whisper_flamingo_b200 import _lib  # noqa: E402
from gpu_util import ptr, stream  # noqa: E402

L = _lib.lib()
R = int(sys.argv[1]) if len(sys.argv) > 1 else 160


def timeit(fn, n=200, warm=20):
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


for name, N, K, epi in (("Q / out-proj", 768, 768, 0), ("out-proj gated", 768, 768, 2), ("ff.0 + GELU", 3072, 768, 1),
                        ("ff.2 gated", 768, 3072, 2)):
    X = torch.randn(R, K, device="cuda").to(torch.bfloat16)
    W = torch.randn(N, K, device="cuda").to(torch.bfloat16)
    bias = torch.randn(N, device="cuda")
    res = torch.randn(R, N, device="cuda")
    gate = torch.tensor([0.5], device="cuda")
    out = torch.empty(R, N, device="cuda", dtype=torch.float32 if epi == 2 else torch.bfloat16)
    out2 = torch.empty(R, N, device="cuda", dtype=torch.bfloat16)
    t_s = timeit(lambda: _lib.check(L.gxa_linear_few_rows_bf16(ptr(X), K, ptr(W), K, R, N, K, epi, ptr(out), N, ptr(bias),
                                                                 ptr(res), N, ptr(gate), stream()), "skinny"))
    # the tile GEMM with the closest epilogue (0 bf16 / 3 bias+GELU / 2 gated residual)
    tepi = {0: 0, 1: 3, 2: 2}[epi]
    outg = torch.empty(R, N, device="cuda", dtype=torch.float32 if epi == 2 else torch.bfloat16)
    t_g = timeit(lambda: _lib.check(L.gxa_gemm_bf16(ptr(X), K, 0, ptr(W), K, 0, R, N, K, tepi, ptr(outg), N, ptr(out2), N,
                                                     ptr(bias), ptr(res) if epi == 2 else None, N, ptr(gate), 0, 0, 0,
                                                     stream()), "gemm"))
    wbytes = N * K * 2
    print(f"{name:16s} R={R} N={N} K={K}: few-rows kernel {t_s:6.1f} us ({wbytes / t_s / 1e3:6.0f} GB/s of weights)   "
          f"tile GEMM {t_g:6.1f} us")
