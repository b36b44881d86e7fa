"""CUDA-event timings of individual kernels at the BASELINE config-3 per-GPU shape (not a bench line)."""
import os
import sys

import torch

sys.path.insert(0, __file__.rsplit("/tools/", 1)[0])
sys.path.insert(0, __file__.rsplit("/tools/", 1)[0] + "/tests")
from mocov2_whisper_flamingo_b200 import _lib  # noqa: E402
from gpu_util import ptr, stream  # noqa: E402

L = _lib.lib()
B, H, Lq, Lk = int(os.environ.get("BB", 64)), 16, 448, 750
D = 64 * H


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


g = torch.Generator(device="cuda").manual_seed(0)
q = torch.randn(B, Lq, D, device="cuda", generator=g).to(torch.bfloat16)
kv = torch.randn(B, Lk, 2 * D, device="cuda", generator=g).to(torch.bfloat16)
dctx = torch.randn(B, Lq, D, device="cuda", generator=g).to(torch.bfloat16)
ctx = torch.empty_like(q)
lse = torch.empty(B, H, Lq, device="cuda")
delta = torch.empty_like(lse)
wsb = torch.empty(L.gxa_attn_bwd_workspace_bytes(B, H, Lq), dtype=torch.uint8, device="cuda")
dq = torch.empty_like(q)
dkv = torch.empty_like(kv)
mask = torch.zeros(B, Lk, dtype=torch.uint8, device="cuda")
kvend = torch.full((B,), Lk, dtype=torch.int32, device="cuda")
k, v = kv[..., :D], kv[..., D:]
fl = 4.0 * B * Lq * Lk * D


def fwd():
    _lib.check(L.gxa_attn_fwd(ptr(q), D, ptr(k), ptr(v), 2 * D, ptr(mask), ptr(kvend), ptr(ctx), D, ptr(lse), B, H, Lq, Lk,
                              stream()), "fwd")


def bwd():
    _lib.check(L.gxa_attn_bwd(ptr(q), D, ptr(k), ptr(v), 2 * D, ptr(mask), ptr(kvend), ptr(ctx), ptr(dctx), D, ptr(lse),
                              ptr(delta), ptr(dq), D, ptr(dkv), ptr(dkv[..., D:]), 2 * D, ptr(wsb), B, H, Lq, Lk, stream()), "bwd")


t = timeit(fwd)
print(f"attn fwd [{os.environ.get('GXA_ATTN', 'tc')}]: {t:8.1f} us  {fl / t / 1e6:7.1f} TFLOP/s (algorithmic 4*B*Lq*Lk*D)")
t = timeit(bwd)
print(f"attn bwd [{os.environ.get('GXA_ATTN', 'tc')}]: {t:8.1f} us  {2.5 * fl / t / 1e6:7.1f} TFLOP/s (algorithmic 10*B*Lq*Lk*D)")
