"""Achieved HBM GB/s / TFLOP/s of the statements either side of the fusion stack in AVNet, each timed
alone with CUDA events: LayerNorm with the positional add folded in, the 3-way sum, and the fat-K 2048 -> d video
projection.  Not a bench line.   python tools/bench_glue.py [B] [T] [d_model]
"""
import json
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import mocov2_whisper_flamingo_b200 as G  # noqa: E402
from mocov2_whisper_flamingo_b200 import _lib, av_glue  # noqa: E402
from mocov2_whisper_flamingo_b200.functional import workspace  # noqa: E402

B = int(sys.argv[1]) if len(sys.argv) > 1 else 64
T = int(sys.argv[2]) if len(sys.argv) > 2 else 400          # reference max_frames (config.py:28)
D = int(sys.argv[3]) if len(sys.argv) > 3 else 512          # reference d_model (config.py:12)
L = _lib.lib()
s = torch.cuda.current_stream().cuda_stream
try:
    HBM = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))["hbm_gbs"]
    src = "measured"
except Exception:
    HBM, src = 6650.0, "fallback"


def timeit(fn, n=50, warm=5):
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


rows = B * T
x = torch.randn(rows, D, device="cuda")
w, b = torch.randn(D, device="cuda"), torch.randn(D, device="cuda")
pe = av_glue.positional_table(D, device="cuda")[0, :T].contiguous()
out = torch.empty_like(x)
mean, rstd = torch.empty(rows, device="cuda"), torch.empty(rows, device="cuda")
flush = torch.empty(256 << 20, dtype=torch.uint8, device="cuda")      # > L2: the inputs below are small enough to sit in it


def ln_pe():
    flush.zero_()
    _lib.check(L.gxa_layernorm_pe_fwd(x.data_ptr(), w.data_ptr(), b.data_ptr(), pe.data_ptr(), T, out.data_ptr(),
                                      mean.data_ptr(), rstd.data_ptr(), rows, D, s), "ln_pe")


def only_flush():
    flush.zero_()


t_flush = timeit(only_flush)
t = timeit(ln_pe) - t_flush
byt = rows * D * 8 + T * D * 4 + rows * 8
print(f"LayerNorm + pe   rows={rows} D={D}: {t:7.1f} us  {byt / t / 1e3:7.0f} GB/s = {byt / t / 1e3 / HBM:.2f} of {src} HBM copy peak")
a, c = torch.randn_like(x), torch.randn_like(x)


def sum3():
    flush.zero_()
    _lib.check(L.gxa_sum3_f32(x.data_ptr(), a.data_ptr(), c.data_ptr(), out.data_ptr(), x.numel(), s), "sum3")


t = timeit(sum3) - t_flush
byt = rows * D * 16
print(f"3-way sum        rows={rows} D={D}: {t:7.1f} us  {byt / t / 1e3:7.0f} GB/s = {byt / t / 1e3 / HBM:.2f} of {src} HBM copy peak")
FV = 2048
xv = torch.randn(rows, FV, device="cuda")
wv, bv = torch.randn(D, FV, device="cuda") / 45, torch.randn(D, device="cuda")
yv = torch.empty(rows, D, device="cuda")
for prec, name in ((_lib.GXA_PREC_BF16, "bf16"), (_lib.GXA_PREC_FP32, "fp32 (3xbf16)")):
    ws = workspace(L.gxa_linear_workspace_bytes(rows, FV, D, prec), x.device)

    def proj():
        _lib.check(L.gxa_linear_fwd(xv.data_ptr(), wv.data_ptr(), bv.data_ptr(), yv.data_ptr(), rows, FV, D, prec,
                                    ws.data_ptr(), s), "linear")
    t = timeit(proj, n=20)
    print(f"video_proj {FV}->{D} rows={rows} {name}: {t:7.1f} us  {2.0 * rows * FV * D / t / 1e6:7.1f} TFLOP/s "
          f"(includes the fp32 -> bf16 operand casts; {rows * FV * 4 / t / 1e3:.0f} GB/s of fp32 input)")
