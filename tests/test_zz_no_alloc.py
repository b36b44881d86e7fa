"""gxa_linear_fwd_static: the allocation-free variant of gxa_linear_fwd (scratch in a static __device__ array inside the
library, csrc/gxa_noalloc.cu).  No counterpart in the reference.

Named test_zz_* so that pytest collects it LAST (a failure here must not hide any
other test).  It passes alone in a fresh process and after the rest of the suite: every buffer is allocated before the
calls under test, both entry points are called through the C ABI (ctypes), and nothing depends on earlier tests.
"""
import pytest
import torch

from mocov2_whisper_flamingo_b200 import _lib

pytestmark = pytest.mark.gpu

GXA_PREC_BF16 = _lib.GXA_PREC_BF16
TOL_VS_EXISTING = 1e-4      # the suite's fp32 tolerance: the two entry points run the same kernels on the same operands
TOL_VS_FP32_MATH = 2e-2     # the suite's bf16 tolerance (north_star), against x W^T + b in fp32


def _rel(a, b):
    den = float(b.abs().max())
    return float((a.double() - b.double()).abs().max()) / (den if den > 0 else 1.0)


LIMIT = 8 << 20             # GXA_LINEAR_FWD_STATIC_MAX_BYTES (include/gxa.h)


def _need(rows, in_f, out_f):
    """bytes of the built-in array a shape uses (include/gxa.h): the two bf16 operand copies"""
    return 2 * (((rows * in_f + 127) // 128) * 128 + out_f * in_f)


def _case(rows, in_f, out_f, seed):
    g = torch.Generator(device="cuda").manual_seed(seed)
    x = torch.randn(rows, in_f, device="cuda", generator=g)
    w = torch.randn(out_f, in_f, device="cuda", generator=g) * in_f ** -0.5
    b = torch.rand(out_f, device="cuda", generator=g) * 0.2 - 0.1
    return x, w, b


def test_static_scratch_linear_matches_workspace_linear_on_two_calls_in_a_row():
    L = _lib.lib()
    s = torch.cuda.current_stream().cuda_stream
    # call 1 (also the per-device set-up) uses half of the array (CTA-pair GEMM kernel); call 2 is smaller and ragged
    # (1-CTA kernel) and runs on top of call 1's left-overs; call 3 repeats call 1's shape with new data
    shapes = [(1024, 1024, 1024), (300, 512, 768), (1024, 1024, 1024)]
    cases = [_case(*shp, seed=100 + i) for i, shp in enumerate(shapes)]
    outs_static = [torch.empty(r, o, device="cuda") for r, _, o in shapes]
    outs_ws = [torch.empty(r, o, device="cuda") for r, _, o in shapes]
    ws = [torch.empty(L.gxa_linear_workspace_bytes(r, i, o, GXA_PREC_BF16), dtype=torch.uint8, device="cuda")
          for r, i, o in shapes]
    for (r, i, o) in shapes:
        assert 0 < _need(r, i, o) <= LIMIT

    # the existing entry point first (also loads every kernel both entry points use)
    for (x, w, b), y, wk, (r, i, o) in zip(cases, outs_ws, ws, shapes):
        _lib.check(L.gxa_linear_fwd(x.data_ptr(), w.data_ptr(), b.data_ptr(), y.data_ptr(), r, i, o, GXA_PREC_BF16,
                                    wk.data_ptr(), s), "gxa_linear_fwd")
    torch.cuda.synchronize()

    # the calls under test, back to back on one stream.  The first one is the set-up of the device (symbol address; with
    # lazy module loading the driver materialises the 8 MiB array here), so the "nothing is allocated" check brackets
    # the calls after it.
    torch_before = free_before = None
    for n, ((x, w, b), y, (r, i, o)) in enumerate(zip(cases, outs_static, shapes)):
        if n == 1:
            torch.cuda.synchronize()
            torch_before = torch.cuda.memory_allocated()
            free_before, _ = torch.cuda.mem_get_info()
        code = L.gxa_linear_fwd_static(x.data_ptr(), w.data_ptr(), b.data_ptr(), y.data_ptr(), r, i, o, s)
        assert code == 0, (code, L.gxa_last_cuda_error().decode())
    torch.cuda.synchronize()
    free_after, _ = torch.cuda.mem_get_info()
    assert torch.cuda.memory_allocated() == torch_before
    assert free_after >= free_before, "device memory was allocated inside gxa_linear_fwd_static"

    for n, ((x, w, b), ys, yw) in enumerate(zip(cases, outs_static, outs_ws)):
        assert torch.isfinite(ys).all()
        e = _rel(ys, yw)
        assert e <= TOL_VS_EXISTING, (n, e)
        e32 = _rel(ys, x @ w.t() + b)
        assert e32 <= TOL_VS_FP32_MATH, (n, e32)


def test_static_scratch_linear_rejects_what_does_not_fit():
    L = _lib.lib()
    rows, in_f, out_f = 4096, 1024, 1024                    # 2 * (4 Mi + 1 Mi) bytes = 10 MiB > 8 MiB
    assert _need(rows, in_f, out_f) > LIMIT
    x = torch.zeros(rows, in_f, device="cuda")
    w = torch.zeros(out_f, in_f, device="cuda")
    y = torch.full((rows, out_f), 7.0, device="cuda")
    code = L.gxa_linear_fwd_static(x.data_ptr(), w.data_ptr(), None, y.data_ptr(), rows, in_f, out_f,
                                   torch.cuda.current_stream().cuda_stream)
    assert code == -1                                        # GXA_ERR_SHAPE
    msg = L.gxa_last_cuda_error().decode()
    assert "gxa_linear_fwd_static" in msg and str(8 << 20) in msg, msg
    torch.cuda.synchronize()
    assert bool((y == 7.0).all())                            # nothing was launched
    assert L.gxa_linear_fwd_static(None, w.data_ptr(), None, y.data_ptr(), 64, 64, 64, None) == -7   # GXA_ERR_ARG
