"""Few-row linear layers of the decode step (csrc/gxa_skinny.cu) through the C ABI, against fp64 math on the same bf16
operands: every epilogue, ragged row counts (1 .. 256), a partial 128-column weight tile, every cluster size the host
picks (K = 128 .. 4096)."""
import pytest
import torch

from mocov2_whisper_flamingo_b200 import _lib
from gpu_util import ptr, rel_err, stream

pytestmark = pytest.mark.gpu


def _run(R, N, K, epi, seed=0):
    L = _lib.lib()
    g = torch.Generator(device="cuda").manual_seed(seed)
    X = torch.randn(R, K, device="cuda", generator=g).to(torch.bfloat16)
    W = (torch.randn(N, K, device="cuda", generator=g) * K ** -0.5).to(torch.bfloat16)
    bias = torch.rand(N, device="cuda", generator=g) - 0.5
    res = torch.randn(R, N, device="cuda", generator=g)
    gate = torch.tensor([0.7], device="cuda")
    out = torch.full((R, N), float("nan"), device="cuda", dtype=torch.float32 if epi == 2 else torch.bfloat16)
    _lib.check(L.gxa_linear_few_rows_bf16(ptr(X), K, ptr(W), K, R, N, K, epi, ptr(out), N, ptr(bias), ptr(res), N,
                                          ptr(gate), stream()), "gxa_linear_few_rows_bf16")
    torch.cuda.synchronize()
    t = X.double() @ W.double().t() + bias.double()
    if epi == 1:
        u = t.float().to(torch.bfloat16).double()
        t = 0.5 * u * (1 + torch.erf(u / 2 ** 0.5))
    elif epi == 2:
        t = res.double() + torch.tanh(gate.double()) * t
    return out, t


@pytest.mark.parametrize("R,N,K", [(1, 128, 128), (5, 384, 384), (32, 768, 768), (160, 768, 768), (160, 3072, 768),
                                   (160, 768, 3072), (256, 1024, 4096), (77, 320, 640), (200, 1280, 1280)])
@pytest.mark.parametrize("epi", [0, 1, 2])
def test_few_rows_linear_matches_fp64(R, N, K, epi):
    out, ref = _run(R, N, K, epi)
    assert torch.isfinite(out.float()).all()
    tol = 2e-5 if epi == 2 else 6e-3          # fp32 output: accumulation order only; bf16 outputs: one rounding (2^-8)
    assert rel_err(out.float(), ref) < tol


def test_few_rows_linear_rejects_unsupported_shapes():
    L = _lib.lib()
    X = torch.zeros(300, 128, device="cuda", dtype=torch.bfloat16)
    W = torch.zeros(128, 128, device="cuda", dtype=torch.bfloat16)
    out = torch.zeros(300, 128, device="cuda", dtype=torch.bfloat16)
    assert L.gxa_linear_few_rows_bf16(ptr(X), 128, ptr(W), 128, 300, 128, 128, 0, ptr(out), 128, None, None, 0, None,
                                      stream()) == -1          # more than 256 rows
    assert L.gxa_linear_few_rows_bf16(ptr(X), 64, ptr(W), 64, 16, 128, 64, 0, ptr(out), 128, None, None, 0, None,
                                      stream()) == -1          # K < 128
