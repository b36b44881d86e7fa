"""Column sums at the BASELINE row counts (config 3: 64 x 750 = 48000 visual rows, 64 x 448 = 28672 query rows) through
the public linear backward (``gxa_linear_bwd``: d bias = column sums of dy, fp32).  The small parity shapes never give a
thread more than a dozen rows; here every thread of the column-sum kernel walks its cp.async ring many times.  The
expected value is exact arithmetic (a float64 sum), so the tolerance is fp32 summation error only."""
import pytest
import torch

import mocov2_whisper_flamingo_b200 as G  # noqa: F401  (fails loudly if the library is missing)
from mocov2_whisper_flamingo_b200._lib import GXA_PREC_BF16
from mocov2_whisper_flamingo_b200.functional import LinearFn

pytestmark = pytest.mark.gpu


@pytest.mark.parametrize("B,L,D", [(64, 750, 1024), (64, 448, 1024), (4, 250, 768)])
def test_linear_bias_gradient_is_the_column_sum(B, L, D):
    g = torch.Generator(device="cuda").manual_seed(5)
    x = torch.randn(B, L, D, device="cuda", generator=g)
    w = (torch.randn(D, D, device="cuda", generator=g) / 32).requires_grad_(True)
    b = torch.zeros(D, device="cuda", requires_grad=True)
    y = LinearFn.apply(x, w, b, GXA_PREC_BF16)
    dy = torch.randn(B, L, D, device="cuda", generator=g) + 0.25      # non-zero mean: the sums grow with the row count
    y.backward(dy)
    torch.cuda.synchronize()
    ref = dy.double().sum((0, 1))
    err = float((b.grad.double() - ref).abs().max()) / float(ref.abs().max())
    assert err < 1e-4, err
    dw_ref = dy.reshape(-1, D).to(torch.bfloat16).float().t() @ x.reshape(-1, D).to(torch.bfloat16).float()
    err_w = float((w.grad - dw_ref).abs().max()) / float(dw_ref.abs().max())
    assert err_w < 2e-2, err_w
