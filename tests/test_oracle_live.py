"""The oracle against the LIVE reference module, at the head counts / widths of every BASELINE config
(Whisper-tiny 384/6, small 768/12, medium 1024/16, large-v2 1280/20) -- shapes whose parameter gradients are too large
to commit as fixtures (12 d^2 floats).  Runs only where /root/reference exists (the build container); on the GPU box the
committed fixtures in tests/golden/ (made by oracle/make_golden.py from the same module) stand in.

Checks, per width: the float64 restatement (oracle/gxa_oracle.py) and the torch CPU port (oracle/torch_port.py) against
reference forward + autograd backward on identical seeded inputs with variable-length padding masks (one sample fully
padded), nonzero gates, non-trivial biases / LayerNorm affine.
"""
from __future__ import annotations

import os
import sys

import numpy as np
import pytest
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)
REF = "/root/reference"
pytestmark = pytest.mark.skipif(not os.path.exists(os.path.join(REF, "models", "gate_cross_attention.py")),
                                reason="live reference not present (GPU box): tests/golden fixtures cover the oracle")

from oracle import gxa_oracle as O  # noqa: E402
from oracle import torch_port as TP  # noqa: E402

WIDTHS = {"tiny": (384, 6), "small": (768, 12), "medium": (1024, 16), "large-v2": (1280, 20)}


def _reference_block(D, H):
    if REF not in sys.path:
        sys.path.insert(0, REF)
    from models.gate_cross_attention import GatedCrossAttentionBlock
    return GatedCrossAttentionBlock(D, H, dropout=0.0)


@pytest.mark.parametrize("name", list(WIDTHS))
def test_oracle_and_port_match_live_reference(name):
    D, H = WIDTHS[name]
    B, Lq, Lk, seed = 3, 24, 40, 4321
    lengths = [40, 17, 0]                                     # full, ragged, fully padded (the reference's zeros case)
    torch.manual_seed(0)
    shapes = O.block_param_shapes(D)
    p = O.make_params(shapes, seed)
    mod = _reference_block(D, H)
    assert list(mod.state_dict().keys()) == list(shapes.keys())
    mod.load_state_dict({k: torch.from_numpy(v) for k, v in p.items()}, strict=True)
    x, xa, mask, dy = O.make_inputs(B, Lq, Lk, D, seed, lengths)
    xt, xat = torch.from_numpy(x).requires_grad_(True), torch.from_numpy(xa).requires_grad_(True)
    y_ref = mod(xt, xat, torch.from_numpy(mask))
    y_ref.backward(torch.from_numpy(dy))
    g_ref = {k: v.grad.numpy() for k, v in mod.named_parameters()}

    # float64 restatement
    y, c = O.block_forward(p, x, xa, mask, H)
    gr, dx, dxa = O.block_backward(p, c, dy)
    assert O.rel_err(y, y_ref.detach().numpy()) < 2e-6
    assert O.rel_err(dx, xt.grad.numpy()) < 2e-5 and O.rel_err(dxa, xat.grad.numpy()) < 2e-5
    assert np.all(dxa[mask] == 0) and np.all(xat.grad.numpy()[mask] == 0)          # padded keys: exact zeros in both
    for k, ref in g_ref.items():
        mine = gr[k]
        if k.endswith("in_proj_bias"):                        # bk third: identically zero / fp32 noise in the reference
            idx = np.r_[0:D, 2 * D:3 * D]
            assert np.abs(mine[D:2 * D]).max() < 1e-9 and np.abs(ref[D:2 * D]).max() < 1e-3
            mine, ref = mine[idx], ref[idx]
        assert O.rel_err(mine, ref) < 5e-5, k

    # torch CPU port (the timed CPU baseline): same ATen kernels as the reference -> fp32-noise agreement
    pt = {k: torch.from_numpy(v).clone().requires_grad_(True) for k, v in p.items()}
    xt2, xat2 = torch.from_numpy(x).requires_grad_(True), torch.from_numpy(xa).requires_grad_(True)
    y2 = TP.block_forward(pt, xt2, xat2, torch.from_numpy(mask), H)
    y2.backward(torch.from_numpy(dy))
    assert O.rel_err(y2.detach().numpy(), y_ref.detach().numpy()) < 2e-6
    assert O.rel_err(xt2.grad.numpy(), xt.grad.numpy()) < 2e-5
    for k, ref in g_ref.items():
        mine = pt[k].grad.numpy()
        if k.endswith("in_proj_bias"):
            idx = np.r_[0:D, 2 * D:3 * D]
            mine, ref = mine[idx], ref[idx]
        assert O.rel_err(mine, ref) < 5e-5, k
