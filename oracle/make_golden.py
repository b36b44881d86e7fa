"""Generate tests/golden/*.npz from the LIVE reference module.  Build-container only.

Imports /root/reference/models/gate_cross_attention.py (it cannot travel to the GPU
box), loads numpy-seeded parameters (oracle.gxa_oracle.make_params) through
``load_state_dict(strict=True)``, runs forward + autograd backward in fp32 on the
CPU and stores outputs and gradients.  Inputs/params are NOT stored: they are
regenerated from the recorded seed by the oracle's numpy generator.

    python oracle/make_golden.py          # rewrites tests/golden/
"""
from __future__ import annotations

import os
import sys

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, "/root/reference")

from oracle import gxa_oracle as O  # noqa: E402
from models.gate_cross_attention import GatedCrossAttentionBlock, GatedCrossModalFusion  # noqa: E402

GOLD = os.path.join(ROOT, "tests", "golden")

# name -> (kind, B, Lq, Lk, D, H, n_layers, lengths, seed, store_grads, gates)
CASES = {
    "block_small":    ("block", 2, 16, 12, 128, 2, 1, None, 11, True, (0.7, -0.4)),
    "block_masked":   ("block", 3, 24, 40, 128, 2, 1, [40, 17, 0], 12, True, (0.7, -0.4)),
    "block_gate0":    ("block", 2, 16, 12, 128, 2, 1, None, 13, True, (0.0, 0.0)),
    "block_noxa":     ("block_noxa", 2, 16, 0, 128, 2, 1, None, 14, True, (0.7, -0.4)),
    "fusion_small":   ("fusion", 2, 20, 28, 128, 2, 2, [28, 9], 15, True, (0.7, -0.4)),
    "block_c1":       ("block", 2, 448, 250, 384, 6, 1, None, 1234, False, (0.7, -0.4)),
}


def run_case(name, spec):
    kind, B, Lq, Lk, D, H, nl, lengths, seed, store_grads, gates = spec
    torch.manual_seed(0)
    torch.set_num_threads(8)
    out = {"spec": np.array([B, Lq, Lk, D, H, nl, seed], np.int64),
           "gates": np.array(gates, np.float64),
           "lengths": np.array(lengths if lengths is not None else [Lk] * B, np.int64)}
    if kind == "fusion":
        shapes = O.fusion_param_shapes(D, nl)
        mod = GatedCrossModalFusion(D, H, nl, dropout=0.0)
    else:
        shapes = O.block_param_shapes(D)
        mod = GatedCrossAttentionBlock(D, H, dropout=0.0)
    p = O.make_params(shapes, seed, *gates)
    assert list(mod.state_dict().keys()) == list(shapes.keys()), "state_dict layout drifted"
    mod.load_state_dict({k: torch.from_numpy(v) for k, v in p.items()}, strict=True)
    x, xa, mask, dy = O.make_inputs(B, Lq, max(Lk, 1), D, seed, lengths)
    xt = torch.from_numpy(x).requires_grad_(True)
    xat = torch.from_numpy(xa).requires_grad_(True)
    mt = torch.from_numpy(mask)
    if kind == "block":
        y = mod(xt, xat, mt)
    elif kind == "block_noxa":
        y = mod(xt)
    else:
        y = mod(xt, xat, video_mask=mt)
    out["y"] = y.detach().numpy()
    # the reference under bf16 autocast (the only reduced-precision mode it supports)
    if kind != "block_noxa":
        with torch.no_grad(), torch.autocast("cpu", dtype=torch.bfloat16):
            yb = mod(xt, xat, mt) if kind == "block" else mod(xt, xat, video_mask=mt)
        out["y_autocast_bf16"] = yb.float().numpy()
    if store_grads:
        y.backward(torch.from_numpy(dy))
        out["dx"] = xt.grad.numpy()
        if kind != "block_noxa":
            out["dxa"] = xat.grad.numpy()
        for k, v in mod.named_parameters():
            if v.grad is not None:
                out["g:" + k] = v.grad.numpy()
    np.savez_compressed(os.path.join(GOLD, name + ".npz"), **out)
    print(name, "y", out["y"].shape, "keys", len(out))


if __name__ == "__main__":
    os.makedirs(GOLD, exist_ok=True)
    for n, s in CASES.items():
        run_case(n, s)
