"""Fixture for the statements either side of the fusion stack in AVNet.  Build-container only.

Runs the reference's own classes on the CPU -- ``LayerNorm`` (models/gate_cross_attention.py:7-10), ``PositionalEncoding``
(models/utils.py:8-28) and ``nn.Linear`` wired exactly as models/av_net.py:83-84 / :88-89 / :106 -- and stores inputs,
parameters, outputs and autograd gradients.    python oracle/make_avglue_golden.py   # rewrites tests/golden/avglue.npz
"""
from __future__ import annotations

import os
import sys

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, "/root/reference")
from models.gate_cross_attention import LayerNorm  # noqa: E402
from models.utils import PositionalEncoding  # noqa: E402


def main():
    torch.manual_seed(77)
    B, T, FA, FV, D = 2, 21, 192, 256, 128
    out = {"spec": np.array([B, T, FA, FV, D], np.int64)}
    feats = {}
    for tag, F in (("audio", FA), ("video", FV)):
        proj, ln, pos = torch.nn.Linear(F, D), LayerNorm(D), PositionalEncoding(D)
        with torch.no_grad():
            ln.weight.copy_(1 + 0.1 * torch.randn(D))
            ln.bias.uniform_(-0.1, 0.1)
        x = torch.randn(B, T, F, requires_grad=True)
        y = pos(ln(proj(x)))                               # av_net.py:83-84 / :88-89
        y.retain_grad()
        feats[tag] = y
        out.update({f"{tag}/x": x.detach().numpy(), f"{tag}/proj.weight": proj.weight.detach().numpy(),
                    f"{tag}/proj.bias": proj.bias.detach().numpy(), f"{tag}/ln.weight": ln.weight.detach().numpy(),
                    f"{tag}/ln.bias": ln.bias.detach().numpy(), f"{tag}/pe_head": pos.pe[0, :T].numpy()})
        feats[tag + "_mods"] = (proj, ln, x)
    fused = torch.randn(B, T, D, requires_grad=True)
    total = fused + feats["audio"] + feats["video"]        # av_net.py:106
    dy = torch.randn(B, T, D)
    total.backward(dy)
    out.update({"fused": fused.detach().numpy(), "out": total.detach().numpy(), "d_out": dy.numpy(),
                "d_fused": fused.grad.numpy()})
    for tag in ("audio", "video"):
        proj, ln, x = feats[tag + "_mods"]
        out.update({f"{tag}/y": feats[tag].detach().numpy(), f"{tag}/dx": x.grad.numpy(),
                    f"{tag}/d_proj.weight": proj.weight.grad.numpy(), f"{tag}/d_proj.bias": proj.bias.grad.numpy(),
                    f"{tag}/d_ln.weight": ln.weight.grad.numpy(), f"{tag}/d_ln.bias": ln.bias.grad.numpy()})
    path = os.path.join(ROOT, "tests", "golden", "avglue.npz")
    np.savez_compressed(path, **out)
    print("wrote", path, os.path.getsize(path), "bytes")


if __name__ == "__main__":
    main()
