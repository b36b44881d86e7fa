"""the statements either side of the fusion stack in AVNet (av_net.py:83-84, :88-89, :106) against outputs of
the reference's own LayerNorm / PositionalEncoding / nn.Linear (tests/golden/avglue.npz, oracle/make_avglue_golden.py)."""
import os

import numpy as np
import pytest
import torch

from conftest import GOLDEN
import mocov2_whisper_flamingo_b200 as G
from mocov2_whisper_flamingo_b200 import av_glue


def _load():
    return np.load(os.path.join(GOLDEN, "avglue.npz"))


def _rel(a, b):
    a, b = a.double().cpu(), torch.as_tensor(b).double()
    return float((a - b).abs().max() / b.abs().max().clamp_min(1e-30))      # max-norm, as everywhere else


def test_positional_table_matches_reference_buffer():
    g = _load()
    B, T, FA, FV, D = (int(v) for v in g["spec"])
    pe = av_glue.positional_table(D)
    assert pe.shape == (1, 5000, D)
    assert np.array_equal(pe[0, :T].numpy(), g["audio/pe_head"])           # same fp32 recipe -> same bits
    assert np.array_equal(pe[0, :T].numpy(), g["video/pe_head"])


def test_cpu_input_fails_loudly():
    g = _load()
    B, T, FA, FV, D = (int(v) for v in g["spec"])
    proj, ln = torch.nn.Linear(FA, D), G.LayerNorm(D)
    with pytest.raises(RuntimeError, match="no CPU fallback"):
        av_glue.proj_ln_pe(torch.from_numpy(g["audio/x"]), proj, ln, av_glue.positional_table(D))
    with pytest.raises(AssertionError):
        av_glue.proj_ln_pe(torch.zeros(3, FA), proj, ln, av_glue.positional_table(D))


@pytest.mark.gpu
@pytest.mark.parametrize("precision,tol_y,tol_g", [("fp32", 1e-4, 1e-4), ("bf16", 2e-2, 2e-2)])
def test_glue_matches_reference(precision, tol_y, tol_g):
    g = _load()
    B, T, FA, FV, D = (int(v) for v in g["spec"])
    G.set_default_precision(precision)
    try:
        feats, mods = {}, {}
        for tag, F in (("audio", FA), ("video", FV)):
            proj, ln = torch.nn.Linear(F, D), G.LayerNorm(D)
            with torch.no_grad():
                proj.weight.copy_(torch.from_numpy(g[f"{tag}/proj.weight"]))
                proj.bias.copy_(torch.from_numpy(g[f"{tag}/proj.bias"]))
                ln.weight.copy_(torch.from_numpy(g[f"{tag}/ln.weight"]))
                ln.bias.copy_(torch.from_numpy(g[f"{tag}/ln.bias"]))
            proj.cuda(), ln.cuda()
            x = torch.from_numpy(g[f"{tag}/x"]).cuda().requires_grad_(True)
            feats[tag] = av_glue.proj_ln_pe(x, proj, ln, av_glue.positional_table(D, device="cuda"))
            mods[tag] = (proj, ln, x)
            assert feats[tag].dtype == torch.float32
            assert _rel(feats[tag].detach(), g[f"{tag}/y"]) < tol_y, tag
        fused = torch.from_numpy(g["fused"]).cuda().requires_grad_(True)
        out = av_glue.fuse_sum(fused, feats["audio"], feats["video"])
        assert _rel(out.detach(), g["out"]) < tol_y
        out.backward(torch.from_numpy(g["d_out"]).cuda())
        assert _rel(fused.grad, g["d_fused"]) < 1e-6
        for tag in ("audio", "video"):
            proj, ln, x = mods[tag]
            assert _rel(x.grad, g[f"{tag}/dx"]) < tol_g, tag
            assert _rel(proj.weight.grad, g[f"{tag}/d_proj.weight"]) < tol_g, tag
            assert _rel(proj.bias.grad, g[f"{tag}/d_proj.bias"]) < tol_g, tag
            assert _rel(ln.weight.grad, g[f"{tag}/d_ln.weight"]) < tol_g, tag
            assert _rel(ln.bias.grad, g[f"{tag}/d_ln.bias"]) < tol_g, tag
    finally:
        G.set_default_precision(None)
