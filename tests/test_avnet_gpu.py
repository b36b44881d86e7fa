"""Drop-in check at the boundary the reference calls (models/av_net.py:51-57 constructor, :99-104 call): a
reference-produced checkpoint (the ``fusion_module.*`` slice of ``AVNet.state_dict()``) is loaded with strict=True and
the exact tensors the reference's AVNet handed to its fusion module are replayed; outputs and every gradient are
compared with what the reference's autograd produced.  Fixture: tests/golden/avnet_fusion.npz, written by
oracle/make_avnet_golden.py from the live reference (full AVNet, tiny random Whisper, MoCo backbone)."""
import os

import numpy as np
import pytest
import torch

from conftest import GOLDEN
import mocov2_whisper_flamingo_b200 as G

FIX = os.path.join(GOLDEN, "avnet_fusion.npz")


def _load():
    g = np.load(FIX)
    B, L, D, H, nl = (int(v) for v in g["spec"])
    prefix = "fusion_module."
    ckpt = {str(k): torch.from_numpy(g["ckpt/" + str(k)]) for k in g["ckpt_keys"]}
    return g, (B, L, D, H, nl), prefix, ckpt


def test_reference_checkpoint_layout_matches():
    """CPU: same keys, same ORDER, same shapes as the reference's state_dict; strict load succeeds."""
    g, (B, L, D, H, nl), prefix, ckpt = _load()
    mod = G.GatedCrossModalFusion(d_model=D, n_heads=H, n_layers=nl, dropout=0.1, enable_logging=False)
    mine = mod.state_dict()
    assert [prefix + k for k in mine.keys()] == list(ckpt.keys())
    for k, v in mine.items():
        assert tuple(v.shape) == tuple(ckpt[prefix + k].shape), k
        assert v.dtype == ckpt[prefix + k].dtype == torch.float32, k
    missing, unexpected = mod.load_state_dict({k[len(prefix):]: v for k, v in ckpt.items()}, strict=True)
    assert not missing and not unexpected
    assert float(mod.layers[0].attn_gate.item()) == pytest.approx(0.6)          # what monitor.py / the notebook read


@pytest.mark.gpu
@pytest.mark.parametrize("prec,tol", [("fp32", 1e-4), ("bf16", 2e-2)])
def test_avnet_call_replayed(prec, tol):
    from gpu_util import rel_err
    g, (B, L, D, H, nl), prefix, ckpt = _load()
    mod = G.GatedCrossModalFusion(d_model=D, n_heads=H, n_layers=nl, dropout=0.0, enable_logging=False)
    mod.load_state_dict({k[len(prefix):]: v for k, v in ckpt.items()}, strict=True)
    mod.cuda().eval()
    a = torch.from_numpy(g["audio_features"]).cuda().requires_grad_(True)
    v = torch.from_numpy(g["video_features"]).cuda().requires_grad_(True)
    vm = torch.from_numpy(g["video_mask"]).cuda()
    am = torch.zeros(B, L, dtype=torch.bool, device="cuda")                      # av_net.py:101 passes it; it is ignored
    G.set_default_precision(prec)
    try:
        fused = mod(a, v, audio_mask=am, video_mask=vm)                          # the call of av_net.py:99-104
        fused.backward(torch.from_numpy(g["d_fused"]).cuda())
        torch.cuda.synchronize()
    finally:
        G.set_default_precision(None)
    assert fused.dtype == torch.float32 and tuple(fused.shape) == (B, L, D)
    assert rel_err(fused.detach().cpu(), torch.from_numpy(g["fused"])) < tol
    gtol = tol
    assert rel_err(a.grad.cpu(), torch.from_numpy(g["d_audio"])) < gtol
    assert rel_err(v.grad.cpu(), torch.from_numpy(g["d_video"])) < gtol
    for n, p in mod.named_parameters():
        ref = torch.from_numpy(g["grad/" + n]).double()
        mine = p.grad.double().cpu()
        if n.endswith("in_proj_bias"):           # the bk third is ~0 (softmax shift invariance): absolute tolerance
            Dm = ref.shape[0] // 3
            assert mine[Dm:2 * Dm].abs().max() <= gtol * max(1.0, float(mine.abs().max())), n
            idx = torch.cat([torch.arange(0, Dm), torch.arange(2 * Dm, 3 * Dm)])
            mine, ref = mine[idx], ref[idx]
        den = float(ref.abs().max())
        assert float((mine - ref).abs().max()) / (den if den > 0 else 1.0) < gtol, n
