"""Integration fixture at the boundary the reference actually calls.  Build-container only.

Runs the reference's full ``AVNet`` (models/av_net.py) on the CPU -- ``WhisperModel.from_pretrained`` monkeypatched to a
random tiny ``WhisperConfig`` (no network), ``MoCofile=None`` -- with a forward hook on ``fusion_module`` and stores

  * the exact tensors AVNet hands to ``fusion_module`` (audio_features, video_features, video_mask) and its output,
  * the gradient AVNet's own loss sends back into the fusion output and the gradients autograd produces for the fusion
    inputs and for every fusion parameter,
  * the ``fusion_module.*`` slice of ``AVNet.state_dict()`` -- a reference-produced checkpoint.

tests/test_avnet_gpu.py loads that checkpoint into the drop-in ``GatedCrossModalFusion`` (strict=True) and replays the
call.    python oracle/make_avnet_golden.py      # rewrites tests/golden/avnet_fusion.npz
"""
from __future__ import annotations

import os
import sys

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, "/root/reference")

import transformers  # noqa: E402
from transformers import WhisperConfig, WhisperModel  # noqa: E402

D_MODEL, N_HEADS, N_LAYERS_CFG, VOCAB = 128, 2, 4, 50          # fusion: d=128, 2 heads of 64, n_layers = 4 // 2 = 2


def tiny_whisper(*_a, **_k):
    cfg = WhisperConfig(d_model=64, encoder_layers=1, decoder_layers=1, encoder_attention_heads=2,
                        decoder_attention_heads=2, encoder_ffn_dim=128, decoder_ffn_dim=128, num_mel_bins=80,
                        max_source_positions=1500, vocab_size=100, pad_token_id=0, bos_token_id=1, eos_token_id=2,
                        decoder_start_token_id=1)
    return WhisperModel(cfg)


def main():
    torch.manual_seed(2024)
    torch.set_num_threads(8)
    transformers.WhisperModel.from_pretrained = staticmethod(tiny_whisper)
    from models.av_net import AVNet
    import models.av_net as av_mod
    av_mod.WhisperModel.from_pretrained = staticmethod(tiny_whisper)
    modelargs = (D_MODEL, N_HEADS, N_LAYERS_CFG, 0, 0, 0.0)       # modelargs[5] = dropout (0: deterministic fixture)
    net = AVNet("AV", None, 0, modelargs, VOCAB)
    net.eval()                                                   # BatchNorm of the frozen frontends in eval mode
    with torch.no_grad():                                        # open the gates (init 0 = identity)
        for layer in net.fusion_module.layers:
            layer.attn_gate.fill_(0.6)
            layer.ff_gate.fill_(-0.5)
            layer.attn.in_proj_bias.uniform_(-0.1, 0.1)
            layer.attn.out_proj.bias.uniform_(-0.1, 0.1)
    B, T = 2, 24
    audio = torch.randn(B, 3000, 80)
    audio_mask = torch.zeros(B, 1500, dtype=torch.bool)
    video = 0.5 * torch.randn(B, T, 3, 32, 32)
    video_len = torch.tensor([T, 13])
    cap = {}

    def pre(mod, args, kwargs):
        cap["audio"], cap["video"] = args[0], args[1]
        cap["video_mask"] = kwargs["video_mask"]
        cap["audio"].retain_grad()
        cap["video"].retain_grad()

    def post(mod, args, kwargs, out):
        cap["out"] = out
        out.retain_grad()

    net.fusion_module.register_forward_pre_hook(pre, with_kwargs=True)
    net.fusion_module.register_forward_hook(post, with_kwargs=True)
    logits = net((audio, audio_mask, video, None, video_len))
    loss = (logits.float().log_softmax(-1)[..., 3]).sum()
    loss.backward()
    out = {"spec": np.array([B, cap["audio"].shape[1], D_MODEL, N_HEADS, N_LAYERS_CFG // 2], np.int64),
           "audio_features": cap["audio"].detach().numpy(), "video_features": cap["video"].detach().numpy(),
           "video_mask": cap["video_mask"].numpy(), "fused": cap["out"].detach().numpy(),
           "d_fused": cap["out"].grad.numpy(), "d_audio": cap["audio"].grad.numpy(), "d_video": cap["video"].grad.numpy()}
    # d_audio / d_video include the skip path `out = fused + audio + video` (av_net.py:106): subtract it to get the
    # gradient that flows through fusion_module alone
    out["d_audio"] = out["d_audio"] - out["d_fused"]
    out["d_video"] = out["d_video"] - out["d_fused"]
    sd = {k: v for k, v in net.state_dict().items() if k.startswith("fusion_module.")}
    out["ckpt_keys"] = np.array(list(sd.keys()))
    for k, v in sd.items():
        out["ckpt/" + k] = v.numpy()
    for n, p in net.fusion_module.named_parameters():
        out["grad/" + n] = p.grad.numpy()
    path = os.path.join(ROOT, "tests", "golden", "avnet_fusion.npz")
    np.savez_compressed(path, **out)
    print("wrote", path, os.path.getsize(path), "bytes; min_len", cap["audio"].shape[1], "mask sums",
          cap["video_mask"].sum(1).tolist())


if __name__ == "__main__":
    main()
