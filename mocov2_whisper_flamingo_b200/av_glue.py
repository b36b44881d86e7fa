"""The step either side of the fusion stack inside the reference's AVNet, on the path's own kernels.

Reference, models/av_net.py:
    :83-84   audio_features = pos_enc_audio(audio_ln(audio_proj(whisper_out)))
    :88-89   video_features = pos_enc_video(video_ln(video_proj(visual_out)))
    :106     out = fused + audio_features + video_features
``proj`` is an ``nn.Linear``, ``ln`` the reference's ``LayerNorm`` (fp32 statistics, eps 1e-5) and ``pe`` the buffer of
``models/utils.py:8-28 PositionalEncoding`` ([1, max_len, d], sin on even / cos on odd channels).

The functions take the reference's OWN modules / buffers, so AVNet keeps its parameters and state_dict keys
(``audio_proj.weight``, ``audio_ln.weight``, ``pos_enc_audio.pe`` ...) and only the three statements change:

    audio_features = proj_ln_pe(whisper_out, self.audio_proj, self.audio_ln, self.pos_enc_audio.pe)
    video_features = proj_ln_pe(visual_out, self.video_proj, self.video_ln, self.pos_enc_video.pe)
    out = fuse_sum(fused, audio_features, video_features)

Everything runs in the library: GEMM (+bias) on the tcgen05 kernels (tensor cores under autocast, split-bf16
fp32-accurate otherwise), LayerNorm with the positional add folded into its store (``gxa_layernorm_pe_fwd``: one HBM
pass instead of two), and the 3-way sum as one pass (``gxa_sum3_f32``: 12 bytes read + 4 written per element instead
of two binary adds' 24).  Achieved GB/s: tools/bench_glue.py, profiles/.
"""
from __future__ import annotations

import math
from typing import Optional

import torch
from torch import nn

from .functional import LayerNormFn, LinearFn, Sum3Fn, resolve_precision


def positional_table(d_model: int, max_len: int = 5000, device=None) -> torch.Tensor:
    """The ``pe`` buffer of models/utils.py:8-19, [1, max_len, d_model] fp32."""
    pe = torch.zeros(max_len, d_model)
    position = torch.arange(0, max_len, dtype=torch.float).unsqueeze(1)
    denominator = torch.exp(torch.arange(0, d_model, 2).float() * (-math.log(10000.0) / d_model))
    pe[:, 0::2] = torch.sin(position * denominator)
    pe[:, 1::2] = torch.cos(position * denominator)
    pe = pe.unsqueeze(0)
    return pe.to(device) if device is not None else pe


def proj_ln_pe(x: torch.Tensor, proj: nn.Linear, ln: nn.Module, pe: torch.Tensor,
               precision: Optional[int] = None) -> torch.Tensor:
    """``pos_enc(ln(proj(x)))`` for x [B, T, in_features] -> [B, T, d_model] fp32 (av_net.py:83-84 / :88-89)."""
    if x.dim() != 3:
        raise AssertionError(f"expected [B, T, F] features, got {tuple(x.shape)}")
    T = x.size(1)
    if T > pe.size(1):
        raise RuntimeError(f"The size of tensor a ({T}) must match the size of tensor b ({pe.size(1)}) at non-singleton "
                           "dimension 1")        # what the reference's broadcast add raises
    prec = resolve_precision(precision)
    h = LinearFn.apply(x, proj.weight, proj.bias, prec)
    return LayerNormFn.apply(h, ln.weight, ln.bias, pe[0, :T] if pe.dim() == 3 else pe[:T])


def fuse_sum(fused: torch.Tensor, audio_features: torch.Tensor, video_features: torch.Tensor) -> torch.Tensor:
    """``fused + audio_features + video_features`` (av_net.py:106)."""
    return Sum3Fn.apply(fused, audio_features, video_features)
