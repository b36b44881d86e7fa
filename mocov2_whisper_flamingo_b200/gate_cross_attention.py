"""Drop-in replacements for the three classes of the reference's models/gate_cross_attention.py.

Same class names, constructor arguments, ``forward`` signatures, padding-mask semantics
(bool, True = padded key) and ``state_dict`` layout, so that

    from mocov2_whisper_flamingo_b200.gate_cross_attention import GatedCrossModalFusion, LayerNorm

can replace ``from .gate_cross_attention import ...`` in the reference's models/av_net.py:9 and a
checkpoint written by the reference loads with ``strict=True``.  The torch sub-modules created here
(``nn.MultiheadAttention``, ``nn.Sequential``) are *parameter containers only* -- they give the
reference's exact parameter names, shapes, registration order and default initialisation; their
``forward`` is never called.  All arithmetic runs in libgxa_b200.so (hand-written sm_100a CUDA).

Precision: ``torch.autocast`` active (the reference's training mode, train.py:316) -> bf16 tensor-core
contractions with fp32 LayerNorm/residual; otherwise fp32-accurate contractions (3xbf16 split) that
match the reference's fp32 module to <= 1e-4.  Override with ``set_default_precision``.
"""
from __future__ import annotations

import logging
from typing import Optional

import torch
import torch.nn as nn

from ._lib import BlockPtrs, GXA_PREC_BF16
from .functional import (BlockFn, BlockState, LayerNormFn, LinearFn, StackShared, _dims, draw_dropout_seeds, resolve_precision,
                         set_default_precision)

logger = logging.getLogger(__name__)

__all__ = ["LayerNorm", "GatedCrossAttentionBlock", "GatedCrossModalFusion", "set_default_precision"]


def _log_shapes(enabled: bool, where: str, **tensors) -> None:
    """``enable_logging=True`` (reference :29-35, :41-49, :61-72): tensor shapes at the module boundaries, INFO level."""
    if enabled:
        logger.info("%s: %s", where, ", ".join(f"{k}{tuple(t.shape)}" for k, t in tensors.items() if t is not None))


class LayerNorm(nn.LayerNorm):
    """reference: models/gate_cross_attention.py:8-10 (fp32 statistics, result cast back to x.dtype)."""

    def forward(self, x: torch.Tensor) -> torch.Tensor:
        return LayerNormFn.apply(x, self.weight, self.bias)


class GatedCrossAttentionBlock(nn.Module):
    """reference: models/gate_cross_attention.py:12-50."""

    def __init__(self, d_model, n_heads, dropout=0.1, enable_logging=False):
        super().__init__()
        self.enable_logging = enable_logging
        hidden = 4 * d_model            # the reference's FFN width (:21), independent of any fc_hidden_size
        # Parameter containers only (their forward is never called), created and registered in the reference's order
        # (:16-27): same default initialisation stream under torch.manual_seed and the same state_dict() ordering.
        for name, holder in (
                ("attn", nn.MultiheadAttention(d_model, n_heads, dropout=dropout, batch_first=True)),
                ("attn_ln", LayerNorm(d_model)),
                ("ff_ln", LayerNorm(d_model)),
                ("attn_gate", nn.Parameter(torch.zeros(1))),
                ("ff_gate", nn.Parameter(torch.zeros(1))),
                ("ff", nn.Sequential(nn.Linear(d_model, hidden), nn.GELU(), nn.Linear(hidden, d_model),
                                     nn.Dropout(dropout)))):
            setattr(self, name, holder)
        self.d_model, self.n_heads, self.dropout = d_model, n_heads, dropout
        self._state = BlockState()
        self.precision: Optional[int] = None   # None: follow autocast / set_default_precision

    def _params(self):
        # order of _lib.PARAM_FIELDS.  Read straight from the registration dicts: nn.Module.__getattr__ costs ~0.4 us
        # per hop and this runs twice per layer and step (same objects `self.attn.in_proj_weight` etc. would return).
        own, mods = self._parameters, self._modules
        attn, ff = mods["attn"], mods["ff"]._modules
        ap, op = attn._parameters, attn._modules["out_proj"]._parameters
        l1, l2, f0, f2 = mods["attn_ln"]._parameters, mods["ff_ln"]._parameters, ff["0"]._parameters, ff["2"]._parameters
        return (own["attn_gate"], own["ff_gate"], ap["in_proj_weight"], ap["in_proj_bias"], op["weight"], op["bias"],
                l1["weight"], l1["bias"], l2["weight"], l2["bias"], f0["weight"], f0["bias"], f2["weight"], f2["bias"])

    def forward(self, x, xa=None, mask=None, _shared: Optional[StackShared] = None):
        _log_shapes(self.enable_logging, "block in", x=x, xa=xa, mask=mask)
        if xa is not None and xa.dim() == 4 and xa.size(1) == 1:          # [B, 1, Lk, D] visual memory (:35-36)
            xa = xa.squeeze(1)
        if mask is not None and mask.dim() == 3 and mask.size(1) == 1:    # [B, 1, Lk] padding mask (:37-38)
            mask = mask.squeeze(1)
        # dropout (reference :16 attention probabilities, :26 after ff.2) is active in training mode only
        p_drop = float(self.dropout) if self.training else 0.0
        shared = _shared
        if shared is None:
            shared = StackShared(xa, mask if xa is not None else None, resolve_precision(self.precision))
        y = BlockFn.apply(x, xa, shared, self._state, self.n_heads, self, p_drop, *self._params())
        _log_shapes(self.enable_logging, "block out", y=y)
        return y


class GatedCrossModalFusion(nn.Module):
    """reference: models/gate_cross_attention.py:52-73 (audio_mask is accepted and ignored there too)."""

    def __init__(self, d_model, n_heads, n_layers, dropout=0.1, enable_logging=False):
        super().__init__()
        self.enable_logging = enable_logging
        self.audio_proj = nn.Linear(d_model, d_model)
        self.video_proj = nn.Linear(d_model, d_model)
        self.layers = nn.ModuleList([GatedCrossAttentionBlock(d_model, n_heads, dropout,
                                                              enable_logging=self.enable_logging)
                                     for _ in range(n_layers)])
        self.ln_post = LayerNorm(d_model)
        self.precision: Optional[int] = None
        self.grad_hook = None    # set by parallel.DataParallelGrads: per-layer gradient all-reduce
        self._cast_streams = {}  # device index -> stream the weight casts are prefetched on

    def invalidate_weight_cache(self):
        """Force the bf16 operand cache of every layer to be rebuilt on the next forward (the cache is
        otherwise refreshed automatically when a weight's version counter changes, e.g. after optimizer.step())."""
        for layer in self.layers:
            layer._state.key = None

    def _prefetch_weight_casts(self, prec: int, device: torch.device) -> None:
        """bf16 mode: the fp32 -> bf16 operand cast of every layer whose weights changed since the last call (all of them
        after an optimizer step) is enqueued on a side stream at the start of the stack, ordered after the work already on
        the caller's stream (the weight update), instead of in front of each layer's first GEMM: 24 HBM-bound passes of
        75 MB leave the critical path (1.25 ms of a 15 ms step at 8 samples per GPU)."""
        if prec != GXA_PREC_BF16 or device.type != "cuda":
            return
        stale = []
        for layer in self.layers:
            pd = layer._params()
            st = layer._state
            if st.wcache is None or st.key != (prec, layer.d_model) + tuple((p.data_ptr(), p._version) for p in pd):
                stale.append((layer, pd))
        if not stale:
            return
        side = self._cast_streams.get(device.index)
        if side is None:
            side = self._cast_streams[device.index] = torch.cuda.Stream(device=device)
        side.wait_stream(torch.cuda.current_stream(device))
        for layer, pd in stale:
            if any(not p.is_cuda for p in pd):
                return                     # BlockFn.forward raises the proper error
            dims = _dims(1, 1, 1, layer.d_model, layer.n_heads, prec, True)
            layer._state.refresh(dims, pd, layer._state.pointers(pd), stream=side)

    def forward(self, audio_features, video_features, audio_mask=None, video_mask=None):
        _log_shapes(self.enable_logging, "fusion in", audio_features=audio_features, video_features=video_features)
        prec = resolve_precision(self.precision)
        self._prefetch_weight_casts(prec, audio_features.device)
        x = LinearFn.apply(audio_features, self.audio_proj.weight, self.audio_proj.bias, prec)
        xa = LinearFn.apply(video_features, self.video_proj.weight, self.video_proj.bias, prec)
        _log_shapes(self.enable_logging, "fusion projected", x=x, xa=xa)
        shared = StackShared(xa, video_mask, prec, grad_hook=self.grad_hook)
        if self.training and any(layer.dropout > 0.0 for layer in self.layers):
            shared.seeds = draw_dropout_seeds(len(self.layers))
        for layer in self.layers:
            x = layer(x, xa, video_mask, _shared=shared)
        x = LayerNormFn.apply(x, self.ln_post.weight, self.ln_post.bias)
        _log_shapes(self.enable_logging, "fusion out", y=x)
        return x
