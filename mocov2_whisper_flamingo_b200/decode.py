"""Decode-time use of the gated cross-attention stack (BASELINE config 5).

The reference has no decode loop of its own; during autoregressive decoding its fusion stack would be called with the
same visual memory at every step.  K|V of that memory do not change, so they are projected once per layer
(``gxa_block_kv_prefill``) and every step runs the blocks on the new query rows only (``gxa_block_decode``): the result
equals ``GatedCrossModalFusion.forward`` restricted to those rows (tests/test_decode_gpu.py).

Beams: the block has no interaction between query positions, so the beams of a sample are passed as the Lq dimension
``[B, beams, D]`` and share the sample's K|V cache -- the cache is read once per sample, not once per beam.

    state = DecodeState(fusion, video_features, video_mask)      # once per utterance batch
    y_t = state.step(audio_step)                                 # [B, Lq_step, D] -> [B, Lq_step, D], every step
    state.capture(audio_step.shape)                              # optional: replay the step as one CUDA graph
"""
from __future__ import annotations

import ctypes as C
from typing import Optional

import torch

from . import _lib
from ._lib import BlockPtrs, GXA_PREC_BF16, check
from .functional import LayerNormFn, LinearFn, _dims, _f32c, _p, _require_cuda, _stream, StackShared


class DecodeState:
    """Per-batch decode state of a GatedCrossModalFusion: projected visual memory, key mask and one K|V cache per layer."""

    def __init__(self, fusion, video_features: torch.Tensor, video_mask: Optional[torch.Tensor] = None):
        L = _lib.lib()
        _require_cuda(video_features, "video_features")
        self.fusion = fusion
        with torch.no_grad():
            xa = LinearFn.apply(video_features, fusion.video_proj.weight, fusion.video_proj.bias, GXA_PREC_BF16)
        self.B, self.Lk, self.D = xa.shape
        self.H = fusion.layers[0].n_heads
        self.shared = StackShared(xa, video_mask, GXA_PREC_BF16)           # validates / prepares the bool mask
        dims = _dims(self.B, 1, self.Lk, self.D, self.H, GXA_PREC_BF16, True)
        xa_op = torch.empty(L.gxa_xa_operand_bytes(C.byref(dims)), dtype=torch.uint8, device=xa.device)
        check(L.gxa_xa_prepare(C.byref(dims), _f32c(xa).data_ptr(), xa_op.data_ptr(), _stream()), "gxa_xa_prepare")
        self.caches, self.ptrs, self.wcaches = [], [], []
        nbytes = L.gxa_block_kv_cache_bytes(C.byref(dims))
        if nbytes == 0:
            raise RuntimeError(f"gxa_b200: unsupported decode shape B={self.B} Lk={self.Lk} D={self.D} H={self.H}")
        for layer in fusion.layers:
            pd = [p.detach() for p in layer._params()]
            pp = BlockPtrs(*[p.data_ptr() for p in pd])
            wcache = layer._state.refresh(dims, pd, pp)
            kv = torch.empty(nbytes, dtype=torch.uint8, device=xa.device)
            check(L.gxa_block_kv_prefill(C.byref(dims), C.byref(pp), wcache.data_ptr(), xa_op.data_ptr(), kv.data_ptr(),
                                         _stream()), "gxa_block_kv_prefill")
            self.caches.append(kv)
            self.ptrs.append((pd, pp))
            self.wcaches.append(wcache)
        self._ws = None
        self._graph = None
        self._static_in = self._static_out = None

    def kv_cache_bytes(self) -> int:
        return sum(c.numel() for c in self.caches)

    def _run(self, audio_step: torch.Tensor) -> torch.Tensor:
        L = _lib.lib()
        f = self.fusion
        B, Lq, D = audio_step.shape
        if B != self.B or D != self.D:
            raise RuntimeError(f"gxa_b200: decode step shape {tuple(audio_step.shape)} does not match the cache (B={self.B}, D={self.D})")
        dims = _dims(B, Lq, self.Lk, D, self.H, GXA_PREC_BF16, True)
        n = L.gxa_block_decode_workspace_bytes(C.byref(dims))
        if self._ws is None or self._ws.numel() < n:
            self._ws = torch.empty(n, dtype=torch.uint8, device=audio_step.device)
        x = LinearFn.apply(audio_step, f.audio_proj.weight, f.audio_proj.bias, GXA_PREC_BF16)
        x = _f32c(x)
        y = torch.empty_like(x)
        for (pd, pp), wcache, kv in zip(self.ptrs, self.wcaches, self.caches):
            check(L.gxa_block_decode(C.byref(dims), C.byref(pp), wcache.data_ptr(), x.data_ptr(), kv.data_ptr(),
                                     _p(self.shared.mask_u8), _p(self.shared.kv_end), y.data_ptr(), self._ws.data_ptr(),
                                     _stream()), "gxa_block_decode")
            x, y = y, x
        return LayerNormFn.apply(x, f.ln_post.weight, f.ln_post.bias)

    @torch.no_grad()
    def step(self, audio_step: torch.Tensor) -> torch.Tensor:
        """One decoding step: audio_step [B, Lq_step, D] (Lq_step = 1, or the beams of each sample) -> same shape."""
        _require_cuda(audio_step, "audio_step")
        if self._graph is not None and tuple(audio_step.shape) == tuple(self._static_in.shape):
            self._static_in.copy_(audio_step)
            self._graph.replay()
            return self._static_out
        return self._run(audio_step)

    @torch.no_grad()
    def capture(self, shape, device: Optional[torch.device] = None) -> None:
        """Record the whole step (projection, n_layers blocks, ln_post) as ONE CUDA graph: at decode shapes the step is
        launch-bound (8 small kernels per layer), and there is no dropout seed or host-side state to bake in."""
        device = device or self.caches[0].device
        self._static_in = torch.zeros(*shape, device=device)
        side = torch.cuda.Stream()
        side.wait_stream(torch.cuda.current_stream())
        with torch.cuda.stream(side):
            for _ in range(3):                      # warm-up: workspaces, tensor maps, kernel attributes
                self._run(self._static_in)
        torch.cuda.current_stream().wait_stream(side)
        g = torch.cuda.CUDAGraph()
        with torch.cuda.graph(g):
            self._static_out = self._run(self._static_in)
        self._graph = g
