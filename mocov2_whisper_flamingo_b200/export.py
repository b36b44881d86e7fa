"""ONNX export of the gated cross-attention stack (reference: export_model.py:15-47, :86-105).

The library's ops are opaque to ``torch.onnx.export`` (they are ctypes calls into libgxa_b200.so), and the reference's
own export script is broken as committed (stale AVNet constructor / forward signature).  This module is the
repaired equivalent for the path this repository owns:

* ``ExportableFusion`` / ``ExportableBlock`` wrap a ``GatedCrossModalFusion`` / ``GatedCrossAttentionBlock``.  Their
  forward runs the SAME CUDA kernels as the wrapped module (through autograd Functions that take the padding mask as a
  tensor argument, so that it becomes a graph input), and each Function carries a ``symbolic`` that emits the reference's
  arithmetic (models/gate_cross_attention.py:8-10, :39-46, :64-70; nn.MultiheadAttention's masked softmax,
  torch/nn/functional.py:6409-6415, :6676-6684) as STANDARD opset-14 ONNX operators.  The symbolic is graph emission
  only -- it never computes anything, and nothing in the product's compute path goes through it.
* ``export_onnx`` is ``export_model.py:export_to_onnx`` for the stack: eval mode, opset 14, constant folding, named
  inputs ``audio_features`` / ``video_features`` / ``video_mask``, dynamic batch and length axes.  Serialising needs the
  ``onnx`` package (torch's TorchScript exporter imports it in its last step); where it is missing the function says so.
* ``onnx_graph`` stops one step earlier and returns torch's in-memory ONNX graph plus the initialisers; tests/test_export_gpu.py
  evaluates that graph node by node and compares it with the CUDA path and the reference goldens, so the emitted graph is checked
  even where ``onnx`` / ``onnxruntime`` are not installed (this image).

Operators used: Add Sub Mul Div Sqrt Tanh Erf ReduceMean MatMul Transpose Reshape Slice Softmax Where IsNaN Unsqueeze
Constant (all in the default domain, opset <= 14; LayerNormalization / Gelu only exist from opset 17 / 20 and are
decomposed).
"""
from __future__ import annotations

import math
from typing import Optional

import torch
import torch.nn as nn

from .functional import LayerNormFn, LinearFn
from .gate_cross_attention import GatedCrossAttentionBlock, GatedCrossModalFusion

__all__ = ["ExportableBlock", "ExportableFusion", "export_onnx", "onnx_graph"]

_EPS = 1e-5          # nn.LayerNorm default (models/gate_cross_attention.py:8-10)


# ------------------------------------------------------------------------------------------------ graph emission helpers
def _const(g, value, dtype=torch.float32):
    return g.op("Constant", value_t=torch.tensor(value, dtype=dtype))


def _layer_norm(g, x, w, b):
    """fp32 LayerNorm over the last axis, biased variance (reference :8-10)."""
    mean = g.op("ReduceMean", x, axes_i=[-1], keepdims_i=1)
    d = g.op("Sub", x, mean)
    var = g.op("ReduceMean", g.op("Mul", d, d), axes_i=[-1], keepdims_i=1)
    rstd = g.op("Sqrt", g.op("Add", var, _const(g, _EPS)))
    return g.op("Add", g.op("Mul", g.op("Div", d, rstd), w), b)


def _linear(g, x, w, b):
    """x W^T + b with W stored [out, in] (nn.Linear)."""
    y = g.op("MatMul", x, g.op("Transpose", w, perm_i=[1, 0]))
    return y if b is None else g.op("Add", y, b)


def _rows(g, t, start, end):
    """t[start:end] along axis 0 (opset >= 10: starts / ends / axes are inputs)."""
    i64 = torch.int64
    return g.op("Slice", t, _const(g, [start], i64), _const(g, [end], i64), _const(g, [0], i64))


def _split_heads(g, t, n_heads):
    """[B, L, D] -> [B, H, L, D/H]"""
    t = g.op("Reshape", t, _const(g, [0, 0, n_heads, -1], torch.int64))
    return g.op("Transpose", t, perm_i=[0, 2, 1, 3])


def _block_graph(g, x, xa, mask, n_heads, p):
    """GatedCrossAttentionBlock.forward in eval mode (reference :39-46) as standard operators.  p: the 14 parameters in
    the order of GatedCrossAttentionBlock._params()."""
    (attn_gate, ff_gate, in_w, in_b, out_w, out_b, ln1_w, ln1_b, ln2_w, ln2_b, ff0_w, ff0_b, ff2_w, ff2_b) = p
    d_model = in_w.type().sizes()[1]
    head_dim = d_model // n_heads
    x1 = x
    if xa is not None:
        h = _layer_norm(g, x, ln1_w, ln1_b)
        q = _linear(g, h, _rows(g, in_w, 0, d_model), _rows(g, in_b, 0, d_model))
        k = _linear(g, xa, _rows(g, in_w, d_model, 2 * d_model), _rows(g, in_b, d_model, 2 * d_model))
        v = _linear(g, xa, _rows(g, in_w, 2 * d_model, 3 * d_model), _rows(g, in_b, 2 * d_model, 3 * d_model))
        q, k, v = (_split_heads(g, t, n_heads) for t in (q, k, v))
        s = g.op("MatMul", q, g.op("Transpose", k, perm_i=[0, 1, 3, 2]))
        s = g.op("Mul", s, _const(g, 1.0 / math.sqrt(head_dim)))
        if mask is not None:
            # key_padding_mask: True = padded key -> -inf on that key for every head and query (functional.py:6409-6415)
            m4 = g.op("Unsqueeze", g.op("Unsqueeze", mask, _const(g, [1], torch.int64)), _const(g, [1], torch.int64))
            s = g.op("Add", s, g.op("Where", m4, _const(g, float("-inf")), _const(g, 0.0)))
        prob = g.op("Softmax", s, axis_i=-1)
        if mask is not None:
            # a sample whose keys are ALL padded: softmax of -inf everywhere is NaN; the library (and its CPU checker) define
            # the attention output of such rows as 0
            prob = g.op("Where", g.op("IsNaN", prob), _const(g, 0.0), prob)
        ctx = g.op("Transpose", g.op("MatMul", prob, v), perm_i=[0, 2, 1, 3])
        ctx = g.op("Reshape", ctx, _const(g, [0, 0, -1], torch.int64))
        o = _linear(g, ctx, out_w, out_b)
        x1 = g.op("Add", x, g.op("Mul", o, g.op("Tanh", attn_gate)))
    h2 = _layer_norm(g, x1, ln2_w, ln2_b)
    u = _linear(g, h2, ff0_w, ff0_b)
    erf = g.op("Erf", g.op("Mul", u, _const(g, 1.0 / math.sqrt(2.0))))
    a = g.op("Mul", g.op("Mul", u, _const(g, 0.5)), g.op("Add", erf, _const(g, 1.0)))       # exact (erf) GELU, nn.GELU()
    f = _linear(g, a, ff2_w, ff2_b)
    return g.op("Add", x1, g.op("Mul", f, g.op("Tanh", ff_gate)))


# ------------------------------------------------------------------------------------------------ export-time Functions
# forward = the product's CUDA path (inference, no autograd graph); symbolic = graph emission
class _no_trace:
    """The tracer keeps recording aten calls made INSIDE an autograd Function's forward (they are dead code once the
    Function's node takes over its outputs), and chokes on some of the host-side ones the library makes (dtype views of
    workspace buffers).  Nothing inside these forwards belongs in the graph: tracing is paused for their duration."""

    def __enter__(self):
        self.state = torch._C._get_tracing_state()
        torch._C._set_tracing_state(None)

    def __exit__(self, *exc):
        torch._C._set_tracing_state(self.state)


class _BlockExportFn(torch.autograd.Function):
    @staticmethod
    def forward(ctx, x, xa, mask, n_heads, block, *params):
        with _no_trace(), torch.no_grad():
            return block(x, xa, mask)

    @staticmethod
    def symbolic(g, x, xa, mask, n_heads, block, *params):
        return _block_graph(g, x, xa, mask, n_heads, params)


class _LinearExportFn(torch.autograd.Function):
    @staticmethod
    def forward(ctx, x, w, b, precision):
        with _no_trace(), torch.no_grad():
            return LinearFn.apply(x, w, b, precision)

    @staticmethod
    def symbolic(g, x, w, b, precision):
        return _linear(g, x, w, b)


class _LayerNormExportFn(torch.autograd.Function):
    @staticmethod
    def forward(ctx, x, w, b):
        with _no_trace(), torch.no_grad():
            return LayerNormFn.apply(x, w, b)

    @staticmethod
    def symbolic(g, x, w, b):
        return _layer_norm(g, x, w, b)


# ------------------------------------------------------------------------------------------------ exportable wrappers
def _squeeze_inputs(xa, mask):
    if xa is not None and xa.dim() == 4 and xa.size(1) == 1:          # [B, 1, Lk, D] visual memory (reference :35-36)
        xa = xa.squeeze(1)
    if mask is not None and mask.dim() == 3 and mask.size(1) == 1:    # [B, 1, Lk] padding mask (reference :37-38)
        mask = mask.squeeze(1)
    return xa, mask


class ExportableBlock(nn.Module):
    """``GatedCrossAttentionBlock`` with an ONNX-exportable forward (same signature: x, xa=None, mask=None)."""

    def __init__(self, block: GatedCrossAttentionBlock):
        super().__init__()
        self.block = block

    def forward(self, x, xa=None, mask=None):
        xa, mask = _squeeze_inputs(xa, mask)
        if xa is None:
            mask = None
        return _BlockExportFn.apply(x, xa, mask, self.block.n_heads, self.block, *self.block._params())


class ExportableFusion(nn.Module):
    """``GatedCrossModalFusion`` with an ONNX-exportable forward (reference :61-72; audio_mask is ignored there too)."""

    def __init__(self, fusion: GatedCrossModalFusion):
        super().__init__()
        self.fusion = fusion

    def forward(self, audio_features, video_features, video_mask=None):
        from .functional import resolve_precision
        f = self.fusion
        prec = resolve_precision(f.precision)
        x = _LinearExportFn.apply(audio_features, f.audio_proj.weight, f.audio_proj.bias, prec)
        xa = _LinearExportFn.apply(video_features, f.video_proj.weight, f.video_proj.bias, prec)
        for layer in f.layers:
            x = _BlockExportFn.apply(x, xa, video_mask, layer.n_heads, layer, *layer._params())
        return _LayerNormExportFn.apply(x, f.ln_post.weight, f.ln_post.bias)


def _wrap(module):
    if isinstance(module, (ExportableBlock, ExportableFusion)):
        return module
    if isinstance(module, GatedCrossModalFusion):
        return ExportableFusion(module)
    if isinstance(module, GatedCrossAttentionBlock):
        return ExportableBlock(module)
    raise TypeError(f"cannot export {type(module).__name__}: expected GatedCrossModalFusion or GatedCrossAttentionBlock")


def _names(wrapped, n_inputs):
    if isinstance(wrapped, ExportableFusion):
        names = ["audio_features", "video_features", "video_mask"]
    else:
        names = ["x", "xa", "mask"]
    names = names[:n_inputs]
    # batch and both sequence lengths are dynamic, as in export_model.py:41-46
    axes = {names[0]: {0: "batch", 1: "query_length"}, "output": {0: "batch", 1: "query_length"}}
    if n_inputs > 1:
        axes[names[1]] = {0: "batch", 1: "key_length"}
    if n_inputs > 2:
        axes[names[2]] = {0: "batch", 1: "key_length"}
    return names, axes


def onnx_graph(module, example_inputs, opset_version: int = 14):
    """torch's in-memory ONNX graph of the module (every node an ``onnx::`` operator) and its initialisers.

    Returns ``(graph, params)``: ``graph`` is a ``torch._C.Graph`` whose inputs are the named module inputs followed by the
    initialisers, ``params`` maps initialiser names to tensors.  Needs a CUDA device (tracing runs the real forward), not
    the ``onnx`` package."""
    from torch.onnx._internal.torchscript_exporter import utils as _ts_utils
    from torch.onnx._internal.torchscript_exporter._globals import GLOBALS
    wrapped = _wrap(module).eval()
    example_inputs = tuple(example_inputs)
    names, axes = _names(wrapped, len(example_inputs))
    old = GLOBALS.export_onnx_opset_version
    GLOBALS.export_onnx_opset_version = opset_version
    try:
        with torch.no_grad():
            graph, params, _ = _ts_utils._model_to_graph(wrapped, example_inputs, input_names=names,
                                                         output_names=["output"], dynamic_axes=axes,
                                                         do_constant_folding=False)
    finally:
        GLOBALS.export_onnx_opset_version = old
    return graph, params


def export_onnx(module, path, example_inputs, opset_version: int = 14) -> None:
    """``export_model.py:export_to_onnx`` for the stack / a block: writes an ONNX file with dynamic batch and length axes.

    ``example_inputs``: CUDA tensors ``(audio_features, video_features[, video_mask])`` for a fusion stack or
    ``(x[, xa[, mask]])`` for a block."""
    wrapped = _wrap(module).eval()
    example_inputs = tuple(example_inputs)
    names, axes = _names(wrapped, len(example_inputs))
    try:
        import onnx  # noqa: F401  (torch's TorchScript exporter needs it to serialise)
    except ImportError as e:
        raise RuntimeError("export_onnx: the `onnx` package is required to write the file (torch.onnx imports it when "
                           "serialising); onnx_graph() builds and returns the same graph without it") from e
    with torch.no_grad():
        torch.onnx.export(wrapped, example_inputs, path, export_params=True, opset_version=opset_version,
                          do_constant_folding=True, input_names=names, output_names=["output"], dynamic_axes=axes,
                          dynamo=False)
