"""autograd.Function wrappers over the C ABI (include/gxa.h).  PyTorch is plumbing here: it owns
device memory, streams and the autograd graph; all arithmetic happens in libgxa_b200.so.

Backward of every Function runs on the autograd engine thread; the entry points take the
stream explicitly and keep no thread-local state.
"""
from __future__ import annotations

import ctypes as C
from typing import Optional

import torch

from . import _lib
from ._lib import BlockPtrs, Dims, GXA_PREC_BF16, GXA_PREC_FP32, check

_default_precision: Optional[int] = None
_workspaces: dict = {}


def set_default_precision(p: Optional[str]) -> None:
    """'bf16' | 'fp32' | None (None: bf16 under torch.autocast, else fp32 -- the reference's own rule)."""
    global _default_precision
    _default_precision = {None: None, "bf16": GXA_PREC_BF16, "fp32": GXA_PREC_FP32}[p]


def resolve_precision(pref: Optional[int] = None) -> int:
    if pref is not None:
        return pref
    if _default_precision is not None:
        return _default_precision
    return GXA_PREC_BF16 if torch.is_autocast_enabled("cuda") else GXA_PREC_FP32


_raw_stream = getattr(torch._C, "_cuda_getCurrentRawStream", None)
_raw_device = getattr(torch._C, "_cuda_getDevice", None)


def _stream() -> int:
    """cudaStream_t of torch's current stream on the current device.  The raw accessor is ~20x cheaper than building a
    torch.cuda.Stream object, and this is called ~130 times per step (host time matters at small batches)."""
    if _raw_stream is not None and _raw_device is not None:
        return _raw_stream(_raw_device())
    return torch.cuda.current_stream().cuda_stream


def _p(t: Optional[torch.Tensor]):
    return None if t is None else t.data_ptr()


def _require_cuda(t: torch.Tensor, what: str) -> None:
    if not t.is_cuda:
        raise RuntimeError(f"gxa_b200: {what} must be a CUDA tensor (this path has no CPU fallback)")


def _f32c(t: torch.Tensor) -> torch.Tensor:
    return t.detach().to(torch.float32).contiguous()


def workspace(nbytes: int, device: torch.device) -> torch.Tensor:
    """Grow-only scratch buffer per (device, stream); kernels on one stream run in order so sharing is safe."""
    key = (device.index, _stream())
    w = _workspaces.get(key)
    if w is None or w.numel() < nbytes:
        w = torch.empty(max(nbytes, 1 << 20), dtype=torch.uint8, device=device)
        _workspaces[key] = w
    return w


# ---------------------------------------------------------------------------------------------
# per-layer gradient bucket: the 14 gradients of a block as views of ONE flat fp32 buffer, so the
# data-parallel all-reduce of a layer is a single collective.  Segments are padded to
# 64 floats so every view is 256-byte aligned (vector atomics in the wgrad epilogues).
# ---------------------------------------------------------------------------------------------
def bucket_layout(D: int):
    sizes = [1, 1, 3 * D * D, 3 * D, D * D, D, D, D, D, D, 4 * D * D, 4 * D, 4 * D * D, D]
    offs, o = [], 0
    for s in sizes:
        offs.append(o)
        o += (s + 63) // 64 * 64
    return sizes, offs, o


PARAM_SHAPES = lambda D: [(1,), (1,), (3 * D, D), (3 * D,), (D, D), (D,), (D,), (D,), (D,), (D,), (4 * D, D), (4 * D,),
                          (D, 4 * D), (D,)]

# Host time per layer matters at small per-GPU batches (8 ranks on one host: the step is host-bound, tools/dp_profile.py):
# everything that depends on the shape only is computed once.
_LAYOUTS: dict = {}
_SIZES: dict = {}


def _layout(D: int):
    """(sizes, byte offsets, total floats, padded sizes, shapes) of a layer's flat gradient bucket."""
    lay = _LAYOUTS.get(D)
    if lay is None:
        sizes, offs, total = bucket_layout(D)
        padded = [(s + 63) // 64 * 64 for s in sizes]
        lay = _LAYOUTS[D] = (sizes, [4 * o for o in offs], total, padded, PARAM_SHAPES(D))
    return lay


def bucket_views(bucket: torch.Tensor, D: int):
    """The 14 gradient views of a flat bucket: one split call, then only the gates (padded to 64 floats) are sliced and
    only the four matrices are reshaped."""
    sizes, _, _, padded, shapes = _layout(D)
    parts = torch.split_with_sizes(bucket, padded)
    out = []
    for part, s, pad, shp in zip(parts, sizes, padded, shapes):
        if s != pad:
            part = part[:s]
        out.append(part.view(shp) if len(shp) > 1 else part)
    return out


def _block_sizes(dims: "Dims"):
    """(saved bytes, workspace bytes) of a block call, cached by shape."""
    key = (dims.B, dims.Lq, dims.Lk, dims.D, dims.H, dims.precision, dims.has_xa)
    v = _SIZES.get(key)
    if v is None:
        L = _lib.lib()
        v = _SIZES[key] = (int(L.gxa_block_saved_bytes(C.byref(dims))), int(L.gxa_block_workspace_bytes(C.byref(dims))))
    return v


class BlockState:
    """Per-module device state: the bf16 operand cache of the weights, refreshed when a weight changes."""

    def __init__(self):
        self.wcache = None
        self.key = None
        self.ready = None        # event of a cast enqueued on another stream (prefetch), not yet waited for
        self._event = None
        self.ptrs = None         # data pointers of the 14 parameters last validated (fp32, contiguous) ...
        self.pp = None           # ... and their BlockPtrs struct

    @staticmethod
    def _key(dims: Dims, params):
        return (dims.precision, dims.D) + tuple((p.data_ptr(), p._version) for p in params)

    def pointers(self, params) -> BlockPtrs:
        """BlockPtrs of the parameters; dtype / layout are validated when a pointer changed, not on every call."""
        ptrs = tuple(p.data_ptr() for p in params)
        if ptrs != self.ptrs:
            for p in params:
                if p.dtype != torch.float32 or not p.is_contiguous():
                    raise RuntimeError("gxa_b200: block parameters must be contiguous fp32 (the reference's only "
                                       "supported storage, SURVEY A.1 M8)")
            self.ptrs, self.pp = ptrs, BlockPtrs(*ptrs)
        return self.pp

    def refresh(self, dims: Dims, params, pp: BlockPtrs, stream: Optional[torch.cuda.Stream] = None) -> torch.Tensor:
        """(Re)build the bf16 operand cache if a weight changed.  With ``stream`` the cast is enqueued there (the stack
        prefetches every layer's cast off the critical path) and ``wait_ready()`` orders the consumer after it."""
        L = _lib.lib()
        key = self._key(dims, params)
        if self.wcache is None or self.key != key:
            n = L.gxa_block_wcache_bytes(C.byref(dims))
            if self.wcache is None or self.wcache.numel() != n:
                self.wcache = torch.empty(n, dtype=torch.uint8, device=params[0].device)
            check(L.gxa_block_prepare_weights(C.byref(dims), C.byref(pp), self.wcache.data_ptr(),
                                              stream.cuda_stream if stream is not None else _stream()),
                  "gxa_block_prepare_weights")
            self.key = key
            if stream is not None:
                if self._event is None:
                    self._event = torch.cuda.Event()
                self._event.record(stream)
                self.ready = self._event
            else:
                self.ready = None
        return self.wcache

    def wait_ready(self) -> None:
        if self.ready is not None:
            torch.cuda.current_stream().wait_event(self.ready)
            self.ready = None


class StackShared:
    """State shared by the layers of one GatedCrossModalFusion call: xa is the same tensor for every
    layer, so its operand form, the mask extents and the dxa accumulator exist once per stack."""

    def __init__(self, xa: Optional[torch.Tensor], mask: Optional[torch.Tensor], precision: int, grad_hook=None):
        self.precision = precision
        self.mask_u8 = None
        self.kv_end = None
        self.xa_op = None
        self.dxa = None
        self.dxa_wanted = False      # some layer saw xa.requires_grad
        self.first_token = None      # identity of the first block call (its backward runs last)
        self.grad_hook = grad_hook   # callable(layer_token, flat_bucket) -> None (data-parallel all-reduce)
        self.seeds = None            # dropout seeds drawn for the whole stack (one per layer), consumed by BlockFn.forward
        L = _lib.lib()
        if xa is not None:
            if xa.dim() == 4 and xa.size(1) == 1:
                xa = xa.squeeze(1)
            if mask is not None and mask.dim() == 3 and mask.size(1) == 1:
                mask = mask.squeeze(1)
            B, Lk, D = xa.shape
            if mask is not None:
                if mask.dtype != torch.bool:
                    raise AssertionError("only bool key_padding_mask is supported (True = padded key)")
                assert mask.shape == (B, Lk), f"expecting key_padding_mask shape of {(B, Lk)}, but got {tuple(mask.shape)}"
                self.mask_u8 = mask.contiguous().view(torch.uint8)
                self.kv_end = torch.empty(B, dtype=torch.int32, device=xa.device)
                check(L.gxa_mask_prepare(self.mask_u8.data_ptr(), self.kv_end.data_ptr(), B, Lk, _stream()),
                      "gxa_mask_prepare")


def _dims(B, Lq, Lk, D, H, precision, has_xa, p_drop=0.0, seed=0) -> Dims:
    return Dims(B=B, Lq=Lq, Lk=max(Lk, 0), D=D, H=H, precision=precision, has_xa=int(has_xa), reserved=0,
                p_drop=float(p_drop), reserved2=0, seed=int(seed) & 0xFFFFFFFFFFFFFFFF)


# Dropout seeds: the masks are a pure function of (seed, element) (csrc/gxa_rng.cuh), so all the host has to supply is
# one fresh 64-bit seed per (forward call, layer).  It is drawn from torch's default CPU generator: reproducible under
# torch.manual_seed(), advances with every call, and costs no kernel launch or device sync.
def next_dropout_seed() -> int:
    hi, lo = torch.randint(0, 1 << 31, (2,), dtype=torch.int64).tolist()
    return ((hi << 33) ^ (lo << 2) ^ 0x9E3779B97F4A7C15) & 0xFFFFFFFFFFFFFFFF


def draw_dropout_seeds(n: int) -> list:
    """n seeds from ONE draw of the same generator (a stack draws its layers' seeds together: one small torch call per
    forward instead of one per layer)."""
    v = torch.randint(0, 1 << 31, (2 * n,), dtype=torch.int64).tolist()
    return [((v[2 * i] << 33) ^ (v[2 * i + 1] << 2) ^ 0x9E3779B97F4A7C15) & 0xFFFFFFFFFFFFFFFF for i in range(n)]


class BlockFn(torch.autograd.Function):
    """GatedCrossAttentionBlock.forward / backward (reference: models/gate_cross_attention.py:28-50)."""

    @staticmethod
    def forward(ctx, x, xa, shared: StackShared, state: BlockState, n_heads: int, token, p_drop: float, *params):
        L = _lib.lib()
        _require_cuda(x, "x")
        if x.dim() != 3:
            raise AssertionError(f"query should be batched 3D tensor [B, Lq, D] but received {x.dim()}-D query tensor")
        B, Lq, D = x.shape
        xf = _f32c(x)
        has_xa = xa is not None
        if has_xa:
            # the checks F.multi_head_attention_forward makes on key / value (torch/nn/functional.py:6229-6241, :6478)
            _require_cuda(xa, "xa")
            if xa.dim() != 3:
                raise AssertionError(f"key and value should be 3D tensors [B, Lk, D] but received a {xa.dim()}-D tensor")
            if xa.device != x.device:
                raise RuntimeError(f"gxa_b200: x is on {x.device} but xa is on {xa.device}")
            if xa.shape[0] != B:
                raise AssertionError(f"key / value batch size {xa.shape[0]} does not match the query batch size {B}")
            if xa.shape[2] != D:
                raise RuntimeError(f"key / value feature width {xa.shape[2]} does not match embed_dim {D} "
                                   "(mat1 and mat2 shapes cannot be multiplied)")
        Lk = xa.shape[1] if has_xa else 0
        seed = 0
        if p_drop > 0.0:
            seed = shared.seeds.pop() if shared.seeds else next_dropout_seed()
        dims = _dims(B, Lq, Lk, D, n_heads, shared.precision, has_xa, p_drop, seed)
        pd = params                        # inside Function.forward these are plain tensors (no graph is recorded)
        pp = state.pointers(pd)
        wcache = state.refresh(dims, pd, pp)
        state.wait_ready()
        if has_xa and shared.xa_op is None:
            n = L.gxa_xa_operand_bytes(C.byref(dims))
            shared.xa_op = torch.empty(n, dtype=torch.uint8, device=x.device)
            check(L.gxa_xa_prepare(C.byref(dims), _f32c(xa).data_ptr(), shared.xa_op.data_ptr(), _stream()),
                  "gxa_xa_prepare")
        if shared.first_token is None:
            shared.first_token = token
        nsaved, nws = _block_sizes(dims)
        if nsaved == 0:
            raise RuntimeError(f"gxa_b200: unsupported shape B={B} Lq={Lq} Lk={Lk} D={D} H={n_heads} "
                               "(need head_dim 64 and d_model % 128 == 0)")
        saved = torch.empty(nsaved, dtype=torch.uint8, device=x.device)
        ws = workspace(nws, x.device)
        y = torch.empty_like(xf)
        check(L.gxa_block_fwd(C.byref(dims), C.byref(pp), wcache.data_ptr(), xf.data_ptr(), _p(shared.xa_op),
                              _p(shared.mask_u8), _p(shared.kv_end), y.data_ptr(), saved.data_ptr(), ws.data_ptr(),
                              _stream()), "gxa_block_fwd")
        ctx.dims, ctx.shared, ctx.state, ctx.token = dims, shared, state, token
        ctx.wcache, ctx.pp = wcache, pp
        ctx.xa_shape = tuple(xa.shape) if has_xa else None
        ctx.xa_needs_grad = bool(has_xa and ctx.needs_input_grad[1])
        if ctx.xa_needs_grad:
            shared.dxa_wanted = True
        ctx.x_dtype = x.dtype
        # the activation stash is an ordinary saved tensor: autograd releases it after this node's backward
        # (layer by layer during the step) and keeps it under retain_graph=True
        ctx.save_for_backward(xf, saved, *pd)
        return y

    @staticmethod
    def backward(ctx, dy):
        L = _lib.lib()
        xf, saved, *pd = ctx.saved_tensors
        dims, shared = ctx.dims, ctx.shared
        D = dims.D
        dyf = _f32c(dy)
        _, boffs, total, _, _ = _layout(D)
        # the data-parallel hook may hand out the bucket itself (slices of one flat buffer per group of layers)
        alloc = getattr(getattr(shared.grad_hook, "__self__", None), "alloc_bucket", None)
        bucket = alloc(ctx.token, total, xf.device) if alloc is not None else \
            torch.zeros(total, dtype=torch.float32, device=xf.device)
        views = bucket_views(bucket, D)
        pp = ctx.pp                                # saved tensors are the parameters themselves: same storage as in forward
        base = bucket.data_ptr()
        gp = BlockPtrs(*[base + o for o in boffs])
        dx = torch.empty_like(xf)
        dxa_ptr = None
        if dims.has_xa and shared.dxa_wanted:      # xa without requires_grad: the d(xa) GEMM is not run at all
            if shared.dxa is None:
                shared.dxa = torch.zeros(ctx.xa_shape, dtype=torch.float32, device=xf.device)
            dxa_ptr = shared.dxa.data_ptr()
        ws = workspace(_block_sizes(dims)[1], xf.device)
        check(L.gxa_block_bwd(C.byref(dims), C.byref(pp), ctx.wcache.data_ptr(), xf.data_ptr(), _p(shared.xa_op),
                              _p(shared.mask_u8), _p(shared.kv_end), saved.data_ptr(), dyf.data_ptr(),
                              dx.data_ptr(), dxa_ptr, C.byref(gp), ws.data_ptr(), _stream()), "gxa_block_bwd")
        owned = False
        if shared.grad_hook is not None:
            # data-parallel hook (parallel.py): installs the layer's gradients as p.grad itself and orders the all-reduce;
            # autograd must then not see them (AccumulateGrad would touch the bucket while NCCL reduces it in place)
            owned = bool(shared.grad_hook(ctx.token, bucket, views))
        dxa = None
        if dims.has_xa and shared.dxa_wanted and ctx.token is shared.first_token:
            dxa = shared.dxa          # every layer has accumulated into it by now
            shared.dxa = None
        grads = [None] * len(views) if owned else list(views)
        if not dims.has_xa and not owned:           # attention parameters did not take part (reference: grad None)
            for i in (0, 2, 3, 4, 5, 6, 7):
                grads[i] = None
        return (dx.to(ctx.x_dtype) if ctx.x_dtype != torch.float32 else dx, dxa, None, None, None, None, None, *grads)


class LayerNormFn(torch.autograd.Function):
    """LayerNorm.forward (reference: models/gate_cross_attention.py:8-10): fp32 statistics, output in x.dtype."""

    @staticmethod
    def forward(ctx, x, weight, bias, pe=None):
        """pe: optional [T, D] fp32 table added after the affine (out[b, t] = LN(x[b, t]) + pe[t]); the result is then
        fp32 like the reference's ``ln(x) + pe`` (fp32 buffer, models/utils.py:9)."""
        L = _lib.lib()
        _require_cuda(x, "x")
        D = x.shape[-1]
        xf = _f32c(x).view(-1, D)
        rows = xf.shape[0]
        out = torch.empty_like(xf)
        mean = torch.empty(rows, dtype=torch.float32, device=x.device)
        rstd = torch.empty_like(mean)
        w, b = _f32c(weight), _f32c(bias)
        if pe is None:
            check(L.gxa_layernorm_fwd(xf.data_ptr(), w.data_ptr(), b.data_ptr(), out.data_ptr(), 1, mean.data_ptr(),
                                      rstd.data_ptr(), rows, D, _stream()), "gxa_layernorm_fwd")
        else:
            T = x.shape[-2]
            pef = _f32c(pe).view(-1, D)
            if pef.shape[0] != T or pef.device != x.device:
                raise RuntimeError(f"gxa_b200: positional table {tuple(pe.shape)} does not match [T={T}, D={D}] on {x.device}")
            check(L.gxa_layernorm_pe_fwd(xf.data_ptr(), w.data_ptr(), b.data_ptr(), pef.data_ptr(), T, out.data_ptr(),
                                         mean.data_ptr(), rstd.data_ptr(), rows, D, _stream()), "gxa_layernorm_pe_fwd")
        ctx.save_for_backward(xf, w, mean, rstd)
        ctx.shape, ctx.dtype = x.shape, x.dtype
        ctx.has_pe = pe is not None
        return out.view(x.shape) if pe is not None else out.view(x.shape).to(x.dtype)

    @staticmethod
    def backward(ctx, g):
        L = _lib.lib()
        xf, w, mean, rstd = ctx.saved_tensors
        rows, D = xf.shape
        gf = _f32c(g).view(rows, D)
        dx = torch.empty_like(xf)
        dw = torch.zeros(D, dtype=torch.float32, device=xf.device)
        db = torch.zeros_like(dw)
        check(L.gxa_layernorm_bwd(gf.data_ptr(), xf.data_ptr(), w.data_ptr(), mean.data_ptr(), rstd.data_ptr(), None,
                                  dx.data_ptr(), dw.data_ptr(), db.data_ptr(), rows, D, _stream()), "gxa_layernorm_bwd")
        return (dx.view(ctx.shape).to(ctx.dtype), dw, db) + ((None,) if ctx.has_pe else ())


class Sum3Fn(torch.autograd.Function):
    """``a + b + c`` in one pass (reference: models/av_net.py:106 ``fused + audio_features + video_features``)."""

    @staticmethod
    def forward(ctx, a, b, c):
        L = _lib.lib()
        for t, n in ((a, "fused"), (b, "audio_features"), (c, "video_features")):
            _require_cuda(t, n)
        if not (a.shape == b.shape == c.shape):
            raise RuntimeError(f"gxa_b200: fuse_sum shapes differ: {tuple(a.shape)}, {tuple(b.shape)}, {tuple(c.shape)}")
        af, bf, cf = _f32c(a), _f32c(b), _f32c(c)
        out = torch.empty_like(af)
        n = af.numel()
        if n % 4:
            raise RuntimeError("gxa_b200: fuse_sum needs a multiple of 4 elements")
        check(L.gxa_sum3_f32(af.data_ptr(), bf.data_ptr(), cf.data_ptr(), out.data_ptr(), n, _stream()), "gxa_sum3_f32")
        ctx.dtypes = (a.dtype, b.dtype, c.dtype)
        return out

    @staticmethod
    def backward(ctx, g):
        return tuple(g.to(dt) for dt in ctx.dtypes)


class LinearFn(torch.autograd.Function):
    """nn.Linear at the stack boundary (audio_proj / video_proj, gate_cross_attention.py:56-57, 64-65)."""

    @staticmethod
    def forward(ctx, x, weight, bias, precision: int):
        L = _lib.lib()
        _require_cuda(x, "x")
        out_f, in_f = weight.shape
        xf = _f32c(x).view(-1, in_f)
        rows = xf.shape[0]
        w, b = _f32c(weight), _f32c(bias)
        y = torch.empty(rows, out_f, dtype=torch.float32, device=x.device)
        ws = workspace(L.gxa_linear_workspace_bytes(rows, in_f, out_f, precision), x.device)
        check(L.gxa_linear_fwd(xf.data_ptr(), w.data_ptr(), b.data_ptr(), y.data_ptr(), rows, in_f, out_f, precision,
                               ws.data_ptr(), _stream()), "gxa_linear_fwd")
        ctx.save_for_backward(xf, w)
        ctx.precision, ctx.in_shape, ctx.x_dtype = precision, x.shape, x.dtype
        ctx.need_dx = x.requires_grad
        return y.view(*x.shape[:-1], out_f)

    @staticmethod
    def backward(ctx, dy):
        L = _lib.lib()
        xf, w = ctx.saved_tensors
        rows, in_f = xf.shape
        out_f = w.shape[0]
        dyf = _f32c(dy).view(rows, out_f)
        dx = torch.empty_like(xf) if ctx.need_dx else None
        dw = torch.zeros_like(w)
        db = torch.zeros(out_f, dtype=torch.float32, device=xf.device)
        ws = workspace(L.gxa_linear_workspace_bytes(rows, in_f, out_f, ctx.precision), xf.device)
        check(L.gxa_linear_bwd(dyf.data_ptr(), xf.data_ptr(), w.data_ptr(), _p(dx), dw.data_ptr(), db.data_ptr(), rows,
                               in_f, out_f, ctx.precision, ws.data_ptr(), _stream()), "gxa_linear_bwd")
        return (dx.view(ctx.in_shape).to(ctx.x_dtype) if dx is not None else None), dw, db, None
