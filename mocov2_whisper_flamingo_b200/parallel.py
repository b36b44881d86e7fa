"""Batch-sharded data parallelism for the fusion stack: one process per GPU, full replica, and ONE
collective on the path -- the gradient all-reduce (reference: train.py:319 DDPStrategy).

Each layer's 14 gradients live in one flat fp32 bucket (functional.bucket_layout).  As soon as a layer's
backward call has been enqueued, its bucket is all-reduced with NCCL on a dedicated communication
stream (ordered after the layer's kernels by an event), so the NVLink/NVSwitch traffic of layer i
overlaps the backward kernels of layers i-1 .. 0.  ``finish()`` makes the compute stream wait for the
communication stream and reduces the few stack-level parameters (audio_proj, video_proj, ln_post).

Ownership of the layer gradients.  While a DataParallelGrads is attached and a gradient is already
present, ``BlockFn.backward`` does NOT hand the parameter gradients of a layer to autograd (it returns None for them):
the hook below installs them as ``p.grad`` itself.  Autograd's AccumulateGrad would otherwise run ``p.grad += view`` on the compute stream whenever a gradient is
already present (gradient accumulation -- the reference trains with accumulate_grad_batches = 4, config.py:42 -- or
``zero_grad(set_to_none=False)``), reading the bucket while NCCL reduces it in place on the other stream.  Here:

* first backward after ``zero_grad(set_to_none=True)`` (no gradient present): autograd adopts the views of the fresh flat
  bucket as ``p.grad`` -- it only stores them, no kernel runs, so nothing can race with the collective;
* a gradient is already present and is a view of the layer's flat bucket (accumulation, or zeroed in place): the new
  bucket is added to that flat buffer with ONE kernel on the compute stream, before the collective is ordered;
* a gradient is present but lives elsewhere (someone assigned it): it is folded into the new bucket, which then
  becomes the home of the layer's gradients.

``no_sync()`` (as torch DDP's) skips the collective for the micro-batches inside it; the all-reduce of the last
micro-batch then carries the accumulated sum.  Call ``finish()`` before ``optimizer.step()`` / ``zero_grad``.

Grouping.  Every collective costs the host ~115 us of torch.distributed / NCCL enqueue work.  ``alloc_bucket`` can hand
``group_layers`` consecutive layers slices of ONE flat buffer, reduced once when the group's lowest layer has produced
its gradients: fewer collectives, at the price of a larger last collective behind the end of backward.  Measured with
tools/dp_profile.py at 8 samples per GPU (config 3's shard on 8 GPUs), device ms per step / host enqueue ms per step:
    N = 2:  1 layer per collective 14.25 / 13.6,  2: 14.50 / 12.2,  4: 14.65 / 11.9   -> device-bound, grouping costs
    N = 8:  1 layer per collective 16.12 / 13.9,  2: 14.91 / 13.6,  4: 15.14 / 13.0   -> eight processes on one host
            enqueue ~30 % slower each (13.65 ms without any collective against 10.6 at N = 2) and the step is host-bound
Default: 2 layers per collective on 4 or more ranks (N = 4, 16 samples per GPU: 24.04 vs 24.21 ms), 1 below (GXA_DP_GROUP / ``group_layers`` override it).  Only the
common path is grouped (no gradient present at backward: the first micro-batch after zero_grad(set_to_none=True)); layers
that accumulate into existing gradients are reduced one by one as before.

SM budget: the NCCL kernel keeps its CTAs (NCCL_MAX_CTAS) resident while a bucket is in flight.  The library's
persistent kernels partition their work statically over the CTAs they launch, so a CTA that has to wait for an SM held
by NCCL doubles that kernel's time.  From the first bucket until ``finish()`` the library is therefore told to size its
grids for ``SMs - comm_sms`` (``gxa_set_sm_limit``); forward and the first backward layer use the whole device.
"""
from __future__ import annotations

import contextlib
import os
from typing import Dict, List, Optional

import torch
import torch.distributed as dist

from .functional import PARAM_SHAPES, bucket_layout


class DataParallelGrads:
    def __init__(self, fusion, process_group: Optional[dist.ProcessGroup] = None, average: bool = True,
                 comm_sms: Optional[int] = None, grad_dtype: Optional[torch.dtype] = None,
                 group_layers: Optional[int] = None):
        self.fusion = fusion
        self.pg = process_group
        self.world = dist.get_world_size(process_group) if dist.is_initialized() else 1
        self.average = average
        self.comm_stream = torch.cuda.Stream() if torch.cuda.is_available() else None
        self.pending: List = []
        self.require_sync = True
        self._home: Dict[int, tuple] = {}              # id(layer) -> (flat bucket its p.grad are views of, their data_ptrs)
        self._touched: List[torch.Tensor] = []
        if group_layers is None:               # measured default: 2 layers per collective on >= 4 ranks, 1 below (see above)
            group_layers = int(os.environ.get("GXA_DP_GROUP", "2" if self.world >= 4 else "1"))
        self.group = max(1, group_layers)
        self._index = {id(layer): i for i, layer in enumerate(fusion.layers)}
        self._gbuf: Optional[torch.Tensor] = None      # flat buffer of the group of layers in flight
        self._gid = self._gnext = -1
        self._gn = 0
        self._gadopted: List[torch.Tensor] = []
        if comm_sms is None:
            comm_sms = int(os.environ.get("GXA_DP_COMM_SMS", os.environ.get("NCCL_MAX_CTAS", "16")))
        self.comm_sms = max(0, comm_sms)
        self._limited = False
        # Optional compressed exchange (as torch DDP's bf16_compress_hook): the layer buckets travel as bf16 and come back
        # into the fp32 bucket.  Halves the NVLink bytes and the time NCCL's CTAs hold their SMs; costs one rounding to
        # 8 mantissa bits per rank's contribution (the sum itself is carried out by NCCL in bf16).  Default: fp32, the
        # reference's DDP behaviour (train.py:319).
        if grad_dtype is None and os.environ.get("GXA_DP_GRAD_DTYPE", "") in ("bf16", "bfloat16"):
            grad_dtype = torch.bfloat16
        if grad_dtype not in (None, torch.float32, torch.bfloat16):
            raise ValueError("grad_dtype must be None, torch.float32 or torch.bfloat16")
        self.grad_dtype = None if grad_dtype == torch.float32 else grad_dtype
        # diagnostics only (tools/, never bench.py's reported line): time the step without the collective
        self._diag_skip = os.environ.get("GXA_DP_DIAG_NO_ALLREDUCE") == "1"
        fusion.grad_hook = self._on_layer_grads

    @contextlib.contextmanager
    def no_sync(self):
        """Gradient accumulation: backward passes inside accumulate locally; the next backward outside all-reduces."""
        old, self.require_sync = self.require_sync, False
        try:
            yield
        finally:
            self.require_sync = old

    # called from BlockFn.backward before the layer's kernels are enqueued: the zeroed flat bucket of the layer
    def alloc_bucket(self, layer, total: int, device) -> torch.Tensor:
        i = self._index.get(id(layer))
        # (reducing the lowest group's layers one by one, to shorten the collective left behind the end of backward, was
        # measured at N = 8 and is no better: 14.73-14.81 vs 14.20-14.58 ms per step)
        if self.group <= 1 or self.world == 1 or i is None:
            return torch.zeros(total, dtype=torch.float32, device=device)
        gid, k = divmod(i, self.group)
        if self._gbuf is None or gid != self._gid or k != self._gnext or self._gbuf.device != torch.device(device) \
                or self._gbuf.numel() != self._gn * total:
            # first layer of a group in backward order (the group's highest index present)
            self._gn, self._gid = k + 1, gid
            self._gbuf = torch.zeros(self._gn * total, dtype=torch.float32, device=device)
            self._gadopted = []
        self._gnext = k - 1
        return self._gbuf[k * total:(k + 1) * total]

    def _group_member(self, bucket: torch.Tensor) -> bool:
        g = self._gbuf
        return g is not None and bucket.numel() > 0 and g.device == bucket.device and \
            g.data_ptr() <= bucket.data_ptr() < g.data_ptr() + g.numel() * 4

    # called from BlockFn.backward (autograd thread) right after the layer's kernels were enqueued.  Returns True when it
    # installed the layer's parameter gradients itself ("do not pass them to autograd"), False when autograd may adopt
    # the views (no gradient present: AccumulateGrad then only stores the view, no kernel touches the bucket).
    # Host cost matters here: at 8 samples per GPU the step is ~2000 launches in ~14 ms, so the common case (all
    # p.grad None) does 14 attribute reads and nothing else.
    def _on_layer_grads(self, layer, bucket: torch.Tensor, views) -> bool:
        params = layer._params()
        owned = False
        if all(p.grad is None for p in params):
            target = bucket                                     # autograd adopts the views of this bucket
        else:
            home = self._home.get(id(layer))
            at_home = home is not None and home[0].device == bucket.device and all(
                p.grad is not None and p.grad.data_ptr() == ptr for p, ptr in zip(params, home[1]))
            if at_home:
                home[0].add_(bucket)                            # accumulate: one kernel over the flat buffer
                target = home[0]
            else:
                with torch.no_grad():
                    for p, v in zip(params, views):
                        if p.grad is not None:                  # a gradient that lives elsewhere: fold it in, re-home
                            v.add_(p.grad.to(v.dtype).view_as(v))
                        p.grad = v
                target = bucket
            owned = True
        if target is bucket:
            self._home[id(layer)] = (bucket, [v.data_ptr() for v in views])
        member = self._group_member(bucket)
        last_of_group = member and self._gnext < 0
        targets = [target]
        if member:
            if not owned:
                self._gadopted.append(bucket)
                targets = []
            if last_of_group:                       # the group is complete: one collective over the whole flat buffer
                if len(self._gadopted) == self._gn:
                    targets = [self._gbuf]
                else:                               # some layers took the accumulation path and were reduced on their own
                    targets = targets + self._gadopted
                self._gbuf, self._gadopted = None, []
        if self.world == 1 or self._diag_skip or not self.require_sync:
            return owned
        for t in targets:
            self._all_reduce(t)
        return owned

    def _all_reduce(self, target: torch.Tensor) -> None:
        if self.comm_stream is None:           # CPU/gloo (tests): synchronous
            if self.average:
                target.div_(self.world)
            if self.grad_dtype is not None:
                wire = target.to(self.grad_dtype)
                dist.all_reduce(wire, group=self.pg)
                target.copy_(wire)
            else:
                dist.all_reduce(target, group=self.pg)
            return
        ready = torch.cuda.Event()
        ready.record(torch.cuda.current_stream())
        if not self._limited and self.comm_sms > 0:      # kernels enqueued from here on share the device with NCCL
            from . import _lib
            sms = torch.cuda.get_device_properties(target.device).multi_processor_count
            _lib.set_sm_limit(max(2, sms - self.comm_sms))
            self._limited = True
        with torch.cuda.stream(self.comm_stream):
            self.comm_stream.wait_event(ready)
            # NCCL averages inside the collective: no separate pass over the bucket on the communication stream
            op = dist.ReduceOp.AVG if self.average else dist.ReduceOp.SUM
            if self.grad_dtype is not None:
                wire = target.to(self.grad_dtype)                  # both casts run on the communication stream
                dist.all_reduce(wire, op=op, group=self.pg, async_op=True).wait()
                target.copy_(wire)
            else:
                self.pending.append(dist.all_reduce(target, op=op, group=self.pg, async_op=True))
        target.record_stream(self.comm_stream)
        self._touched.append(target)

    def broadcast_parameters(self, src: int = 0) -> None:
        if self.world == 1:
            return
        for p in self.fusion.parameters():
            dist.broadcast(p.data, src, group=self.pg)

    def finish(self) -> None:
        """Call after backward (of a synchronising micro-batch): reduces the stack-level grads and joins the
        communication stream.  After it returns, every ``p.grad`` is ordered on the current stream."""
        if self.world == 1 or not self.require_sync:
            return
        rest = [p.grad for n, p in self.fusion.named_parameters() if not n.startswith("layers.") and p.grad is not None]
        if rest and not self._diag_skip:
            flat = torch.cat([g.reshape(-1) for g in rest])
            if self.comm_stream is None:
                if self.average:
                    flat.div_(self.world)
                dist.all_reduce(flat, group=self.pg)
            else:
                dist.all_reduce(flat, op=dist.ReduceOp.AVG if self.average else dist.ReduceOp.SUM, group=self.pg)
            o = 0
            for g in rest:
                g.copy_(flat[o:o + g.numel()].view_as(g))
                o += g.numel()
        if self._limited:
            from . import _lib
            _lib.set_sm_limit(0)
            self._limited = False
        for w in self.pending:
            w.wait()                           # makes the current stream wait for the collective
        if self.comm_stream is not None:
            torch.cuda.current_stream().wait_stream(self.comm_stream)
        self.pending.clear()
        self._touched.clear()


def shard_batch(global_batch: int, rank: int, world: int):
    """Contiguous batch shard of rank `rank` (reference: DistributedSamplerWrapper, data_module.py:84-106)."""
    per = global_batch // world
    rem = global_batch % world
    start = rank * per + min(rank, rem)
    return start, start + per + (1 if rank < rem else 0)


def balanced_shards(lengths, world: int):
    """Sample indices per rank for variable-length visual memories (BASELINE config 4).

    The K|V projection and the attention core scale with the sum of the samples' frame counts, so contiguous shards of a
    length-skewed batch leave ranks unevenly loaded (up to ~30 % at 60..750 frames) and the slowest rank sets the step.
    Samples are sorted by length, longest first, and dealt to the ranks in a serpentine (0..N-1, N-1..0, ...): every rank
    gets the same number of samples (+-1) and nearly the same total length.  The reference buckets by frame count for the
    same reason (``ByFrameCountSampler``, datamodule/data_module.py:18-62).  Pure host logic, no collective: every rank
    computes the same assignment from the same ``lengths``.
    """
    if world < 1:
        raise ValueError("world must be >= 1")
    order = sorted(range(len(lengths)), key=lambda i: (-int(lengths[i]), i))
    shards = [[] for _ in range(world)]
    for pos, idx in enumerate(order):
        lap, k = divmod(pos, world)
        shards[k if lap % 2 == 0 else world - 1 - k].append(idx)
    return shards
