"""The N > 1 path on CPU: two processes over ``gloo`` (reference train.py:319 DDPStrategy).

The product has no CPU compute path, so the per-shard gradients come from the oracle's torch port (the checker) and are
packed exactly as ``BlockFn.backward`` packs them: one flat fp32 bucket per layer (``functional.bucket_layout``) whose views
become the parameters' ``.grad``, handed to ``fusion.grad_hook`` layer by layer in backward order.  What is under test is
the product's host logic around the single collective of the path: parameter broadcast, the per-layer bucket all-reduce
(average and sum), ``finish()`` for the stack-level parameters, contiguous and length-balanced sharding.  The check is
the data-parallel identity: the mean over ranks of the shard gradients equals the full-batch gradient / world.
"""
from __future__ import annotations

import os
import socket
import sys

import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

D, H, LAYERS, GB, LQ, LK = 128, 2, 4, 4, 16, 12
LAYER_KEYS = ("attn_gate", "ff_gate", "attn.in_proj_weight", "attn.in_proj_bias", "attn.out_proj.weight",
              "attn.out_proj.bias", "attn_ln.weight", "attn_ln.bias", "ff_ln.weight", "ff_ln.bias", "ff.0.weight",
              "ff.0.bias", "ff.2.weight", "ff.2.bias")          # order of _lib.PARAM_FIELDS / bucket_layout


def _free_port() -> int:
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


def _global_batch():
    g = torch.Generator().manual_seed(99)
    audio = torch.randn(GB, LQ, D, generator=g)
    video = torch.randn(GB, LK, D, generator=g)
    lengths = torch.tensor([12, 5, 9, 1])
    mask = torch.arange(LK)[None, :] >= lengths[:, None]
    dy = torch.randn(GB, LQ, D, generator=g)
    return audio, video, mask, dy


def _port_grads(state, audio, video, mask, dy):
    """Gradients of sum(y * dy) over the given samples, from the oracle's torch port (the checker)."""
    from oracle import torch_port as TP
    p = {k: v.detach().clone().requires_grad_(True) for k, v in state.items()}
    y = TP.fusion_forward(p, audio, video, mask, H, LAYERS)
    (y * dy).sum().backward()
    return {k: v.grad for k, v in p.items()}


def _worker(rank: int, world: int, port: int, average: bool, wire: str = "fp32"):
    import mocov2_whisper_flamingo_b200 as G
    from mocov2_whisper_flamingo_b200.functional import PARAM_SHAPES, bucket_layout
    from mocov2_whisper_flamingo_b200.parallel import DataParallelGrads, balanced_shards, shard_batch
    dist.init_process_group("gloo", init_method=f"tcp://127.0.0.1:{port}", rank=rank, world_size=world)
    try:
        torch.manual_seed(100 + rank)                       # replicas start out different ...
        model = G.GatedCrossModalFusion(D, H, LAYERS, dropout=0.0)
        with torch.no_grad():
            for n, p in model.named_parameters():
                if n.endswith("_gate"):
                    p.fill_(0.5 + 0.1 * rank)
                elif n.endswith("bias"):
                    p.uniform_(-0.1, 0.1)
        # wire = "bf16": the layer buckets travel as bf16 (torch DDP's bf16_compress_hook equivalent); each contribution is
        # rounded to 8 mantissa bits once and the sum is formed in bf16 -> 2^-8 relative per element at world 2
        ddp = DataParallelGrads(model, average=average, grad_dtype=torch.bfloat16 if wire == "bf16" else None,
                                group_layers=2 if average else 1)           # two layers per collective, and per-layer
        tol = 2e-5 if wire == "fp32" else 8e-3
        ddp.broadcast_parameters(src=0)                     # ... and are made identical to rank 0's
        state = {k: v.detach().clone() for k, v in model.state_dict().items()}
        flat = torch.cat([v.reshape(-1) for v in state.values()])
        gathered = [torch.empty_like(flat) for _ in range(world)]
        dist.all_gather(gathered, flat)
        assert all(torch.equal(gathered[0], t) for t in gathered), "broadcast_parameters left the replicas different"

        audio, video, mask, dy = _global_batch()
        b0, b1 = shard_batch(GB, rank, world)
        local = _port_grads(state, audio[b0:b1], video[b0:b1], mask[b0:b1], dy[b0:b1])

        # what BlockFn.backward does, layer by layer in backward order
        sizes, offs, total = bucket_layout(D)
        calls = []
        real_all_reduce = dist.all_reduce
        dist.all_reduce = lambda t, *a, **k: (calls.append(t.numel()), real_all_reduce(t, *a, **k))[1]
        for i in reversed(range(LAYERS)):
            bucket = ddp.alloc_bucket(model.layers[i], total, torch.device("cpu"))    # as BlockFn.backward does
            assert bucket.numel() == total and not bool(bucket.any())
            views = [bucket[o:o + s].view(shp) for s, o, shp in zip(sizes, offs, PARAM_SHAPES(D))]
            for v, key in zip(views, LAYER_KEYS):
                v.copy_(local[f"layers.{i}.{key}"])
            # no gradient present: the hook leaves the views to autograd (which only stores them) ...
            assert model.grad_hook(model.layers[i], bucket, views) is False
            for v, prm in zip(views, model.layers[i]._params()):
                prm.grad = v                                                  # ... as AccumulateGrad does
        dist.all_reduce = real_all_reduce
        # grouped: layers (3, 2) and (1, 0) are slices of one flat buffer each, reduced by ONE collective per pair;
        # ungrouped: one collective per layer
        assert calls == ([2 * total, 2 * total] if average else [total] * LAYERS), calls
        for n, prm in model.named_parameters():
            if not n.startswith("layers."):
                prm.grad = local[n].clone()
        ddp.finish()

        full = _port_grads(state, audio, video, mask, dy)   # every rank can check itself against the full batch
        scale = 1.0 / world if average else 1.0
        for n, prm in model.named_parameters():
            ref = full[n] * scale
            den = float(ref.abs().max())
            if n.endswith("_gate"):     # judged against the size of the shard contributions it sums, which may cancel
                den = max(den, float(local[n].abs().max()) * scale)
            err = float((prm.grad - ref).abs().max()) / (den if den > 0 else 1.0)
            # the two scalar gate gradients are sums of signed terms that cancel: fp32 summation
            # order (shard sums vs one full-batch sum) shows at 1e-4 relative to the cancelled total
            lim = max(tol, 2e-3) if n.endswith("_gate") else (tol if n.startswith("layers.") else 2e-5)
            assert err < lim, (n, err)

        # ---- gradient accumulation (reference: accumulate_grad_batches, config.py:42): two micro-batches, the first one
        # inside no_sync(); p.grad is already present in the second backward, so the new bucket has to be ADDED to the
        # flat buffer behind p.grad before the single all-reduce.  Layer 0 starts from a "foreign" gradient (assigned by
        # someone else, not a view of a bucket) to cover the re-homing path.
        model.zero_grad(set_to_none=True)
        foreign = {}
        for prm, key in zip(model.layers[0]._params(), LAYER_KEYS):
            foreign[key] = torch.full_like(prm, 0.25)
            prm.grad = foreign[key].clone()
        half = (b1 - b0 + 1) // 2
        micro = [(b0, b0 + half), (b0 + half, b1)]
        for step, (m0, m1) in enumerate(micro):
            if m1 <= m0:
                loc = {k: torch.zeros_like(v) for k, v in state.items()}
            else:
                loc = _port_grads(state, audio[m0:m1], video[m0:m1], mask[m0:m1], dy[m0:m1])
            ctxm = ddp.no_sync() if step == 0 else __import__("contextlib").nullcontext()
            with ctxm:
                for i in reversed(range(LAYERS)):
                    bucket = torch.zeros(total)
                    views = [bucket[o:o + s].view(shp) for s, o, shp in zip(sizes, offs, PARAM_SHAPES(D))]
                    for v, key in zip(views, LAYER_KEYS):
                        v.copy_(loc[f"layers.{i}.{key}"])
                    had = [prm.grad is not None for prm in model.layers[i]._params()]
                    owned = model.grad_hook(model.layers[i], bucket, views)
                    assert owned is any(had)          # gradients present: the hook owns them; none: autograd adopts the views
                    if not owned:
                        for v, prm in zip(views, model.layers[i]._params()):
                            prm.grad = v
                for n, prm in model.named_parameters():
                    if not n.startswith("layers."):
                        prm.grad = loc[n].clone() if prm.grad is None else prm.grad + loc[n]
                ddp.finish()                                    # a no-op inside no_sync()
        for n, prm in model.named_parameters():
            ref = full[n] * scale
            if n.startswith("layers.0."):
                # every rank folded the same foreign 0.25 in BEFORE the collective: AVG keeps it, SUM multiplies it
                ref = ref + 0.25 * (1.0 if average else world)
            den = float(ref.abs().max())
            if n.endswith("_gate"):
                den = max(den, float(local[n].abs().max()) * scale)
            err = float((prm.grad - ref).abs().max()) / (den if den > 0 else 1.0)
            lim = max(tol, 2e-3) if n.endswith("_gate") else (tol if n.startswith("layers.") else 2e-5)
            assert err < lim, ("accumulate", n, err)

        # sharding helpers: every rank derives the same partition from the same lengths
        lengths = [750, 60, 400, 399, 90, 700, 61, 333]
        mine = balanced_shards(lengths, world)
        box = [None] * world
        dist.all_gather_object(box, mine)
        assert all(b == box[0] for b in box)
        assert sorted(i for s in mine for i in s) == list(range(len(lengths)))
        spans = [None] * world
        dist.all_gather_object(spans, (b0, b1))
        assert spans[0][0] == 0 and spans[-1][1] == GB and all(spans[k][1] == spans[k + 1][0] for k in range(world - 1))
    finally:
        dist.destroy_process_group()


@pytest.mark.parametrize("average", [True, False])
def test_world_size_2_gradient_allreduce_matches_full_batch(average):
    world = 2
    mp.spawn(_worker, args=(world, _free_port(), average), nprocs=world, join=True)


def test_world_size_2_bf16_gradient_exchange():
    mp.spawn(_worker, args=(2, _free_port(), True, "bf16"), nprocs=2, join=True)
