"""Where a data-parallel step goes (not a bench line): run under torchrun with N >= 2 ranks, each with PER_GPU_BATCH samples
of config 3 (default 8 = the shard of the 8-GPU run), and time on every rank
    plain        no gradient exchange object at all
    ddp-skip     DataParallelGrads attached, collective skipped (GXA_DP_DIAG_NO_ALLREDUCE): hook + SM budget only
    ddp-fp32/gK  fp32 exchange, one collective per K layers (g1 = the bench configuration)
    ddp-bf16     bf16 gradient exchange
device time per step (CUDA events, max over ranks) and host time to enqueue a step; then a CUPTI timeline of rank 0 for
ddp-fp32: per-kernel totals, NCCL kernel durations, idle gaps.
    python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29531 tools/dp_profile.py
"""
import collections
import os
import sys
import time

import torch
import torch.distributed as dist

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
os.environ.setdefault("NCCL_MAX_CTAS", "16")
if os.environ.get("NCCL_DEBUG", "VERSION") == "VERSION":
    os.environ["NCCL_DEBUG"] = "WARN"
import bench  # noqa: E402
import mocov2_whisper_flamingo_b200 as G  # noqa: E402
from mocov2_whisper_flamingo_b200.parallel import DataParallelGrads  # noqa: E402

wl = os.environ.get("WORKLOAD", "whisper-medium-c3")
D, H, layers, GB, Lq, Lk, _variable = bench.WORKLOADS[wl]
B = int(os.environ.get("PER_GPU_BATCH", "8"))
rank, world, local = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
torch.cuda.set_device(local)
dist.init_process_group("nccl", device_id=torch.device("cuda", local))
torch.manual_seed(1234)
model = G.GatedCrossModalFusion(D, H, layers, dropout=0.1)
bench.init_like_bench(model)
model.cuda().train()
G.set_default_precision("bf16")
g = torch.Generator(device="cuda").manual_seed(1234 + rank)
a = torch.randn(B, Lq, D, device="cuda", generator=g, requires_grad=True)
v = torch.randn(B, Lk, D, device="cuda", generator=g, requires_grad=True)
mask = torch.zeros(B, Lk, dtype=torch.bool, device="cuda")
dy = torch.randn(B, Lq, D, device="cuda", generator=g)
probe = torch.zeros(sum(p.numel() for p in model.layers[0].parameters()), device="cuda")
for _ in range(4):
    dist.all_reduce(probe, op=dist.ReduceOp.AVG)
torch.cuda.synchronize()
ddp = None

plist = list(model.parameters())


def step():
    model.invalidate_weight_cache()
    bench.drop_grads(plist)
    a.grad = v.grad = None
    y = model(a, v, video_mask=mask)
    (y * dy).sum().backward()
    if ddp:
        ddp.finish()


def measure(name, steps=10):
    for _ in range(4):
        step()
    dist.barrier()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    t0 = time.perf_counter()
    for _ in range(steps):
        step()
    t_host = (time.perf_counter() - t0) / steps * 1e3
    e1.record()
    torch.cuda.synchronize()
    t = torch.tensor([e0.elapsed_time(e1) / steps, t_host], device="cuda")
    dist.all_reduce(t, op=dist.ReduceOp.MAX)
    if rank == 0:
        print(f"{name:10s} N={world} B/GPU={B}: device {float(t[0]):7.2f} ms/step   host enqueue {float(t[1]):7.2f} ms/step "
              f"(max over ranks; NCCL_MAX_CTAS={os.environ.get('NCCL_MAX_CTAS')})", flush=True)


# DP_CONFIGS=plain,skip,bf16,g1,g2,g4,g2/sms8,...   gK = K layers per collective, /smsS = SMs taken from the library's
# persistent grids while buckets are in flight (default NCCL_MAX_CTAS)
configs = os.environ.get("DP_CONFIGS", "plain,skip,bf16,g2,g4,g1").split(",")
for cfg in configs:
    if cfg == "plain":
        ddp = None
        model.grad_hook = None
    elif cfg == "skip":
        os.environ["GXA_DP_DIAG_NO_ALLREDUCE"] = "1"
        ddp = DataParallelGrads(model)
        os.environ.pop("GXA_DP_DIAG_NO_ALLREDUCE")
    elif cfg == "bf16":
        ddp = DataParallelGrads(model, grad_dtype=torch.bfloat16)
    else:
        grp, _, sms = cfg.partition("/sms")
        ddp = DataParallelGrads(model, group_layers=int(grp[1:]), comm_sms=int(sms) if sms else None)
    measure("ddp-" + cfg if cfg != "plain" else cfg)

if os.environ.get("TIMELINE", "1") == "1":
    from torch.profiler import ProfilerActivity, profile
    dist.barrier()
    torch.cuda.synchronize()
    with profile(activities=[ProfilerActivity.CUDA]) as prof:
        for _ in range(2):
            step()
        torch.cuda.synchronize()
    if rank == 0:
        evs = sorted((e for e in prof.events() if e.device_type == torch.autograd.DeviceType.CUDA),
                     key=lambda e: e.time_range.start)
        tot = collections.defaultdict(lambda: [0.0, 0])
        for e in evs:
            k = e.name[:70]
            tot[k][0] += e.time_range.elapsed_us()
            tot[k][1] += 1
        span = (max(e.time_range.end for e in evs) - evs[0].time_range.start) / 2e3
        print(f"timeline rank 0, ddp-fp32: span {span:.2f} ms/step; kernel totals per step:")
        for k, (us, n) in sorted(tot.items(), key=lambda kv: -kv[1][0])[:14]:
            print(f"  {us / 2e3:8.3f} ms  n/step={n // 2:4d}  avg={us / n:8.1f} us  {k}")
        # compute-stream idle: gaps between consecutive non-NCCL kernels
        comp = [e for e in evs if "nccl" not in e.name.lower()]
        idle, end = 0.0, comp[0].time_range.end
        gaps = []
        for e in comp[1:]:
            if e.time_range.start > end:
                idle += e.time_range.start - end
                gaps.append((e.time_range.start - end, e.name[:50]))
            end = max(end, e.time_range.end)
        print(f"  idle between library kernels: {idle / 2e3:.2f} ms/step; largest gaps (us, before kernel):")
        for gap, nm in sorted(gaps, reverse=True)[:8]:
            print(f"    {gap:8.1f}  {nm}")
        nc = [e for e in evs if "nccl" in e.name.lower()]
        if nc:
            first = nc[0].time_range.start - evs[0].time_range.start
            print(f"  NCCL kernels: {len(nc) // 2}/step, total {sum(e.time_range.elapsed_us() for e in nc) / 2e3:.2f} ms/step, "
                  f"first starts {first / 1e3:.2f} ms into the span")
dist.barrier()
dist.destroy_process_group()
