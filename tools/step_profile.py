"""Per-kernel time inside a real bench step (CUPTI activity records via torch.profiler; kernels run back to back,
warm caches -- unlike the serialised ncu launch list).  Not a bench line.
    python tools/step_profile.py [workload] [per_gpu_batch]
LAYERS=n overrides the layer count; PLAIN=1 only runs a few steps (the command to put under `ncu --set full -k regex:...`).
"""
import collections
import os
import re
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import bench  # noqa: E402
import mocov2_whisper_flamingo_b200 as G  # noqa: E402

wl = sys.argv[1] if len(sys.argv) > 1 else "whisper-medium-c3"
D, H, layers, GB, Lq, Lk, _variable = bench.WORKLOADS[wl]
B = int(sys.argv[2]) if len(sys.argv) > 2 else GB
layers = int(os.environ.get("LAYERS", layers))
torch.manual_seed(1234)
model = G.GatedCrossModalFusion(D, H, layers, dropout=float(os.environ.get("DROPOUT", "0.0")))
with torch.no_grad():
    for n, p in model.named_parameters():
        if n.endswith("attn_gate"):
            p.fill_(0.7)
        elif n.endswith("ff_gate"):
            p.fill_(-0.4)
model.cuda().train()
G.set_default_precision("bf16")
a = torch.randn(B, Lq, D, device="cuda", requires_grad=True)
v = torch.randn(B, Lk, D, device="cuda", requires_grad=True)
mask = torch.zeros(B, Lk, dtype=torch.bool, device="cuda")
dy = torch.randn(B, Lq, D, device="cuda")

plist = list(model.parameters())


def step():
    model.invalidate_weight_cache()
    bench.drop_grads(plist)
    a.grad = v.grad = None
    y = model(a, v, video_mask=mask)
    (y * dy).sum().backward()


# HIPRI=1: run the step on a high-priority torch stream (the library's side stream has the default = lowest priority, so
# its weight-gradient GEMMs then only take SMs the dependency chain does not want)
if os.environ.get("HIPRI"):
    _hp = torch.cuda.Stream(priority=-1)
    _hp.wait_stream(torch.cuda.current_stream())
    torch.cuda.set_stream(_hp)
for _ in range(3):
    step()
torch.cuda.synchronize()
# host time to ENQUEUE one step (no synchronisation inside): if it approaches the device time the step is launch-bound
import time  # noqa: E402
torch.cuda.synchronize()
t0 = time.perf_counter()
for _ in range(3):
    step()
t_host = (time.perf_counter() - t0) / 3
torch.cuda.synchronize()
t_dev = (time.perf_counter() - t0) / 3
print(f"{wl} B={B}: host enqueue {t_host * 1e3:.2f} ms/step, wall {t_dev * 1e3:.2f} ms/step (3 steps back to back)")
if os.environ.get("PLAIN"):
    print("plain run done")
    sys.exit(0)
from torch.profiler import ProfilerActivity, profile  # noqa: E402

N = 2
with profile(activities=[ProfilerActivity.CUDA]) as prof:
    for _ in range(N):
        step()
    torch.cuda.synchronize()
agg = collections.defaultdict(lambda: [0, 0.0])
t_min, t_max = None, None
for e in prof.events():
    if e.device_type != torch.autograd.DeviceType.CUDA:
        continue
    name = re.sub(r"\(.*", "", e.name)[:90]
    agg[name][0] += 1
    agg[name][1] += e.device_time if hasattr(e, "device_time") else e.cuda_time
    s, t = e.time_range.start, e.time_range.end
    t_min = s if t_min is None else min(t_min, s)
    t_max = t if t_max is None else max(t_max, t)
tot = sum(v_[1] for v_ in agg.values())
print(f"{wl} B={B}: sum of kernel time {tot / N / 1e3:.2f} ms/step; wall span {(t_max - t_min) / N / 1e3:.2f} ms/step "
      f"(gap = launch/idle)")
for k, v_ in sorted(agg.items(), key=lambda kv: -kv[1][1]):
    print(f"{v_[1] / N / 1e3:9.3f} ms {100 * v_[1] / tot:5.1f}%  n/step={v_[0] // N:5d}  avg={v_[1] / v_[0]:8.1f} us  {k}")

# ---- idle gaps on the device timeline: intervals where no kernel of any stream is running
iv = sorted((e.time_range.start, e.time_range.end, re.sub(r"\(.*", "", e.name)[:60]) for e in prof.events()
            if e.device_type == torch.autograd.DeviceType.CUDA)
gaps, cur_end, cur_name = [], iv[0][1], iv[0][2]
for s_, t_, nm in iv[1:]:
    if s_ > cur_end:
        gaps.append((s_ - cur_end, cur_name, nm, cur_end - iv[0][0]))
    if t_ > cur_end:
        cur_end, cur_name = t_, nm
idle = sum(g_[0] for g_ in gaps)
print(f"device idle inside the {N}-step span: {idle / N / 1e3:.2f} ms/step in {len(gaps) // N} gaps/step; largest:")
for g_ in sorted(gaps, key=lambda g_: -g_[0])[:14]:
    print(f"  {g_[0]:8.1f} us at +{g_[3] / 1e3:8.2f} ms  after {g_[1]}  before {g_[2]}")
