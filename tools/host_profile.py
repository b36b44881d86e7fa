"""Where the HOST time of a step goes at a small per-GPU batch (cProfile of the enqueue path, no synchronisation inside).
    python tools/host_profile.py [workload] [per_gpu_batch]
"""
import cProfile
import os
import pstats
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import bench  # noqa: E402
import mocov2_whisper_flamingo_b200 as G  # noqa: E402

wl = sys.argv[1] if len(sys.argv) > 1 else "whisper-medium-c3"
D, H, layers, GB, Lq, Lk, _v = bench.WORKLOADS[wl]
B = int(sys.argv[2]) if len(sys.argv) > 2 else 8
model = G.GatedCrossModalFusion(D, H, layers, dropout=0.1)
bench.init_like_bench(model)
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


for _ in range(3):
    step()
torch.cuda.synchronize()
pr = cProfile.Profile()
pr.enable()
for _ in range(5):
    step()
pr.disable()
torch.cuda.synchronize()
st = pstats.Stats(pr)
st.sort_stats("tottime").print_stats(18)
