"""Decode-step timing at BASELINE config 5 (Whisper-small, batch 32, beam 5, 750 cached visual frames, 1 GPU).
Not a bench.py line (the bench metric is quoted on c3); prints one JSON line for DESIGN.md / profiles/.
    python tools/bench_decode.py [B beams Lk D H layers]
"""
import json
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import mocov2_whisper_flamingo_b200 as G  # noqa: E402

B, beams, Lk, D, H, layers = (int(a) for a in (sys.argv[1:7] + ["32", "5", "750", "768", "12", "12"][len(sys.argv) - 1:]))
torch.manual_seed(0)
mod = G.GatedCrossModalFusion(D, H, layers, dropout=0.1).cuda().eval()
with torch.no_grad():
    for l in mod.layers:
        l.attn_gate.fill_(0.5)
        l.ff_gate.fill_(0.5)
video = torch.randn(B, Lk, D, device="cuda")
lens = torch.randint(60, Lk + 1, (B,), device="cuda")
lens[0] = Lk
mask = torch.arange(Lk, device="cuda")[None, :] >= lens[:, None]
x = torch.randn(B, beams, D, device="cuda")
G.set_default_precision("bf16")


def timeit(fn, n=200, warm=20):
    for _ in range(warm):
        fn()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    torch.cuda.synchronize()
    e0.record()
    for _ in range(n):
        fn()
    e1.record()
    torch.cuda.synchronize()
    return e0.elapsed_time(e1) / n * 1e3


with torch.no_grad():
    t_full = timeit(lambda: mod(x, video, video_mask=mask), n=20, warm=3)     # what the reference does: re-project K|V
    st = G.DecodeState(mod, video, mask)
    n0 = G.launch_count()
    st.step(x)
    launches = G.launch_count() - n0
    t_eager = timeit(lambda: st.step(x))
    st.capture(x.shape)
    t_graph = timeit(lambda: st.step(x))
live = int(lens.sum().item())
kv_bytes = layers * live * 2 * D * 2                      # live K|V rows actually attended to
w_bytes = layers * 10 * D * D * 2 + 2 * D * D * 2           # bf16 operand copies: Q, out, ff.0, ff.2 (+ audio_proj)
floor_us = (kv_bytes + w_bytes) / 7.0e3 / 1e3             # at ~7 TB/s
print(json.dumps({"workload": f"decode-c5 B={B} beams={beams} Lk={Lk} D={D} H={H} layers={layers}",
                  "us_per_step_no_cache": round(t_full, 1), "us_per_step_cached_eager": round(t_eager, 1),
                  "us_per_step_cached_graph": round(t_graph, 1), "launches_per_step": launches,
                  "kv_cache_bytes": st.kv_cache_bytes(), "algorithmic_bytes_per_step": kv_bytes + w_bytes,
                  "hbm_floor_us_at_7TBs": round(floor_us, 1),
                  "achieved_GBs_graph": round((kv_bytes + w_bytes) / t_graph / 1e3, 1),
                  "tokens_per_s_graph": round(B * beams / t_graph * 1e6)}))
if os.environ.get("PROFILE"):
    from torch.profiler import profile, ProfilerActivity
    st2 = G.DecodeState(mod, video, mask)
    with torch.no_grad(), profile(activities=[ProfilerActivity.CUDA]) as prof:
        for _ in range(10):
            st2.step(x)
        torch.cuda.synchronize()
    print(prof.key_averages().table(sort_by="cuda_time_total", row_limit=20, max_name_column_width=70))
