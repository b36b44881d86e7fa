#!/usr/bin/env python
"""bench.py -- gated cross-attention fwd+bwd query-tokens/s at the Whisper-medium shape (BASELINE.json
configs[2]: d=1024, 16 heads, 24 gated x-attn layers, global batch 64, Lq=448 x Lk=750, bf16), batch-sharded
over N GPUs (strong scaling: the global batch is fixed).

    python bench.py [--gpus N --steps K --warmup W]          # product arm (CUDA, libgxa_b200.so)
    python bench.py --impl reference ...                      # reference's CPU path (oracle port) on host cores

A step = H2D of the step's inputs (e2e only) -> GatedCrossModalFusion forward (24 layers) -> loss ->
backward (all dgrad + wgrad) -> per-layer gradient all-reduce (N>1) -> D2H of the loss (e2e only).
The fp32->bf16 weight cast is redone every step (a real optimizer step would change the weights).
Prints ONE JSON line on rank 0.
"""
from __future__ import annotations

import argparse
import json
import os

os.environ.setdefault("NCCL_MAX_CTAS", "16")   # bound the SMs the gradient all-reduce may hold (parallel.py: SM budget)
import statistics
import subprocess
import sys
import threading
import time

import torch

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

WORKLOADS = {
    # name: (D, H, layers, global batch, Lq, Lk)
    "whisper-medium-c3": (1024, 16, 24, 64, 448, 750),
    "whisper-small-c2": (768, 12, 12, 16, 448, 750),
    "whisper-tiny-c1": (384, 6, 1, 2, 448, 250),
}
METRIC = "gated-xattn fwd+bwd tokens/s at Whisper-medium"


def block_fwd_flops(B, Lq, Lk, D):
    """20 B Lq D^2 + 4 (sum len) D^2 + 4 Lq (sum len) D, full-length keys."""
    return 20 * B * Lq * D * D + 4 * B * Lk * D * D + 4 * B * Lq * Lk * D


def peaks():
    try:
        with open(os.path.join(ROOT, "MEASURED_PEAKS.json")) as f:
            p = json.load(f)
        return dict(burst=float(p["bf16_tflops"]), sustained=float(p["bf16_tflops_sustained"]),
                    hbm=float(p["hbm_gbs"]), source="measured")
    except Exception:
        return dict(burst=1590.0, sustained=1400.0, hbm=6650.0, source="fallback")


class ClockSampler:
    """nvidia-smi clocks + throttle reasons DURING the timed region (B200_PROFILING.md recipe)."""
    Q = ("clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,"
         "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,"
         "clocks_event_reasons.sw_power_cap")

    def __init__(self, index: int):
        self.index, self.rows, self.proc = index, [], None

    def start(self):
        try:
            self.proc = subprocess.Popen(["nvidia-smi", f"--query-gpu={self.Q}", "--format=csv,noheader,nounits",
                                          "-i", str(self.index), "-lms", "100"], stdout=subprocess.PIPE, text=True)
            threading.Thread(target=self._read, daemon=True).start()
        except Exception:
            self.proc = None

    def _read(self):
        for line in self.proc.stdout:
            self.rows.append([c.strip() for c in line.split(",")])

    def stop(self):
        if self.proc is None:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        time.sleep(0.15)
        self.proc.terminate()
        sm, mx, reasons = [], None, set()
        for r in self.rows:
            try:
                sm.append(float(r[0])); mx = float(r[1])
            except Exception:
                continue
            for name, v in zip(("hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"), r[3:7]):
                if v.lower().startswith("active"):
                    reasons.add(name)
        return {"sm_mhz": statistics.median(sm) if sm else None, "sm_max_mhz": mx, "reasons": sorted(reasons),
                "samples": len(sm)}


def cpu_port_tokens_per_s(D, H, layers, Lq, Lk, threads, sample_B=2, sample_layers=2, reps=3, dropout=0.1):
    """The reference's CPU path (oracle/torch_port.py: same ATen CPU kernels as the reference module) on a bounded
    sample: `sample_layers` layers of the stack at batch `sample_B`, fwd+bwd, extrapolated linearly to `layers`."""
    from oracle import gxa_oracle as O
    from oracle import torch_port as TP
    torch.set_num_threads(threads)
    p = TP.to_torch(O.make_params(O.fusion_param_shapes(D, sample_layers), 1234), requires_grad=True)
    x, xa, mask, dy = O.make_inputs(sample_B, Lq, Lk, D, 1234)
    xt, xat = torch.from_numpy(x).requires_grad_(True), torch.from_numpy(xa).requires_grad_(True)
    mt, dyt = torch.from_numpy(mask), torch.from_numpy(dy)
    times = []
    for i in range(reps + 1):
        t0 = time.perf_counter()
        y = TP.fusion_forward(p, xt, xat, mt, H, sample_layers, dropout, True)
        y.backward(dyt)
        times.append(time.perf_counter() - t0)
    t = statistics.median(times[1:])
    t_full = t * layers / sample_layers
    return sample_B * Lq / t_full, t, f"{sample_layers} of {layers} layers at batch {sample_B} (Lq={Lq}, Lk={Lk}, D={D}), fp32 fwd+bwd, " \
                                      f"median of {reps} after 1 warm-up, extrapolated linearly in layers"


def run_reference(args, D, H, layers, GB, Lq, Lk):
    """--impl reference: the reference's own CPU implementation of the path on the host cores.  The reference is a
    Python module that cannot travel to the GPU box, so this is the oracle port (same ATen CPU kernels)."""
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    threads = os.cpu_count() or 1
    sample_B, sample_layers = 2, 2
    from oracle import gxa_oracle as O
    from oracle import torch_port as TP
    torch.set_num_threads(threads)
    p = TP.to_torch(O.make_params(O.fusion_param_shapes(D, sample_layers), 1234), requires_grad=True)
    x, xa, mask, dy = O.make_inputs(sample_B, Lq, Lk, D, 1234)
    xt, xat = torch.from_numpy(x).requires_grad_(True), torch.from_numpy(xa).requires_grad_(True)
    mt, dyt = torch.from_numpy(mask), torch.from_numpy(dy)

    def step():
        TP.fusion_forward(p, xt, xat, mt, H, sample_layers, args.dropout, True).backward(dyt)

    for _ in range(args.warmup):
        step()
    t0 = time.perf_counter()
    for _ in range(args.steps):
        step()
    t = (time.perf_counter() - t0) / args.steps
    t_full = t * layers / sample_layers           # seconds for sample_B samples through all layers
    value = sample_B * Lq / t_full
    sample = (f"{sample_layers} of {layers} layers at batch {sample_B} per step, fp32 fwd+bwd on {threads} host threads, "
              "extrapolated linearly in layers; reference module is Python and cannot travel -> oracle/torch_port.py "
              "(same ATen CPU kernels)")
    print(json.dumps({
        "impl": "reference", "metric": METRIC, "value": value, "unit": "tokens/s", "n_gpus": args.gpus,
        "steps": args.steps, "warmup": args.warmup, "ms_per_step": t * 1e3, "higher_is_better": True,
        "scaling": "strong", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
        "config": {"workload": args.workload, "d_model": D, "heads": H, "layers": layers, "global_batch": GB,
                   "Lq": Lq, "Lk": Lk, "dropout": args.dropout, "mode": "train"},
        "cpu_baseline": {"value": value, "unit": "tokens/s", "cores": threads, "kind": "port", "sample": sample},
        "e2e": {"value": value, "unit": "tokens/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
    }))


def time_dominant_kernel(M, D, pk):
    """FFN1 GEMM with the fused bias + erf-GELU epilogue at this rank's shape, CUDA events on the launch stream."""
    import ctypes as C  # noqa: F401
    from mocov2_whisper_flamingo_b200 import _lib
    L = _lib.lib()
    N, K = 4 * D, D
    A = torch.randn(M, K, device="cuda").to(torch.bfloat16)
    W = (torch.randn(N, K, device="cuda") / K ** 0.5).to(torch.bfloat16)
    bias = torch.randn(N, device="cuda")
    U = torch.empty(M, N, device="cuda", dtype=torch.bfloat16)
    Act = torch.empty_like(U)
    s = torch.cuda.current_stream().cuda_stream

    def launch():
        _lib.check(L.gxa_gemm_bf16(A.data_ptr(), K, 0, W.data_ptr(), K, 0, M, N, K, _lib.EPI_BIAS_GELU, U.data_ptr(), N,
                                   Act.data_ptr(), N, bias.data_ptr(), None, 0, None, 0, 0, 0, s), "gemm")
    for _ in range(5):
        launch()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    torch.cuda.synchronize()
    n = 20
    e0.record()
    for _ in range(n):
        launch()
    e1.record()
    torch.cuda.synchronize()
    ms = e0.elapsed_time(e1) / n
    tf = 2.0 * M * N * K / (ms * 1e-3) / 1e12
    return {"bound": "tensor", "kernel": "gemm_tc2_kernel<EPI_BIAS_GELU> (CTA-pair 256x256 tiles; FFN1 + bias + erf-GELU and its derivative, M=%d N=%d K=%d)" % (M, N, K),
            "achieved": tf, "peak": pk["burst"], "unit": "TFLOP/s", "frac": tf / pk["burst"],
            "peak_source": pk["source"] + " bf16_tflops (burst: kernel timed alone)", "ms_per_launch": ms,
            # dram__bytes_read.sum + dram__bytes_write.sum of this kernel at this shape, one ncu --set full capture
            # (profiles/r01_c_ncu_gemm_pair.txt: 67.7 MB read + 417.6 MB written); other shapes were not captured
            "traffic": 485.3e6 if (M, N, K) == (28672, 4096, 1024) else None,
            "algorithmic_bytes": 2.0 * (M * K + N * K + 2 * M * N)}


def run_product(args, D, H, layers, GB, Lq, Lk):
    import torch.distributed as dist
    import mocov2_whisper_flamingo_b200 as G
    from mocov2_whisper_flamingo_b200.parallel import DataParallelGrads, shard_batch
    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if not torch.cuda.is_available():
        raise SystemExit("bench.py: no CUDA device (the product path has no CPU fallback)")
    torch.cuda.set_device(local)
    if world > 1:
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    b0, b1 = shard_batch(GB, rank, world)
    B = b1 - b0
    torch.manual_seed(1234)
    model = G.GatedCrossModalFusion(D, H, layers, dropout=args.dropout)   # reference default: config.py:21 fusion_dropout 0.1
    with torch.no_grad():                      # non-trivial biases / LN affine, nonzero gates
        for n, p in model.named_parameters():
            if n.endswith("attn_gate"):
                p.fill_(0.7)
            elif n.endswith("ff_gate"):
                p.fill_(-0.4)
            elif "ln" in n and n.endswith("weight"):
                p.copy_(1 + 0.1 * torch.randn_like(p))
            elif n.endswith("bias"):
                p.uniform_(-0.1, 0.1)
    model.cuda().train()
    G.set_default_precision("bf16")
    ddp = DataParallelGrads(model) if world > 1 else None
    if ddp:
        ddp.broadcast_parameters()
        # communicator set-up, not part of any step: NCCL connects its channels lazily on the first large collectives of
        # each size class, which otherwise lands in the warm-up / first timed steps of a fresh process
        probe = torch.zeros(sum(p.numel() for p in model.layers[0].parameters()), device="cuda")
        for _ in range(4):
            dist.all_reduce(probe, op=dist.ReduceOp.AVG)
        torch.cuda.synchronize()
        del probe
    g = torch.Generator().manual_seed(1234 + rank)
    audio_h = torch.randn(B, Lq, D, generator=g).pin_memory()
    video_h = torch.randn(B, Lk, D, generator=g).pin_memory()
    mask_h = torch.zeros(B, Lk, dtype=torch.bool).pin_memory()       # len_b = Lk, mask still passed (8d2)
    loss_h = torch.zeros(1).pin_memory()
    audio_d, video_d, mask_d = audio_h.cuda(), video_h.cuda(), mask_h.cuda()
    dy = torch.randn(B, Lq, D, device="cuda")
    audio_d.requires_grad_(True)
    video_d.requires_grad_(True)

    # e2e: the H2D copy of a step's inputs is part of every timed step.  Like any input pipeline (pinned memory +
    # non_blocking copies) it is double-buffered: the copy for step i+1 runs on a copy stream while step i computes.
    bufs = [(audio_d, video_d, mask_d),
            (audio_d.detach().clone().requires_grad_(True), video_d.detach().clone().requires_grad_(True), mask_d.clone())]
    copy_stream = torch.cuda.Stream()
    ready = [torch.cuda.Event(), torch.cuda.Event()]
    free = [torch.cuda.Event(), torch.cuda.Event()]
    state = {"i": 0, "prefetched": None}

    def prefetch(k):
        with torch.cuda.stream(copy_stream), torch.no_grad():
            copy_stream.wait_event(free[k])                 # the step that last read this buffer has been enqueued & done
            bufs[k][0].copy_(audio_h, non_blocking=True)
            bufs[k][1].copy_(video_h, non_blocking=True)
            bufs[k][2].copy_(mask_h, non_blocking=True)
            ready[k].record(copy_stream)
        state["prefetched"] = k

    def step(e2e: bool, more: bool = True):
        k = state["i"] & 1
        state["i"] += 1
        a_d, v_d, m_d = bufs[k] if e2e else bufs[0]
        if e2e:
            if state["prefetched"] != k:
                prefetch(k)                                 # first step of a timed region: nothing to hide behind
            torch.cuda.current_stream().wait_event(ready[k])
            if more:
                prefetch(k ^ 1)                             # next step's inputs, overlapped with this step's kernels
        model.invalidate_weight_cache()
        model.zero_grad(set_to_none=True)
        a_d.grad = v_d.grad = None
        y = model(a_d, v_d, video_mask=m_d)
        loss = (y * dy).sum()
        loss.backward()
        if ddp:
            ddp.finish()
        if e2e:
            free[k].record(torch.cuda.current_stream())
            loss_h.copy_(loss.detach().view(1), non_blocking=True)
            torch.cuda.current_stream().synchronize()      # the user reads the loss every step
            return float(loss_h[0])
        return None

    def timed(e2e: bool, steps: int, sampler=None):
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        if sampler:
            sampler.start()
        e0.record()
        state["prefetched"] = None
        for i in range(steps):
            step(e2e, more=i + 1 < steps)
        e1.record()
        torch.cuda.synchronize()
        if world > 1:
            dist.barrier()
        clocks = sampler.stop() if sampler else None
        ms = torch.tensor([e0.elapsed_time(e1) / steps], device="cuda")
        if world > 1:
            dist.all_reduce(ms, op=dist.ReduceOp.MAX)
        return float(ms[0]), clocks

    for _ in range(max(args.warmup, 3)):
        step(False)
    l0 = G.launch_count()
    ms, clocks = timed(False, args.steps, ClockSampler(local) if rank == 0 else None)
    launches = G.launch_count() - l0
    step(True)
    ms_e2e, clocks_e2e = timed(True, args.steps, ClockSampler(local) if rank == 0 else None)
    # diagnostic, outside every timed region: what the box's host->device link gives for this step's pinned inputs alone
    # (e2e hides the copy behind the previous step only if 314 MB move in less than a step)
    h0, h1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    torch.cuda.synchronize()
    h0.record()
    bufs[1][0].detach().copy_(audio_h, non_blocking=True)
    bufs[1][1].detach().copy_(video_h, non_blocking=True)
    h1.record()
    torch.cuda.synchronize()
    h2d_gbs = (audio_h.numel() + video_h.numel()) * 4 / (h0.elapsed_time(h1) * 1e-3) / 1e9
    pk = peaks()
    tokens = GB * Lq
    flops_gpu = 3 * layers * block_fwd_flops(B, Lq, Lk, D)
    roof = time_dominant_kernel(B * Lq, D, pk)
    out = None
    if rank == 0:
        step_tf = flops_gpu / (ms * 1e-3) / 1e12
        out = {
            "metric": METRIC, "value": tokens / (ms * 1e-3), "unit": "tokens/s", "n_gpus": world, "steps": args.steps,
            "warmup": max(args.warmup, 3), "ms_per_step": ms, "higher_is_better": True, "scaling": "strong",
            "vs_baseline": None, "dtype": "bf16", "data": "synthetic",
            "config": {"workload": args.workload, "d_model": D, "heads": H, "layers": layers, "global_batch": GB,
                       "per_gpu_batch": B, "Lq": Lq, "Lk": Lk, "parallelism": f"dp{world}", "dropout": args.dropout, "mode": "train",
                       "step": "fwd + bwd (dgrad+wgrad) + weight bf16 re-cast" + (" + per-layer NCCL grad all-reduce" if world > 1 else ""),
                       "l2": "per-step working set (>30 GB of activations) far exceeds the 126 MB L2; no flush needed"},
            "clocks": clocks, "gpu_launches": launches,
            "e2e": {"value": tokens / (ms_e2e * 1e-3), "unit": "tokens/s", "ms_per_step": ms_e2e,
                    "h2d_bytes_per_step": (audio_h.numel() + video_h.numel()) * 4 + mask_h.numel(),
                    "d2h_bytes_per_step": 4, "h2d_link_gbs_measured": h2d_gbs, "clocks": clocks_e2e},
            "roofline": roof,
            "step_roofline": {"bound": "tensor", "achieved": step_tf, "peak": pk["sustained"], "unit": "TFLOP/s",
                              "frac": step_tf / pk["sustained"], "frac_of_2250_spec": step_tf / 2250.0,
                              "flops_per_step_per_gpu": flops_gpu,
                              "peak_source": pk["source"] + " bf16_tflops_sustained (kernels timed inside a long step)"},
        }
    if world > 1:
        dist.barrier()
    if rank == 0:
        if world == 1 and not args.no_cpu_baseline:
            threads = os.cpu_count() or 1
            v, t, sample = cpu_port_tokens_per_s(D, H, layers, Lq, Lk, threads, dropout=args.dropout)
            out["cpu_baseline"] = {"value": v, "unit": "tokens/s", "cores": threads, "kind": "port", "sample": sample,
                                   "seconds_per_sample": t}
        print(json.dumps(out))
    if world > 1:
        dist.destroy_process_group()


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=5)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="gxa", choices=["gxa", "reference"])
    ap.add_argument("--workload", default="whisper-medium-c3", choices=sorted(WORKLOADS))
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--dropout", type=float, default=0.1, help="training-mode dropout of the stack (reference config.py: 0.1)")
    args = ap.parse_args()
    D, H, layers, GB, Lq, Lk = WORKLOADS[args.workload]
    if args.impl == "reference":
        run_reference(args, D, H, layers, GB, Lq, Lk)
    else:
        run_product(args, D, H, layers, GB, Lq, Lk)


if __name__ == "__main__":
    main()
