"""SURVEY 8 d6, second half: the "existing Blackwell path" -- the reference's architecture built from stock torch modules
(nn.MultiheadAttention, nn.LayerNorm, nn.Linear, nn.GELU), eager, torch.autocast(bf16), on one B200, at the bench's
workload (config 3 by default).  Measurement tool, not a test and not part of the product: nothing here touches
libgxa_b200.so.  Prints one JSON line; run it next to `python bench.py` on the same box.

    python tests/bench_torch_eager_gpu.py [--layers 24 --batch 64 --steps 5 --warmup 3]
"""
from __future__ import annotations

import argparse
import json

import torch
import torch.nn as nn


class EagerBlock(nn.Module):
    """Same statements as the reference block (models/gate_cross_attention.py:13-50), stock torch ops only."""

    def __init__(self, d, h, p):
        super().__init__()
        self.attn = nn.MultiheadAttention(d, h, dropout=p, batch_first=True)
        self.attn_ln, self.ff_ln = nn.LayerNorm(d), nn.LayerNorm(d)
        self.attn_gate = nn.Parameter(torch.tensor([0.7]))
        self.ff_gate = nn.Parameter(torch.tensor([-0.4]))
        self.ff = nn.Sequential(nn.Linear(d, 4 * d), nn.GELU(), nn.Linear(4 * d, d), nn.Dropout(p))

    def forward(self, x, xa, mask):
        h = self.attn_ln(x.float()).type(x.dtype)
        o = self.attn(h, xa, xa, key_padding_mask=mask, need_weights=False)[0]
        x = x + o * self.attn_gate.tanh()
        h = self.ff_ln(x.float()).type(x.dtype)
        return x + self.ff(h) * self.ff_gate.tanh()


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--d", type=int, default=1024)
    ap.add_argument("--heads", type=int, default=16)
    ap.add_argument("--layers", type=int, default=24)
    ap.add_argument("--batch", type=int, default=64)
    ap.add_argument("--lq", type=int, default=448)
    ap.add_argument("--lk", type=int, default=750)
    ap.add_argument("--dropout", type=float, default=0.1)
    ap.add_argument("--steps", type=int, default=5)
    ap.add_argument("--warmup", type=int, default=3)
    a = ap.parse_args()
    dev = torch.device("cuda:0")
    torch.manual_seed(1234)
    layers = a.layers
    while True:
        try:
            mods = nn.ModuleList([EagerBlock(a.d, a.heads, a.dropout) for _ in range(layers)]).to(dev).train()
            x = torch.randn(a.batch, a.lq, a.d, device=dev, requires_grad=True)
            xa = torch.randn(a.batch, a.lk, a.d, device=dev, requires_grad=True)
            mask = torch.zeros(a.batch, a.lk, dtype=torch.bool, device=dev)
            dy = torch.randn(a.batch, a.lq, a.d, device=dev)

            def step():
                for p in mods.parameters():
                    p.grad = None
                x.grad = xa.grad = None
                with torch.autocast("cuda", dtype=torch.bfloat16):
                    y = x
                    for m in mods:
                        y = m(y, xa, mask)
                y.backward(dy)

            for _ in range(a.warmup):
                step()
            torch.cuda.synchronize()
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record()
            for _ in range(a.steps):
                step()
            e1.record()
            torch.cuda.synchronize()
            break
        except torch.OutOfMemoryError:
            del mods
            torch.cuda.empty_cache()
            if layers <= 2:
                raise
            layers //= 2
    ms = e0.elapsed_time(e1) / a.steps * (a.layers / layers)
    F = 3 * a.layers * (20 * a.batch * a.lq * a.d ** 2 + 4 * a.batch * a.lk * a.d ** 2 + 4 * a.batch * a.lq * a.lk * a.d)
    print(json.dumps({"impl": "torch-eager-autocast-bf16 (stock nn.MultiheadAttention / LayerNorm / Linear / GELU)",
                      "metric": "gated-xattn fwd+bwd tokens/s", "value": a.batch * a.lq / (ms * 1e-3), "unit": "tokens/s",
                      "ms_per_step": ms, "tflops": F / (ms * 1e-3) / 1e12, "layers_timed": layers, "layers": a.layers,
                      "extrapolated": layers != a.layers, "batch": a.batch, "Lq": a.lq, "Lk": a.lk, "d_model": a.d,
                      "dropout": a.dropout, "steps": a.steps, "warmup": a.warmup, "torch": torch.__version__,
                      "peak_mem_gb": torch.cuda.max_memory_allocated() / 2 ** 30}))


if __name__ == "__main__":
    main()
