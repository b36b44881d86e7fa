"""CPU port of the reference hot path with torch fp32 functional ops.  TEST INFRASTRUCTURE ONLY.

Same op sequence as /root/reference/models/gate_cross_attention.py:28-50 and the
``F.multi_head_attention_forward`` it reaches (torch/nn/functional.py:6244-6691):
layer_norm -> packed in-projection -> additive -inf key-padding mask -> SDPA ->
out-proj -> tanh-gated residual -> layer_norm -> Linear/GELU(erf)/Linear -> gated
residual.  It therefore dispatches to the same ATen CPU kernels (MKL GEMM,
_scaled_dot_product_flash_attention_for_cpu, vectorised GELU/LN) the reference
module does, which makes it a fair timed CPU baseline (bench.py cpu_baseline kind
"port") on the GPU box where /root/reference is absent.  Autograd provides the
backward exactly as it does for the reference.  Never imported by the product.
"""
from __future__ import annotations

import torch
import torch.nn.functional as F


def block_forward(p: dict, x, xa, mask, n_heads: int, dropout_p: float = 0.0, training: bool = False, taps=None):
    """p maps the 14 state_dict keys of a block to torch tensors (fp32, CPU).  taps: optional dict that receives the
    attention output `o` and the FFN output `f` with retain_grad() set (tests: natural scale of the gate gradients)."""
    B, Lq, D = x.shape
    hd = D // n_heads
    if xa is not None:
        if xa.dim() == 4 and xa.size(1) == 1:
            xa = xa.squeeze(1)
        if mask is not None and mask.dim() == 3 and mask.size(1) == 1:
            mask = mask.squeeze(1)
        Lk = xa.shape[1]
        W, bi = p["attn.in_proj_weight"], p["attn.in_proj_bias"]
        h = F.layer_norm(x.float(), (D,), p["attn_ln.weight"], p["attn_ln.bias"]).type(x.dtype)
        q = F.linear(h, W[:D], bi[:D])
        kv = F.linear(xa, W[D:], bi[D:])
        k, v = kv[..., :D], kv[..., D:]
        q = q.view(B, Lq, n_heads, hd).transpose(1, 2)
        k = k.reshape(B, Lk, n_heads, hd).transpose(1, 2)
        v = v.reshape(B, Lk, n_heads, hd).transpose(1, 2)
        am = None
        if mask is not None:
            am = torch.zeros(B, 1, 1, Lk, dtype=q.dtype).masked_fill_(mask.view(B, 1, 1, Lk), float("-inf"))
        ctx = F.scaled_dot_product_attention(q, k, v, am, dropout_p if training else 0.0)
        ctx = ctx.transpose(1, 2).reshape(B, Lq, D)
        o = F.linear(ctx, p["attn.out_proj.weight"], p["attn.out_proj.bias"])
        if taps is not None and o.requires_grad:
            o.retain_grad()
            taps["o"] = o
        x = x + o * p["attn_gate"].tanh()
    h2 = F.layer_norm(x.float(), (D,), p["ff_ln.weight"], p["ff_ln.bias"]).type(x.dtype)
    f = F.linear(F.gelu(F.linear(h2, p["ff.0.weight"], p["ff.0.bias"])), p["ff.2.weight"], p["ff.2.bias"])
    f = F.dropout(f, dropout_p, training)
    if taps is not None and f.requires_grad:
        f.retain_grad()
        taps["f"] = f
    return x + f * p["ff_gate"].tanh()


def gate_scales(p: dict, taps: list) -> dict:
    """After backward: 2-norm of the terms each gate gradient sums (see oracle.gxa_oracle.block_backward)."""
    out = {}
    for i, t in enumerate(taps):
        for key, name in (("o", "attn_gate"), ("f", "ff_gate")):
            if key in t and t[key].grad is not None:
                g = float(p[f"layers.{i}.{name}"].detach().tanh())
                # d(o) = tanh(g) * dx1  =>  terms of d(gate) = (1 - tanh^2) * dx1 * o = (1 - tanh^2) / tanh * d(o) * o
                out[f"layers.{i}.{name}"] = abs((1 - g * g) / g) * float((t[key].grad * t[key].detach()).norm())
    return out


def fusion_forward(p: dict, audio, video, video_mask, n_heads: int, n_layers: int,
                   dropout_p: float = 0.0, training: bool = False, taps=None):
    D = audio.shape[-1]
    x = F.linear(audio, p["audio_proj.weight"], p["audio_proj.bias"])
    xa = F.linear(video, p["video_proj.weight"], p["video_proj.bias"])
    for i in range(n_layers):
        pre = f"layers.{i}."
        sub = {k[len(pre):]: v for k, v in p.items() if k.startswith(pre)}
        t = None
        if taps is not None:
            t = {}
            taps.append(t)
        x = block_forward(sub, x, xa, video_mask, n_heads, dropout_p, training, taps=t)
    return F.layer_norm(x.float(), (D,), p["ln_post.weight"], p["ln_post.bias"]).type(x.dtype)


def to_torch(p_np: dict, requires_grad: bool = False) -> dict:
    return {k: torch.from_numpy(v.copy()).requires_grad_(requires_grad) for k, v in p_np.items()}
