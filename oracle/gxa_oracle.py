"""CPU oracle for the gated cross-attention hot path.  TEST INFRASTRUCTURE ONLY.

This file restates, on the CPU, the arithmetic of the reference's
``GatedCrossAttentionBlock`` / ``GatedCrossModalFusion``
(/root/reference/models/gate_cross_attention.py:12-73) and of the
``nn.MultiheadAttention`` call it delegates to
(torch/nn/functional.py:5848-5865 in-projection, :6608-6619 mask merge,
:6676-6691 SDPA + out-proj).  Only ``tests/``, ``__graft_entry__.smoke()`` and
``bench.py``'s cpu_baseline / ``--impl reference`` leg may import it; the product
package never does.

Two restatements live here:

* ``block_forward`` / ``block_backward`` / ``fusion_forward`` / ``fusion_backward``:
  numpy float64, explicit formulas, no autograd.  This is
  the parity checker.
* ``TorchPort``: the same op sequence written with torch CPU fp32 functional ops
  (layer_norm, linear, scaled_dot_product_attention, gelu).  It dispatches to the
  same ATen CPU kernels the reference module reaches and is what ``bench.py``
  times as the CPU baseline (kind "port") on the GPU box, where /root/reference
  does not exist.

Pinning: the reference ships no tests or golden vectors, so
the oracle is pinned against outputs of the *live reference module* generated in
the build container by ``oracle/make_golden.py`` and committed under
``tests/golden/``; ``tests/test_oracle.py`` replays them.
"""
from __future__ import annotations

import math

import numpy as np

try:  # scipy is in the image; keep a slow fallback so the oracle never silently changes
    from scipy.special import erf as _erf
except Exception:  # pragma: no cover
    _erf = np.vectorize(math.erf, otypes=[np.float64])

LN_EPS = 1e-5           # torch/nn/modules/normalization.py:192 default
HEAD_DIM = 64           # every config of BASELINE.json has d_model / n_heads == 64

BLOCK_KEYS = (
    # registration order of the reference block (gate_cross_attention.py:16-27)
    "attn_gate", "ff_gate",
    "attn.in_proj_weight", "attn.in_proj_bias",
    "attn.out_proj.weight", "attn.out_proj.bias",
    "attn_ln.weight", "attn_ln.bias",
    "ff_ln.weight", "ff_ln.bias",
    "ff.0.weight", "ff.0.bias", "ff.2.weight", "ff.2.bias",
)


def block_param_shapes(d: int) -> dict:
    return {
        "attn_gate": (1,), "ff_gate": (1,),
        "attn.in_proj_weight": (3 * d, d), "attn.in_proj_bias": (3 * d,),
        "attn.out_proj.weight": (d, d), "attn.out_proj.bias": (d,),
        "attn_ln.weight": (d,), "attn_ln.bias": (d,),
        "ff_ln.weight": (d,), "ff_ln.bias": (d,),
        "ff.0.weight": (4 * d, d), "ff.0.bias": (4 * d,),
        "ff.2.weight": (d, 4 * d), "ff.2.bias": (d,),
    }


def fusion_param_shapes(d: int, n_layers: int) -> dict:
    # gate_cross_attention.py:56-59
    out = {"audio_proj.weight": (d, d), "audio_proj.bias": (d,),
           "video_proj.weight": (d, d), "video_proj.bias": (d,)}
    for i in range(n_layers):
        for k, s in block_param_shapes(d).items():
            out[f"layers.{i}.{k}"] = s
    out["ln_post.weight"] = (d,)
    out["ln_post.bias"] = (d,)
    return out


# ----------------------------------------------------------------------------
# synthetic inputs; numpy RNG so that they are identical on
# every machine and independent of torch's generator
# ----------------------------------------------------------------------------
def make_params(shapes: dict, seed: int, attn_gate: float = 0.7, ff_gate: float = -0.4) -> dict:
    rng = np.random.RandomState(seed)
    p = {}
    for k, s in shapes.items():
        if k.endswith("attn_gate"):
            p[k] = np.full(s, attn_gate, np.float32)
        elif k.endswith("ff_gate"):
            p[k] = np.full(s, ff_gate, np.float32)
        elif k.endswith("ln.weight") or k.endswith("ln_post.weight"):
            p[k] = (1.0 + 0.1 * rng.standard_normal(s)).astype(np.float32)
        elif k.endswith("ln.bias") or k.endswith("ln_post.bias"):
            p[k] = (0.1 * rng.standard_normal(s)).astype(np.float32)
        elif k.endswith("bias"):
            p[k] = rng.uniform(-0.1, 0.1, s).astype(np.float32)
        else:  # weight matrix [out, in]: uniform with the xavier/kaiming scale of the default init
            bound = 1.0 / math.sqrt(s[1])
            p[k] = rng.uniform(-bound, bound, s).astype(np.float32)
    return p


def make_inputs(B, Lq, Lk, D, seed, lengths=None, garbage=True):
    """x, xa, mask(bool, True = padded key), dy.  Padded xa rows hold 1e3-scale garbage."""
    rng = np.random.RandomState(seed + 7919)
    x = rng.standard_normal((B, Lq, D)).astype(np.float32)
    xa = rng.standard_normal((B, Lk, D)).astype(np.float32)
    dy = rng.standard_normal((B, Lq, D)).astype(np.float32)
    if lengths is None:
        lengths = [Lk] * B
    lengths = np.asarray(lengths, np.int64)
    mask = np.arange(Lk)[None, :] >= lengths[:, None]      # av_net.py:111-114
    if garbage and mask.any():
        g = (1e3 * rng.standard_normal((B, Lk, D))).astype(np.float32)
        xa = np.where(mask[:, :, None], g, xa)
    return x, xa, mask, dy


# ----------------------------------------------------------------------------
# float64 restatement
# ----------------------------------------------------------------------------
def _f64(a):
    return np.asarray(a, dtype=np.float64)


def ln_fwd(x, w, b):
    mu = x.mean(-1, keepdims=True)
    var = ((x - mu) ** 2).mean(-1, keepdims=True)          # biased variance
    rstd = 1.0 / np.sqrt(var + LN_EPS)
    xhat = (x - mu) * rstd
    return xhat * w + b, (xhat, rstd)


def ln_bwd(g, cache, w):
    xhat, rstd = cache
    gw = g * w
    dx = rstd * (gw - gw.mean(-1, keepdims=True) - xhat * (gw * xhat).mean(-1, keepdims=True))
    red = tuple(range(g.ndim - 1))
    return dx, (g * xhat).sum(red), g.sum(red)


def gelu_fwd(u):
    return 0.5 * u * (1.0 + _erf(u / math.sqrt(2.0)))       # exact erf GELU (activation.py:808)


def gelu_grad(u):
    cdf = 0.5 * (1.0 + _erf(u / math.sqrt(2.0)))
    pdf = np.exp(-0.5 * u * u) / math.sqrt(2.0 * math.pi)
    return cdf + u * pdf


def _split_heads(t, H):
    B, L, D = t.shape
    return t.reshape(B, L, H, D // H).transpose(0, 2, 1, 3)   # [B,H,L,hd]


def _merge_heads(t):
    B, H, L, hd = t.shape
    return t.transpose(0, 2, 1, 3).reshape(B, L, H * hd)


def block_forward(p, x, xa, mask, n_heads, drop=None):
    """One GatedCrossAttentionBlock forward (gate_cross_attention.py:28-50).

    drop: None (eval / dropout = 0) or (attn_m [B,H,Lq,Lk], ffn_m [B,Lq,D]): the training-mode dropout multipliers
    (0 or 1/(1-p)) applied to the attention probabilities inside SDPA (:16) and to the ff.2 output (:26).

    p: dict of the 14 block tensors.  mask: bool [B,Lk], True = key ignored, or None.
    xa None => FFN only (:39).  Returns (y, cache)."""
    P = {k: _f64(v) for k, v in p.items()}
    x = _f64(x)
    D = x.shape[-1]
    c = {"x": x, "H": n_heads}
    if xa is not None:
        xa = _f64(xa)
        if xa.ndim == 4 and xa.shape[1] == 1:              # :35-36
            xa = xa[:, 0]
        if mask is not None and mask.ndim == 3 and mask.shape[1] == 1:   # :37-38
            mask = mask[:, 0]
        B, Lk, _ = xa.shape
        if mask is None:
            mask = np.zeros((B, Lk), bool)
        W, bi = P["attn.in_proj_weight"], P["attn.in_proj_bias"]
        h, c["ln1"] = ln_fwd(x, P["attn_ln.weight"], P["attn_ln.bias"])
        q = h @ W[:D].T + bi[:D]
        k = xa @ W[D:2 * D].T + bi[D:2 * D]
        v = xa @ W[2 * D:].T + bi[2 * D:]
        qh, kh, vh = _split_heads(q, n_heads), _split_heads(k, n_heads), _split_heads(v, n_heads)
        S = np.einsum("bhqd,bhkd->bhqk", qh, kh) / math.sqrt(D // n_heads)
        S = np.where(mask[:, None, None, :], -np.inf, S)
        m = S.max(-1, keepdims=True)
        m = np.where(np.isfinite(m), m, 0.0)
        e = np.exp(S - m)
        den = e.sum(-1, keepdims=True)
        # fully masked row: torch 2.11 SDPA yields a zero context
        Pm = np.where(den > 0, e / np.where(den > 0, den, 1.0), 0.0)
        attn_m = None if drop is None else _f64(drop[0])
        Pd = Pm if attn_m is None else Pm * attn_m
        ctx = _merge_heads(np.einsum("bhqk,bhkd->bhqd", Pd, vh))
        o = ctx @ P["attn.out_proj.weight"].T + P["attn.out_proj.bias"]
        ga = math.tanh(float(P["attn_gate"][0]))
        x1 = x + ga * o
        c.update(xa=xa, mask=mask, h=h, qh=qh, kh=kh, vh=vh, P=Pm, Pd=Pd, attn_m=attn_m, ctx=ctx, o=o, ga=ga)
    else:
        x1 = x
    h2, c["ln2"] = ln_fwd(x1, P["ff_ln.weight"], P["ff_ln.bias"])
    u = h2 @ P["ff.0.weight"].T + P["ff.0.bias"]
    a = gelu_fwd(u)
    f = a @ P["ff.2.weight"].T + P["ff.2.bias"]
    ffn_m = None if drop is None else _f64(drop[1])
    if ffn_m is not None:
        f = f * ffn_m
    gf = math.tanh(float(P["ff_gate"][0]))
    y = x1 + gf * f
    c.update(x1=x1, h2=h2, u=u, a=a, f=f, gf=gf, ffn_m=ffn_m, has_attn=xa is not None)
    return y, c


def block_backward(p, c, dy):
    """Explicit backward.  Returns (grads dict, dx, dxa or None)."""
    P = {k: _f64(v) for k, v in p.items()}
    dy = _f64(dy)
    D = dy.shape[-1]
    H = c["H"]
    g = {}
    flat = lambda t: t.reshape(-1, t.shape[-1])
    gf = c["gf"]
    g["ff_gate"] = np.array([(1 - gf * gf) * (dy * c["f"]).sum()])
    # A gate gradient is ONE scalar: the sum of B*Lq*D signed terms.  Its natural magnitude is the 2-norm of the terms
    # (what a random-sign sum of them has); the total itself can be cancelled far below that, which makes "relative to
    # the total" an ill-conditioned yardstick for a single number.  Kept beside the gradients for the parity tests.
    c["gate_scale"] = {"ff_gate": (1 - gf * gf) * float(np.sqrt(((dy * c["f"]) ** 2).sum()))}
    df = gf * dy
    if c.get("ffn_m") is not None:
        df = df * c["ffn_m"]
    g["ff.2.weight"] = flat(df).T @ flat(c["a"])
    g["ff.2.bias"] = flat(df).sum(0)
    da = df @ P["ff.2.weight"]
    du = da * gelu_grad(c["u"])
    g["ff.0.weight"] = flat(du).T @ flat(c["h2"])
    g["ff.0.bias"] = flat(du).sum(0)
    dh2 = du @ P["ff.0.weight"]
    dln, g["ff_ln.weight"], g["ff_ln.bias"] = ln_bwd(dh2, c["ln2"], P["ff_ln.weight"])
    dx1 = dy + dln
    if not c["has_attn"]:
        for k in ("attn_gate", "attn.in_proj_weight", "attn.in_proj_bias", "attn.out_proj.weight",
                  "attn.out_proj.bias", "attn_ln.weight", "attn_ln.bias"):
            g[k] = None
        return g, dx1, None
    ga = c["ga"]
    g["attn_gate"] = np.array([(1 - ga * ga) * (dx1 * c["o"]).sum()])
    c["gate_scale"]["attn_gate"] = (1 - ga * ga) * float(np.sqrt(((dx1 * c["o"]) ** 2).sum()))
    do = ga * dx1
    g["attn.out_proj.weight"] = flat(do).T @ flat(c["ctx"])
    g["attn.out_proj.bias"] = flat(do).sum(0)
    dctx = _split_heads(do @ P["attn.out_proj.weight"], H)
    Pm, qh, kh, vh = c["P"], c["qh"], c["kh"], c["vh"]
    scale = 1.0 / math.sqrt(D // H)
    dv = np.einsum("bhqk,bhqd->bhkd", c.get("Pd", Pm), dctx)
    dP = np.einsum("bhqd,bhkd->bhqk", dctx, vh)
    if c.get("attn_m") is not None:
        dP = dP * c["attn_m"]
    delta = (dP * Pm).sum(-1, keepdims=True)
    dS = Pm * (dP - delta)
    dq = _merge_heads(np.einsum("bhqk,bhkd->bhqd", dS, kh)) * scale
    dk = _merge_heads(np.einsum("bhqk,bhqd->bhkd", dS, qh)) * scale
    dv = _merge_heads(dv)
    W = P["attn.in_proj_weight"]
    xa = c["xa"]
    g["attn.in_proj_weight"] = np.concatenate(
        [flat(dq).T @ flat(c["h"]), flat(dk).T @ flat(xa), flat(dv).T @ flat(xa)], 0)
    g["attn.in_proj_bias"] = np.concatenate([flat(dq).sum(0), flat(dk).sum(0), flat(dv).sum(0)])
    dh = dq @ W[:D]
    dxa = dk @ W[D:2 * D] + dv @ W[2 * D:]
    dln1, g["attn_ln.weight"], g["attn_ln.bias"] = ln_bwd(dh, c["ln1"], P["attn_ln.weight"])
    return g, dx1 + dln1, dxa


def _sub(p, prefix):
    n = len(prefix)
    return {k[n:]: v for k, v in p.items() if k.startswith(prefix)}


def fusion_forward(p, audio, video, video_mask, n_heads, n_layers):
    """GatedCrossModalFusion.forward (gate_cross_attention.py:60-73); audio_mask is ignored there."""
    audio, video = _f64(audio), _f64(video)
    x = audio @ _f64(p["audio_proj.weight"]).T + _f64(p["audio_proj.bias"])
    xa = video @ _f64(p["video_proj.weight"]).T + _f64(p["video_proj.bias"])
    caches = []
    for i in range(n_layers):
        x, c = block_forward(_sub(p, f"layers.{i}."), x, xa, video_mask, n_heads)
        caches.append(c)
    y, lnc = ln_fwd(x, _f64(p["ln_post.weight"]), _f64(p["ln_post.bias"]))
    return y, dict(caches=caches, ln=lnc, audio=audio, video=video, n_layers=n_layers)


def fusion_backward(p, c, dy):
    g = {}
    flat = lambda t: t.reshape(-1, t.shape[-1])
    dx, g["ln_post.weight"], g["ln_post.bias"] = ln_bwd(_f64(dy), c["ln"], _f64(p["ln_post.weight"]))
    dxa = 0.0
    for i in reversed(range(c["n_layers"])):
        gi, dx, dxa_i = block_backward(_sub(p, f"layers.{i}."), c["caches"][i], dx)
        dxa = dxa + dxa_i
        for k, v in gi.items():
            g[f"layers.{i}.{k}"] = v
    g["audio_proj.weight"] = flat(dx).T @ flat(c["audio"])
    g["audio_proj.bias"] = flat(dx).sum(0)
    g["video_proj.weight"] = flat(dxa).T @ flat(c["video"])
    g["video_proj.bias"] = flat(dxa).sum(0)
    d_audio = dx @ _f64(p["audio_proj.weight"])
    d_video = dxa @ _f64(p["video_proj.weight"])
    return g, d_audio, d_video


def gate_scales(c) -> dict:
    """name -> 2-norm of the terms a gate gradient sums (after block_backward / fusion_backward ran on cache c)."""
    if "caches" in c:
        return {f"layers.{i}.{k}": v for i, ci in enumerate(c["caches"]) for k, v in ci.get("gate_scale", {}).items()}
    return dict(c.get("gate_scale", {}))


def rel_err(a, b):
    """max|a-b| / max|b|  -- the relative error every parity test in tests/ uses."""
    a, b = _f64(a), _f64(b)
    den = np.abs(b).max()
    return float(np.abs(a - b).max() / (den if den > 0 else 1.0))
