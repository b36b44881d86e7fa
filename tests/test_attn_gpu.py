"""Masked-softmax attention kernels (gxa_attn_fwd / gxa_attn_bwd) vs fp32 torch math on the same bf16
inputs.  Mask handling must be exact: padded keys contribute exactly nothing and get exactly zero grad.
Tolerance 1.5e-2 relative (bf16 output rounding and bf16 probabilities in the PV product)."""

import pytest
import torch

from mocov2_whisper_flamingo_b200 import _lib
from gpu_util import ptr, rel_err, stream

pytestmark = pytest.mark.gpu


def ref_attn(q, k, v, mask, H):
    B, Lq, D = q.shape
    Lk = k.shape[1]
    qh = q.float().view(B, Lq, H, 64).transpose(1, 2)
    kh = k.float().view(B, Lk, H, 64).transpose(1, 2)
    vh = v.float().view(B, Lk, H, 64).transpose(1, 2)
    S = qh @ kh.transpose(-1, -2) / 8.0
    if mask is not None:
        S = S.masked_fill(mask[:, None, None, :], float("-inf"))
    P = torch.softmax(S, -1)
    P = torch.nan_to_num(P, nan=0.0)          # fully masked sample -> zero context (torch 2.11 SDPA behaviour)
    return (P @ vh).transpose(1, 2).reshape(B, Lq, D)


def run(B, H, Lq, Lk, lengths, seed=0, garbage=True):
    L = _lib.lib()
    D = 64 * H
    g = torch.Generator(device="cuda").manual_seed(seed)
    q = torch.randn(B, Lq, D, device="cuda", generator=g).to(torch.bfloat16)
    kv = torch.randn(B, Lk, 2 * D, device="cuda", generator=g).to(torch.bfloat16)
    dctx = torch.randn(B, Lq, D, device="cuda", generator=g).to(torch.bfloat16)
    mask = kvend = None
    if lengths is not None:
        ln = torch.tensor(lengths, device="cuda")
        mask = torch.arange(Lk, device="cuda")[None, :] >= ln[:, None]
        if garbage:
            kv = torch.where(mask[:, :, None], (kv.float() * 1e3).to(torch.bfloat16), kv)
        kvend = torch.empty(B, dtype=torch.int32, device="cuda")
        m8 = mask.contiguous().view(torch.uint8)
        _lib.check(L.gxa_mask_prepare(ptr(m8), ptr(kvend), B, Lk, stream()), "mask_prepare")
        assert kvend.tolist() == [min(int(x), Lk) for x in lengths]
    else:
        m8 = None
    k, v = kv[..., :D], kv[..., D:]
    ctx = torch.full((B, Lq, D), float("nan"), device="cuda", dtype=torch.bfloat16)
    lse = torch.empty(B, H, Lq, device="cuda")
    _lib.check(L.gxa_attn_fwd(ptr(q), D, ptr(k), ptr(v), 2 * D, ptr(m8), ptr(kvend), ptr(ctx), D, ptr(lse), B, H, Lq, Lk,
                              stream()), "attn_fwd")
    dq = torch.full_like(q, float("nan"))
    dkv = torch.full_like(kv, float("nan"))
    delta = torch.empty(B, H, Lq, device="cuda")
    wsb = torch.empty(L.gxa_attn_bwd_workspace_bytes(B, H, Lq), dtype=torch.uint8, device="cuda")
    _lib.check(L.gxa_attn_bwd(ptr(q), D, ptr(k), ptr(v), 2 * D, ptr(m8), ptr(kvend), ptr(ctx), ptr(dctx), D, ptr(lse),
                              ptr(delta), ptr(dq), D, ptr(dkv), ptr(dkv[..., D:]), 2 * D, ptr(wsb), B, H, Lq, Lk, stream()),
               "attn_bwd")
    torch.cuda.synchronize()
    qr, kr, vr = (t.float().clone().requires_grad_(True) for t in (q, k, v))
    ref = ref_attn(qr, kr, vr, mask, H)
    ref.backward(dctx.float())
    return dict(ctx=ctx.float(), ref=ref.detach(), dq=dq.float(), dk=dkv[..., :D].float(), dv=dkv[..., D:].float(),
                rdq=qr.grad, rdk=kr.grad, rdv=vr.grad, mask=mask, lse=lse)


@pytest.mark.parametrize("B,H,Lq,Lk,lengths", [
    (2, 2, 64, 64, None),
    (2, 6, 448, 250, None),
    (3, 2, 100, 200, [200, 77, 1]),
    (4, 4, 448, 750, [750, 60, 333, 64]),
    (2, 2, 50, 130, [130, 0]),
    # BASELINE widths (configs 2, 3, 4): every head index up to 11 / 15 / 19, six key blocks, ragged and empty samples
    (2, 12, 448, 750, [750, 60]),
    (3, 16, 448, 750, [750, 333, 0]),
    (2, 20, 448, 750, [61, 750]),
])
def test_attention_fwd_bwd(B, H, Lq, Lk, lengths):
    r = run(B, H, Lq, Lk, lengths)
    assert torch.isfinite(r["ctx"]).all()
    assert rel_err(r["ctx"], r["ref"]) < 1.5e-2
    for a, b in (("dq", "rdq"), ("dk", "rdk"), ("dv", "rdv")):
        assert torch.isfinite(r[a]).all(), a
        assert rel_err(r[a], r[b]) < 1.5e-2, a
    if r["mask"] is not None:
        m = r["mask"]
        assert (r["dk"][m] == 0).all() and (r["dv"][m] == 0).all()      # bit-exact zero grad on padded keys
        for b in range(B):
            if lengths[b] == 0:                                           # fully masked sample
                assert (r["ctx"][b] == 0).all() and (r["dq"][b] == 0).all()


def test_padded_keys_have_bit_exact_zero_influence():
    a = run(3, 2, 100, 200, [200, 77, 1], seed=5, garbage=True)
    b = run(3, 2, 100, 200, [200, 77, 1], seed=5, garbage=False)
    assert torch.equal(a["ctx"], b["ctx"]) and torch.equal(a["dq"], b["dq"])
    assert torch.equal(a["lse"], b["lse"])


# ---------------------------------------------------------------------------------------------- dropout
def run_dropout(B, H, Lq, Lk, lengths, p, seed64, seed=0):
    """gxa_attn_fwd_dropout / gxa_attn_bwd_dropout vs torch math with the SAME mask (tests/dropout_ref.py)."""
    import numpy as np
    import dropout_ref as DR
    L = _lib.lib()
    D = 64 * H
    g = torch.Generator(device="cuda").manual_seed(seed)
    q = torch.randn(B, Lq, D, device="cuda", generator=g).to(torch.bfloat16)
    kv = torch.randn(B, Lk, 2 * D, device="cuda", generator=g).to(torch.bfloat16)
    dctx = torch.randn(B, Lq, D, device="cuda", generator=g).to(torch.bfloat16)
    mask = kvend = m8 = None
    if lengths is not None:
        ln = torch.tensor(lengths, device="cuda")
        mask = torch.arange(Lk, device="cuda")[None, :] >= ln[:, None]
        kvend = torch.empty(B, dtype=torch.int32, device="cuda")
        m8 = mask.contiguous().view(torch.uint8)
        _lib.check(L.gxa_mask_prepare(ptr(m8), ptr(kvend), B, Lk, stream()), "mask_prepare")
    k, v = kv[..., :D], kv[..., D:]
    ctx = torch.full((B, Lq, D), float("nan"), device="cuda", dtype=torch.bfloat16)
    lse = torch.empty(B, H, Lq, device="cuda")
    _lib.check(L.gxa_attn_fwd_dropout(ptr(q), D, ptr(k), ptr(v), 2 * D, ptr(m8), ptr(kvend), ptr(ctx), D, ptr(lse), B, H,
                                      Lq, Lk, p, seed64, stream()), "attn_fwd_dropout")
    dq = torch.full_like(q, float("nan"))
    dkv = torch.full_like(kv, float("nan"))
    delta = torch.empty(B, H, Lq, device="cuda")
    wsb = torch.empty(L.gxa_attn_bwd_workspace_bytes(B, H, Lq), dtype=torch.uint8, device="cuda")
    _lib.check(L.gxa_attn_bwd_dropout(ptr(q), D, ptr(k), ptr(v), 2 * D, ptr(m8), ptr(kvend), ptr(ctx), ptr(dctx), D,
                                      ptr(lse), ptr(delta), ptr(dq), D, ptr(dkv), ptr(dkv[..., D:]), 2 * D, ptr(wsb), B,
                                      H, Lq, Lk, p, seed64, stream()), "attn_bwd_dropout")
    torch.cuda.synchronize()
    mult = torch.from_numpy(DR.attn_multipliers(B, H, Lq, Lk, p, seed64)).cuda()
    qr, kr, vr = (t.float().clone().requires_grad_(True) for t in (q, k, v))
    qh = qr.view(B, Lq, H, 64).transpose(1, 2)
    kh = kr.view(B, Lk, H, 64).transpose(1, 2)
    vh = vr.view(B, Lk, H, 64).transpose(1, 2)
    S = qh @ kh.transpose(-1, -2) / 8.0
    if mask is not None:
        S = S.masked_fill(mask[:, None, None, :], float("-inf"))
    P = torch.nan_to_num(torch.softmax(S, -1), nan=0.0) * mult
    ref = (P @ vh).transpose(1, 2).reshape(B, Lq, D)
    ref.backward(dctx.float())
    return dict(ctx=ctx.float(), ref=ref.detach(), dq=dq.float(), dk=dkv[..., :D].float(), dv=dkv[..., D:].float(),
                rdq=qr.grad, rdk=kr.grad, rdv=vr.grad, mult=mult)


@pytest.mark.parametrize("B,H,Lq,Lk,lengths,p", [
    (2, 2, 64, 64, None, 0.1),
    (2, 6, 448, 250, None, 0.1),
    (3, 2, 100, 201, [201, 77, 30], 0.25),         # odd Lk: pair counters restart per row
    (2, 4, 448, 750, [750, 333], 0.1),
    (2, 16, 448, 750, [750, 60], 0.1),             # config 3's heads and lengths
    (2, 2, 128, 130, [130, 97], 0.01),             # threshold below the smallest normal fp16 pattern (packed compare, gxa_rng.cuh)
    (2, 2, 128, 130, None, 0.9),
])
def test_attention_dropout_matches_reference_with_same_mask(B, H, Lq, Lk, lengths, p):
    r = run_dropout(B, H, Lq, Lk, lengths, p, seed64=0x1234ABCD5678EF01)
    keep = (r["mult"] > 0).float().mean().item()
    assert abs(keep - (1 - p)) < 0.01                                     # the mask really drops ~p of the entries
    assert rel_err(r["ctx"], r["ref"]) < 1.5e-2
    for a, b in (("dq", "rdq"), ("dk", "rdk"), ("dv", "rdv")):
        assert torch.isfinite(r[a]).all(), a
        assert rel_err(r[a], r[b]) < 1.5e-2, a


def test_attention_dropout_seed_and_p0():
    a = run_dropout(2, 2, 64, 64, None, 0.1, seed64=7)
    b = run_dropout(2, 2, 64, 64, None, 0.1, seed64=7)
    c = run_dropout(2, 2, 64, 64, None, 0.1, seed64=8)
    z = run_dropout(2, 2, 64, 64, None, 0.0, seed64=7)
    n = run(2, 2, 64, 64, None)
    assert torch.equal(a["ctx"], b["ctx"]) and torch.equal(a["dq"], b["dq"])          # same seed: same bits
    assert not torch.equal(a["ctx"], c["ctx"])                                        # another seed: another mask
    assert torch.equal(z["ctx"], n["ctx"]) and torch.equal(z["dk"], n["dk"])          # p = 0 is the plain kernel
