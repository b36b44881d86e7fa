"""Oracle parity of the CUDA path at every BASELINE width (configs 2, 3, 4: d_model / heads = 768/12, 1024/16, 1280/20)
at the real sequence lengths (Lq = 448 query tokens x Lk = 750 visual frames), two layers of the stack, ragged and
fully padded samples -- forward AND every gradient, fp32 and bf16 mode.

Checker: oracle/torch_port.py run here on the host CPU in fp32 (the reference's op sequence on the same ATen CPU kernels,
autograd backward; pinned against the live reference module at these widths by tests/test_oracle_live.py in the build
container), plus the float64 numpy oracle at the 768 width.  A mis-indexed head >= 6, a descriptor bug that only shows at
N = 4096 / K = 1280, or a wrong log-sum-exp at the sixth key block would show up here at O(1), not at the tolerance.

Tolerances (BASELINE.json north_star): fp32 mode 1e-4, bf16 mode 2e-2, rel = max|a - b| / max|b|; padded keys exact.
The measured errors are written to gpurun_out/parity_errors.json (tests/gpu_util.py).
"""
import numpy as np
import pytest
import torch

from oracle import gxa_oracle as O
from oracle import torch_port as TP
import mocov2_whisper_flamingo_b200 as G
from gpu_util import np_dev, record_err, rel_err, scalar_sum_err

pytestmark = pytest.mark.gpu
TOL = {"fp32": 1e-4, "bf16": 2e-2}
WIDTHS = [(768, 12), (1024, 16), (1280, 20)]
LENGTHS = [[750, 60], [333, 0]]
B, LQ, LK, NL = 2, 448, 750, 2


def _cpu_reference(p, x, xa, mask, dy, H):
    torch.set_num_threads(max(1, torch.get_num_threads()))
    pt = TP.to_torch(p, requires_grad=True)
    xt, xat = torch.from_numpy(x).requires_grad_(True), torch.from_numpy(xa).requires_grad_(True)
    taps = []
    y = TP.fusion_forward(pt, xt, xat, torch.from_numpy(mask), H, NL, taps=taps)
    y.backward(torch.from_numpy(dy))
    return y.detach(), xt.grad, xat.grad, {k: v.grad for k, v in pt.items()}, TP.gate_scales(pt, taps)


def _gpu(mod, x, xa, mask, dy, prec):
    G.set_default_precision(prec)
    try:
        xt, xat = np_dev(x).requires_grad_(True), np_dev(xa).requires_grad_(True)
        mod.zero_grad(set_to_none=True)
        y = mod(xt, xat, video_mask=torch.from_numpy(mask).cuda())
        y.backward(np_dev(dy))
        torch.cuda.synchronize()
        return y.detach().cpu(), xt.grad.cpu(), xat.grad.cpu(), {k: v.grad.cpu() for k, v in mod.named_parameters()}
    finally:
        G.set_default_precision(None)


def _grad_err(name, mine, ref, scales=None):
    """Relative error of one gradient tensor.  The bk third of in_proj_bias is mathematically zero;
    the scalar gate gradients are measured against the natural size of the sum they are (gpu_util.scalar_sum_err)."""
    mine, ref = mine.double(), ref.double()
    if name.endswith("_gate") and scales is not None and name in scales:
        return scalar_sum_err(mine.reshape(-1)[0], ref.reshape(-1)[0], scales[name])
    if name.endswith("in_proj_bias"):
        D = ref.shape[0] // 3
        zero_third = float(mine[D:2 * D].abs().max()) / max(1.0, float(mine.abs().max()))
        idx = torch.cat([torch.arange(0, D), torch.arange(2 * D, 3 * D)])
        return max(rel_err(mine[idx], ref[idx]), zero_third)
    return rel_err(mine, ref)


@pytest.mark.parametrize("lengths", LENGTHS, ids=["ragged750_60", "ragged333_0"])
@pytest.mark.parametrize("D,H", WIDTHS)
def test_stack_parity_at_baseline_widths(D, H, lengths):
    seed = 1234 + D + lengths[1]
    p = O.make_params(O.fusion_param_shapes(D, NL), seed)
    x, xa, mask, dy = O.make_inputs(B, LQ, LK, D, seed, lengths)            # padded xa rows hold 1e3-scale garbage
    yr, dxr, dxar, gr, scales = _cpu_reference(p, x, xa, mask, dy, H)
    assert torch.isfinite(yr).all()
    mod = G.GatedCrossModalFusion(D, H, NL, dropout=0.0)
    mod.load_state_dict({k: torch.from_numpy(v) for k, v in p.items()}, strict=True)
    mod.cuda()
    pad = torch.from_numpy(mask)
    for prec in ("fp32", "bf16"):
        case = f"stack_D{D}_H{H}_len{lengths[0]}_{lengths[1]}_{prec}"
        y, dx, dxa, g = _gpu(mod, x, xa, mask, dy, prec)
        tol = TOL[prec]
        errs = {"y": rel_err(y, yr), "dx": rel_err(dx, dxr), "dxa": rel_err(dxa, dxar)}
        for k, ref in gr.items():
            errs["g:" + k] = _grad_err(k, g[k], ref, scales)
            if k.endswith("_gate"):                          # the plain relative error is recorded, not asserted
                record_err(case, "raw:" + k, rel_err(g[k], ref))
        for k, v in errs.items():
            record_err(case, k, v)
        assert torch.isfinite(y).all() and torch.isfinite(dx).all() and torch.isfinite(dxa).all()
        assert (dxa[pad] == 0).all(), "padded keys must get exactly zero gradient"
        bad = {k: v for k, v in errs.items() if not v < tol}
        assert not bad, f"{case}: above {tol:g}: " + ", ".join(f"{k} {v:.3e}" for k, v in sorted(bad.items(), key=lambda kv: -kv[1]))


def test_float64_oracle_agrees_at_768():
    """The numpy float64 restatement (explicit backward formulas) at a BASELINE width: second opinion on the checker
    itself and on the CUDA fp32 mode, one layer."""
    D, H, lengths, seed = 768, 12, [750, 60], 4321
    p = O.make_params(O.block_param_shapes(D), seed)
    x, xa, mask, dy = O.make_inputs(B, LQ, LK, D, seed, lengths)
    yo, c = O.block_forward(p, x, xa, mask, H)
    go, dxo, dxao = O.block_backward(p, c, dy)
    blk = G.GatedCrossAttentionBlock(D, H, dropout=0.0)
    blk.load_state_dict({k: torch.from_numpy(v) for k, v in p.items()}, strict=True)
    blk.cuda()
    for prec in ("fp32", "bf16"):
        G.set_default_precision(prec)
        try:
            xt, xat = np_dev(x).requires_grad_(True), np_dev(xa).requires_grad_(True)
            blk.zero_grad(set_to_none=True)
            y = blk(xt, xat, torch.from_numpy(mask).cuda())
            y.backward(np_dev(dy))
            torch.cuda.synchronize()
        finally:
            G.set_default_precision(None)
        tol = TOL[prec]
        case = f"block_f64_D{D}_{prec}"
        errs = {"y": rel_err(y.detach().cpu(), torch.from_numpy(yo)), "dx": rel_err(xt.grad.cpu(), torch.from_numpy(dxo)),
                "dxa": rel_err(xat.grad.cpu(), torch.from_numpy(dxao))}
        for k, v in blk.named_parameters():
            errs["g:" + k] = _grad_err(k, v.grad.cpu(), torch.from_numpy(np.asarray(go[k])), O.gate_scales(c))
        for k, v in errs.items():
            record_err(case, k, v)
        bad = {k: v for k, v in errs.items() if not v < tol}
        assert not bad, f"{case}: " + ", ".join(f"{k} {v:.3e}" for k, v in bad.items())


# ------------------------------------------------------------------------------------------ a9: the block's input forms
def test_block_input_forms_4d_xa_3d_mask_noncontiguous_bf16():
    """reference forward :35-38 squeezes a [B,1,Lk,D] visual memory and a [B,1,Lk] mask; inputs need not be contiguous;
    a bf16 query stream still gives an fp32 result (the [1]-shaped fp32 gate promotes the residual) and
    a bf16 gradient for x."""
    Bq, Lq, Lk, D, H, seed = 3, 40, 72, 256, 4, 77
    p = O.make_params(O.block_param_shapes(D), seed)
    x, xa, mask, dy = O.make_inputs(Bq, Lq, Lk, D, seed, [72, 5, 31])
    yo, c = O.block_forward(p, x, xa, mask, H)
    go, dxo, dxao = O.block_backward(p, c, dy)
    blk = G.GatedCrossAttentionBlock(D, H, dropout=0.0)
    blk.load_state_dict({k: torch.from_numpy(v) for k, v in p.items()}, strict=True)
    blk.cuda()
    for prec in ("fp32", "bf16"):
        G.set_default_precision(prec)
        try:
            # (1) plain call
            xt, xat, mt = np_dev(x).requires_grad_(True), np_dev(xa).requires_grad_(True), torch.from_numpy(mask).cuda()
            y0 = blk(xt, xat, mt)
            y0.backward(np_dev(dy))
            # (2) 4-D xa, 3-D mask, non-contiguous x (a transposed view of a [Lq,B,D] buffer) and xa (every second row)
            xs = np_dev(x).transpose(0, 1).contiguous().transpose(0, 1).requires_grad_(True)
            assert not xs.is_contiguous()
            wide = torch.zeros(Bq, 2 * Lk, D, device="cuda")
            wide[:, ::2] = np_dev(xa)
            xa4 = wide[:, ::2].unsqueeze(1).requires_grad_(True)
            assert xa4.dim() == 4 and not xa4.is_contiguous()
            y1 = blk(xs, xa4, mt.unsqueeze(1))
            y1.backward(np_dev(dy))
            torch.cuda.synchronize()
            tol = TOL[prec]
            rerun = 1e-5 if prec == "fp32" else 5e-3       # fp32 atomics order (dQ key-block partials, split-K) may differ
            assert torch.equal(y0, y1) and rel_err(xs.grad, xt.grad) < rerun
            assert xa4.grad.shape == xa4.shape and rel_err(xa4.grad.squeeze(1), xat.grad) < rerun
            assert rel_err(y1.detach().cpu(), torch.from_numpy(yo)) < tol
            assert rel_err(xs.grad.cpu(), torch.from_numpy(dxo)) < tol
            assert rel_err(xa4.grad.squeeze(1).cpu(), torch.from_numpy(dxao)) < tol
            # (3) bf16 query stream; xa without requires_grad (no d(xa) GEMM, no gradient returned for it)
            xb = np_dev(x).to(torch.bfloat16).requires_grad_(True)
            xa_ng = np_dev(xa)
            yb = blk(xb, xa_ng, mt)
            assert yb.dtype == torch.float32
            yb.backward(np_dev(dy))
            assert xb.grad is not None and xb.grad.dtype == torch.bfloat16 and xa_ng.grad is None
            yo_b, cb = O.block_forward(p, xb.detach().float().cpu().numpy(), xa, mask, H)
            assert rel_err(yb.cpu(), torch.from_numpy(yo_b)) < tol
            _, dxo_b, _ = O.block_backward(p, cb, dy)
            assert rel_err(xb.grad.float().cpu(), torch.from_numpy(dxo_b)) < max(tol, 8e-3)   # dx rounded to bf16
        finally:
            G.set_default_precision(None)


def test_block_rejects_mismatched_visual_memory():
    """reference: F.multi_head_attention_forward asserts on the key/value and mask shapes (torch/nn/functional.py:6229-6241)."""
    blk = G.GatedCrossAttentionBlock(128, 2, dropout=0.0).cuda()
    x = torch.randn(2, 8, 128, device="cuda")
    with pytest.raises((AssertionError, RuntimeError)):
        blk(x, torch.randn(3, 8, 128, device="cuda"))                # batch mismatch
    with pytest.raises((AssertionError, RuntimeError)):
        blk(x, torch.randn(2, 8, 64, device="cuda"))                 # feature width mismatch
    with pytest.raises(AssertionError):
        blk(x, torch.randn(2, 8, 128, device="cuda"), torch.zeros(2, 9, dtype=torch.bool, device="cuda"))
    with pytest.raises(RuntimeError):
        blk(x.cpu(), torch.randn(2, 8, 128))                          # no CPU fallback
