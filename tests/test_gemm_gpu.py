"""tcgen05 GEMM (gxa_gemm_bf16) vs torch fp32 matmul of the same bf16 operands.
bf16 x bf16 products are exact in fp32, so only the summation order differs: tolerance 2e-5 relative."""
import math

import pytest
import torch

from mocov2_whisper_flamingo_b200 import _lib
from gpu_util import err_map, gemm, rel_err

pytestmark = pytest.mark.gpu
TOL = 2e-5


def operands(M, N, K, seed=0):
    g = torch.Generator(device="cuda").manual_seed(seed)
    A = torch.randn(M, K, device="cuda", generator=g).to(torch.bfloat16)
    B = torch.randn(N, K, device="cuda", generator=g).to(torch.bfloat16)
    return A, B


def ref_mm(A, B):
    return A.float() @ B.float().t()


@pytest.mark.parametrize("a_mn,b_mn", [(False, False), (False, True), (True, True), (True, False)])
@pytest.mark.parametrize("M,N,K,bn", [(128, 128, 64, 128), (128, 256, 128, 256), (500, 384, 384, 0),
                                      (1000, 1536, 200, 0), (896, 512, 1024, 256), (3000, 1024, 520, 0)])
def test_layouts_f32(a_mn, b_mn, M, N, K, bn):
    if (a_mn and M % 8) or (b_mn and N % 8) or ((not a_mn or not b_mn) and K % 8):
        # TMA needs 16-byte row pitch: the contiguous dimension of each operand must be a multiple of 8
        if (a_mn and M % 8) or (not a_mn and K % 8) or (b_mn and N % 8) or (not b_mn and K % 8):
            pytest.skip("row pitch not 16-byte aligned for this layout")
    A, B = operands(M, N, K)
    Cm, _ = gemm(A, B, a_mn, b_mn, bn=bn)
    R = ref_mm(A, B)
    e = rel_err(Cm, R)
    assert e < TOL, f"rel err {e:.3e}\n{err_map(Cm, R)}"


@pytest.mark.parametrize("ks", [2, 3, 4])
def test_split_k_and_accumulate(ks):
    A, B = operands(256, 256, 1000, seed=1)
    R = ref_mm(A, B)
    Cm, _ = gemm(A, B, True, True, k_splits=ks)
    assert rel_err(Cm, R) < TOL
    C0 = torch.randn(256, 256, device="cuda")
    Cm, _ = gemm(A, B, False, True, accumulate=1, C_init=C0, k_splits=1)
    assert rel_err(Cm, R + C0) < TOL


def test_auto_split_wgrad_shape():
    # wgrad-like: few output tiles, long contraction (tokens) -> auto split-K
    A, B = operands(384, 384, 5000, seed=2)
    Cm, _ = gemm(A, B, True, True)
    assert rel_err(Cm, ref_mm(A, B)) < TOL


def test_epilogue_bias_bf16():
    A, B = operands(300, 768, 384, seed=3)
    bias = torch.randn(768, device="cuda")
    Cm, _ = gemm(A, B, epi=_lib.EPI_BF16, bias=bias)
    R = (ref_mm(A, B) + bias).to(torch.bfloat16)
    assert rel_err(Cm.float(), R.float()) < 8e-3        # one bf16 ulp
    assert (Cm.float() - R.float()).abs().max() <= 2 * 2 ** -8 * R.float().abs().max()


def test_epilogue_gate_residual():
    A, B = operands(300, 384, 384, seed=4)
    bias = torch.randn(384, device="cuda")
    X = torch.randn(300, 384, device="cuda")
    gate = torch.tensor([0.7], device="cuda")
    Y, O = gemm(A, B, epi=_lib.EPI_GATE_RES, bias=bias, aux=X, gate=gate)
    o_ref = (ref_mm(A, B) + bias).to(torch.bfloat16)
    assert rel_err(O.float(), o_ref.float()) < 8e-3
    y_ref = X + math.tanh(0.7) * O.float()               # exact given the stored O
    assert rel_err(Y, y_ref) < 1e-6


def test_epilogue_bias_gelu_and_backward():
    A, B = operands(260, 1536, 384, seed=5)
    bias = torch.randn(1536, device="cuda")
    # forward epilogue: u = bf16(A B^T + bias);  C2 = gelu(u);  C = gelu'(u) (what the backward needs of u)
    G, Act = gemm(A, B, epi=_lib.EPI_BIAS_GELU, bias=bias)
    u_ref = (ref_mm(A, B) + bias).to(torch.bfloat16).float().requires_grad_(True)
    a_ref = torch.nn.functional.gelu(u_ref)
    a_ref.sum().backward()
    assert rel_err(Act.float(), a_ref.detach().to(torch.bfloat16).float()) < 8e-3
    assert rel_err(G.float(), u_ref.grad.to(torch.bfloat16).float()) < 8e-3
    # GELU backward epilogue: dU = bf16(dA) * gelu'(u), with dA = dF @ W2 (B operand MN-major as in dgrad)
    dF, W2t = operands(260, 1536, 384, seed=6)           # dF [260,384], W2^T logical [1536,384]
    dU, _ = gemm(dF, W2t, False, True, epi=_lib.EPI_GELU_BWD, aux=G)
    dA = ref_mm(dF, W2t).to(torch.bfloat16).float()
    assert rel_err(dU.float(), dA * u_ref.grad) < 1e-2


# ---- CTA-pair (cta_group::2) 256x256 kernel, selected with bn=512 ---------------------------------
@pytest.mark.parametrize("a_mn,b_mn", [(False, False), (False, True), (True, True), (True, False)])
@pytest.mark.parametrize("M,N,K", [(256, 256, 64), (256, 256, 512), (1000, 1536, 200), (896, 384, 1024),
                                   (3000, 1024, 520), (4096, 4096, 1024)])
def test_pair_kernel_layouts_f32(a_mn, b_mn, M, N, K):
    if (a_mn and M % 8) or (b_mn and N % 8):
        pytest.skip("row pitch not 16-byte aligned for this layout")
    A, B = operands(M, N, K, seed=7)
    Cm, _ = gemm(A, B, a_mn, b_mn, bn=512)
    R = ref_mm(A, B)
    e = rel_err(Cm, R)
    assert e < TOL, f"rel err {e:.3e}\n{err_map(Cm, R)}"


@pytest.mark.parametrize("ks", [1, 2, 5])
def test_pair_kernel_split_k_accumulate(ks):
    A, B = operands(1024, 512, 3000, seed=8)
    R = ref_mm(A, B)
    C0 = torch.randn(1024, 512, device="cuda")
    Cm, _ = gemm(A, B, True, True, accumulate=1, C_init=C0, k_splits=ks, bn=512)
    assert rel_err(Cm, R + C0) < TOL


def test_pair_kernel_epilogues():
    A, B = operands(700, 1536, 384, seed=9)
    bias = torch.randn(1536, device="cuda")
    Cm, _ = gemm(A, B, epi=_lib.EPI_BF16, bias=bias, bn=512)
    R = (ref_mm(A, B) + bias).to(torch.bfloat16)
    assert rel_err(Cm.float(), R.float()) < 8e-3
    G, Act = gemm(A, B, epi=_lib.EPI_BIAS_GELU, bias=bias, bn=512)      # C = gelu'(u), C2 = gelu(u), u = bf16(R)
    u = R.float().requires_grad_(True)
    a = torch.nn.functional.gelu(u)
    a.sum().backward()
    a_ref = a.detach().to(torch.bfloat16)
    assert rel_err(Act.float(), a_ref.float()) < 8e-3
    assert (Act.float() - a_ref.float()).abs().max() < 2e-2 * a_ref.float().abs().max()
    assert rel_err(G.float(), u.grad.to(torch.bfloat16).float()) < 8e-3
    dF, W2t = operands(700, 1536, 384, seed=10)
    dU, _ = gemm(dF, W2t, False, True, epi=_lib.EPI_GELU_BWD, aux=G, bn=512)
    dA = ref_mm(dF, W2t).to(torch.bfloat16).float()
    assert rel_err(dU.float(), dA * u.grad) < 1e-2
    A2, B2 = operands(700, 512, 1536, seed=11)
    b2 = torch.randn(512, device="cuda")
    X = torch.randn(700, 512, device="cuda")
    gate = torch.tensor([-0.4], device="cuda")
    Y, O = gemm(A2, B2, epi=_lib.EPI_GATE_RES, bias=b2, aux=X, gate=gate, bn=512)
    o_ref = (ref_mm(A2, B2) + b2).to(torch.bfloat16)
    assert rel_err(O.float(), o_ref.float()) < 8e-3
    assert rel_err(Y, X + math.tanh(-0.4) * O.float()) < 1e-6
