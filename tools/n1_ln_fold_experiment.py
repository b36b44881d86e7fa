"""n1 (north_star "fused LayerNorm + projection"): what folding the LayerNorm INTO the tensor-core projection costs in
accuracy.  CPU experiment, torch only -- no kernels involved.

Design (i) of VERDICT r1 #6: the GEMM runs on bf16(x) * bf16(gamma o W)^T and the epilogue applies
    q = rstd * (acc - mu * sum_k gamma_k W_nk) + (beta W^T + b)
so that no normalised tensor h = LN(x) is ever written.  The rounding point moves: x is rounded to bf16 BEFORE the mean
is removed.  Compared against the fp32 reference  LN(x) W^T + b  are
    (a) the shipped path:  bf16(LN(x)) @ bf16(W)^T, fp32 accumulate  (fp32 statistics, rounding after normalisation)
    (b) design (i) as above, statistics in fp32 from the unrounded x
for rows whose mean is |mu| / sigma = 0, 1, 3, 10 standard deviations away from zero (the residual stream of the block is
not centred: x1 = x + tanh(a) * o carries whatever offset the stream has).  rel = max|a - b| / max|b|, as in the tests.
    python tools/n1_ln_fold_experiment.py
"""
import torch

torch.manual_seed(0)
M, N = 2048, 1024


def bf(t):
    return t.to(torch.bfloat16).to(torch.float32)


print(f"{'D':>5} {'|mu|/sigma':>10} {'(a) shipped':>12} {'(b) folded':>12} {'(b)/(a)':>8}")
for D in (384, 768, 1024, 1280):
    W = torch.randn(N, D) / D ** 0.5
    b = torch.rand(N) * 0.2 - 0.1
    gamma = 1 + 0.1 * torch.randn(D)
    beta = 0.1 * torch.randn(D)
    for ratio in (0.0, 1.0, 3.0, 10.0):
        x = torch.randn(M, D) + ratio
        mu = x.mean(1, keepdim=True)
        rstd = (x.var(1, unbiased=False, keepdim=True) + 1e-5).rsqrt()
        h = (x - mu) * rstd * gamma + beta
        ref = (h.double() @ W.double().T + b.double()).float()
        a = bf(h) @ bf(W).T + b
        Wg = bf(gamma * W)
        acc = bf(x) @ Wg.T                                           # what the tensor cores would accumulate (fp32)
        corr = (gamma * W).sum(1)                                    # sum_k gamma_k W_nk, fp32, once per weight update
        bq = rstd * (acc - mu * corr) + (beta @ W.T + b)
        den = ref.abs().max()
        ea, eb = float((a - ref).abs().max() / den), float((bq - ref).abs().max() / den)
        print(f"{D:5d} {ratio:10.1f} {ea:12.2e} {eb:12.2e} {eb / ea:8.1f}")
