"""Bring-up probe for the CTA-pair GEMM (bn=512): one process, prints error per shape/layout."""
import sys

import torch

sys.path.insert(0, __file__.rsplit("/tests/", 1)[0])
sys.path.insert(0, __file__.rsplit("/", 1)[0])
from gpu_util import err_map, gemm, rel_err  # noqa: E402

a_mn, b_mn = bool(int(sys.argv[1])), bool(int(sys.argv[2]))
for (M, N, K) in [(256, 256, 64), (256, 256, 512), (1024, 512, 200), (4096, 4096, 1024)]:
    g = torch.Generator(device="cuda").manual_seed(0)
    A = torch.randn(M, K, device="cuda", generator=g).to(torch.bfloat16)
    B = torch.randn(N, K, device="cuda", generator=g).to(torch.bfloat16)
    try:
        Cm, _ = gemm(A, B, a_mn, b_mn, bn=512)
    except Exception as e:  # noqa: BLE001
        print(f"PAIR a_mn={a_mn} b_mn={b_mn} M={M} N={N} K={K}: EXC {e}", flush=True)
        break
    R = A.float() @ B.float().t()
    e = rel_err(Cm, R)
    print(f"PAIR a_mn={a_mn} b_mn={b_mn} M={M} N={N} K={K}: rel_err={e:.3e} nan={int(torch.isnan(Cm).sum())}", flush=True)
    if not e < 2e-5:
        print(err_map(Cm, R, 64, 64), flush=True)
