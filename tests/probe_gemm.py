"""One process per operand layout: prints the error of the tcgen05 GEMM vs torch for a few shapes, with a
coarse error map.  Used for bring-up on the GPU box (a trapped kernel poisons only its own process)."""
import sys

import torch

sys.path.insert(0, __file__.rsplit("/tests/", 1)[0])
sys.path.insert(0, __file__.rsplit("/", 1)[0])
from gpu_util import err_map, gemm, rel_err  # noqa: E402

a_mn, b_mn = bool(int(sys.argv[1])), bool(int(sys.argv[2]))
for (M, N, K, bn) in [(128, 128, 64, 128), (128, 128, 256, 128), (256, 256, 128, 256), (1000, 1536, 200, 0),
                      (4096, 4096, 1024, 0)]:
    g = torch.Generator(device="cuda").manual_seed(0)
    A = torch.randn(M, K, device="cuda", generator=g).to(torch.bfloat16)
    B = torch.randn(N, K, device="cuda", generator=g).to(torch.bfloat16)
    try:
        Cm, _ = gemm(A, B, a_mn, b_mn, bn=bn)
    except Exception as e:  # noqa: BLE001
        print(f"a_mn={a_mn} b_mn={b_mn} M={M} N={N} K={K} bn={bn}: EXC {e}", flush=True)
        break
    R = A.float() @ B.float().t()
    e = rel_err(Cm, R)
    print(f"a_mn={a_mn} b_mn={b_mn} M={M} N={N} K={K} bn={bn}: rel_err={e:.3e} nan={int(torch.isnan(Cm).sum())}",
          flush=True)
    if not e < 2e-5:
        print(err_map(Cm, R), flush=True)
