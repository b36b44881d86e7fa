"""Launch the dominant kernels a few times at the BASELINE config-3 shape (for ncu --set full captures)."""
import sys

import torch

sys.path.insert(0, __file__.rsplit("/tools/", 1)[0])
sys.path.insert(0, __file__.rsplit("/tools/", 1)[0] + "/tests")
from mocov2_whisper_flamingo_b200 import _lib  # noqa: E402
from gpu_util import gemm  # noqa: E402

which = sys.argv[1] if len(sys.argv) > 1 else "gemm"
M, D = 64 * 448, 1024
g = torch.Generator(device="cuda").manual_seed(0)
if which == "gemm":
    A = torch.randn(M, D, device="cuda", generator=g).to(torch.bfloat16)
    W1 = (torch.randn(4 * D, D, device="cuda", generator=g) / 32).to(torch.bfloat16)
    bias = torch.randn(4 * D, device="cuda")
    for _ in range(3):
        gemm(A, W1, epi=_lib.EPI_BIAS_GELU, bias=bias)          # FFN1 + bias + GELU
    Wq = (torch.randn(D, D, device="cuda", generator=g) / 32).to(torch.bfloat16)
    for _ in range(3):
        gemm(A, Wq, epi=_lib.EPI_BF16, bias=bias[:D].contiguous())  # plain projection
    dY = torch.randn(M, D, device="cuda", generator=g).to(torch.bfloat16)
    for _ in range(3):
        gemm(dY.t().contiguous(), A.t().contiguous(), True, True)    # wgrad [D, D] over M tokens
print("done")
