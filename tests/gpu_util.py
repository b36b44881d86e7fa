"""Helpers shared by the -m gpu tests: they call the product through its C ABI (ctypes) and compare
against torch fp32/fp64 math or the oracle."""

import numpy as np
import torch

from mocov2_whisper_flamingo_b200 import _lib


def stream():
    return torch.cuda.current_stream().cuda_stream


def ptr(t):
    return None if t is None else t.data_ptr()


def rel_err(a, b):
    a, b = a.double(), b.double()
    den = b.abs().max().item()
    return (a - b).abs().max().item() / (den if den > 0 else 1.0)


def scalar_sum_err(mine, ref, scale):
    """Error of a scalar that is a sum of many signed terms (the two gate gradients), relative to
    max(|ref|, 2-norm of the terms): the natural size of such a sum.  The total itself may be cancelled far below it, and
    a relative error against a cancelled total says nothing about the kernels (the reference's own bf16-autocast run is
    1.7e-2 off its fp32 run on such a scalar)."""
    mine, ref = float(mine), float(ref)
    return abs(mine - ref) / max(abs(ref), float(scale), 1e-30)


def err_map(a, b, br=32, bc=32):
    """coarse map of where two matrices disagree (for diagnosing tile/descriptor bugs from one run)"""
    d = (a.double() - b.double()).abs()
    M, N = d.shape
    rows = []
    for r in range(0, min(M, 8 * br), br):
        rows.append(" ".join(f"{d[r:r + br, c:c + bc].max().item():8.2e}" for c in range(0, min(N, 8 * bc), bc)))
    return "\n".join(rows)


def gemm(A_log, B_log, a_mn=False, b_mn=False, epi=_lib.EPI_F32, bias=None, aux=None, gate=None, accumulate=0,
         C_init=None, bn=0, k_splits=0):
    """C = A_log @ B_log.T through gxa_gemm_bf16.  A_log [M,K], B_log [N,K] bf16 (logical); storage is
    transposed when the *_mn flag is set.  Returns (C, C2)."""
    L = _lib.lib()
    M, K = A_log.shape
    N = B_log.shape[0]
    A = A_log.t().contiguous() if a_mn else A_log.contiguous()
    B = B_log.t().contiguous() if b_mn else B_log.contiguous()
    lda, ldb = A.shape[1], B.shape[1]
    dev = A.device
    out_f32 = epi in (_lib.EPI_F32, _lib.EPI_GATE_RES)
    if C_init is not None:
        Cmat = C_init.clone()
    else:
        Cmat = torch.full((M, N), float("nan"), dtype=torch.float32 if out_f32 else torch.bfloat16, device=dev)
    C2 = None
    if epi in (_lib.EPI_GATE_RES, _lib.EPI_BIAS_GELU):
        C2 = torch.full((M, N), float("nan"), dtype=torch.bfloat16, device=dev)
    ld_aux = aux.shape[1] if aux is not None else 0
    _lib.check(L.gxa_gemm_bf16(ptr(A), lda, int(a_mn), ptr(B), ldb, int(b_mn), M, N, K, epi, ptr(Cmat), N, ptr(C2), N,
                               ptr(bias), ptr(aux), ld_aux, ptr(gate), accumulate, bn, k_splits, stream()),
               "gxa_gemm_bf16")
    torch.cuda.synchronize()
    return Cmat, C2


def np_dev(a, dtype=torch.float32):
    return torch.from_numpy(np.ascontiguousarray(a)).to("cuda", dtype)


# ---- measured parity errors, written next to the test log so that a tolerance is a measurement, not an argument
_ERRLOG = {}


def record_err(case: str, name: str, value: float) -> None:
    _ERRLOG.setdefault(case, {})[name] = float(value)


def dump_errlog() -> None:
    import json
    import os
    if not _ERRLOG:
        return
    out = os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "gpurun_out")
    try:
        os.makedirs(out, exist_ok=True)
        path = os.path.join(out, "parity_errors.json")
        old = {}
        if os.path.exists(path):
            with open(path) as f:
                old = json.load(f)
        old.update(_ERRLOG)
        with open(path, "w") as f:
            json.dump(old, f, indent=1, sort_keys=True)
    except OSError:
        pass
