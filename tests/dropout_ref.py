"""numpy restatement of csrc/gxa_rng.cuh: the dropout multipliers (0 or 1/(1-p)) the CUDA kernels apply.

Used by the -m gpu parity tests to hand the oracle the SAME masks the kernels generate from (seed, element)."""
import numpy as np

STREAM_ATTN, STREAM_FFN = 1, 2
M32 = 0xFFFFFFFF


def _mix(ctr, seed_lo, seed_hi):
    x = (ctr ^ np.uint32(seed_lo)).astype(np.uint64)
    t = x * np.uint64(0x9E3779B1)
    y = ((t >> np.uint64(32)) ^ (t & np.uint64(M32)) ^ np.uint64(seed_hi)) & np.uint64(M32)
    t = y * np.uint64(0x85EBCA77)
    return (((t >> np.uint64(32)) ^ (t & np.uint64(M32))) & np.uint64(M32)).astype(np.uint32)


def drop_cfg(p, seed, stream):
    thresh = min(int(p * 65536.0 + 0.5), 65535)
    scale = np.float32(1.0) / (np.float32(1.0) - np.float32(p))
    seed_lo = (seed & M32) ^ ((stream * 0x85EBCA6B) & M32)
    seed_hi = ((seed >> 32) + stream * 0xC2B2AE35) & M32
    return thresh, float(scale), seed_lo, seed_hi


def multipliers(rows, cols, p, seed, stream):
    """[rows, cols] float32: 0 where dropped, 1/(1-p) where kept; row r, column c <-> counter r*ceil(cols/2) + c//2."""
    if p <= 0:
        return np.ones((rows, cols), np.float32)
    thresh, scale, lo, hi = drop_cfg(p, seed, stream)
    half = (cols + 1) // 2
    ctr = ((np.arange(rows, dtype=np.uint64)[:, None] * np.uint64(half) + np.arange(half, dtype=np.uint64)[None, :])
           & np.uint64(M32)).astype(np.uint32)
    x = _mix(ctr, lo, hi)
    u = np.stack([x & np.uint32(0xFFFF), x >> np.uint32(16)], -1).reshape(rows, 2 * half)[:, :cols]
    return np.where(u >= thresh, np.float32(scale), np.float32(0.0)).astype(np.float32)


def attn_multipliers(B, H, Lq, Lk, p, seed):
    return multipliers(B * H * Lq, Lk, p, seed, STREAM_ATTN).reshape(B, H, Lq, Lk)


def ffn_multipliers(B, Lq, D, p, seed):
    return multipliers(B * Lq, D, p, seed, STREAM_FFN).reshape(B, Lq, D)
