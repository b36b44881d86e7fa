"""numpy restatement of csrc/gxa_rng.cuh: the dropout multipliers (0 or 1/(1-p)) the CUDA kernels apply.

Used by the -m gpu parity tests to hand the oracle the SAME masks the kernels generate from (seed, element)."""
import numpy as np

STREAM_ATTN, STREAM_FFN = 1, 2
M32 = 0xFFFFFFFF


def _fold(t):
    return ((t >> np.uint64(32)) ^ (t & np.uint64(M32))) & np.uint64(M32)


def _words(ctr, seed_lo, seed_hi):
    """the two 2 x 15-bit words of a counter (gxa_rng.cuh drop_words)"""
    x = (ctr ^ np.uint32(seed_lo)).astype(np.uint64)
    y = _fold(x * np.uint64(0x9E3779B1)) ^ np.uint64(seed_hi)
    w0 = _fold(y * np.uint64(0x85EBCA77)) & np.uint64(0x7FFF7FFF)
    w1 = _fold(y * np.uint64(0xC2B2AE3D)) & np.uint64(0x7FFF7FFF)
    return w0.astype(np.uint32), w1.astype(np.uint32)


def drop_cfg(p, seed, stream):
    thresh = max(1, min(int(p * 32768.0 + 0.5), 0x7C00))
    scale = np.float32(1.0) / (np.float32(1.0) - np.float32(p))
    seed_lo = (seed & M32) ^ ((stream * 0x85EBCA6B) & M32)
    seed_hi = ((seed >> 32) + stream * 0xC2B2AE35) & M32
    return thresh, float(scale), seed_lo, seed_hi


def multipliers(rows, cols, p, seed, stream):
    """[rows, cols] float32: 0 where dropped, 1/(1-p) where kept; row r, column c <-> counter r*ceil(cols/4) + c//4,
    15-bit field c % 4 of that counter's two words."""
    if p <= 0:
        return np.ones((rows, cols), np.float32)
    thresh, scale, lo, hi = drop_cfg(p, seed, stream)
    quarter = (cols + 3) // 4
    ctr = ((np.arange(rows, dtype=np.uint64)[:, None] * np.uint64(quarter) + np.arange(quarter, dtype=np.uint64)[None, :])
           & np.uint64(M32)).astype(np.uint32)
    w0, w1 = _words(ctr, lo, hi)
    u = np.stack([w0 & np.uint32(0xFFFF), w0 >> np.uint32(16), w1 & np.uint32(0xFFFF), w1 >> np.uint32(16)],
                 -1).reshape(rows, 4 * quarter)[:, :cols]
    return np.where(u >= thresh, np.float32(scale), np.float32(0.0)).astype(np.float32)


def attn_multipliers(B, H, Lq, Lk, p, seed):
    return multipliers(B * H * Lq, Lk, p, seed, STREAM_ATTN).reshape(B, H, Lq, Lk)


def ffn_multipliers(B, Lq, D, p, seed):
    return multipliers(B * Lq, D, p, seed, STREAM_FFN).reshape(B, Lq, D)
