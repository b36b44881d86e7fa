// Counter-based dropout bits shared by every kernel that applies or re-creates a dropout mask
// (reference: nn.MultiheadAttention(dropout=p) -> dropout on the attention probabilities inside SDPA, and
// nn.Dropout(p) after ff.2 -- models/gate_cross_attention.py:16, :26).  Nothing is stored: forward and backward
// evaluate the same function of (seed, element), so the mask costs no HBM traffic.
//
//   keep(row, col) = u16(row, col) >= thresh,   thresh = round(p * 65536)   (drop probability thresh / 65536)
//   u16(row, col)  = 16-bit half (col & 1) of mix(seed, row * ceil(ncols / 2) + (col >> 1))
//
// One 32-bit mix yields the decision for two neighbouring columns; the counter wraps modulo 2^32.
// tests/dropout_ref.py holds the numpy restatement used by the parity tests.
#pragma once
#include <stdint.h>

namespace gxa {

struct DropCfg {
  uint32_t thresh;     // 0 = dropout off
  float scale;         // 1 / (1 - p)
  uint32_t seed_lo, seed_hi;
};

// two multiply-fold rounds (64-bit product, high ^ low): 2 IMAD.WIDE + 2 LOP3 per 32 random bits
__host__ __device__ __forceinline__ uint32_t drop_mix(uint32_t ctr, uint32_t seed_lo, uint32_t seed_hi) {
  uint64_t t = static_cast<uint64_t>(ctr ^ seed_lo) * 0x9E3779B1ULL;
  const uint32_t y = static_cast<uint32_t>(t >> 32) ^ static_cast<uint32_t>(t) ^ seed_hi;
  t = static_cast<uint64_t>(y) * 0x85EBCA77ULL;
  return static_cast<uint32_t>(t >> 32) ^ static_cast<uint32_t>(t);
}
// keep decisions for the column pair (2*pair, 2*pair + 1):  u16 >= thresh, evaluated on the full word
// -DGXA_DROP_DPX (GXA_NVCC_EXTRA, A/B experiment, same mask): both half-word compares in one VIMNMX.U16x2 with two
// predicate outputs instead of a shift and two ISETP.
__device__ __forceinline__ void drop_keep(const DropCfg& c, uint32_t ctr, bool& k0, bool& k1) {
  const uint32_t x = drop_mix(ctr, c.seed_lo, c.seed_hi);
#ifdef GXA_DROP_DPX
  __vibmax_u16x2(x, c.thresh | (c.thresh << 16), &k1, &k0);      // pred = (x.half >= thresh), high half first
#else
  const uint32_t t = c.thresh << 16;
  k0 = (x << 16) >= t;
  k1 = x >= t;
#endif
}
// multipliers (0 or cfg.scale) for the same pair
__device__ __forceinline__ void drop_pair(const DropCfg& c, uint32_t ctr, float& m0, float& m1) {
  bool k0, k1;
  drop_keep(c, ctr, k0, k1);
  m0 = k0 ? c.scale : 0.f;
  m1 = k1 ? c.scale : 0.f;
}
inline DropCfg make_drop_cfg(float p, uint64_t seed, uint32_t stream) {
  DropCfg c;
  c.thresh = 0; c.scale = 1.f; c.seed_lo = 0; c.seed_hi = 0;
  if (p > 0.f) {
    double t = double(p) * 65536.0 + 0.5;
    c.thresh = t >= 65535.0 ? 65535u : uint32_t(t);
    c.scale = 1.0f / (1.0f - p);
    c.seed_lo = uint32_t(seed) ^ (stream * 0x85ebca6bU);
    c.seed_hi = uint32_t(seed >> 32) + stream * 0xc2b2ae35U;
  }
  return c;
}
constexpr uint32_t kDropStreamAttn = 1, kDropStreamFfn = 2;

}  // namespace gxa
