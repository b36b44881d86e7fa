// Counter-based dropout bits shared by every kernel that applies or re-creates a dropout mask
// (reference: nn.MultiheadAttention(dropout=p) -> dropout on the attention probabilities inside SDPA, and
// nn.Dropout(p) after ff.2 -- models/gate_cross_attention.py:16, :26).  Nothing is stored: forward and backward
// evaluate the same function of (seed, element), so the mask costs no HBM traffic.
//
//   keep(row, col) = u15(row, col) >= thresh,   thresh = round(p * 32768)   (drop probability thresh / 32768)
//   one counter per FOUR neighbouring columns:  ctr = row * ceil(ncols / 4) + (col >> 2)     (modulo 2^32)
//     t  = (ctr ^ seed_lo) * 0x9E3779B1                  64-bit product
//     y  = hi(t) ^ lo(t) ^ seed_hi
//     w0 = fold(y * 0x85EBCA77) & 0x7FFF7FFF             fold(x) = hi(x) ^ lo(x)
//     w1 = fold(y * 0xC2B2AE3D) & 0x7FFF7FFF
//     u15 of column 4k+0, 4k+1 = low / high half of w0;  of column 4k+2, 4k+3 = low / high half of w1
//
// Why this shape (round 2; ncu of the attention forward: the first version -- one 2-round hash per TWO columns, a shift,
// two integer compares and two selects -- was 6.4 of the 11.7 instructions per element of the softmax loop):
//   * three multiplies yield four decisions (2 instructions per element instead of 3);
//   * the 15-bit fields sit in the two halves of a 32-bit word with the sign bits cleared, which is a pair of
//     non-negative fp16 bit patterns: ONE packed half-precision compare (HSET2.GEU) turns a word into the 0xFFFF / 0
//     masks of two columns, and one AND applies them to a packed bf16x2 pair of probabilities.  For non-negative fp16
//     patterns the floating-point order IS the integer order (denormals included, no flush: the compare is not .ftz);
//     NaN patterns (0x7C01..0x7FFF) are "unordered" and count as kept by GEU -- as integers they are above every
//     admissible threshold as well (thresh <= 0x7C00, i.e. p <= 0.968), so the two definitions agree bit for bit.
// tests/dropout_ref.py holds the numpy restatement used by the parity tests; gxa_dropout_multipliers_host evaluates
// the functions below on the host.
#pragma once
#include <stdint.h>
#ifdef __CUDACC__
#include <cuda_fp16.h>
#endif

namespace gxa {

struct DropCfg {
  uint32_t thresh;     // 0 = dropout off; 15-bit threshold replicated in both halves of the word (t | t << 16)
  float scale;         // 1 / (1 - p)
  uint32_t seed_lo, seed_hi;
};

// the two 2x15-bit words of one counter (four columns)
__host__ __device__ __forceinline__ void drop_words(uint32_t ctr, uint32_t seed_lo, uint32_t seed_hi, uint32_t& w0, uint32_t& w1) {
  const uint64_t t = static_cast<uint64_t>(ctr ^ seed_lo) * 0x9E3779B1ULL;
  const uint32_t y = static_cast<uint32_t>(t >> 32) ^ static_cast<uint32_t>(t) ^ seed_hi;
  const uint64_t a = static_cast<uint64_t>(y) * 0x85EBCA77ULL;
  const uint64_t b = static_cast<uint64_t>(y) * 0xC2B2AE3DULL;
  w0 = (static_cast<uint32_t>(a >> 32) ^ static_cast<uint32_t>(a)) & 0x7FFF7FFFu;
  w1 = (static_cast<uint32_t>(b >> 32) ^ static_cast<uint32_t>(b)) & 0x7FFF7FFFu;
}
// host / scalar form: keep decision of column `col` (0..3) of the counter
__host__ __device__ __forceinline__ bool drop_keep_scalar(const DropCfg& c, uint32_t ctr, int col) {
  uint32_t w0, w1;
  drop_words(ctr, c.seed_lo, c.seed_hi, w0, w1);
  const uint32_t w = (col & 2) ? w1 : w0;
  const uint32_t u = (col & 1) ? (w >> 16) : (w & 0xFFFFu);
  return u >= (c.thresh & 0xFFFFu);
}
#ifdef __CUDACC__
// 0xFFFF / 0 per 16-bit half: columns (4k, 4k+1) in m01, (4k+2, 4k+3) in m23 -- AND them onto packed bf16x2 pairs
__device__ __forceinline__ void drop_masks4(const DropCfg& c, uint32_t ctr, uint32_t& m01, uint32_t& m23) {
  uint32_t w0, w1;
  drop_words(ctr, c.seed_lo, c.seed_hi, w0, w1);
  const __half2 t = *reinterpret_cast<const __half2*>(&c.thresh);
  m01 = __hgeu2_mask(*reinterpret_cast<const __half2*>(&w0), t);
  m23 = __hgeu2_mask(*reinterpret_cast<const __half2*>(&w1), t);
}
// multipliers (0 or cfg.scale) of the four columns of the counter
__device__ __forceinline__ void drop_mult4(const DropCfg& c, uint32_t ctr, float (&m)[4]) {
  uint32_t m01, m23;
  drop_masks4(c, ctr, m01, m23);
  const uint32_t s = __float_as_uint(c.scale);
  m[0] = __uint_as_float((m01 & 0xFFFFu) ? s : 0u);
  m[1] = __uint_as_float((m01 >> 16) ? s : 0u);
  m[2] = __uint_as_float((m23 & 0xFFFFu) ? s : 0u);
  m[3] = __uint_as_float((m23 >> 16) ? s : 0u);
}
#endif
inline DropCfg make_drop_cfg(float p, uint64_t seed, uint32_t stream) {
  DropCfg c;
  c.thresh = 0; c.scale = 1.f; c.seed_lo = 0; c.seed_hi = 0;
  if (p > 0.f) {
    double t = double(p) * 32768.0 + 0.5;
    uint32_t t15 = t >= double(0x7C00) ? 0x7C00u : uint32_t(t);     // fp16 patterns above 0x7C00 are NaN (see above)
    if (t15 == 0) t15 = 1;
    c.thresh = t15 | (t15 << 16);
    c.scale = 1.0f / (1.0f - p);
    c.seed_lo = uint32_t(seed) ^ (stream * 0x85ebca6bU);
    c.seed_hi = uint32_t(seed >> 32) + stream * 0xc2b2ae35U;
  }
  return c;
}
constexpr uint32_t kDropStreamAttn = 1, kDropStreamFfn = 2;

}  // namespace gxa
