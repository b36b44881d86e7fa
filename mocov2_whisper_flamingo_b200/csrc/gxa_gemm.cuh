// Persistent, warp-specialised tcgen05 GEMM for sm_100a with fused epilogues.
//
//   C[M,N] = A[M,K] * B[N,K]^T      bf16 operands, fp32 accumulation in TMEM
//
// Operand layouts (both supported for A and for B, selected at compile time):
//   K-major  : operand stored [rows, K] row-major (the contraction index is contiguous)
//   MN-major : operand stored [K, rows] row-major (the row index is contiguous) -- used by
//              dgrad (B = W as stored) and wgrad (A = dY^T, B = X^T as stored), so no
//              transposed copies of weights or activations are ever materialised.
//
// CTA = 320 threads: warps 0-7 epilogue (TMEM -> registers -> fused math -> HBM),
// warp 8 = TMA producer, warp 9 = TMEM allocator + single-thread tcgen05.mma issuer.
// Tile 128 x BN x 64, 128B-swizzled smem ring, two TMEM accumulator stages so the
// epilogue of tile i overlaps the main loop of tile i+1.  Grid = min(#work, #SMs).
#pragma once
#include "gxa_ptx.cuh"
#include "gxa_rng.cuh"

namespace gxa {

enum Epilogue : int {
  EPI_BF16 = 0,       // C(bf16) = acc + bias
  EPI_F32 = 1,        // C(f32)  = acc + bias          (atomic add when accumulate / split-K); with `gate` (1-CTA kernel only):
                      // C += tanh(gate) * (acc + bias) -- the split-K form of EPI_GATE_RES for few-row (decode) problems
  EPI_GATE_RES = 2,   // t = bf16(acc + bias) -> C2(bf16);  C(f32) = aux(f32) + tanh(gate) * t
  EPI_BIAS_GELU = 3,  // u = bf16(acc + bias);   C2(bf16) = gelu_erf(u);   C(bf16) = gelu_erf'(u), all the backward needs of u
  EPI_GELU_BWD = 4,   // C(bf16) = bf16(acc) * aux(bf16),  aux = gelu'(u) as stored by EPI_BIAS_GELU
};

struct GemmParams {
  int M, N, K;
  int k_splits;
  void* C;
  int ldc;
  void* C2;
  int ldc2;
  const float* bias;
  const void* aux;
  int ld_aux;
  const float* gate;   // device scalar, tanh applied in-kernel
  int accumulate;
  DropCfg drop;        // EPI_GATE_RES: dropout on t before it is stored / gated (nn.Dropout after ff.2); thresh 0 = off
  float* colsum;       // EPI_GELU_BWD: colsum[n] += sum over rows of the (bf16-rounded) output, i.e. the bias gradient
};

constexpr int GEMM_BM = 128;
constexpr int GEMM_BK = 64;
constexpr int GEMM_THREADS = 320;
constexpr int GEMM_EPI_WARPS = 8;

template <int BN>
struct GemmCfg {
  static constexpr int A_BYTES = GEMM_BM * GEMM_BK * 2;
  static constexpr int B_BYTES = BN * GEMM_BK * 2;
  static constexpr int STAGE_BYTES = A_BYTES + B_BYTES;
  static constexpr int STAGES = (BN == 256) ? 4 : 6;
  static constexpr int TMEM_COLS = (2 * BN <= 256) ? 256 : 512;
  static constexpr int SMEM_BYTES = STAGES * STAGE_BYTES + 1024 /*align slack*/ + 256 /*barriers*/;
};

__device__ __forceinline__ float gelu_erf(float u) { return 0.5f * u * (1.0f + erff(u * 0.70710678118654752f)); }
__device__ __forceinline__ float gelu_erf_grad(float u) {
  float cdf = 0.5f * (1.0f + erff(u * 0.70710678118654752f));
  float pdf = 0.3989422804014327f * __expf(-0.5f * u * u);
  return cdf + u * pdf;
}
__device__ __forceinline__ float bf16r(float v) { return __bfloat162float(__float2bfloat16_rn(v)); }
__device__ __forceinline__ uint32_t pack_bf16(float lo, float hi) {
  __nv_bfloat162 t = __floats2bfloat162_rn(lo, hi);
  return *reinterpret_cast<uint32_t*>(&t);
}
__device__ __forceinline__ void store_bf16x32(__nv_bfloat16* dst, const float (&v)[32]) {
  uint4* d4 = reinterpret_cast<uint4*>(dst);
#pragma unroll
  for (int i = 0; i < 4; ++i) {
    uint4 q;
    q.x = pack_bf16(v[8 * i + 0], v[8 * i + 1]);
    q.y = pack_bf16(v[8 * i + 2], v[8 * i + 3]);
    q.z = pack_bf16(v[8 * i + 4], v[8 * i + 5]);
    q.w = pack_bf16(v[8 * i + 6], v[8 * i + 7]);
    d4[i] = q;
  }
}

// One thread owns 32 consecutive columns [col, col+32) of output row `row`.
template <int EPI>
__device__ __forceinline__ void epilogue_chunk(const GemmParams& p, int row, int col, const uint32_t (&r)[32],
                                               float gate_t, bool first_split) {
  float v[32];
#pragma unroll
  for (int i = 0; i < 32; ++i) v[i] = __uint_as_float(r[i]);
  if (p.bias != nullptr && first_split) {
    const float4* b4 = reinterpret_cast<const float4*>(p.bias + col);
#pragma unroll
    for (int i = 0; i < 8; ++i) {
      float4 b = __ldg(b4 + i);
      v[4 * i + 0] += b.x; v[4 * i + 1] += b.y; v[4 * i + 2] += b.z; v[4 * i + 3] += b.w;
    }
  }
  if (row >= p.M) return;
  const size_t off = static_cast<size_t>(row) * p.ldc + col;
  if constexpr (EPI == EPI_BF16) {
    store_bf16x32(reinterpret_cast<__nv_bfloat16*>(p.C) + off, v);
  } else if constexpr (EPI == EPI_F32) {
    float* c = reinterpret_cast<float*>(p.C) + off;
    if (p.gate != nullptr) {
#pragma unroll
      for (int i = 0; i < 32; ++i) v[i] *= gate_t;
    }
    if (p.accumulate || p.k_splits > 1) {
#pragma unroll
      for (int i = 0; i < 8; ++i)
        atomicAdd(reinterpret_cast<float4*>(c) + i, make_float4(v[4 * i], v[4 * i + 1], v[4 * i + 2], v[4 * i + 3]));
    } else {
#pragma unroll
      for (int i = 0; i < 8; ++i)
        reinterpret_cast<float4*>(c)[i] = make_float4(v[4 * i], v[4 * i + 1], v[4 * i + 2], v[4 * i + 3]);
    }
  } else if constexpr (EPI == EPI_GATE_RES) {
#pragma unroll
    for (int i = 0; i < 32; ++i) v[i] = bf16r(v[i]);
    if (p.drop.thresh != 0) {
      const uint32_t ctr = static_cast<uint32_t>(row) * static_cast<uint32_t>(p.N >> 2) + (col >> 2);
#pragma unroll
      for (int i = 0; i < 8; ++i) {
        float m[4];
        drop_mult4(p.drop, ctr + i, m);
#pragma unroll
        for (int j = 0; j < 4; ++j) v[4 * i + j] = bf16r(v[4 * i + j] * m[j]);
      }
    }
    store_bf16x32(reinterpret_cast<__nv_bfloat16*>(p.C2) + static_cast<size_t>(row) * p.ldc2 + col, v);
    const float4* x4 = reinterpret_cast<const float4*>(reinterpret_cast<const float*>(p.aux) +
                                                       static_cast<size_t>(row) * p.ld_aux + col);
    float4* y4 = reinterpret_cast<float4*>(reinterpret_cast<float*>(p.C) + off);
#pragma unroll
    for (int i = 0; i < 8; ++i) {
      float4 x = __ldg(x4 + i);
      y4[i] = make_float4(x.x + gate_t * v[4 * i], x.y + gate_t * v[4 * i + 1], x.z + gate_t * v[4 * i + 2],
                          x.w + gate_t * v[4 * i + 3]);
    }
  } else if constexpr (EPI == EPI_BIAS_GELU) {
    float dg[32];
#pragma unroll
    for (int i = 0; i < 32; ++i) {
      v[i] = bf16r(v[i]);
      dg[i] = gelu_erf_grad(v[i]);
      v[i] = gelu_erf(v[i]);
    }
    store_bf16x32(reinterpret_cast<__nv_bfloat16*>(p.C) + off, dg);
    store_bf16x32(reinterpret_cast<__nv_bfloat16*>(p.C2) + static_cast<size_t>(row) * p.ldc2 + col, v);
  } else if constexpr (EPI == EPI_GELU_BWD) {
    const uint4* u4 = reinterpret_cast<const uint4*>(reinterpret_cast<const __nv_bfloat16*>(p.aux) +
                                                     static_cast<size_t>(row) * p.ld_aux + col);
#pragma unroll
    for (int i = 0; i < 4; ++i) {
      uint4 q = __ldg(u4 + i);
      const __nv_bfloat162* h = reinterpret_cast<const __nv_bfloat162*>(&q);
#pragma unroll
      for (int j = 0; j < 4; ++j) {
        float2 dg = __bfloat1622float2(h[j]);
        v[8 * i + 2 * j] = bf16r(v[8 * i + 2 * j]) * dg.x;
        v[8 * i + 2 * j + 1] = bf16r(v[8 * i + 2 * j + 1]) * dg.y;
      }
    }
    store_bf16x32(reinterpret_cast<__nv_bfloat16*>(p.C) + off, v);
  }
}


// ---------------------------------------------------------------- epilogue v2 (coalesced)
// The TMEM load gives each thread one ROW of a 32x32 fp32 chunk.  The warp transposes it through a
// 4 KB XOR-swizzled staging buffer so that 8 lanes cover one 128-byte row segment: every global
// load/store of the fused epilogue is then fully coalesced, and bias / gate are loaded once per lane.
// erf-GELU on packed f32x2 (FFMA2/FMUL2 on sm_100 halve the issue slots of the polynomials): gelu_both2 below.
__device__ __forceinline__ float ex2_approx(float x) {
  float y;
  asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x));
  return y;
}
__device__ __forceinline__ float2 f2(float a) { return make_float2(a, a); }
// value and derivative from ONE exponential and no reciprocal (round 2).  The Abramowitz-Stegun 7.1.26 form used in round 1 costs two MUFU
// operations per element (rcp + ex2): at 16 / clk / SM that is 4096 cycles for a CTA's 128 x 256 tile -- as long as the
// tile's main loop at K = 1024, so the FFN1 epilogue could not hide under it (1113 vs 1450 TFLOP/s for the plain-bf16
// epilogue).  Here   1 - Phi(a) = 0.5 * exp(-q(a)),  q = c1 a + ... + c5 a^5,  a = |u|   (minimax fit on [0, 8], q keeps
// growing beyond; max |gelu error| 1.8e-6, max |gelu' error| 1.8e-6 against erf in float64 -- three orders of magnitude
// below the bf16 rounding of the two outputs), and the density comes from the same exponential:
//   phi(a) ~= d/da Phi = h * q'(a),  h = 1 - Phi(a)
//   gelu(u)  = relu(u) - a h            gelu'(u) = 0.5 + copysign(0.5 - h + a h q'(a), u)
__device__ __forceinline__ void gelu_both2(float2 u, float2& g, float2& dg) {
  constexpr float L2E = 1.4426950408889634f;
  constexpr float c1 = 0.7978507334384695f, c2 = 0.3183989204813974f, c3 = 0.036381011328425965f,
                  c4 = -0.005117729294495289f, c5 = 0.00036235314137177013f;
  const float2 a = make_float2(fabsf(u.x), fabsf(u.y));
  float2 q = __ffma2_rn(a, f2(-c5 * L2E), f2(-c4 * L2E));          // -q(a) * log2(e) - 1  (the 0.5 rides in the exponent)
  q = __ffma2_rn(a, q, f2(-c3 * L2E));
  q = __ffma2_rn(a, q, f2(-c2 * L2E));
  q = __ffma2_rn(a, q, f2(-c1 * L2E));
  q = __ffma2_rn(a, q, f2(-1.0f));
  float2 dq = __ffma2_rn(a, f2(5.f * c5), f2(4.f * c4));           // q'(a)
  dq = __ffma2_rn(a, dq, f2(3.f * c3));
  dq = __ffma2_rn(a, dq, f2(2.f * c2));
  dq = __ffma2_rn(a, dq, f2(c1));
  const float2 h = make_float2(ex2_approx(q.x), ex2_approx(q.y));
  const float2 ah = __fmul2_rn(a, h);
  g = __fadd2_rn(make_float2(fmaxf(u.x, 0.f), fmaxf(u.y, 0.f)), make_float2(-ah.x, -ah.y));
  const float2 r = __ffma2_rn(ah, dq, __fadd2_rn(f2(0.5f), make_float2(-h.x, -h.y)));     // 0.5 - h + a h q'
  dg = __fadd2_rn(make_float2(copysignf(r.x, u.x), copysignf(r.y, u.y)), f2(0.5f));
}
__device__ __forceinline__ uint2 pack_bf16x4(float4 v) {
  uint2 r;
  r.x = pack_bf16(v.x, v.y);
  r.y = pack_bf16(v.z, v.w);
  return r;
}
__device__ __forceinline__ float4 unpack_bf16x4(uint2 q) {
  return make_float4(__uint_as_float(q.x << 16), __uint_as_float(q.x & 0xffff0000u), __uint_as_float(q.y << 16),
                     __uint_as_float(q.y & 0xffff0000u));
}
__device__ __forceinline__ float4 round_bf16x4(float4 v) { return unpack_bf16x4(pack_bf16x4(v)); }

// Per-lane operands of one 32x32 chunk that come from HBM (x for the gated residual, u for gelu'): fetched one
// chunk AHEAD of their use so the load latency overlaps the previous chunk's TMEM read + math (the loads used to
// be issued one by one inside the row loop: 8 exposed HBM round trips per chunk).
template <int EPI>
struct AuxRegs {};
template <>
struct AuxRegs<EPI_GATE_RES> { float4 v[8]; };
template <>
struct AuxRegs<EPI_GELU_BWD> { uint4 v[4]; };     // bf16-first layout: 4 lanes per row, 8 columns (16 B) per lane

template <int EPI>
__device__ __forceinline__ void aux_prefetch(const GemmParams& p, int row0, int col, int lane, AuxRegs<EPI>& ax) {
  if constexpr (EPI == EPI_GATE_RES) {
    const int c = col + 4 * (lane & 7), rsub = lane >> 3;
#pragma unroll
    for (int it = 0; it < 8; ++it) {
      const int row = row0 + it * 4 + rsub;
      ax.v[it] = make_float4(0.f, 0.f, 0.f, 0.f);
      if (row < p.M)
        ax.v[it] = __ldg(reinterpret_cast<const float4*>(reinterpret_cast<const float*>(p.aux) +
                                                        static_cast<size_t>(row) * p.ld_aux + c));
    }
  } else if constexpr (EPI == EPI_GELU_BWD) {
    const int c = col + 8 * (lane & 3), rsub8 = lane >> 2;
#pragma unroll
    for (int it = 0; it < 4; ++it) {
      const int row = row0 + it * 8 + rsub8;
      ax.v[it] = make_uint4(0u, 0u, 0u, 0u);
      if (row < p.M)
        ax.v[it] = __ldg(reinterpret_cast<const uint4*>(reinterpret_cast<const __nv_bfloat16*>(p.aux) +
                                                       static_cast<size_t>(row) * p.ld_aux + c));
    }
  }
}

template <int EPI>
__device__ __forceinline__ void epilogue_chunk_v2(const GemmParams& p, uint8_t* stage, int row0, int col,
                                                  const uint32_t (&r)[32], float gate_t, bool first_split, int lane,
                                                  const AuxRegs<EPI>& ax, const CUtensorMap* tmC = nullptr) {
  const uint32_t sbase = smem_u32(stage);
  if constexpr (EPI == EPI_BF16 || EPI == EPI_GELU_BWD) {
    // bf16-first: bias, rounding and packing happen in the TMEM layout (thread = row; the 32 bias values of the chunk are
    // the same for every lane), so the transpose moves a 2 KB bf16 tile instead of the 4 KB fp32 one.  The staging
    // traffic goes through the same shared-memory port as the MMA operands and the TMA writes: at
    // K = 1024 the fp32 transposes were +20 % on that port.  Rows are 64 B here; 16-byte slot j of row l sits at
    // j ^ ((l >> 1) & 3), which keeps both the row-wise stores and the 4-lanes-per-row loads conflict-free.
    // EPI_GELU_BWD: the staged tile is bf16(acc) (the rounding point of autocast's bf16 linear output); gelu'(u) (aux,
    // prefetched in the post-transpose layout) multiplies it after the transpose, and the column sums of the rounded
    // products (= d ff.0.bias) are folded over the 8 row lanes of each column group.
    uint32_t pk[16];
    if (EPI == EPI_BF16 && p.bias != nullptr && first_split) {
#pragma unroll
      for (int j = 0; j < 8; ++j) {
        const float4 b = __ldg(reinterpret_cast<const float4*>(p.bias + col) + j);
        pk[2 * j] = pack_bf16(__uint_as_float(r[4 * j]) + b.x, __uint_as_float(r[4 * j + 1]) + b.y);
        pk[2 * j + 1] = pack_bf16(__uint_as_float(r[4 * j + 2]) + b.z, __uint_as_float(r[4 * j + 3]) + b.w);
      }
    } else {
#pragma unroll
      for (int j = 0; j < 16; ++j) pk[j] = pack_bf16(__uint_as_float(r[2 * j]), __uint_as_float(r[2 * j + 1]));
    }
#pragma unroll
    for (int j = 0; j < 4; ++j) {
      const uint32_t a = sbase + lane * 64 + ((j ^ ((lane >> 1) & 3)) << 4);
      asm volatile("st.shared.v4.b32 [%0], {%1, %2, %3, %4};" ::"r"(a), "r"(pk[4 * j]), "r"(pk[4 * j + 1]),
                   "r"(pk[4 * j + 2]), "r"(pk[4 * j + 3])
                   : "memory");
    }
    __syncwarp();
    const int cl4 = lane & 3, rsub8 = lane >> 2;
    float cs8[8] = {0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f};
#pragma unroll
    for (int it = 0; it < 4; ++it) {
      const int rr = it * 8 + rsub8;
      const int row = row0 + rr;
      uint4 v;
      asm volatile("ld.shared.v4.b32 {%0, %1, %2, %3}, [%4];"
                   : "=r"(v.x), "=r"(v.y), "=r"(v.z), "=r"(v.w)
                   : "r"(sbase + rr * 64 + ((cl4 ^ ((rr >> 1) & 3)) << 4)));
      if constexpr (EPI == EPI_GELU_BWD) {
        const uint4 g = ax.v[it];              // gelu'(u), stored by the forward epilogue (zero for rows past M)
        const float4 a0 = unpack_bf16x4(make_uint2(v.x, v.y)), a1 = unpack_bf16x4(make_uint2(v.z, v.w));
        const float4 g0 = unpack_bf16x4(make_uint2(g.x, g.y)), g1 = unpack_bf16x4(make_uint2(g.z, g.w));
        const uint2 o0 = pack_bf16x4(make_float4(a0.x * g0.x, a0.y * g0.y, a0.z * g0.z, a0.w * g0.w));
        const uint2 o1 = pack_bf16x4(make_float4(a1.x * g1.x, a1.y * g1.y, a1.z * g1.z, a1.w * g1.w));
        v = make_uint4(o0.x, o0.y, o1.x, o1.y);
        if (row < p.M) {
          const float4 s0 = unpack_bf16x4(o0), s1 = unpack_bf16x4(o1);
          cs8[0] += s0.x; cs8[1] += s0.y; cs8[2] += s0.z; cs8[3] += s0.w;
          cs8[4] += s1.x; cs8[5] += s1.y; cs8[6] += s1.z; cs8[7] += s1.w;
        }
      }
      if (row < p.M)
        *reinterpret_cast<uint4*>(reinterpret_cast<__nv_bfloat16*>(p.C) + static_cast<size_t>(row) * p.ldc + col + cl4 * 8) = v;
    }
    if constexpr (EPI == EPI_GELU_BWD) {
      if (p.colsum != nullptr) {      // 32 rows of this chunk -> one vector reduction per 8 columns
#pragma unroll
        for (int o = 4; o <= 16; o <<= 1) {
#pragma unroll
          for (int k = 0; k < 8; ++k) cs8[k] += __shfl_xor_sync(0xffffffffu, cs8[k], o);
        }
        if (rsub8 == 0) {
          float* dst = p.colsum + col + cl4 * 8;
          atomicAdd(reinterpret_cast<float4*>(dst), make_float4(cs8[0], cs8[1], cs8[2], cs8[3]));
          atomicAdd(reinterpret_cast<float4*>(dst + 4), make_float4(cs8[4], cs8[5], cs8[6], cs8[7]));
        }
      }
    }
    __syncwarp();          // the next chunk re-uses the staging buffer
  } else {
  if constexpr (EPI == EPI_F32) {
    // the previous chunk's TMA reduction must have finished READING this staging buffer
    if (tmC != nullptr) {
      if (lane == 0) bulk_wait_group_read0();
      __syncwarp();
    }
  }
#pragma unroll
  for (int j = 0; j < 8; ++j) {
    const uint32_t a = sbase + lane * 128 + ((j ^ (lane & 7)) << 4);
    asm volatile("st.shared.v4.b32 [%0], {%1, %2, %3, %4};" ::"r"(a), "r"(r[4 * j]), "r"(r[4 * j + 1]),
                 "r"(r[4 * j + 2]), "r"(r[4 * j + 3])
                 : "memory");
  }
  if constexpr (EPI == EPI_F32) {
    // wgrad / split-K accumulation: the staging buffer IS a 128B-swizzled [32 x 32] fp32 box -> one TMA reduce-add
    // (no LSU atomics; rows past M are clipped by the tensor map)
    if (tmC != nullptr) {
      fence_proxy_async_smem();
      __syncwarp();
      if (lane == 0) {
        tma_reduce_add_2d(tmC, stage, col, row0);
        bulk_commit_group();
      }
      return;
    }
  }
  __syncwarp();
  const int cl = lane & 7, rsub = lane >> 3;
  const int c = col + 4 * cl;
  float4 b = make_float4(0.f, 0.f, 0.f, 0.f);
  if (p.bias != nullptr && first_split) b = __ldg(reinterpret_cast<const float4*>(p.bias + c));
#pragma unroll
  for (int it = 0; it < 8; ++it) {
    const int rr = it * 4 + rsub;
    const int row = row0 + rr;
    float4 a;
    asm volatile("ld.shared.v4.f32 {%0, %1, %2, %3}, [%4];"
                 : "=f"(a.x), "=f"(a.y), "=f"(a.z), "=f"(a.w)
                 : "r"(sbase + rr * 128 + ((cl ^ (rr & 7)) << 4)));
    if (row >= p.M) continue;
    a.x += b.x; a.y += b.y; a.z += b.z; a.w += b.w;
    const size_t off = static_cast<size_t>(row) * p.ldc + c;
    if constexpr (EPI == EPI_F32) {
      float4* dst = reinterpret_cast<float4*>(reinterpret_cast<float*>(p.C) + off);
      if (p.accumulate || p.k_splits > 1) atomicAdd(dst, a);
      else *dst = a;
    } else if constexpr (EPI == EPI_GATE_RES) {
      uint2 pk = pack_bf16x4(a);
      a = unpack_bf16x4(pk);
      if (p.drop.thresh != 0) {      // bf16(linear) -> dropout -> bf16, as autocast does; one counter = these 4 columns
        uint32_t m01, m23;
        drop_masks4(p.drop, static_cast<uint32_t>(row) * static_cast<uint32_t>(p.N >> 2) + (c >> 2), m01, m23);
        const float sc = p.drop.scale;
        pk = pack_bf16x4(make_float4(a.x * sc, a.y * sc, a.z * sc, a.w * sc));
        pk.x &= m01;
        pk.y &= m23;
        a = unpack_bf16x4(pk);
      }
      *reinterpret_cast<uint2*>(reinterpret_cast<__nv_bfloat16*>(p.C2) + static_cast<size_t>(row) * p.ldc2 + c) = pk;
      const float4 x = ax.v[it];
      *reinterpret_cast<float4*>(reinterpret_cast<float*>(p.C) + off) =
          make_float4(fmaf(gate_t, a.x, x.x), fmaf(gate_t, a.y, x.y), fmaf(gate_t, a.z, x.z), fmaf(gate_t, a.w, x.w));
    } else if constexpr (EPI == EPI_BIAS_GELU) {
      a = round_bf16x4(a);          // u as autocast's bf16 linear output
      float2 g0, g1, d0, d1;
      gelu_both2(make_float2(a.x, a.y), g0, d0);
      gelu_both2(make_float2(a.z, a.w), g1, d1);
      *reinterpret_cast<uint2*>(reinterpret_cast<__nv_bfloat16*>(p.C) + off) = pack_bf16x4(make_float4(d0.x, d0.y, d1.x, d1.y));
      *reinterpret_cast<uint2*>(reinterpret_cast<__nv_bfloat16*>(p.C2) + static_cast<size_t>(row) * p.ldc2 + c) =
          pack_bf16x4(make_float4(g0.x, g0.y, g1.x, g1.y));
    }
  }
  __syncwarp();
  }   // fp32-staged epilogues
}

}  // namespace gxa
