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

namespace gxa {

enum Epilogue : int {
  EPI_BF16 = 0,       // C(bf16) = acc + bias
  EPI_F32 = 1,        // C(f32)  = acc + bias          (atomic add when accumulate / split-K)
  EPI_GATE_RES = 2,   // t = bf16(acc + bias) -> C2(bf16);  C(f32) = aux(f32) + tanh(gate) * t
  EPI_BIAS_GELU = 3,  // u = bf16(acc + bias) -> C(bf16);   C2(bf16) = gelu_erf(u)
  EPI_GELU_BWD = 4,   // C(bf16) = bf16(acc) * gelu'(aux(bf16))
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
#pragma unroll
    for (int i = 0; i < 32; ++i) v[i] = bf16r(v[i]);
    store_bf16x32(reinterpret_cast<__nv_bfloat16*>(p.C) + off, v);
#pragma unroll
    for (int i = 0; i < 32; ++i) v[i] = gelu_erf(v[i]);
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
        float2 u = __bfloat1622float2(h[j]);
        v[8 * i + 2 * j] = bf16r(v[8 * i + 2 * j]) * gelu_erf_grad(u.x);
        v[8 * i + 2 * j + 1] = bf16r(v[8 * i + 2 * j + 1]) * gelu_erf_grad(u.y);
      }
    }
    store_bf16x32(reinterpret_cast<__nv_bfloat16*>(p.C) + off, v);
  }
}

}  // namespace gxa
