// Masked-softmax attention over variable-length keys, head_dim 64, bf16 in / fp32 softmax.
// Round-1 implementation: flash-style tiles on the legacy warp-level tensor-core path
// (mma.sync m16n8k16) with cp.async double buffering and 128B-XOR-swizzled smem tiles.
// The 85-90 % of the block's FLOPs that live in the projections/FFN run on tcgen05 (gxa_gemm.cu);
// moving these contractions to tcgen05/TMEM as well is tracked in DESIGN.md ("next").
//
//   forward : one CTA = 64 queries of one (batch, head); loops over 64-key tiles up to kv_end[b]
//   backward: dK/dV kernel (CTA = 64 keys, loops over query tiles, works on S^T so that no
//             register/smem transposes are needed) + dQ kernel (CTA = 64 queries, loops over key tiles).
//             Deterministic: no atomics.
#include <cstdlib>

#include "gxa_host.h"

namespace gxa {

constexpr int AT = 64;            // tile edge (queries, keys) and head_dim
constexpr int ATT_THREADS = 128;  // 4 warps x 16 rows
constexpr float kLog2e = 1.4426950408889634f;
constexpr float kScaleLog2 = 0.125f * kLog2e;   // 1/sqrt(64) folded into the exp2 argument

__device__ __forceinline__ uint32_t s_u32(const void* p) { return static_cast<uint32_t>(__cvta_generic_to_shared(p)); }

__device__ __forceinline__ void cp_async16(void* dst, const void* src, bool valid) {
  const int sz = valid ? 16 : 0;   // src-size 0 => 16 bytes of zeros
  asm volatile("cp.async.cg.shared.global [%0], [%1], 16, %2;" ::"r"(s_u32(dst)), "l"(src), "r"(sz) : "memory");
}
__device__ __forceinline__ void cp_async_commit() { asm volatile("cp.async.commit_group;" ::: "memory"); }
template <int N>
__device__ __forceinline__ void cp_async_wait() { asm volatile("cp.async.wait_group %0;" ::"n"(N) : "memory"); }

// tile: [64 rows][64 bf16] = 128 B rows, 16-byte chunk c of row r lives at chunk (c ^ (r & 7))
__device__ __forceinline__ uint32_t tile_addr(const __nv_bfloat16* tile, int r, int chunk) {
  return s_u32(tile) + r * 128 + ((chunk ^ (r & 7)) << 4);
}
// rows [r0, r0+64) of a [*, ld] bf16 matrix (64 columns starting at g); rows >= rmax are zero-filled
__device__ __forceinline__ void tile_load_async(__nv_bfloat16* tile, const __nv_bfloat16* g, int ld, int r0, int rmax) {
#pragma unroll
  for (int i = 0; i < 4; ++i) {
    const int idx = threadIdx.x + i * ATT_THREADS;
    const int r = idx >> 3, c = idx & 7;
    const bool ok = (r0 + r) < rmax;
    const __nv_bfloat16* src = g + size_t(ok ? (r0 + r) : 0) * ld + c * 8;
    cp_async16(reinterpret_cast<uint8_t*>(tile) + r * 128 + ((c ^ (r & 7)) << 4), src, ok);
  }
}
__device__ __forceinline__ void ldsm_x4(uint32_t addr, uint32_t& r0, uint32_t& r1, uint32_t& r2, uint32_t& r3) {
  asm volatile("ldmatrix.sync.aligned.m8n8.x4.shared.b16 {%0,%1,%2,%3}, [%4];"
               : "=r"(r0), "=r"(r1), "=r"(r2), "=r"(r3) : "r"(addr));
}
__device__ __forceinline__ void ldsm_x4_t(uint32_t addr, uint32_t& r0, uint32_t& r1, uint32_t& r2, uint32_t& r3) {
  asm volatile("ldmatrix.sync.aligned.m8n8.x4.trans.shared.b16 {%0,%1,%2,%3}, [%4];"
               : "=r"(r0), "=r"(r1), "=r"(r2), "=r"(r3) : "r"(addr));
}
__device__ __forceinline__ void mma16816(float (&c)[4], const uint32_t (&a)[4], uint32_t b0, uint32_t b1) {
  asm volatile(
      "mma.sync.aligned.m16n8k16.row.col.f32.bf16.bf16.f32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, {%0,%1,%2,%3};"
      : "+f"(c[0]), "+f"(c[1]), "+f"(c[2]), "+f"(c[3])
      : "r"(a[0]), "r"(a[1]), "r"(a[2]), "r"(a[3]), "r"(b0), "r"(b1));
}
__device__ __forceinline__ uint32_t pack2(float lo, float hi) {
  __nv_bfloat162 t = __floats2bfloat162_rn(lo, hi);
  return *reinterpret_cast<uint32_t*>(&t);
}

// A fragments (16 rows starting at m0, all 64 k) of a tile -> a[ks][4]
__device__ __forceinline__ void load_a_frags(const __nv_bfloat16* tile, int m0, uint32_t (&a)[4][4]) {
  const int l = threadIdx.x & 31;
  const int r = m0 + (l & 7) + ((l >> 3) & 1) * 8;
#pragma unroll
  for (int ks = 0; ks < 4; ++ks) ldsm_x4(tile_addr(tile, r, ks * 2 + (l >> 4)), a[ks][0], a[ks][1], a[ks][2], a[ks][3]);
}
// C[16 x 64] += A[16 x 64k] * T^T where tile T is stored [n][k]  (S = Q K^T, dP = dO V^T, ...)
__device__ __forceinline__ void mma_a_tileT(float (&c)[8][4], const uint32_t (&a)[4][4], const __nv_bfloat16* tile) {
  const int l = threadIdx.x & 31;
#pragma unroll
  for (int ks = 0; ks < 4; ++ks) {
#pragma unroll
    for (int np = 0; np < 4; ++np) {
      uint32_t b0, b1, b2, b3;
      ldsm_x4(tile_addr(tile, np * 16 + (l & 7) + (l >> 4) * 8, ks * 2 + ((l >> 3) & 1)), b0, b1, b2, b3);
      mma16816(c[2 * np], a[ks], b0, b1);
      mma16816(c[2 * np + 1], a[ks], b2, b3);
    }
  }
}
// C[16 x 64] += P[16 x 64k] * T where tile T is stored [k][n] and P comes from accumulator registers
__device__ __forceinline__ void mma_p_tile(float (&c)[8][4], const float (&p)[8][4], const __nv_bfloat16* tile) {
  const int l = threadIdx.x & 31;
#pragma unroll
  for (int kk = 0; kk < 4; ++kk) {
    uint32_t a[4];
    a[0] = pack2(p[2 * kk][0], p[2 * kk][1]);
    a[1] = pack2(p[2 * kk][2], p[2 * kk][3]);
    a[2] = pack2(p[2 * kk + 1][0], p[2 * kk + 1][1]);
    a[3] = pack2(p[2 * kk + 1][2], p[2 * kk + 1][3]);
#pragma unroll
    for (int np = 0; np < 4; ++np) {
      uint32_t b0, b1, b2, b3;
      ldsm_x4_t(tile_addr(tile, kk * 16 + (l & 7) + ((l >> 3) & 1) * 8, np * 2 + (l >> 4)), b0, b1, b2, b3);
      mma16816(c[2 * np], a, b0, b1);
      mma16816(c[2 * np + 1], a, b2, b3);
    }
  }
}
__device__ __forceinline__ void zero_acc(float (&c)[8][4]) {
#pragma unroll
  for (int i = 0; i < 8; ++i) c[i][0] = c[i][1] = c[i][2] = c[i][3] = 0.f;
}
// store a warp's 16 x 64 accumulator as bf16 rows (row g and g+8 of the warp's 16), rows >= rmax skipped
__device__ __forceinline__ void store_acc_bf16(__nv_bfloat16* g, int ld, int r0, int rmax, const float (&c)[8][4],
                                               float scale) {
  const int l = threadIdx.x & 31, gq = l >> 2, t = l & 3;
#pragma unroll
  for (int nt = 0; nt < 8; ++nt) {
    const int col = nt * 8 + 2 * t;
    if (r0 + gq < rmax)
      *reinterpret_cast<uint32_t*>(g + size_t(r0 + gq) * ld + col) = pack2(c[nt][0] * scale, c[nt][1] * scale);
    if (r0 + gq + 8 < rmax)
      *reinterpret_cast<uint32_t*>(g + size_t(r0 + gq + 8) * ld + col) = pack2(c[nt][2] * scale, c[nt][3] * scale);
  }
}

// ================================================================== forward
__global__ void __launch_bounds__(ATT_THREADS)
attn_fwd_kernel(const __nv_bfloat16* __restrict__ q, int ldq, const __nv_bfloat16* __restrict__ k,
                const __nv_bfloat16* __restrict__ v, int ldkv, const uint8_t* __restrict__ mask,
                const int* __restrict__ kv_end, __nv_bfloat16* __restrict__ ctx, int ldc, float* __restrict__ lse,
                int H, int Lq, int Lk) {
  __shared__ __align__(128) __nv_bfloat16 sQ[AT * AT];
  __shared__ __align__(128) __nv_bfloat16 sK[2][AT * AT];
  __shared__ __align__(128) __nv_bfloat16 sV[2][AT * AT];
  __shared__ uint8_t sM[2][AT];
  const int q0 = blockIdx.x * AT, h = blockIdx.y, b = blockIdx.z;
  const int warp = threadIdx.x >> 5, l = threadIdx.x & 31, gq = l >> 2, t = l & 3;
  const int kend = kv_end ? min(kv_end[b], Lk) : Lk;
  const int nblk = (kend + AT - 1) / AT;
  const __nv_bfloat16* qg = q + size_t(b) * Lq * ldq + h * AT;
  const __nv_bfloat16* kg = k + size_t(b) * Lk * ldkv + h * AT;
  const __nv_bfloat16* vg = v + size_t(b) * Lk * ldkv + h * AT;
  const uint8_t* mg = mask ? mask + size_t(b) * Lk : nullptr;

  tile_load_async(sQ, qg, ldq, q0, Lq);
  cp_async_commit();
  auto load_kv = [&](int j, int buf) {
    tile_load_async(sK[buf], kg, ldkv, j * AT, kend);
    tile_load_async(sV[buf], vg, ldkv, j * AT, kend);
    if (threadIdx.x < AT) {
      const int col = j * AT + threadIdx.x;
      sM[buf][threadIdx.x] = (col >= kend) ? 1 : (mg ? mg[col] : 0);
    }
  };
  if (nblk > 0) load_kv(0, 0);
  cp_async_commit();
  cp_async_wait<1>();
  __syncthreads();
  uint32_t aq[4][4];
  load_a_frags(sQ, warp * 16, aq);

  float o[8][4];
  zero_acc(o);
  float m0 = -INFINITY, m1 = -INFINITY, l0 = 0.f, l1 = 0.f;
  for (int j = 0; j < nblk; ++j) {
    const int buf = j & 1;
    if (j + 1 < nblk) load_kv(j + 1, buf ^ 1);
    cp_async_commit();
    cp_async_wait<1>();
    __syncthreads();
    float s[8][4];
    zero_acc(s);
    mma_a_tileT(s, aq, sK[buf]);
    float mx0 = -INFINITY, mx1 = -INFINITY;
#pragma unroll
    for (int nt = 0; nt < 8; ++nt) {
      const bool k0m = sM[buf][nt * 8 + 2 * t] != 0, k1m = sM[buf][nt * 8 + 2 * t + 1] != 0;
      s[nt][0] = k0m ? -INFINITY : s[nt][0] * kScaleLog2;
      s[nt][1] = k1m ? -INFINITY : s[nt][1] * kScaleLog2;
      s[nt][2] = k0m ? -INFINITY : s[nt][2] * kScaleLog2;
      s[nt][3] = k1m ? -INFINITY : s[nt][3] * kScaleLog2;
      mx0 = fmaxf(mx0, fmaxf(s[nt][0], s[nt][1]));
      mx1 = fmaxf(mx1, fmaxf(s[nt][2], s[nt][3]));
    }
    mx0 = fmaxf(mx0, __shfl_xor_sync(0xffffffffu, mx0, 1));
    mx0 = fmaxf(mx0, __shfl_xor_sync(0xffffffffu, mx0, 2));
    mx1 = fmaxf(mx1, __shfl_xor_sync(0xffffffffu, mx1, 1));
    mx1 = fmaxf(mx1, __shfl_xor_sync(0xffffffffu, mx1, 2));
    const float mn0 = fmaxf(m0, mx0), mn1 = fmaxf(m1, mx1);
    const float mu0 = (mn0 == -INFINITY) ? 0.f : mn0, mu1 = (mn1 == -INFINITY) ? 0.f : mn1;
    const float al0 = exp2f(m0 - mu0), al1 = exp2f(m1 - mu1);
    m0 = mn0; m1 = mn1;
    float rs0 = 0.f, rs1 = 0.f;
#pragma unroll
    for (int nt = 0; nt < 8; ++nt) {
      s[nt][0] = exp2f(s[nt][0] - mu0); s[nt][1] = exp2f(s[nt][1] - mu0);
      s[nt][2] = exp2f(s[nt][2] - mu1); s[nt][3] = exp2f(s[nt][3] - mu1);
      rs0 += s[nt][0] + s[nt][1];
      rs1 += s[nt][2] + s[nt][3];
      o[nt][0] *= al0; o[nt][1] *= al0; o[nt][2] *= al1; o[nt][3] *= al1;
    }
    l0 = l0 * al0 + rs0;
    l1 = l1 * al1 + rs1;
    mma_p_tile(o, s, sV[buf]);
    __syncthreads();
  }
  l0 += __shfl_xor_sync(0xffffffffu, l0, 1); l0 += __shfl_xor_sync(0xffffffffu, l0, 2);
  l1 += __shfl_xor_sync(0xffffffffu, l1, 1); l1 += __shfl_xor_sync(0xffffffffu, l1, 2);
  const float i0 = l0 > 0.f ? 1.f / l0 : 0.f, i1 = l1 > 0.f ? 1.f / l1 : 0.f;
#pragma unroll
  for (int nt = 0; nt < 8; ++nt) { o[nt][0] *= i0; o[nt][1] *= i0; o[nt][2] *= i1; o[nt][3] *= i1; }
  store_acc_bf16(ctx + size_t(b) * Lq * ldc + h * AT, ldc, q0 + warp * 16, Lq, o, 1.0f);
  if (lse && t == 0) {
    // natural-log LSE; +inf for a row without any visible key so that backward's exp() gives 0
    const int r = q0 + warp * 16 + gq;
    float* lg = lse + (size_t(b) * H + h) * Lq;
    if (r < Lq) lg[r] = l0 > 0.f ? (m0 / kLog2e + logf(l0)) : INFINITY;
    if (r + 8 < Lq) lg[r + 8] = l1 > 0.f ? (m1 / kLog2e + logf(l1)) : INFINITY;
  }
}

// ================================================================== backward: delta = rowsum(dctx * ctx) per head
__global__ void __launch_bounds__(256)
attn_delta_kernel(const __nv_bfloat16* __restrict__ ctx, const __nv_bfloat16* __restrict__ dctx, int ldc,
                  float* __restrict__ delta, int B, int H, int Lq) {
  pdl_launch_dependents();
  pdl_wait();
  const int row = blockIdx.x * 8 + (threadIdx.x >> 5);
  const int l = threadIdx.x & 31;
  if (row >= B * Lq) return;
  const int b = row / Lq, qi = row % Lq;
  for (int h = 0; h < H; ++h) {
    const size_t off = size_t(row) * ldc + h * AT + 2 * l;
    float2 a = __bfloat1622float2(*reinterpret_cast<const __nv_bfloat162*>(ctx + off));
    float2 d = __bfloat1622float2(*reinterpret_cast<const __nv_bfloat162*>(dctx + off));
    float s = a.x * d.x + a.y * d.y;
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) s += __shfl_xor_sync(0xffffffffu, s, o);
    if (l == 0) delta[(size_t(b) * H + h) * Lq + qi] = s;
  }
}

// ================================================================== backward: dK, dV   (CTA = 64 keys)
struct BwdKVSmem {
  __nv_bfloat16 K[AT * AT];
  __nv_bfloat16 V[AT * AT];
  __nv_bfloat16 Q[2][AT * AT];
  __nv_bfloat16 dO[2][AT * AT];
  float lse[2][AT];
  float delta[2][AT];
};

__global__ void __launch_bounds__(ATT_THREADS)
attn_bwd_dkv_kernel(const __nv_bfloat16* __restrict__ q, int ldq, const __nv_bfloat16* __restrict__ k,
                    const __nv_bfloat16* __restrict__ v, int ldkv, const uint8_t* __restrict__ mask,
                    const int* __restrict__ kv_end, const __nv_bfloat16* __restrict__ dctx, int ldc,
                    const float* __restrict__ lse, const float* __restrict__ delta, __nv_bfloat16* __restrict__ dk,
                    __nv_bfloat16* __restrict__ dv, int lddkv, int H, int Lq, int Lk) {
  extern __shared__ __align__(128) uint8_t smem_dyn[];
  BwdKVSmem& sm = *reinterpret_cast<BwdKVSmem*>(smem_dyn);
  const int k0 = blockIdx.x * AT, h = blockIdx.y, b = blockIdx.z;
  const int warp = threadIdx.x >> 5, l = threadIdx.x & 31, gq = l >> 2, t = l & 3;
  const int kend = kv_end ? min(kv_end[b], Lk) : Lk;
  __nv_bfloat16* dkg = dk + size_t(b) * Lk * lddkv + h * AT;
  __nv_bfloat16* dvg = dv + size_t(b) * Lk * lddkv + h * AT;
  float dka[8][4], dva[8][4];
  zero_acc(dka);
  zero_acc(dva);
  if (k0 >= kend) {   // whole tile is padding: exact zeros (padded keys receive no gradient)
    store_acc_bf16(dkg, lddkv, k0 + warp * 16, Lk, dka, 1.0f);
    store_acc_bf16(dvg, lddkv, k0 + warp * 16, Lk, dva, 1.0f);
    return;
  }
  const __nv_bfloat16* qg = q + size_t(b) * Lq * ldq + h * AT;
  const __nv_bfloat16* dog = dctx + size_t(b) * Lq * ldc + h * AT;
  const float* lg = lse + (size_t(b) * H + h) * Lq;
  const float* dg = delta + (size_t(b) * H + h) * Lq;
  const int nq = (Lq + AT - 1) / AT;

  tile_load_async(sm.K, k + size_t(b) * Lk * ldkv + h * AT, ldkv, k0, kend);
  tile_load_async(sm.V, v + size_t(b) * Lk * ldkv + h * AT, ldkv, k0, kend);
  auto load_q = [&](int i, int buf) {
    tile_load_async(sm.Q[buf], qg, ldq, i * AT, Lq);
    tile_load_async(sm.dO[buf], dog, ldc, i * AT, Lq);
    if (threadIdx.x < AT) {
      const int r = i * AT + threadIdx.x;
      sm.lse[buf][threadIdx.x] = r < Lq ? lg[r] * kLog2e : INFINITY;   // +inf => p = 0 for rows past Lq
      sm.delta[buf][threadIdx.x] = r < Lq ? dg[r] : 0.f;
    }
  };
  load_q(0, 0);
  cp_async_commit();
  cp_async_wait<0>();
  __syncthreads();
  uint32_t ak[4][4], av[4][4];
  load_a_frags(sm.K, warp * 16, ak);
  load_a_frags(sm.V, warp * 16, av);
  // key-side mask of this thread's two rows (keys g and g+8 of the warp's 16)
  const int key0 = k0 + warp * 16 + gq, key1 = key0 + 8;
  const bool km0 = key0 >= kend || (mask && mask[size_t(b) * Lk + key0] != 0);
  const bool km1 = key1 >= kend || (mask && mask[size_t(b) * Lk + key1] != 0);

  for (int i = 0; i < nq; ++i) {
    const int buf = i & 1;
    if (i + 1 < nq) load_q(i + 1, buf ^ 1);
    cp_async_commit();
    cp_async_wait<1>();
    __syncthreads();
    float st[8][4], dp[8][4];
    zero_acc(st);
    zero_acc(dp);
    mma_a_tileT(st, ak, sm.Q[buf]);    // S^T[key, q]
    mma_a_tileT(dp, av, sm.dO[buf]);   // dP^T[key, q]
#pragma unroll
    for (int nt = 0; nt < 8; ++nt) {
      const int c = nt * 8 + 2 * t;
      const float ls0 = sm.lse[buf][c], ls1 = sm.lse[buf][c + 1];
      const float de0 = sm.delta[buf][c], de1 = sm.delta[buf][c + 1];
      const float p0 = km0 ? 0.f : exp2f(st[nt][0] * kScaleLog2 - ls0);
      const float p1 = km0 ? 0.f : exp2f(st[nt][1] * kScaleLog2 - ls1);
      const float p2 = km1 ? 0.f : exp2f(st[nt][2] * kScaleLog2 - ls0);
      const float p3 = km1 ? 0.f : exp2f(st[nt][3] * kScaleLog2 - ls1);
      st[nt][0] = p0; st[nt][1] = p1; st[nt][2] = p2; st[nt][3] = p3;
      dp[nt][0] = p0 * (dp[nt][0] - de0);
      dp[nt][1] = p1 * (dp[nt][1] - de1);
      dp[nt][2] = p2 * (dp[nt][2] - de0);
      dp[nt][3] = p3 * (dp[nt][3] - de1);
    }
    mma_p_tile(dva, st, sm.dO[buf]);   // dV += P^T dO
    mma_p_tile(dka, dp, sm.Q[buf]);    // dK += dS^T Q
    __syncthreads();
  }
  store_acc_bf16(dkg, lddkv, k0 + warp * 16, Lk, dka, 0.125f);
  store_acc_bf16(dvg, lddkv, k0 + warp * 16, Lk, dva, 1.0f);
}

// ================================================================== backward: dQ   (CTA = 64 queries)
__global__ void __launch_bounds__(ATT_THREADS)
attn_bwd_dq_kernel(const __nv_bfloat16* __restrict__ q, int ldq, const __nv_bfloat16* __restrict__ k,
                   const __nv_bfloat16* __restrict__ v, int ldkv, const uint8_t* __restrict__ mask,
                   const int* __restrict__ kv_end, const __nv_bfloat16* __restrict__ dctx, int ldc,
                   const float* __restrict__ lse, const float* __restrict__ delta, __nv_bfloat16* __restrict__ dq,
                   int lddq, int H, int Lq, int Lk) {
  __shared__ __align__(128) __nv_bfloat16 sQ[AT * AT];    // Q, then reused for dO
  __shared__ __align__(128) __nv_bfloat16 sK[2][AT * AT];
  __shared__ __align__(128) __nv_bfloat16 sV[2][AT * AT];
  __shared__ uint8_t sM[2][AT];
  const int q0 = blockIdx.x * AT, h = blockIdx.y, b = blockIdx.z;
  const int warp = threadIdx.x >> 5, l = threadIdx.x & 31, gq = l >> 2, t = l & 3;
  const int kend = kv_end ? min(kv_end[b], Lk) : Lk;
  const int nblk = (kend + AT - 1) / AT;
  const __nv_bfloat16* kg = k + size_t(b) * Lk * ldkv + h * AT;
  const __nv_bfloat16* vg = v + size_t(b) * Lk * ldkv + h * AT;
  const uint8_t* mg = mask ? mask + size_t(b) * Lk : nullptr;

  uint32_t aq[4][4], ado[4][4];
  tile_load_async(sQ, q + size_t(b) * Lq * ldq + h * AT, ldq, q0, Lq);
  cp_async_commit();
  cp_async_wait<0>();
  __syncthreads();
  load_a_frags(sQ, warp * 16, aq);
  __syncthreads();
  tile_load_async(sQ, dctx + size_t(b) * Lq * ldc + h * AT, ldc, q0, Lq);
  cp_async_commit();
  cp_async_wait<0>();
  __syncthreads();
  load_a_frags(sQ, warp * 16, ado);

  const int r0 = q0 + warp * 16 + gq, r1 = r0 + 8;
  const float* lg = lse + (size_t(b) * H + h) * Lq;
  const float* dg = delta + (size_t(b) * H + h) * Lq;
  const float ls0 = r0 < Lq ? lg[r0] * kLog2e : INFINITY, ls1 = r1 < Lq ? lg[r1] * kLog2e : INFINITY;
  const float de0 = r0 < Lq ? dg[r0] : 0.f, de1 = r1 < Lq ? dg[r1] : 0.f;

  auto load_kv = [&](int j, int buf) {
    tile_load_async(sK[buf], kg, ldkv, j * AT, kend);
    tile_load_async(sV[buf], vg, ldkv, j * AT, kend);
    if (threadIdx.x < AT) {
      const int col = j * AT + threadIdx.x;
      sM[buf][threadIdx.x] = (col >= kend) ? 1 : (mg ? mg[col] : 0);
    }
  };
  if (nblk > 0) load_kv(0, 0);
  cp_async_commit();
  float dqa[8][4];
  zero_acc(dqa);
  for (int j = 0; j < nblk; ++j) {
    const int buf = j & 1;
    if (j + 1 < nblk) load_kv(j + 1, buf ^ 1);
    cp_async_commit();
    cp_async_wait<1>();
    __syncthreads();
    float s[8][4], dp[8][4];
    zero_acc(s);
    zero_acc(dp);
    mma_a_tileT(s, aq, sK[buf]);     // S  = Q K^T
    mma_a_tileT(dp, ado, sV[buf]);   // dP = dO V^T
#pragma unroll
    for (int nt = 0; nt < 8; ++nt) {
      const bool k0m = sM[buf][nt * 8 + 2 * t] != 0, k1m = sM[buf][nt * 8 + 2 * t + 1] != 0;
      const float p0 = k0m ? 0.f : exp2f(s[nt][0] * kScaleLog2 - ls0);
      const float p1 = k1m ? 0.f : exp2f(s[nt][1] * kScaleLog2 - ls0);
      const float p2 = k0m ? 0.f : exp2f(s[nt][2] * kScaleLog2 - ls1);
      const float p3 = k1m ? 0.f : exp2f(s[nt][3] * kScaleLog2 - ls1);
      dp[nt][0] = p0 * (dp[nt][0] - de0);
      dp[nt][1] = p1 * (dp[nt][1] - de0);
      dp[nt][2] = p2 * (dp[nt][2] - de1);
      dp[nt][3] = p3 * (dp[nt][3] - de1);
    }
    mma_p_tile(dqa, dp, sK[buf]);    // dQ += dS K
    __syncthreads();
  }
  store_acc_bf16(dq + size_t(b) * Lq * lddq + h * AT, lddq, q0 + warp * 16, Lq, dqa, 0.125f);
}

// ================================================================== host launchers
static inline int chk() { count_launch(); return cudaGetLastError() == cudaSuccess ? 0 : GXA_ERR_CUDA; }

int attn_fwd_tc(const void* q, int ldq, const void* kv, int ldkv, const uint8_t* mask, const int* kv_end, void* ctx,
                int ldc, float* lse, int B, int H, int Lq, int Lk, cudaStream_t s, DropCfg drop);

// GXA_ATTN=legacy selects the mma.sync kernels (A/B comparison); default is the tcgen05 path
static bool use_legacy_attn() {
  static int v = -1;
  if (v < 0) {
    const char* e = getenv("GXA_ATTN");
    v = (e && e[0] == 'l') ? 1 : 0;
  }
  return v == 1;
}

int attn_fwd_bf16(const void* q, int ldq, const void* k, const void* v, int ldkv, const uint8_t* mask, const int* kv_end,
                  void* ctx, int ldc, float* lse, int B, int H, int Lq, int Lk, cudaStream_t s, DropCfg drop) {
  if (B <= 0 || H <= 0 || Lq <= 0 || Lk <= 0) return GXA_ERR_SHAPE;
  if ((ldq % 8) || (ldkv % 8) || (ldc % 8)) return GXA_ERR_ALIGN;
  // tcgen05 path: K|V packed side by side in one [B*Lk, 2D] matrix (the KV projection's natural output)
  if (!use_legacy_attn() && Lk <= 1024 && ldkv == 2 * H * 64 &&
      v == reinterpret_cast<const __nv_bfloat16*>(k) + H * 64 && (reinterpret_cast<uintptr_t>(ctx) & 15) == 0)
    return attn_fwd_tc(q, ldq, k, ldkv, mask, kv_end, ctx, ldc, lse, B, H, Lq, Lk, s, drop);
  if (drop.thresh != 0) return GXA_ERR_UNSUPPORTED;     // dropout exists on the tcgen05 path only
  dim3 grid((Lq + AT - 1) / AT, H, B);
  attn_fwd_kernel<<<grid, ATT_THREADS, 0, s>>>(
      reinterpret_cast<const __nv_bfloat16*>(q), ldq, reinterpret_cast<const __nv_bfloat16*>(k),
      reinterpret_cast<const __nv_bfloat16*>(v), ldkv, mask, kv_end, reinterpret_cast<__nv_bfloat16*>(ctx), ldc, lse, H,
      Lq, Lk);
  return chk();
}

int attn_bwd_tc(const void* q, int ldq, const void* kv, int ldkv, const uint8_t* mask, const int* kv_end,
                const void* dctx, int ldc, const float* lse, const float* delta, float* dq_acc, void* dq, int lddq,
                void* dk, void* dv, int lddkv, int B, int H, int Lq, int Lk, cudaStream_t s);
int attn_bwd_tc2(const void* q, int ldq, const void* kv, int ldkv, const uint8_t* mask, const int* kv_end,
                 const void* dctx, int ldc, const float* lse, const float* delta, float* dq_acc, void* dq, int lddq,
                 void* dk, void* dv, int lddkv, int B, int H, int Lq, int Lk, float* dq_colsum, cudaStream_t s, DropCfg drop);
size_t attn_bwd_tc_workspace(int B, int H, int Lq);

size_t attn_bwd_workspace_bytes(int B, int H, int Lq) { return attn_bwd_tc_workspace(B, H, Lq); }

int attn_bwd_bf16(const void* q, int ldq, const void* k, const void* v, int ldkv, const uint8_t* mask, const int* kv_end,
                  const void* ctx, const void* dctx, int ldc, const float* lse, float* delta, void* dq, int lddq,
                  void* dk, void* dv, int lddkv, void* workspace, int B, int H, int Lq, int Lk, cudaStream_t s,
                  float* dq_colsum, DropCfg drop) {
  if (B <= 0 || H <= 0 || Lq <= 0 || Lk <= 0) return GXA_ERR_SHAPE;
  if ((ldq % 8) || (ldkv % 8) || (ldc % 8) || (lddq % 8) || (lddkv % 8)) return GXA_ERR_ALIGN;
  auto Q = reinterpret_cast<const __nv_bfloat16*>(q);
  auto K = reinterpret_cast<const __nv_bfloat16*>(k);
  auto V = reinterpret_cast<const __nv_bfloat16*>(v);
  auto dO = reinterpret_cast<const __nv_bfloat16*>(dctx);
  launch_pdl(attn_delta_kernel, dim3((B * Lq + 7) / 8), dim3(256), 0, s, reinterpret_cast<const __nv_bfloat16*>(ctx), dO, ldc,
             delta, B, H, Lq);
  if (int e = chk()) return e;
  // tcgen05 path: K|V (and dK|dV) packed side by side, scratch for the fp32 dQ accumulator available
  const bool packed = workspace != nullptr && ldkv == 2 * H * 64 && lddkv == ldkv && V == K + H * 64 &&
                      reinterpret_cast<__nv_bfloat16*>(dv) == reinterpret_cast<__nv_bfloat16*>(dk) + H * 64;
  static const bool bwd_v1 = getenv("GXA_ATTN_BWD") && getenv("GXA_ATTN_BWD")[0] == '1';   // A/B: per-item CTAs
  if (!use_legacy_attn() && packed && !bwd_v1)       // persistent tcgen05 kernel (any Lq, Lk)
    return attn_bwd_tc2(q, ldq, k, ldkv, mask, kv_end, dctx, ldc, lse, delta, reinterpret_cast<float*>(workspace), dq,
                        lddq, dk, dv, lddkv, B, H, Lq, Lk, dq_colsum, s, drop);
  if (drop.thresh != 0) return GXA_ERR_UNSUPPORTED;     // dropout exists on the persistent tcgen05 path only
  if (dq_colsum != nullptr) {     // the other paths have no fused d(bq): one pass over dq afterwards
    const int e = attn_bwd_bf16(q, ldq, k, v, ldkv, mask, kv_end, ctx, dctx, ldc, lse, delta, dq, lddq, dk, dv, lddkv,
                                workspace, B, H, Lq, Lk, s, nullptr, drop);
    return e ? e : colsum(dq, 0, lddq, dq_colsum, B * Lq, H * 64, s);
  }
  if (!use_legacy_attn() && packed && Lk <= 1024 && Lq <= 1024)
    return attn_bwd_tc(q, ldq, k, ldkv, mask, kv_end, dctx, ldc, lse, delta, reinterpret_cast<float*>(workspace), dq,
                       lddq, dk, dv, lddkv, B, H, Lq, Lk, s);
  static bool attr = false;
  if (!attr) {
    if (cudaFuncSetAttribute(attn_bwd_dkv_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize,
                             (int)sizeof(BwdKVSmem)) != cudaSuccess)
      return GXA_ERR_CUDA;
    attr = true;
  }
  attn_bwd_dkv_kernel<<<dim3((Lk + AT - 1) / AT, H, B), ATT_THREADS, sizeof(BwdKVSmem), s>>>(
      Q, ldq, K, V, ldkv, mask, kv_end, dO, ldc, lse, delta, reinterpret_cast<__nv_bfloat16*>(dk),
      reinterpret_cast<__nv_bfloat16*>(dv), lddkv, H, Lq, Lk);
  if (int e = chk()) return e;
  attn_bwd_dq_kernel<<<dim3((Lq + AT - 1) / AT, H, B), ATT_THREADS, 0, s>>>(
      Q, ldq, K, V, ldkv, mask, kv_end, dO, ldc, lse, delta, reinterpret_cast<__nv_bfloat16*>(dq), lddq, H, Lq, Lk);
  return chk();
}

}  // namespace gxa
