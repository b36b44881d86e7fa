// tcgen05 / TMEM masked-softmax attention, head_dim 64, bf16 operands, fp32 softmax.
//
// Forward: one CTA = 128 queries of one (batch, head); 128-key blocks up to kv_end[b].
//   warp 4  : TMA producer (Q once; K double-buffered, V single-buffered; 3-D tensor maps so that rows past
//             a sample's length are zero-filled, never another sample's data)
//   warp 5  : TMEM allocator + tcgen05.mma issuer:  S_j = Q K_j^T (128x128x64, TMEM cols 0..127) and
//             O_j = P_j V_j (128x64x128, TMEM cols 128..191; V consumed MN-major straight from its TMA tile)
//   warps 0-3: softmax, one thread per query row: two passes over S_j in TMEM (row max, then exp2 / sum),
//             P_j written as bf16 into a 128B-swizzled K-major smem tile (the A operand of the PV MMA),
//             running output kept in registers: O = (O + O_{j-1}) * alpha_j.
// 96 KB smem and 256 TMEM columns per CTA -> two CTAs per SM overlap each other's softmax and MMA phases.
#include <cstdlib>

#include "gxa_host.h"
#include "gxa_ptx.cuh"

namespace gxa {

constexpr int TC_BQ = 128, TC_BK = 128, TC_HD = 64;
constexpr int TC_TILE = 128 * 64 * 2;                    // one 128-row x 64-col bf16 tile (16 KB)
constexpr float kLog2e_tc = 1.4426950408889634f;
constexpr float kScaleLog2_tc = 0.125f * kLog2e_tc;

constexpr int TC_MAXK = 1024;                           // longest key sequence the tcgen05 path handles
constexpr int TC_SOFTMAX_WARPS = 8;
constexpr int TC_FWD_THREADS = (TC_SOFTMAX_WARPS + 2) * 32;

struct FwdTcSmem {
  uint8_t q[TC_TILE];
  uint8_t k[2][TC_TILE];
  uint8_t v[TC_TILE];
  uint8_t p[2 * TC_TILE];          // [128 q][128 keys] bf16 as two 64-key swizzle atoms
  float kbias[TC_MAXK];            // 0 for a visible key, -inf for a padded / masked one
  float pmax[2][2][128];           // [block parity][column half][row]: partial row maxima
  float psum[2][128];
  int blk_masked[TC_MAXK / TC_BK];
  uint64_t q_full, k_full[2], k_empty[2], v_full, v_empty, s_full, s_free, p_ready, pv_done;
  uint32_t tmem_slot;
};
constexpr int FWD_TC_SMEM = sizeof(FwdTcSmem) + 1024;

__device__ __forceinline__ uint32_t pack_bf16x2(float lo, float hi) {
  __nv_bfloat162 t = __floats2bfloat162_rn(lo, hi);
  return *reinterpret_cast<uint32_t*>(&t);
}
__device__ __forceinline__ float ex2_approx(float x) {
  float y;
  asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x));
  return y;
}
__device__ __forceinline__ void softmax_bar_sync() { asm volatile("bar.sync 1, 256;" ::: "memory"); }

__global__ void __launch_bounds__(TC_FWD_THREADS, 2)
attn_fwd_tc_kernel(const __grid_constant__ CUtensorMap tmQ, const __grid_constant__ CUtensorMap tmKV,
                   const uint8_t* __restrict__ mask, const int* __restrict__ kv_end, __nv_bfloat16* __restrict__ ctx,
                   int ldc, float* __restrict__ lse, int H, int Lq, int Lk, int D, const DropCfg drop) {
  extern __shared__ __align__(1024) uint8_t smem_raw[];   // direct use keeps the shared state space (LDS/STS, not generic)
  FwdTcSmem& sm = *reinterpret_cast<FwdTcSmem*>(smem_raw);
  if ((smem_u32(smem_raw) & 1023u) != 0) __trap();
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int q0 = blockIdx.x * TC_BQ, h = blockIdx.y, b = blockIdx.z;
  const int kend = kv_end ? min(kv_end[b], Lk) : Lk;
  const int nblk = (kend + TC_BK - 1) / TC_BK;

  if (threadIdx.x == 0) {
    mbar_init(&sm.q_full, 1);
    mbar_init(&sm.k_full[0], 1); mbar_init(&sm.k_full[1], 1);
    mbar_init(&sm.k_empty[0], 1); mbar_init(&sm.k_empty[1], 1);
    mbar_init(&sm.v_full, 1); mbar_init(&sm.v_empty, 1);
    mbar_init(&sm.s_full, 1); mbar_init(&sm.s_free, TC_SOFTMAX_WARPS);
    mbar_init(&sm.p_ready, TC_SOFTMAX_WARPS); mbar_init(&sm.pv_done, 1);
    mbar_fence_init();
  }
  if (threadIdx.x < TC_MAXK / TC_BK) sm.blk_masked[threadIdx.x] = 0;
  if (warp == TC_SOFTMAX_WARPS && lane == 0) { tma_prefetch_desc(&tmQ); tma_prefetch_desc(&tmKV); }
  if (warp == TC_SOFTMAX_WARPS + 1) tmem_alloc(&sm.tmem_slot, 256);
  __syncthreads();
  {   // key visibility of this sample, once per CTA: the mask is consumed as bytes, never as a float tensor
    const uint8_t* mg = mask ? mask + size_t(b) * Lk : nullptr;
    for (int key = threadIdx.x; key < nblk * TC_BK; key += TC_FWD_THREADS) {
      const bool dead = key >= kend || (mg && mg[key] != 0);
      sm.kbias[key] = dead ? -INFINITY : 0.f;
      if (dead) sm.blk_masked[key >> 7] = 1;
    }
  }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem = sm.tmem_slot;
  const uint32_t tmem_s = tmem, tmem_o = tmem + 128;

  if (warp == TC_SOFTMAX_WARPS) {
    // ------------------------------------------------------------ TMA producer
    if (lane == 0 && nblk > 0) {
      mbar_arrive_expect_tx(&sm.q_full, TC_TILE);
      tma_load_3d(sm.q, &tmQ, &sm.q_full, h * TC_HD, q0, b);
      for (int j = 0; j < nblk; ++j) {
        const int kb = j & 1;
        mbar_wait(&sm.k_empty[kb], ((j >> 1) & 1) ^ 1);
        mbar_arrive_expect_tx(&sm.k_full[kb], TC_TILE);
        tma_load_3d(sm.k[kb], &tmKV, &sm.k_full[kb], h * TC_HD, j * TC_BK, b);
        mbar_wait(&sm.v_empty, (j & 1) ^ 1);
        mbar_arrive_expect_tx(&sm.v_full, TC_TILE);
        tma_load_3d(sm.v, &tmKV, &sm.v_full, D + h * TC_HD, j * TC_BK, b);
      }
    }
  } else if (warp == TC_SOFTMAX_WARPS + 1) {
    // ------------------------------------------------------------ MMA issuer
    if (nblk > 0) {
      constexpr uint32_t idesc_s = umma_idesc_bf16(128, 128, false, false);
      constexpr uint32_t idesc_pv = umma_idesc_bf16(128, 64, false, true);
      // converged issue (gxa_ptx.cuh): warp-uniform descriptors in all lanes, tcgen05 predicated on the elected lane
      constexpr uint32_t HI = umma_desc_hi(1024);
      const uint32_t el = elect_one_pred();
      const uint32_t qd = umma_desc_lo(smem_u32(sm.q), 16), pd = umma_desc_lo(smem_u32(sm.p), 16);
      const uint32_t vd = umma_desc_lo(smem_u32(sm.v), 8192);
      auto issue_s = [&](int j) {
        mbar_wait(&sm.k_full[j & 1], (j >> 1) & 1);
        tc_fence_after();
        const uint32_t kd = umma_desc_lo(smem_u32(sm.k[j & 1]), 16);
#pragma unroll
        for (int k = 0; k < 4; ++k) umma_bf16_e(tmem_s, qd + k * 2, HI, kd + k * 2, HI, idesc_s, k > 0 ? 1u : 0u, el);
        umma_commit_e(&sm.s_full, el);
        umma_commit_e(&sm.k_empty[j & 1], el);
      };
      mbar_wait(&sm.q_full, 0);
      issue_s(0);
      for (int j = 0; j < nblk; ++j) {
        if (j + 1 < nblk) {                       // next S as soon as the softmax warps have read this one
          mbar_wait(&sm.s_free, j & 1);
          issue_s(j + 1);
        }
        mbar_wait(&sm.p_ready, j & 1);
        mbar_wait(&sm.v_full, j & 1);
        tc_fence_after();
#pragma unroll
        for (int k = 0; k < 8; ++k)
          umma_bf16_e(tmem_o, pd + (k >> 2) * (TC_TILE >> 4) + (k & 3) * 2, HI, vd + k * 128, HI, idesc_pv, k > 0 ? 1u : 0u,
                      el);
        umma_commit_e(&sm.pv_done, el);
        umma_commit_e(&sm.v_empty, el);
      }
    }
  } else {
    // ------------------------------------------------------------ softmax: 2 threads per query row
    // warp w: TMEM lane quarter w&3 (rows), column half w>>2: S columns [64*half, +64), O columns [32*half, +32)
    const int quarter = warp & 3, half = warp >> 2;
    const int r = quarter * 32 + lane;
    const uint32_t lane_addr = static_cast<uint32_t>(quarter * 32) << 16;
    float m = -INFINITY, l = 0.f;
    const float drop_log2 = drop.thresh != 0 ? log2f(drop.scale) : 0.f;
    float o[32];
#pragma unroll
    for (int i = 0; i < 32; ++i) o[i] = 0.f;
    const uint32_t p_row = smem_u32(sm.p) + half * TC_TILE + r * 128;     // this thread's 64 keys = swizzle atom `half`
    // dropout counter of this query row: (b, h, q) * ceil(Lk / 2) + key / 2   (gxa_rng.cuh)
    const uint32_t drop_row = (static_cast<uint32_t>(b * H + h) * Lq + (q0 + r)) * static_cast<uint32_t>((Lk + 1) >> 1);

    for (int j = 0; j < nblk; ++j) {
      const float* kb = sm.kbias + j * TC_BK + half * 64;
      const bool any_masked = sm.blk_masked[j] != 0;
      mbar_wait(&sm.s_full, j & 1);
      tc_fence_after();
      // ---- pass 1: partial row max over this thread's 64 raw scores
      float mx = -INFINITY;
#pragma unroll
      for (int c = 0; c < 64; c += 32) {
        uint32_t sr[32];
        tmem_ld_32x32(tmem_s + lane_addr + half * 64 + c, sr);
        tmem_ld_wait();
        if (!any_masked) {
#pragma unroll
          for (int i = 0; i < 32; i += 2) mx = fmaxf(mx, fmaxf(__uint_as_float(sr[i]), __uint_as_float(sr[i + 1])));
        } else {
#pragma unroll
          for (int i = 0; i < 32; ++i) mx = fmaxf(mx, __uint_as_float(sr[i]) + kb[c + i]);
        }
      }
      sm.pmax[j & 1][half][r] = mx;
      softmax_bar_sync();
      mx = fmaxf(mx, sm.pmax[j & 1][half ^ 1][r]);
      const float m_new = fmaxf(m, mx);
      const float m_use = (m_new == -INFINITY) ? 0.f : m_new;
      const float alpha = ex2_approx((m - m_use) * kScaleLog2_tc);      // m = -inf -> 0
      // ---- fold in the previous block's P V product, then rescale to the new maximum
      if (j > 0) {
        mbar_wait(&sm.pv_done, (j - 1) & 1);
        tc_fence_after();
        uint32_t orr[32];
        tmem_ld_32x32(tmem_o + lane_addr + half * 32, orr);
        tmem_ld_wait();
#pragma unroll
        for (int i = 0; i < 32; ++i) o[i] = (o[i] + __uint_as_float(orr[i])) * alpha;
      }
      l *= alpha;
      m = m_new;
      // ---- pass 2: p = exp2((s - m) * scale*log2e), partial row sum, bf16 P tile
      const float mneg = -m_use * kScaleLog2_tc + drop_log2;    // dropout: p carries the 1/(1-p) factor
#pragma unroll
      for (int c = 0; c < 64; c += 32) {
        uint32_t sr[32];
        tmem_ld_32x32(tmem_s + lane_addr + half * 64 + c, sr);
        tmem_ld_wait();
        float pv[32];
        if (!any_masked) {
#pragma unroll
          for (int i = 0; i < 32; ++i) pv[i] = ex2_approx(fmaf(__uint_as_float(sr[i]), kScaleLog2_tc, mneg));
        } else {
#pragma unroll
          for (int i = 0; i < 32; ++i)
            pv[i] = ex2_approx(fmaf(__uint_as_float(sr[i]) + kb[c + i], kScaleLog2_tc, mneg));   // -inf -> 0
        }
#pragma unroll
        for (int i = 0; i < 32; ++i) l += pv[i];
        if (drop.thresh != 0) {      // dropout on the probabilities: the row sum above stays undropped (softmax first);
          // the 1/(1-p) factor is already in pv (folded into the exponent), so dropping is one select per element
          const uint32_t ctr = drop_row + ((j * TC_BK + half * 64 + c) >> 1);
#pragma unroll
          for (int i = 0; i < 16; ++i) {
            bool k0, k1;
            drop_keep(drop, ctr + i, k0, k1);
            pv[2 * i] = k0 ? pv[2 * i] : 0.f;
            pv[2 * i + 1] = k1 ? pv[2 * i + 1] : 0.f;
          }
        }
#pragma unroll
        for (int g = 0; g < 4; ++g) {
          const int chunk = (c >> 3) + g;
          const uint32_t addr = p_row + ((chunk ^ (r & 7)) << 4);
          asm volatile("st.shared.v4.b32 [%0], {%1, %2, %3, %4};" ::"r"(addr),
                       "r"(pack_bf16x2(pv[8 * g + 0], pv[8 * g + 1])), "r"(pack_bf16x2(pv[8 * g + 2], pv[8 * g + 3])),
                       "r"(pack_bf16x2(pv[8 * g + 4], pv[8 * g + 5])), "r"(pack_bf16x2(pv[8 * g + 6], pv[8 * g + 7]))
                       : "memory");
        }
      }
      tc_fence_before();                 // our TMEM reads of S_j are complete
      fence_proxy_async_smem();          // P_j visible to the tensor core (async proxy)
      __syncwarp();
      if (lane == 0) {
        mbar_arrive(&sm.s_free);
        mbar_arrive(&sm.p_ready);
      }
    }
    // ---- last block's product, total row sum, normalise, store
    sm.psum[half][r] = l;
    softmax_bar_sync();
    l += sm.psum[half ^ 1][r];
    if (drop.thresh != 0) l *= 1.0f / drop.scale;      // back to the plain softmax denominator; O keeps the factor
    const float inv = l > 0.f ? 1.0f / l : 0.f;
    if (nblk > 0) {
      mbar_wait(&sm.pv_done, (nblk - 1) & 1);
      tc_fence_after();
      uint32_t orr[32];
      tmem_ld_32x32(tmem_o + lane_addr + half * 32, orr);
      tmem_ld_wait();
#pragma unroll
      for (int i = 0; i < 32; ++i) o[i] = (o[i] + __uint_as_float(orr[i])) * inv;
    }
    const int row = q0 + r;
    if (row < Lq) {
      uint4* dst = reinterpret_cast<uint4*>(ctx + (size_t(b) * Lq + row) * ldc + h * TC_HD + half * 32);
#pragma unroll
      for (int g = 0; g < 4; ++g)
        dst[g] = make_uint4(pack_bf16x2(o[8 * g], o[8 * g + 1]), pack_bf16x2(o[8 * g + 2], o[8 * g + 3]),
                            pack_bf16x2(o[8 * g + 4], o[8 * g + 5]), pack_bf16x2(o[8 * g + 6], o[8 * g + 7]));
      if (lse && half == 0) lse[(size_t(b) * H + h) * Lq + row] = l > 0.f ? (m * 0.125f + logf(l)) : INFINITY;
    }
  }
  tc_fence_before();
  __syncthreads();
  if (warp == TC_SOFTMAX_WARPS + 1) {
    tc_fence_after();
    tmem_dealloc(tmem, 256);
  }
}

// =====================================================================================================
// Forward, second generation (default).  Same CTA shape and roles as above; what changed is where P and O live:
//   * P_j goes registers -> TMEM (tcgen05.st) and is the TMEM A operand of the P V MMA: no 32 KB shared-memory
//     write + 32 KB operand read per key block (the first kernel was shared-memory-bandwidth bound: 176 KB / block)
//   * O accumulates in TMEM over all key blocks; the running maximum is only raised when a block exceeds it by more
//     than 2^8 (then the warp rescales its O columns in TMEM), so the steady state never reads O back
//   * S_j is read once (64 registers per thread) and released to the MMA warp at once; the two threads of a query
//     row meet at a 64-thread named barrier instead of a CTA-wide one
// TMEM columns: S 0..127, P 128..191 (bf16 pairs), O 192..255.   smem: Q, K x2, V x2 = 80 KB -> two CTAs per SM.
// =====================================================================================================
struct FwdTc2Smem {
  uint8_t q[TC_TILE];
  uint8_t k[2][TC_TILE];
  uint8_t v[2][TC_TILE];
  float kbias[TC_MAXK];
  float pmax[2][2][128];
  float psum[2][128];
  int blk_masked[TC_MAXK / TC_BK];
  uint64_t q_full, k_full[2], k_empty[2], v_full[2], v_empty[2], s_full, s_free, p_ready, pv_done;
  uint32_t tmem_slot;
};
constexpr int FWD_TC2_SMEM = sizeof(FwdTc2Smem) + 1024;
// the two warps that share a lane quarter (column halves of the same 32 query rows): barrier ids 1..4 as immediates
__device__ __forceinline__ void pair_bar_sync(int quarter) {
  switch (quarter) {
    case 0: asm volatile("bar.sync 1, 64;" ::: "memory"); break;
    case 1: asm volatile("bar.sync 2, 64;" ::: "memory"); break;
    case 2: asm volatile("bar.sync 3, 64;" ::: "memory"); break;
    default: asm volatile("bar.sync 4, 64;" ::: "memory"); break;
  }
}
constexpr float kLazyRescaleRaw = 8.0f / kScaleLog2_tc;      // raw-score distance that corresponds to a factor 2^8

__global__ void __launch_bounds__(TC_FWD_THREADS, 2)
attn_fwd_tc2_kernel(const __grid_constant__ CUtensorMap tmQ, const __grid_constant__ CUtensorMap tmKV,
                    const uint8_t* __restrict__ mask, const int* __restrict__ kv_end, __nv_bfloat16* __restrict__ ctx,
                    int ldc, float* __restrict__ lse, int H, int Lq, int Lk, int D, const DropCfg drop) {
  extern __shared__ __align__(1024) uint8_t smem_raw[];
  FwdTc2Smem& sm = *reinterpret_cast<FwdTc2Smem*>(smem_raw);
  if ((smem_u32(smem_raw) & 1023u) != 0) __trap();
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int q0 = blockIdx.x * TC_BQ, h = blockIdx.y, b = blockIdx.z;

  if (threadIdx.x == 0) {
    mbar_init(&sm.q_full, 1);
    for (int i = 0; i < 2; ++i) {
      mbar_init(&sm.k_full[i], 1); mbar_init(&sm.k_empty[i], 1);
      mbar_init(&sm.v_full[i], 1); mbar_init(&sm.v_empty[i], 1);
    }
    mbar_init(&sm.s_full, 1); mbar_init(&sm.s_free, TC_SOFTMAX_WARPS);
    mbar_init(&sm.p_ready, TC_SOFTMAX_WARPS); mbar_init(&sm.pv_done, 1);
    mbar_fence_init();
  }
  if (threadIdx.x < TC_MAXK / TC_BK) sm.blk_masked[threadIdx.x] = 0;
  if (warp == TC_SOFTMAX_WARPS && lane == 0) { tma_prefetch_desc(&tmQ); tma_prefetch_desc(&tmKV); }
  if (warp == TC_SOFTMAX_WARPS + 1) tmem_alloc(&sm.tmem_slot, 256);
  __syncthreads();
  pdl_wait();                           // prologue done; everything below reads / writes global memory
  const int kend = kv_end ? min(kv_end[b], Lk) : Lk;
  const int nblk = (kend + TC_BK - 1) / TC_BK;
  // the first two key blocks need no ring slot to be released: issue Q and those K / V tiles before the key-visibility
  // pass below, so that their HBM latency overlaps it (the first S is on every CTA's critical path)
  auto load_kv = [&](int j) {
    const int st = j & 1;
    mbar_arrive_expect_tx(&sm.k_full[st], TC_TILE);
    tma_load_3d(sm.k[st], &tmKV, &sm.k_full[st], h * TC_HD, j * TC_BK, b);
    mbar_arrive_expect_tx(&sm.v_full[st], TC_TILE);
    tma_load_3d(sm.v[st], &tmKV, &sm.v_full[st], D + h * TC_HD, j * TC_BK, b);
  };
  if (warp == TC_SOFTMAX_WARPS && lane == 0 && nblk > 0) {
    mbar_arrive_expect_tx(&sm.q_full, TC_TILE);
    tma_load_3d(sm.q, &tmQ, &sm.q_full, h * TC_HD, q0, b);
    load_kv(0);
    if (nblk > 1) load_kv(1);
  }
  {
    const uint8_t* mg = mask ? mask + size_t(b) * Lk : nullptr;
    for (int key = threadIdx.x; key < nblk * TC_BK; key += TC_FWD_THREADS) {
      const bool dead = key >= kend || (mg && mg[key] != 0);
      sm.kbias[key] = dead ? -INFINITY : 0.f;
      if (dead) sm.blk_masked[key >> 7] = 1;
    }
  }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem = sm.tmem_slot;
  const uint32_t tmem_s = tmem, tmem_p = tmem + 128, tmem_o = tmem + 192;

  if (warp == TC_SOFTMAX_WARPS) {
    // ------------------------------------------------------------ TMA producer
    if (lane == 0) {
      for (int j = 2; j < nblk; ++j) {           // blocks 0 and 1 were issued in the prologue
        const int st = j & 1;
        const uint32_t ph = ((j >> 1) & 1) ^ 1;
        mbar_wait(&sm.k_empty[st], ph);
        mbar_arrive_expect_tx(&sm.k_full[st], TC_TILE);
        tma_load_3d(sm.k[st], &tmKV, &sm.k_full[st], h * TC_HD, j * TC_BK, b);
        mbar_wait(&sm.v_empty[st], ph);
        mbar_arrive_expect_tx(&sm.v_full[st], TC_TILE);
        tma_load_3d(sm.v[st], &tmKV, &sm.v_full[st], D + h * TC_HD, j * TC_BK, b);
      }
    }
    if (lane == 0) pdl_launch_dependents();      // all loads of this CTA issued: the next kernel may be staged
  } else if (warp == TC_SOFTMAX_WARPS + 1) {
    // ------------------------------------------------------------ MMA issuer
    if (nblk > 0) {
      constexpr uint32_t idesc_s = umma_idesc_bf16(128, 128, false, false);
      constexpr uint32_t idesc_pv = umma_idesc_bf16(128, 64, false, true);
      constexpr uint32_t HI = umma_desc_hi(1024);
      const uint32_t el = elect_one_pred();
      const uint32_t qd = umma_desc_lo(smem_u32(sm.q), 16);
      auto issue_s = [&](int j) {
        mbar_wait(&sm.k_full[j & 1], (j >> 1) & 1);
        tc_fence_after();
        const uint32_t kd = umma_desc_lo(smem_u32(sm.k[j & 1]), 16);
#pragma unroll
        for (int k = 0; k < 4; ++k) umma_bf16_e(tmem_s, qd + k * 2, HI, kd + k * 2, HI, idesc_s, k > 0 ? 1u : 0u, el);
        umma_commit_e(&sm.s_full, el);
        umma_commit_e(&sm.k_empty[j & 1], el);
      };
      mbar_wait(&sm.q_full, 0);
      issue_s(0);
      for (int j = 0; j < nblk; ++j) {
        if (j + 1 < nblk) {                       // next S as soon as the softmax warps hold this one in registers
          mbar_wait(&sm.s_free, j & 1);
          issue_s(j + 1);
        }
        mbar_wait(&sm.p_ready, j & 1);
        mbar_wait(&sm.v_full[j & 1], (j >> 1) & 1);
        tc_fence_after();
        const uint32_t vd = umma_desc_lo(smem_u32(sm.v[j & 1]), 8192);
#pragma unroll
        for (int k = 0; k < 8; ++k)               // 16 keys per step: 8 TMEM columns of P, 16 rows of the V tile
          umma_bf16_ts_e(tmem_o, tmem_p + k * 8, vd + k * 128, HI, idesc_pv, (j > 0 || k > 0) ? 1u : 0u, el);
        umma_commit_e(&sm.pv_done, el);
        umma_commit_e(&sm.v_empty[j & 1], el);
      }
    }
  } else {
    // ------------------------------------------------------------ softmax: 2 threads per query row
    const int quarter = warp & 3, half = warp >> 2;
    const int r = quarter * 32 + lane;
    const uint32_t lane_addr = static_cast<uint32_t>(quarter * 32) << 16;
    float m_ref = -INFINITY, l = 0.f;
    const float drop_log2 = drop.thresh != 0 ? log2f(drop.scale) : 0.f;
    const uint32_t drop_row = (static_cast<uint32_t>(b * H + h) * Lq + (q0 + r)) * static_cast<uint32_t>((Lk + 1) >> 1);

    for (int j = 0; j < nblk; ++j) {
      const float* kb = sm.kbias + j * TC_BK + half * 64;
      const bool any_masked = sm.blk_masked[j] != 0;
      mbar_wait(&sm.s_full, j & 1);
      tc_fence_after();
      uint32_t s0[32], s1[32];
      tmem_ld_32x32(tmem_s + lane_addr + half * 64, s0);
      tmem_ld_32x32(tmem_s + lane_addr + half * 64 + 32, s1);
      tmem_ld_wait();
      tc_fence_before();
      __syncwarp();
      if (lane == 0) mbar_arrive(&sm.s_free);
      if (any_masked) {
#pragma unroll
        for (int i = 0; i < 32; i += 2) {
          const float2 a = __fadd2_rn(make_float2(__uint_as_float(s0[i]), __uint_as_float(s0[i + 1])),
                                      *reinterpret_cast<const float2*>(kb + i));
          const float2 c = __fadd2_rn(make_float2(__uint_as_float(s1[i]), __uint_as_float(s1[i + 1])),
                                      *reinterpret_cast<const float2*>(kb + 32 + i));
          s0[i] = __float_as_uint(a.x); s0[i + 1] = __float_as_uint(a.y);
          s1[i] = __float_as_uint(c.x); s1[i + 1] = __float_as_uint(c.y);
        }
      }
      float mx = -INFINITY, mx2 = -INFINITY;
#pragma unroll
      for (int i = 0; i < 32; i += 2) {
        mx = fmax3(mx, __uint_as_float(s0[i]), __uint_as_float(s0[i + 1]));
        mx2 = fmax3(mx2, __uint_as_float(s1[i]), __uint_as_float(s1[i + 1]));
      }
      mx = fmaxf(mx, mx2);
      sm.pmax[j & 1][half][r] = mx;
      pair_bar_sync(quarter);
      mx = fmaxf(mx, sm.pmax[j & 1][half ^ 1][r]);
      // ---- raise the reference maximum only when this block exceeds it by more than 2^8 (always on the first finite one)
      const bool grow = mx > m_ref + kLazyRescaleRaw;
      float alpha = 1.f;
      if (grow) {
        alpha = ex2_approx((m_ref - mx) * kScaleLog2_tc);     // m_ref = -inf -> 0 (nothing accumulated yet)
        m_ref = mx;
      }
      l *= alpha;
      if (j > 0) {       // every P V issued so far has landed in O and has finished reading the P columns we overwrite
        mbar_wait(&sm.pv_done, (j - 1) & 1);
        tc_fence_after();
      }
      if (j > 0 && __any_sync(0xffffffffu, grow)) {            // warp-uniform: the TMEM accesses are .aligned
#pragma unroll 1
        for (int c = 0; c < 32; c += 16) {                     // rare path: small chunks, no pressure on the S registers
          uint32_t orr[16];
          tmem_ld_32x16(tmem_o + lane_addr + half * 32 + c, orr);
          tmem_ld_wait();
#pragma unroll
          for (int i = 0; i < 16; ++i) orr[i] = __float_as_uint(__uint_as_float(orr[i]) * alpha);
          tmem_st_32x16(tmem_o + lane_addr + half * 32 + c, orr);
        }
      }
      const float m_use = (m_ref == -INFINITY) ? 0.f : m_ref;
      const float mneg = -m_use * kScaleLog2_tc + drop_log2;    // dropout: p carries the 1/(1-p) factor
      float2 la = make_float2(0.f, 0.f), lb = make_float2(0.f, 0.f);
      const float2 c2 = make_float2(kScaleLog2_tc, kScaleLog2_tc), mneg2 = make_float2(mneg, mneg);
#pragma unroll
      for (int c = 0; c < 2; ++c) {
        float pv[32];
#pragma unroll
        for (int i = 0; i < 32; i += 2) {      // packed f32x2 FMA for the exponent argument, packed adds for the row sum
          const float2 e = __ffma2_rn(make_float2(__uint_as_float(c == 0 ? s0[i] : s1[i]),
                                                  __uint_as_float(c == 0 ? s0[i + 1] : s1[i + 1])), c2, mneg2);
          pv[i] = ex2_approx(e.x);
          pv[i + 1] = ex2_approx(e.y);
        }
#pragma unroll
        for (int i = 0; i < 32; i += 4) {
          la = __fadd2_rn(la, make_float2(pv[i], pv[i + 1]));
          lb = __fadd2_rn(lb, make_float2(pv[i + 2], pv[i + 3]));
        }
        if (drop.thresh != 0) {      // the row sum above stays undropped (softmax first); dropping is one select per element
          const uint32_t ctr = drop_row + ((j * TC_BK + half * 64 + c * 32) >> 1);
#pragma unroll
          for (int i = 0; i < 16; ++i) {
            bool k0, k1;
            drop_keep(drop, ctr + i, k0, k1);
            pv[2 * i] = k0 ? pv[2 * i] : 0.f;
            pv[2 * i + 1] = k1 ? pv[2 * i + 1] : 0.f;
          }
        }
        uint32_t pk[16];
#pragma unroll
        for (int i = 0; i < 16; ++i) pk[i] = pack_bf16x2(pv[2 * i], pv[2 * i + 1]);
        tmem_st_32x16(tmem_p + lane_addr + half * 32 + c * 16, pk);
      }
      l += (la.x + la.y) + (lb.x + lb.y);
      tmem_st_wait();
      tc_fence_before();
      __syncwarp();
      if (lane == 0) mbar_arrive(&sm.p_ready);
    }
    // ---- total row sum, last product, normalise, store
    sm.psum[half][r] = l;
    pair_bar_sync(quarter);
    l += sm.psum[half ^ 1][r];
    if (drop.thresh != 0) l *= 1.0f / drop.scale;      // back to the plain softmax denominator; O keeps the factor
    const float inv = l > 0.f ? 1.0f / l : 0.f;
    float o[32];
#pragma unroll
    for (int i = 0; i < 32; ++i) o[i] = 0.f;
    if (nblk > 0) {
      mbar_wait(&sm.pv_done, (nblk - 1) & 1);
      tc_fence_after();
      uint32_t orr[32];
      tmem_ld_32x32(tmem_o + lane_addr + half * 32, orr);
      tmem_ld_wait();
#pragma unroll
      for (int i = 0; i < 32; ++i) o[i] = __uint_as_float(orr[i]) * inv;
    }
    const int row = q0 + r;
    if (row < Lq) {
      uint4* dst = reinterpret_cast<uint4*>(ctx + (size_t(b) * Lq + row) * ldc + h * TC_HD + half * 32);
#pragma unroll
      for (int g = 0; g < 4; ++g)
        dst[g] = make_uint4(pack_bf16x2(o[8 * g], o[8 * g + 1]), pack_bf16x2(o[8 * g + 2], o[8 * g + 3]),
                            pack_bf16x2(o[8 * g + 4], o[8 * g + 5]), pack_bf16x2(o[8 * g + 6], o[8 * g + 7]));
      if (lse && half == 0) lse[(size_t(b) * H + h) * Lq + row] = l > 0.f ? (m_ref * 0.125f + logf(l)) : INFINITY;
    }
  }
  tc_fence_before();
  __syncthreads();
  if (warp == TC_SOFTMAX_WARPS + 1) {
    tc_fence_after();
    tmem_dealloc(tmem, 256);
  }
}

int attn_fwd_tc(const void* q, int ldq, const void* kv, int ldkv, const uint8_t* mask, const int* kv_end, void* ctx,
                int ldc, float* lse, int B, int H, int Lq, int Lk, cudaStream_t s, DropCfg drop) {
  const int D = H * TC_HD;
  if (Lk > TC_MAXK) return GXA_ERR_UNSUPPORTED;
  CUtensorMap tmQ, tmKV;
  if (int e = make_tmap_bf16_3d(&tmQ, q, D, Lq, B, ldq, size_t(Lq) * ldq, 64, TC_BQ)) return e;
  if (int e = make_tmap_bf16_3d(&tmKV, kv, 2 * D, Lk, B, ldkv, size_t(Lk) * ldkv, 64, TC_BK)) return e;
  static bool attr = false;
  static int gen = 2;               // GXA_ATTN_FWD=1 selects the first-generation kernel (P through shared memory) for A/B
  if (!attr) {
    if (const char* e = getenv("GXA_ATTN_FWD")) gen = atoi(e) == 1 ? 1 : 2;
    if (cudaFuncSetAttribute(attn_fwd_tc_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, FWD_TC_SMEM) != cudaSuccess ||
        cudaFuncSetAttribute(attn_fwd_tc2_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, FWD_TC2_SMEM) != cudaSuccess)
      return GXA_ERR_CUDA;
    attr = true;
  }
  dim3 grid((Lq + TC_BQ - 1) / TC_BQ, H, B);
  if (gen == 1)
    attn_fwd_tc_kernel<<<grid, TC_FWD_THREADS, FWD_TC_SMEM, s>>>(tmQ, tmKV, mask, kv_end, reinterpret_cast<__nv_bfloat16*>(ctx),
                                                                 ldc, lse, H, Lq, Lk, D, drop);
  else
    launch_pdl(attn_fwd_tc2_kernel, grid, dim3(TC_FWD_THREADS), FWD_TC2_SMEM, s, tmQ, tmKV, mask, kv_end,
               reinterpret_cast<__nv_bfloat16*>(ctx), ldc, lse, H, Lq, Lk, D, drop);
  count_launch();
  return cudaGetLastError() == cudaSuccess ? 0 : GXA_ERR_CUDA;
}


// =====================================================================================================
// Backward.  One CTA = one 128-key block of one (batch, head); loops over the 128-query blocks.
//   S = Q_i K^T, dP = dO_i V^T            (TMEM cols 0..127, 128..255; single-buffered, refilled as soon as read)
//   P = exp2(S*c - lse), dS = P*(dP - delta)   16 warps, thread = (row, 32 key columns): no cross-thread reductions
//   dV += P^T dO_i, dK += dS^T Q_i         (TMEM cols 256..319, 320..383; accumulate over all query blocks; the
//                                           transposes are free: P/dS/dO/Q are read MN-major through descriptors)
//   dQ_i = dS K                            (TMEM cols 384..447) -> 4 drain warps -> fp32 smem tile -> TMA bulk
//                                           reduce-add into the [B,H,LqPad,64] fp32 accumulator (keys are split
//                                           over CTAs, so dQ needs a cross-CTA sum; no register-level atomics)
// warps 0-15 compute, 16-19 dQ drain, 20 TMA producer, 21 MMA issuer.
// =====================================================================================================
constexpr int TCB_COMPUTE_WARPS = 16;
constexpr int TCB_DRAIN_WARPS = 4;
constexpr int TCB_THREADS = (TCB_COMPUTE_WARPS + TCB_DRAIN_WARPS + 2) * 32;
constexpr int TCB_MAXQ = 1024;

struct BwdTcSmem {
  uint8_t k[TC_TILE];
  uint8_t v[TC_TILE];
  uint8_t q[2][TC_TILE];
  uint8_t dO[2][TC_TILE];
  uint8_t p[2 * TC_TILE];
  uint8_t ds[2 * TC_TILE];
  float dq_stage[128 * 64];
  float lse2[TCB_MAXQ];        // lse * log2(e); +inf for rows >= Lq so that P = 0 there
  float delta[TCB_MAXQ];
  float kbias[TC_BK];
  uint64_t kv_full, q_full[2], q_empty[2], sdp_full, sdp_free, pds_ready, pds_free, dq_full, dq_free, dkv_done;
  uint32_t tmem_slot;
  int any_masked;
};
constexpr int BWD_TC_SMEM = sizeof(BwdTcSmem) + 1024;

__global__ void __launch_bounds__(TCB_THREADS, 1)
attn_bwd_tc_kernel(const __grid_constant__ CUtensorMap tmQ, const __grid_constant__ CUtensorMap tmKV,
                   const __grid_constant__ CUtensorMap tmDO, const uint8_t* __restrict__ mask,
                   const int* __restrict__ kv_end, const float* __restrict__ lse, const float* __restrict__ delta,
                   float* __restrict__ dq_acc, __nv_bfloat16* __restrict__ dk, __nv_bfloat16* __restrict__ dv, int lddkv,
                   int H, int Lq, int Lk, int D, int LqPad, int dbg) {
  extern __shared__ __align__(1024) uint8_t smem_raw[];
  BwdTcSmem& sm = *reinterpret_cast<BwdTcSmem*>(smem_raw);
  if ((smem_u32(smem_raw) & 1023u) != 0) __trap();
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int k0 = blockIdx.x * TC_BK, h = blockIdx.y, b = blockIdx.z;
  const int kend = kv_end ? min(kv_end[b], Lk) : Lk;
  const int nq = (Lq + TC_BQ - 1) / TC_BQ;

  if (k0 >= kend) {      // the whole key block is padding: exact zeros, nothing else to do
    for (int idx = threadIdx.x; idx < TC_BK * 8; idx += TCB_THREADS) {
      const int key = k0 + (idx >> 3);
      if (key < Lk) {
        const size_t off = (size_t(b) * Lk + key) * lddkv + h * TC_HD + (idx & 7) * 8;
        *reinterpret_cast<uint4*>(dk + off) = make_uint4(0, 0, 0, 0);
        *reinterpret_cast<uint4*>(dv + off) = make_uint4(0, 0, 0, 0);
      }
    }
    return;
  }

  if (threadIdx.x == 0) {
    mbar_init(&sm.kv_full, 1);
    for (int i = 0; i < 2; ++i) { mbar_init(&sm.q_full[i], 1); mbar_init(&sm.q_empty[i], 1); }
    mbar_init(&sm.sdp_full, 1); mbar_init(&sm.sdp_free, TCB_COMPUTE_WARPS);
    mbar_init(&sm.pds_ready, TCB_COMPUTE_WARPS); mbar_init(&sm.pds_free, 1);
    mbar_init(&sm.dq_full, 1); mbar_init(&sm.dq_free, TCB_DRAIN_WARPS);
    mbar_init(&sm.dkv_done, 1);
    mbar_fence_init();
    sm.any_masked = 0;
  }
  if (warp == TCB_COMPUTE_WARPS + TCB_DRAIN_WARPS && lane == 0) {
    tma_prefetch_desc(&tmQ); tma_prefetch_desc(&tmKV); tma_prefetch_desc(&tmDO);
  }
  if (warp == TCB_COMPUTE_WARPS + TCB_DRAIN_WARPS + 1) tmem_alloc(&sm.tmem_slot, 512);
  __syncthreads();
  {
    const float* lg = lse + (size_t(b) * H + h) * Lq;
    const float* dg = delta + (size_t(b) * H + h) * Lq;
    for (int r = threadIdx.x; r < nq * TC_BQ; r += TCB_THREADS) {
      sm.lse2[r] = r < Lq ? lg[r] * kLog2e_tc : INFINITY;
      sm.delta[r] = r < Lq ? dg[r] : 0.f;
    }
    if (threadIdx.x < TC_BK) {
      const int key = k0 + threadIdx.x;
      const bool dead = key >= kend || (mask && mask[size_t(b) * Lk + key] != 0);
      sm.kbias[threadIdx.x] = dead ? -INFINITY : 0.f;
      if (dead) sm.any_masked = 1;
    }
  }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem = sm.tmem_slot;
  const uint32_t tm_s = tmem, tm_dp = tmem + 128, tm_dv = tmem + 256, tm_dk = tmem + 320, tm_dq = tmem + 384;

  if (warp == TCB_COMPUTE_WARPS + TCB_DRAIN_WARPS) {
    // ------------------------------------------------------------ TMA producer
    if (lane == 0) {
      mbar_arrive_expect_tx(&sm.kv_full, 2 * TC_TILE);
      tma_load_3d(sm.k, &tmKV, &sm.kv_full, h * TC_HD, k0, b);
      tma_load_3d(sm.v, &tmKV, &sm.kv_full, D + h * TC_HD, k0, b);
      for (int i = 0; i < nq; ++i) {
        const int buf = i & 1;
        mbar_wait(&sm.q_empty[buf], ((i >> 1) & 1) ^ 1);
        mbar_arrive_expect_tx(&sm.q_full[buf], 2 * TC_TILE);
        tma_load_3d(sm.q[buf], &tmQ, &sm.q_full[buf], h * TC_HD, i * TC_BQ, b);
        tma_load_3d(sm.dO[buf], &tmDO, &sm.q_full[buf], h * TC_HD, i * TC_BQ, b);
      }
    }
  } else if (warp == TCB_COMPUTE_WARPS + TCB_DRAIN_WARPS + 1) {
    // ------------------------------------------------------------ MMA issuer
    constexpr uint32_t id_s = umma_idesc_bf16(128, 128, false, false);   // S, dP
    constexpr uint32_t id_t = umma_idesc_bf16(128, 64, true, true);      // dV, dK (A = P^T / dS^T)
    constexpr uint32_t id_q = umma_idesc_bf16(128, 64, false, true);     // dQ
    const uint32_t kb_ = smem_u32(sm.k), vb_ = smem_u32(sm.v), pb_ = smem_u32(sm.p), sb_ = smem_u32(sm.ds);
    auto issue_sdp = [&](int i) {
      mbar_wait(&sm.q_full[i & 1], (i >> 1) & 1);
      tc_fence_after();
      if (lane == 0) {
        const uint32_t qb = smem_u32(sm.q[i & 1]), ob = smem_u32(sm.dO[i & 1]);
#pragma unroll
        for (int k = 0; k < 4; ++k)
          umma_bf16(tm_s, umma_desc_sw128(qb + k * 32, 16, 1024), umma_desc_sw128(kb_ + k * 32, 16, 1024), id_s, k > 0);
#pragma unroll
        for (int k = 0; k < 4; ++k)
          umma_bf16(tm_dp, umma_desc_sw128(ob + k * 32, 16, 1024), umma_desc_sw128(vb_ + k * 32, 16, 1024), id_s, k > 0);
        umma_commit(&sm.sdp_full);
      }
      __syncwarp();
    };
    mbar_wait(&sm.kv_full, 0);
    issue_sdp(0);
    for (int i = 0; i < nq; ++i) {
      if (i + 1 < nq) {
        mbar_wait(&sm.sdp_free, i & 1);
        issue_sdp(i + 1);
      }
      mbar_wait(&sm.pds_ready, i & 1);
      if (i > 0) mbar_wait(&sm.dq_free, (i - 1) & 1);
      tc_fence_after();
      if (lane == 0) {
        const uint32_t qb = smem_u32(sm.q[i & 1]), ob = smem_u32(sm.dO[i & 1]);
#pragma unroll
        for (int k = 0; k < 8; ++k) {   // contraction over the 128 queries, 16 per step (rows of P / dS / dO / Q)
          umma_bf16(tm_dv, umma_desc_sw128(pb_ + k * 2048, TC_TILE, 1024), umma_desc_sw128(ob + k * 2048, 8192, 1024),
                    id_t, (i > 0 || k > 0) ? 1u : 0u);
        }
#pragma unroll
        for (int k = 0; k < 8; ++k) {
          umma_bf16(tm_dk, umma_desc_sw128(sb_ + k * 2048, TC_TILE, 1024), umma_desc_sw128(qb + k * 2048, 8192, 1024),
                    id_t, (i > 0 || k > 0) ? 1u : 0u);
        }
#pragma unroll
        for (int k = 0; k < 8; ++k) {   // contraction over the 128 keys
          umma_bf16(tm_dq, umma_desc_sw128(sb_ + (k >> 2) * TC_TILE + (k & 3) * 32, 16, 1024),
                    umma_desc_sw128(kb_ + k * 2048, 8192, 1024), id_q, k > 0 ? 1u : 0u);
        }
        umma_commit(&sm.pds_free);
        umma_commit(&sm.q_empty[i & 1]);
        umma_commit(&sm.dq_full);
        if (i == nq - 1) umma_commit(&sm.dkv_done);
      }
      __syncwarp();
    }
  } else if (warp < TCB_COMPUTE_WARPS) {
    // ------------------------------------------------------------ compute: thread = (query row, 32 key columns)
    const int quarter = warp & 3, cq = warp >> 2;
    const int r = quarter * 32 + lane;
    const uint32_t lane_addr = static_cast<uint32_t>(quarter * 32) << 16;
    const bool any_masked = sm.any_masked != 0;
    const uint32_t p_row = smem_u32(sm.p) + (cq >> 1) * TC_TILE + r * 128;
    const uint32_t s_row = smem_u32(sm.ds) + (cq >> 1) * TC_TILE + r * 128;
    for (int i = 0; i < nq; ++i) {
      const float ls = sm.lse2[i * TC_BQ + r], de = sm.delta[i * TC_BQ + r];
      mbar_wait(&sm.sdp_full, i & 1);
      tc_fence_after();
      uint32_t sr[32], dr[32];
      tmem_ld_32x32(tm_s + lane_addr + cq * 32, sr);
      tmem_ld_32x32(tm_dp + lane_addr + cq * 32, dr);
      tmem_ld_wait();
      tc_fence_before();
      __syncwarp();
      if (lane == 0) mbar_arrive(&sm.sdp_free);
      float pv[32], dsv[32];
      if (!any_masked) {
#pragma unroll
        for (int c = 0; c < 32; ++c) pv[c] = ex2_approx(fmaf(__uint_as_float(sr[c]), kScaleLog2_tc, -ls));
      } else {
#pragma unroll
        for (int c = 0; c < 32; ++c)
          pv[c] = ex2_approx(fmaf(__uint_as_float(sr[c]) + sm.kbias[cq * 32 + c], kScaleLog2_tc, -ls));
      }
#pragma unroll
      for (int c = 0; c < 32; ++c) dsv[c] = pv[c] * (__uint_as_float(dr[c]) - de);
      if (i > 0) mbar_wait(&sm.pds_free, (i - 1) & 1);     // previous dV/dK/dQ MMAs no longer read P / dS
#pragma unroll
      for (int g = 0; g < 4; ++g) {
        const int chunk = (cq & 1) * 4 + g;
        const uint32_t sw = (chunk ^ (r & 7)) << 4;
        asm volatile("st.shared.v4.b32 [%0], {%1, %2, %3, %4};" ::"r"(p_row + sw),
                     "r"(pack_bf16x2(pv[8 * g + 0], pv[8 * g + 1])), "r"(pack_bf16x2(pv[8 * g + 2], pv[8 * g + 3])),
                     "r"(pack_bf16x2(pv[8 * g + 4], pv[8 * g + 5])), "r"(pack_bf16x2(pv[8 * g + 6], pv[8 * g + 7]))
                     : "memory");
        asm volatile("st.shared.v4.b32 [%0], {%1, %2, %3, %4};" ::"r"(s_row + sw),
                     "r"(pack_bf16x2(dsv[8 * g + 0], dsv[8 * g + 1])), "r"(pack_bf16x2(dsv[8 * g + 2], dsv[8 * g + 3])),
                     "r"(pack_bf16x2(dsv[8 * g + 4], dsv[8 * g + 5])), "r"(pack_bf16x2(dsv[8 * g + 6], dsv[8 * g + 7]))
                     : "memory");
      }
      fence_proxy_async_smem();
      __syncwarp();
      if (lane == 0) mbar_arrive(&sm.pds_ready);
    }
    // ---- epilogue: dV (warps 0-3) and dK (warps 4-7), thread = key row
    if (warp < 8) {
      mbar_wait(&sm.dkv_done, 0);
      tc_fence_after();
      const bool is_dk = warp >= 4;
      const int key = k0 + r;
      __nv_bfloat16* dst = (is_dk ? dk : dv) + (size_t(b) * Lk + key) * lddkv + h * TC_HD;
      const float scale = is_dk ? 0.125f : 1.0f;
#pragma unroll
      for (int c = 0; c < TC_HD; c += 32) {
        uint32_t a[32];
        tmem_ld_32x32((is_dk ? tm_dk : tm_dv) + lane_addr + c, a);
        tmem_ld_wait();
        if (key < Lk) {
#pragma unroll
          for (int g = 0; g < 4; ++g)
            reinterpret_cast<uint4*>(dst + c)[g] =
                make_uint4(pack_bf16x2(__uint_as_float(a[8 * g]) * scale, __uint_as_float(a[8 * g + 1]) * scale),
                           pack_bf16x2(__uint_as_float(a[8 * g + 2]) * scale, __uint_as_float(a[8 * g + 3]) * scale),
                           pack_bf16x2(__uint_as_float(a[8 * g + 4]) * scale, __uint_as_float(a[8 * g + 5]) * scale),
                           pack_bf16x2(__uint_as_float(a[8 * g + 6]) * scale, __uint_as_float(a[8 * g + 7]) * scale));
        }
      }
    }
  } else {
    // ------------------------------------------------------------ dQ drain warps (TMEM -> smem -> bulk reduce-add)
    const int quarter = warp & 3;
    const int r = quarter * 32 + lane;
    const uint32_t lane_addr = static_cast<uint32_t>(quarter * 32) << 16;
    const bool issuer = (warp == TCB_COMPUTE_WARPS) && lane == 0;
    float* gbase = dq_acc + (size_t(b) * H + h) * LqPad * TC_HD;
    for (int i = 0; i < nq; ++i) {
      mbar_wait(&sm.dq_full, i & 1);
      tc_fence_after();
      uint32_t a0[32], a1[32];
      tmem_ld_32x32(tm_dq + lane_addr, a0);
      tmem_ld_32x32(tm_dq + lane_addr + 32, a1);
      tmem_ld_wait();
      tc_fence_before();
      __syncwarp();
      if (lane == 0) mbar_arrive(&sm.dq_free);
      if (i > 0) {                        // the previous bulk reduce must have finished READING the staging tile
        if (issuer) bulk_wait_group_read0();
        asm volatile("bar.sync 2, 128;" ::: "memory");
      }
      const uint32_t row_addr = smem_u32(sm.dq_stage) + r * 256;
#pragma unroll
      for (int g = 0; g < 8; ++g) {
        asm volatile("st.shared.v4.b32 [%0], {%1, %2, %3, %4};" ::"r"(row_addr + g * 16), "r"(a0[4 * g]),
                     "r"(a0[4 * g + 1]), "r"(a0[4 * g + 2]), "r"(a0[4 * g + 3]) : "memory");
        asm volatile("st.shared.v4.b32 [%0], {%1, %2, %3, %4};" ::"r"(row_addr + 128 + g * 16), "r"(a1[4 * g]),
                     "r"(a1[4 * g + 1]), "r"(a1[4 * g + 2]), "r"(a1[4 * g + 3]) : "memory");
      }
      fence_proxy_async_smem();
      asm volatile("bar.sync 2, 128;" ::: "memory");
      if (issuer && !(dbg & 1)) {
        bulk_reduce_add_f32(gbase + size_t(i) * TC_BQ * TC_HD, sm.dq_stage, TC_BQ * TC_HD * 4);
        bulk_commit_group();
      }
    }
    if (issuer) bulk_wait_group0();       // global writes complete before the CTA retires its smem
  }
  tc_fence_before();
  __syncthreads();
  if (warp == TCB_COMPUTE_WARPS + TCB_DRAIN_WARPS + 1) {
    tc_fence_after();
    tmem_dealloc(tmem, 512);
  }
}

// dq(bf16)[b, q, h*64 + d] = 0.125 * acc[b, h, q, d]
__global__ void
attn_dq_finalize_kernel(const float* __restrict__ acc, __nv_bfloat16* __restrict__ dq, int lddq, int H, int Lq,
                        int LqPad, size_t total4) {
  for (size_t i = blockIdx.x * size_t(blockDim.x) + threadIdx.x; i < total4; i += size_t(gridDim.x) * blockDim.x) {
    const int d4 = int(i & 15);
    size_t t = i >> 4;
    const int qi = int(t % Lq);
    t /= Lq;
    const int h = int(t % H);
    const size_t b = t / H;
    const float4 v = *reinterpret_cast<const float4*>(acc + ((b * H + h) * LqPad + qi) * TC_HD + d4 * 4);
    uint2 o;
    o.x = pack_bf16x2(v.x * 0.125f, v.y * 0.125f);
    o.y = pack_bf16x2(v.z * 0.125f, v.w * 0.125f);
    *reinterpret_cast<uint2*>(dq + (b * Lq + qi) * lddq + h * TC_HD + d4 * 4) = o;
  }
}

size_t attn_bwd_tc_workspace(int B, int H, int Lq) {
  const size_t LqPad = size_t((Lq + TC_BQ - 1) / TC_BQ) * TC_BQ;
  return size_t(B) * H * LqPad * TC_HD * sizeof(float);
}

int attn_bwd_tc(const void* q, int ldq, const void* kv, int ldkv, const uint8_t* mask, const int* kv_end,
                const void* dctx, int ldc, const float* lse, const float* delta, float* dq_acc, void* dq, int lddq,
                void* dk, void* dv, int lddkv, int B, int H, int Lq, int Lk, cudaStream_t s) {
  const int D = H * TC_HD;
  if (Lq > TCB_MAXQ) return GXA_ERR_UNSUPPORTED;
  const int LqPad = (Lq + TC_BQ - 1) / TC_BQ * TC_BQ;
  CUtensorMap tmQ, tmKV, tmDO;
  if (int e = make_tmap_bf16_3d(&tmQ, q, D, Lq, B, ldq, size_t(Lq) * ldq, 64, TC_BQ)) return e;
  if (int e = make_tmap_bf16_3d(&tmKV, kv, 2 * D, Lk, B, ldkv, size_t(Lk) * ldkv, 64, TC_BK)) return e;
  if (int e = make_tmap_bf16_3d(&tmDO, dctx, D, Lq, B, ldc, size_t(Lq) * ldc, 64, TC_BQ)) return e;
  static bool attr = false;
  if (!attr) {
    if (cudaFuncSetAttribute(attn_bwd_tc_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, BWD_TC_SMEM) != cudaSuccess)
      return GXA_ERR_CUDA;
    attr = true;
  }
  if (cudaMemsetAsync(dq_acc, 0, attn_bwd_tc_workspace(B, H, Lq), s) != cudaSuccess) return GXA_ERR_CUDA;
  dim3 grid((Lk + TC_BK - 1) / TC_BK, H, B);
  attn_bwd_tc_kernel<<<grid, TCB_THREADS, BWD_TC_SMEM, s>>>(tmQ, tmKV, tmDO, mask, kv_end, lse, delta, dq_acc,
                                                           reinterpret_cast<__nv_bfloat16*>(dk),
                                                           reinterpret_cast<__nv_bfloat16*>(dv), lddkv, H, Lq, Lk, D, LqPad,
                                                           getenv("GXA_DBG") ? atoi(getenv("GXA_DBG")) : 0);
  count_launch();
  if (cudaGetLastError() != cudaSuccess) return GXA_ERR_CUDA;
  const size_t total4 = size_t(B) * H * Lq * 16;
  size_t blocks = (total4 + 255) / 256;
  if (blocks > size_t(num_sms()) * 8) blocks = size_t(num_sms()) * 8;
  attn_dq_finalize_kernel<<<(unsigned)blocks, 256, 0, s>>>(dq_acc, reinterpret_cast<__nv_bfloat16*>(dq), lddq, H, Lq,
                                                           LqPad, total4);
  count_launch();
  return cudaGetLastError() == cudaSuccess ? 0 : GXA_ERR_CUDA;
}

}  // namespace gxa
