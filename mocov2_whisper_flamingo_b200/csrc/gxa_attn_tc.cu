// tcgen05 / TMEM masked-softmax attention forward, head_dim 64, bf16 operands, fp32 softmax.
//
// One CTA = 128 queries of one (batch, head); 128-key blocks up to kv_end[b]; two CTAs per SM overlap each other's
// softmax and MMA phases.
//   warp 8  : TMA producer (Q once; K and V double-buffered; 3-D tensor maps so that rows past a sample's length are
//             zero-filled, never another sample's data)
//   warp 9  : TMEM allocator + tcgen05.mma issuer:  S_j = Q K_j^T (128x128x64) and O += P_j V_j (128x64x128, P read
//             from TMEM as the A operand, V consumed MN-major straight from its TMA tile)
//   warps 0-7: softmax, two threads per query row (see the kernel's header below)
// The persistent backward lives in gxa_attn_bwd_tc.cu.
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

__device__ __forceinline__ uint32_t pack_bf16x2(float lo, float hi) {
  __nv_bfloat162 t = __floats2bfloat162_rn(lo, hi);
  return *reinterpret_cast<uint32_t*>(&t);
}
__device__ __forceinline__ float ex2_approx(float x) {
  float y;
  asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x));
  return y;
}

// =====================================================================================================
// Where P and O live:
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
    // dropout counter of this query row: (b, h, q) * ceil(Lk / 4) + key / 4   (gxa_rng.cuh)
    const uint32_t drop_row = (static_cast<uint32_t>(b * H + h) * Lq + (q0 + r)) * static_cast<uint32_t>((Lk + 3) >> 2);

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
        uint32_t pk[16];
#pragma unroll
        for (int i = 0; i < 16; ++i) pk[i] = pack_bf16x2(pv[2 * i], pv[2 * i + 1]);
        if (drop.thresh != 0) {      // the row sum above stays undropped (softmax first); dropping = one AND per packed pair
          const uint32_t ctr = drop_row + ((j * TC_BK + half * 64 + c * 32) >> 2);
#pragma unroll
          for (int i = 0; i < 8; ++i) {
            uint32_t m01, m23;
            drop_masks4(drop, ctr + i, m01, m23);
            pk[2 * i] &= m01;
            pk[2 * i + 1] &= m23;
          }
        }
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
  if (int e = set_max_smem_once(attn_fwd_tc2_kernel, FWD_TC2_SMEM)) return e;
  dim3 grid((Lq + TC_BQ - 1) / TC_BQ, H, B);
  launch_pdl(attn_fwd_tc2_kernel, grid, dim3(TC_FWD_THREADS), FWD_TC2_SMEM, s, tmQ, tmKV, mask, kv_end,
               reinterpret_cast<__nv_bfloat16*>(ctx), ldc, lse, H, Lq, Lk, D, drop);
  count_launch();
  return cudaGetLastError() == cudaSuccess ? 0 : GXA_ERR_CUDA;
}


size_t attn_bwd_tc_workspace(int B, int H, int Lq) {      // fp32 dQ accumulator of the persistent backward (gxa_attn_bwd_tc.cu)
  const size_t LqPad = size_t((Lq + TC_BQ - 1) / TC_BQ) * TC_BQ;
  return size_t(B) * H * LqPad * TC_HD * sizeof(float);
}

}  // namespace gxa
