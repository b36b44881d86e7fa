// tcgen05 / TMEM attention backward, head_dim 64 -- PERSISTENT version.
//
// One CTA per SM walks a list of work items (batch b, head h, 128-key block kb); the pipelines run ACROSS items,
// so the prologue of item n+1 (K/V load, first S/dP products) and the epilogue of item n (dK/dV drain) overlap
// the steady state.  (The first version launched one CTA per item: 4 query iterations per CTA with every TMA,
// TMEM-allocation and drain latency exposed -- 15 % tensor-pipe activity, profiles/r01_c_*.)
//
// Per item:  for each 128-query block i
//   S = Q_i K^T, dP = dO_i V^T            (TMEM cols 0..127, 128..255; refilled as soon as the compute warps read them)
//   P = exp2(S*c - lse), dS = P*(dP - delta)   16 warps, thread = (query row, 32 key columns) in two 16-column halves
//                                           (packed f32x2 math, packed bf16 results held in registers: no spills --
//                                           local memory shares the L1/smem pipe with the MMA operand fetches, which
//                                           bound this kernel: tools/mma_rate.cu)
//   dV += P^T dO_i, dK += dS^T Q_i         (TMEM cols 256..319, 320..383; accumulate over the query blocks; operands
//                                           are read MN-major through descriptors -> the transposes are free)
//   dQ_i = dS K                            (TMEM cols 384..447 / 448..511, double-buffered) -> 4 drain warps -> staged
//                                           in the freed Q/dO ring slot -> TMA bulk reduce-add into the fp32
//                                           [B,H,Lq/128,16,128,4] accumulator (attn_dq_finalize2_kernel converts)
// warps 0-15 compute, 16-19 drain (dQ every iteration, dK/dV + next key mask at item end), 20 TMA producer
// (K/V double-buffered per item, Q/dO 3-stage ring), 21 MMA issuer.
#include <cstdlib>

#include "gxa_host.h"
#include "gxa_ptx.cuh"

namespace gxa {

namespace {

constexpr int BQ = 128, BK = 128, HD = 64;
constexpr int TILE = 128 * 64 * 2;                      // one 128-row x 64-col bf16 tile (16 KB)
constexpr float kLog2e = 1.4426950408889634f;
constexpr float kScaleLog2 = 0.125f * kLog2e;
constexpr int COMPUTE_WARPS = 16, DRAIN_WARPS = 4;
constexpr int THREADS = (COMPUTE_WARPS + DRAIN_WARPS + 2) * 32;
constexpr int QSTAGES = 3;
constexpr int MAX_ITEMS = 1024;                      // work items per CTA (the host sizes the grid accordingly)

struct Smem {
  uint8_t k[2][TILE];
  uint8_t v[2][TILE];
  uint8_t qdo[QSTAGES][2 * TILE];  // ring stage = Q tile | dO tile; once its MMAs are done the same 32 KB stage the
                                   // fp32 dQ tile for the TMA bulk reduce-add before going back to the producer
  uint8_t p[2 * TILE];             // [128 q][128 keys] bf16 as two 64-key swizzle atoms
  uint8_t ds[2 * TILE];
  float kbias[2][BK];              // 0 for a visible key, -inf for a padded / masked one
  int kb_any[2][DRAIN_WARPS];
  uint64_t kv_full[2], kv_empty[2], q_full[QSTAGES], q_empty[QSTAGES];
  uint64_t sdp_full, sdp_free, pds_ready, p_free, pds_free, dq_full[2], dq_free[2], dkv_full, dkv_free, kb_full[2];
  uint32_t tmem_slot;
  uint8_t live[MAX_ITEMS];         // this CTA's items: 0 = the whole key block is padding
};
constexpr int SMEM_BYTES = sizeof(Smem);   // the kernel uses the (1024-aligned) dynamic smem base directly

__device__ __forceinline__ uint32_t pk2(float lo, float hi) {
  __nv_bfloat162 t = __floats2bfloat162_rn(lo, hi);
  return *reinterpret_cast<uint32_t*>(&t);
}
__device__ __forceinline__ float ex2a(float x) {
  float y;
  asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x));
  return y;
}
__device__ __forceinline__ void red_add_v4(float* dst, float a, float b, float c, float d) {
  asm volatile("red.relaxed.gpu.global.add.v4.f32 [%0], {%1, %2, %3, %4};" ::"l"(dst), "f"(a), "f"(b), "f"(c), "f"(d)
               : "memory");
}

struct Item {
  int b, h, k0;
};

}  // namespace

__global__ void __launch_bounds__(THREADS, 1)
attn_bwd_tc2_kernel(const __grid_constant__ CUtensorMap tmQ, const __grid_constant__ CUtensorMap tmKV,
                    const __grid_constant__ CUtensorMap tmDO, const uint8_t* __restrict__ mask,
                    const int* __restrict__ kv_end, const float* __restrict__ lse, const float* __restrict__ delta,
                    float* __restrict__ dq_acc, __nv_bfloat16* __restrict__ dk, __nv_bfloat16* __restrict__ dv, int lddkv,
                    int B, int H, int Lq, int Lk, int D, int LqPad, const DropCfg drop, float* __restrict__ dv_colsum,
                    long long* dbg) {
#ifdef GXA_BWD_STAMPS   // clock stamps of the pipeline events of CTA 0 (tools/dbg_attn_bwd.py); off in the product build
#define STAMP(role, k) do { if (dbg && blockIdx.x == 0 && g < 16 && lane == 0) dbg[(role) * 128 + g * 8 + (k)] = clock64(); } while (0)
#else
#define STAMP(role, k) do { } while (0)
#endif
  extern __shared__ __align__(1024) uint8_t smem_raw[];
  Smem& sm = *reinterpret_cast<Smem*>(smem_raw);
  if ((smem_u32(smem_raw) & 1023u) != 0) __trap();
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int nq = (Lq + BQ - 1) / BQ;
  const int nkb = (Lk + BK - 1) / BK;
  const int total = B * H * nkb;

  auto decode = [&](int w, Item& it) -> bool {      // false: the whole key block is padding (no pipeline work)
    const int kb = w % nkb, bh = w / nkb;
    it.h = bh % H;
    it.b = bh / H;
    it.k0 = kb * BK;
    return sm.live[(w - blockIdx.x) / gridDim.x] != 0;      // no global load on the item-boundary critical path
  };

  if (threadIdx.x == 0) {
    for (int i = 0; i < 2; ++i) {
      mbar_init(&sm.kv_full[i], 1); mbar_init(&sm.kv_empty[i], 1);
      mbar_init(&sm.dq_full[i], 1); mbar_init(&sm.dq_free[i], DRAIN_WARPS);
      mbar_init(&sm.kb_full[i], DRAIN_WARPS);
    }
    for (int i = 0; i < QSTAGES; ++i) { mbar_init(&sm.q_full[i], 1); mbar_init(&sm.q_empty[i], 1); }
    mbar_init(&sm.sdp_full, 1); mbar_init(&sm.sdp_free, COMPUTE_WARPS);
    mbar_init(&sm.pds_ready, COMPUTE_WARPS); mbar_init(&sm.pds_free, 1); mbar_init(&sm.p_free, 1);
    mbar_init(&sm.dkv_full, 1); mbar_init(&sm.dkv_free, DRAIN_WARPS);
    mbar_fence_init();
  }
  if (warp == COMPUTE_WARPS + DRAIN_WARPS && lane == 0) {
    tma_prefetch_desc(&tmQ); tma_prefetch_desc(&tmKV); tma_prefetch_desc(&tmDO);
  }
  if (warp == COMPUTE_WARPS + DRAIN_WARPS + 1) tmem_alloc(&sm.tmem_slot, 512);
  pdl_wait();                           // everything below reads / writes global memory
  for (int t = threadIdx.x; t < MAX_ITEMS; t += THREADS) {
    const int w = blockIdx.x + t * gridDim.x;
    uint8_t lv = 0;
    if (w < total) {
      const int kb = w % nkb, b = (w / nkb) / H;
      const int kend = kv_end ? min(__ldg(kv_end + b), Lk) : Lk;
      lv = kb * BK < kend;
    }
    sm.live[t] = lv;
  }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem = sm.tmem_slot;
  const uint32_t tm_s = tmem, tm_dp = tmem + 128, tm_dv = tmem + 256, tm_dk = tmem + 320, tm_dq = tmem + 384;

  if (warp == COMPUTE_WARPS + DRAIN_WARPS) {
    // ------------------------------------------------------------ TMA producer
    if (lane == 0) {
      int n = 0, qs = 0;
      uint32_t qphase = 0;
      auto load_kv = [&](const Item& it, int nn) {
        const int st = nn & 1;
        mbar_wait(&sm.kv_empty[st], ((nn >> 1) & 1) ^ 1);
        mbar_arrive_expect_tx(&sm.kv_full[st], 2 * TILE);
        tma_load_3d(sm.k[st], &tmKV, &sm.kv_full[st], it.h * HD, it.k0, it.b);
        tma_load_3d(sm.v[st], &tmKV, &sm.kv_full[st], D + it.h * HD, it.k0, it.b);
      };
      // find the first live item
      int w = blockIdx.x;
      Item cur, nxt;
      while (w < total && !decode(w, cur)) w += gridDim.x;
      if (w < total) load_kv(cur, 0);
      while (w < total) {
        int w2 = w + gridDim.x;
        while (w2 < total && !decode(w2, nxt)) w2 += gridDim.x;
        for (int i = 0; i < nq; ++i) {
          mbar_wait(&sm.q_empty[qs], qphase ^ 1);
          mbar_arrive_expect_tx(&sm.q_full[qs], 2 * TILE);
          tma_load_3d(sm.qdo[qs], &tmQ, &sm.q_full[qs], cur.h * HD, i * BQ, cur.b);
          tma_load_3d(sm.qdo[qs] + TILE, &tmDO, &sm.q_full[qs], cur.h * HD, i * BQ, cur.b);
          if (++qs == QSTAGES) { qs = 0; qphase ^= 1; }
          // next item's K/V early (hidden behind this item); placed after the second Q/dO load because its buffer is
          // released by the same MMA (last of the previous item) that releases the third ring stage
          if (i == (nq > 1 ? 1 : 0) && w2 < total) load_kv(nxt, n + 1);
        }
        ++n;
        w = w2;
        cur = nxt;
      }
      pdl_launch_dependents();                   // all loads of this CTA issued: the next kernel may be staged
    }
  } else if (warp == COMPUTE_WARPS + DRAIN_WARPS + 1) {
    // ------------------------------------------------------------ MMA issuer
    constexpr uint32_t id_s = umma_idesc_bf16(128, 128, false, false);   // S, dP
    constexpr uint32_t id_t = umma_idesc_bf16(128, 64, true, true);      // dV, dK (A = P^T / dS^T)
    constexpr uint32_t id_q = umma_idesc_bf16(128, 64, false, true);     // dQ
    int g = 0;              // global iteration counter (all items)
    int qs = 0;             // Q/dO ring stage of iteration g
    uint32_t qphase = 0;
    int qs_n = 0;           // ring stage / phase of the NEXT S/dP issue
    uint32_t qphase_n = 0;
    constexpr uint32_t HI = umma_desc_hi(1024);
    const uint32_t el = elect_one_pred();
    const uint32_t p_t = umma_desc_lo(smem_u32(sm.p), TILE), s_t = umma_desc_lo(smem_u32(sm.ds), TILE);   // MN-major A
    const uint32_t s_k = umma_desc_lo(smem_u32(sm.ds), 16);                                               // K-major A
    auto issue_sdp = [&](int kvst) {
      mbar_wait(&sm.q_full[qs_n], qphase_n);
      tc_fence_after();
      const uint32_t qd = umma_desc_lo(smem_u32(sm.qdo[qs_n]), 16), od = umma_desc_lo(smem_u32(sm.qdo[qs_n]) + TILE, 16);
      const uint32_t kd = umma_desc_lo(smem_u32(sm.k[kvst]), 16), vd = umma_desc_lo(smem_u32(sm.v[kvst]), 16);
#pragma unroll
      for (int k = 0; k < 4; ++k) umma_bf16_e(tm_s, qd + k * 2, HI, kd + k * 2, HI, id_s, k > 0, el);
#pragma unroll
      for (int k = 0; k < 4; ++k) umma_bf16_e(tm_dp, od + k * 2, HI, vd + k * 2, HI, id_s, k > 0, el);
      umma_commit_e(&sm.sdp_full, el);
      if (++qs_n == QSTAGES) { qs_n = 0; qphase_n ^= 1; }
    };
    int w = blockIdx.x, n = 0;
    Item cur, nxt;
    while (w < total && !decode(w, cur)) w += gridDim.x;
    if (w < total) {
      mbar_wait(&sm.kv_full[0], 0);
      issue_sdp(0);
    }
    while (w < total) {
      int w2 = w + gridDim.x;
      while (w2 < total && !decode(w2, nxt)) w2 += gridDim.x;
      const int kvst = n & 1;
      for (int i = 0; i < nq; ++i, ++g) {
        // next S/dP as soon as the compute warps hold this one in registers
        if (i + 1 < nq) {
          mbar_wait(&sm.sdp_free, g & 1);
          issue_sdp(kvst);
        } else if (w2 < total) {
          mbar_wait(&sm.sdp_free, g & 1);
          mbar_wait(&sm.kv_full[(n + 1) & 1], ((n + 1) >> 1) & 1);
          issue_sdp((n + 1) & 1);
        }
        STAMP(0, 0);
        mbar_wait(&sm.pds_ready, g & 1);
        STAMP(0, 1);
        if (g >= 2) mbar_wait(&sm.dq_free[g & 1], ((g >> 1) - 1) & 1);
        STAMP(0, 2);
        if (i == 0 && n > 0) mbar_wait(&sm.dkv_free, (n - 1) & 1);
        STAMP(0, 3);
        tc_fence_after();
        {
          const uint32_t qm = umma_desc_lo(smem_u32(sm.qdo[qs]), 8192), om = umma_desc_lo(smem_u32(sm.qdo[qs]) + TILE, 8192);
          const uint32_t km = umma_desc_lo(smem_u32(sm.k[kvst]), 8192);                                  // MN-major B
          const uint32_t acc0 = i > 0 ? 1u : 0u;
#pragma unroll
          for (int k = 0; k < 8; ++k)     // contraction over the 128 queries, 16 per step (rows of P / dS / dO / Q)
            umma_bf16_e(tm_dv, p_t + k * 128, HI, om + k * 128, HI, id_t, k > 0 ? 1u : acc0, el);
          umma_commit_e(&sm.p_free, el);  // P is read by dV only: the compute warps may store the next P while dK / dQ run
#pragma unroll
          for (int k = 0; k < 8; ++k)
            umma_bf16_e(tm_dk, s_t + k * 128, HI, qm + k * 128, HI, id_t, k > 0 ? 1u : acc0, el);
#pragma unroll
          for (int k = 0; k < 8; ++k)     // contraction over the 128 keys
            umma_bf16_e(tm_dq + (g & 1) * 64, s_k + (k >> 2) * (TILE >> 4) + (k & 3) * 2, HI, km + k * 128, HI, id_q,
                        k > 0 ? 1u : 0u, el);
          umma_commit_e(&sm.pds_free, el);
          umma_commit_e(&sm.dq_full[g & 1], el);
          if (i == nq - 1) {
            umma_commit_e(&sm.dkv_full, el);
            umma_commit_e(&sm.kv_empty[kvst], el);
          }
        }
        STAMP(0, 4);
        if (++qs == QSTAGES) { qs = 0; qphase ^= 1; }
      }
      ++n;
      w = w2;
      cur = nxt;
    }
    (void)qphase;
  } else if (warp < COMPUTE_WARPS) {
    // ------------------------------------------------------------ compute: thread = (query row, 32 key columns)
    const int quarter = warp & 3, cq = warp >> 2;
    const int r = quarter * 32 + lane;
    const uint32_t lane_addr = static_cast<uint32_t>(quarter * 32) << 16;
    const uint32_t p_row = smem_u32(sm.p) + (cq >> 1) * TILE + r * 128;
    const uint32_t s_row = smem_u32(sm.ds) + (cq >> 1) * TILE + r * 128;
    // the compute warps keep only (b*H + h, first key) of an item: every register that stays live across their loop is
    // one they cannot use in it (80-register cap; ncu r02: 13 % of their stall samples were waits on re-loads of spilled
    // item fields at the top of each iteration)
    auto decode_c = [&](int w, int& bh, int& k0) -> bool {
      bh = w / nkb;
      k0 = (w - bh * nkb) * BK;
      return sm.live[(w - blockIdx.x) / gridDim.x] != 0;
    };
    auto row_stats = [&](int bh, int i, float& ls, float& de) {
      const int row = i * BQ + r;
      ls = INFINITY;            // rows past Lq: P = exp2(-inf) = 0 (ls is pre-multiplied by log2 e below)
      de = 0.f;
      if (row < Lq) {
        const size_t o = size_t(bh) * Lq + row;
        // asm volatile: keeps the loads HERE (one iteration ahead of their use); plain __ldg was sunk to the use
        asm volatile("ld.global.nc.f32 %0, [%1];" : "=f"(ls) : "l"(lse + o));
        asm volatile("ld.global.nc.f32 %0, [%1];" : "=f"(de) : "l"(delta + o));
      }
    };
    int w = blockIdx.x, n = 0, g = 0;
    int cur_bh = 0, cur_k0 = 0, nxt_bh = 0, nxt_k0 = 0;
    while (w < total && !decode_c(w, cur_bh, cur_k0)) w += gridDim.x;
    float ls = 0.f, de = 0.f, ls_n, de_n;
    if (w < total) row_stats(cur_bh, 0, ls, de);
    while (w < total) {
      int w2 = w + gridDim.x;
      while (w2 < total && !decode_c(w2, nxt_bh, nxt_k0)) w2 += gridDim.x;
      mbar_wait(&sm.kb_full[n & 1], (n >> 1) & 1);
      const float* kbias = sm.kbias[n & 1] + cq * 32;
      const bool any_masked = (sm.kb_any[n & 1][0] | sm.kb_any[n & 1][1] | sm.kb_any[n & 1][2] | sm.kb_any[n & 1][3]) != 0;
      for (int i = 0; i < nq; ++i, ++g) {
        if (warp == 0) STAMP(1, 5);
        ls_n = 0.f; de_n = 0.f;
        if (i + 1 < nq) row_stats(cur_bh, i + 1, ls_n, de_n);       // one iteration ahead: latency hidden by the math
        else if (w2 < total) row_stats(nxt_bh, 0, ls_n, de_n);
        if (warp == 0) STAMP(1, 0);
        mbar_wait(&sm.sdp_full, g & 1);
        if (warp == 0) STAMP(1, 1);
        tc_fence_after();
        // two 16-column halves: the packed bf16 results of the first half (16 registers) wait in registers while the second
        // half is loaded and computed, so the working set stays below the 80-register cap of this 704-thread CTA
        // (one 32-column pass needed 64 raw + 32 result registers and spilled ~40 local-memory accesses per iteration,
        // which go through the same L1/shared-memory pipe the MMA operands are fetched from)
        // dropout (forward: ctx = (m * P) V with m in {0, 1/(1-p)}): the factor 1/(1-p) rides in the exponent, so
        //   pe = P / (1-p),  P~ = bf16(pe) & mask (the A operand of dV, exactly what the forward multiplied V with),
        //   dS = P * (m * dP - delta) = P~ * dP - pe * (delta * (1-p))
        const float nls = -ls * kLog2e + (drop.thresh != 0 ? log2f(drop.scale) : 0.f);
        const float nde = drop.thresh != 0 ? -de / drop.scale : -de;
        const float2 c2 = make_float2(kScaleLog2, kScaleLog2), nls2 = make_float2(nls, nls), nde2 = make_float2(nde, nde);
        uint32_t ppk[16], dpk[16];
#pragma unroll
        for (int hf = 0; hf < 2; ++hf) {
          uint32_t sr[16], dr[16];
          tmem_ld_32x16(tm_s + lane_addr + cq * 32 + hf * 16, sr);
          tmem_ld_32x16(tm_dp + lane_addr + cq * 32 + hf * 16, dr);
          tmem_ld_wait();
          if (hf == 1) {                  // S / dP are in registers: the MMA warp may overwrite them
            tc_fence_before();
            __syncwarp();
            if (lane == 0) mbar_arrive(&sm.sdp_free);
            if (warp == 0) STAMP(1, 6);
          }
          uint32_t ctr = 0;               // same function of (row, key) as the forward (gxa_rng.cuh): 4 keys per counter
          if (drop.thresh != 0)
            ctr = (static_cast<uint32_t>(cur_bh) * Lq + (i * BQ + r)) * static_cast<uint32_t>((Lk + 3) >> 2) +
                  ((cur_k0 + cq * 32 + hf * 16) >> 2);
#pragma unroll
          for (int w = 0; w < 4; ++w) {   // one counter = 4 keys = 2 packed pairs
            uint32_t mk[2] = {0xffffffffu, 0xffffffffu};
            if (drop.thresh != 0) drop_masks4(drop, ctr + w, mk[0], mk[1]);
#pragma unroll
            for (int cc = 0; cc < 2; ++cc) {   // packed f32x2 arithmetic (FFMA2 / FMUL2 / FADD2): half the issue slots
              const int c = 2 * w + cc;
              float2 sv = make_float2(__uint_as_float(sr[2 * c]), __uint_as_float(sr[2 * c + 1]));
              if (any_masked) sv = __fadd2_rn(sv, *reinterpret_cast<const float2*>(kbias + hf * 16 + 2 * c));
              const float2 e = __ffma2_rn(sv, c2, nls2);
              const float2 pe = make_float2(ex2a(e.x), ex2a(e.y));
              const float2 dp = make_float2(__uint_as_float(dr[2 * c]), __uint_as_float(dr[2 * c + 1]));
              float2 ds;
              uint32_t pp = pk2(pe.x, pe.y);
              if (drop.thresh == 0) {
                ds = __fmul2_rn(pe, __fadd2_rn(dp, nde2));
              } else {
                pp &= mk[cc];
                const float2 pm = make_float2(__uint_as_float(pp << 16), __uint_as_float(pp & 0xffff0000u));
                ds = __ffma2_rn(pm, dp, __fmul2_rn(pe, nde2));
              }
              ppk[hf * 8 + c] = pp;
              dpk[hf * 8 + c] = pk2(ds.x, ds.y);
            }
          }
        }
        if (warp == 0) STAMP(1, 2);
        if (g > 0) mbar_wait(&sm.p_free, (g - 1) & 1);       // previous dV MMAs no longer read P
#pragma unroll
        for (int gg = 0; gg < 4; ++gg) {
          const uint32_t sw = (((cq & 1) * 4 + gg) ^ (r & 7)) << 4;
          asm volatile("st.shared.v4.b32 [%0], {%1, %2, %3, %4};" ::"r"(p_row + sw), "r"(ppk[4 * gg]), "r"(ppk[4 * gg + 1]),
                       "r"(ppk[4 * gg + 2]), "r"(ppk[4 * gg + 3])
                       : "memory");
        }
        if (g > 0) mbar_wait(&sm.pds_free, (g - 1) & 1);     // previous dK / dQ MMAs no longer read dS
#pragma unroll
        for (int gg = 0; gg < 4; ++gg) {
          const uint32_t sw = (((cq & 1) * 4 + gg) ^ (r & 7)) << 4;
          asm volatile("st.shared.v4.b32 [%0], {%1, %2, %3, %4};" ::"r"(s_row + sw), "r"(dpk[4 * gg]), "r"(dpk[4 * gg + 1]),
                       "r"(dpk[4 * gg + 2]), "r"(dpk[4 * gg + 3])
                       : "memory");
        }
        fence_proxy_async_smem();
        __syncwarp();
        if (lane == 0) mbar_arrive(&sm.pds_ready);
        if (warp == 0) STAMP(1, 3);
        asm volatile("" ::"f"(ls_n), "f"(de_n));
        if (warp == 0) STAMP(1, 4);
        ls = ls_n;
        de = de_n;
      }
      ++n;
      w = w2;
      cur_bh = nxt_bh;
      cur_k0 = nxt_k0;
    }
  } else {
    // ------------------------------------------------------------ drain warps: dQ every iteration; dK/dV and the
    // key-visibility vector of the item after next at item end; zeros for fully padded key blocks
    const int quarter = warp & 3;
    const int r = quarter * 32 + lane;
    const uint32_t lane_addr = static_cast<uint32_t>(quarter * 32) << 16;
    Item cur, nx1, nx2;
    for (int w = blockIdx.x; w < total; w += gridDim.x) {
      if (decode(w, cur)) continue;
      const int key = cur.k0 + r;
      if (key < Lk) {
        const size_t off = (size_t(cur.b) * Lk + key) * lddkv + cur.h * HD;
#pragma unroll
        for (int c = 0; c < 8; ++c) {
          reinterpret_cast<uint4*>(dk + off)[c] = make_uint4(0, 0, 0, 0);
          reinterpret_cast<uint4*>(dv + off)[c] = make_uint4(0, 0, 0, 0);
        }
      }
    }
    auto next_live = [&](int w, Item& it) {
      w += gridDim.x;
      while (w < total && !decode(w, it)) w += gridDim.x;
      return w;
    };
    auto fill_kbias = [&](const Item& it, int nn) {
      const int key = it.k0 + r;
      const int kend = kv_end ? min(__ldg(kv_end + it.b), Lk) : Lk;
      const bool dead = key >= kend || (mask && mask[size_t(it.b) * Lk + key] != 0);
      sm.kbias[nn & 1][r] = dead ? -INFINITY : 0.f;
      const bool any = __any_sync(0xffffffffu, dead);
      if (lane == 0) sm.kb_any[nn & 1][quarter] = any ? 1 : 0;
      __syncwarp();
      if (lane == 0) mbar_arrive(&sm.kb_full[nn & 1]);
    };
    int w = blockIdx.x - gridDim.x;
    w = next_live(w, cur);
    int w1 = total, w2 = total;
    if (w < total) {
      fill_kbias(cur, 0);
      w1 = next_live(w, nx1);
      if (w1 < total) fill_kbias(nx1, 1);
    }
    int n = 0, g = 0, qs = 0;
    const bool issuer = warp == COMPUTE_WARPS && lane == 0;
    while (w < total) {
      float* gq = dq_acc + (size_t(cur.b) * H + cur.h) * LqPad * HD;
      for (int i = 0; i < nq; ++i, ++g) {
        if (quarter == 0) STAMP(2, 0);
        mbar_wait(&sm.dq_full[g & 1], (g >> 1) & 1);      // dQ_g complete => every MMA reading ring stage qs is complete
        if (quarter == 0) STAMP(2, 1);
        tc_fence_after();
        // staging layout = accumulator layout [16 column groups][128 rows][4 floats]: conflict-free 16-byte stores
        const uint32_t st_row = smem_u32(sm.qdo[qs]) + r * 16;
        {   // both halves of the 64 accumulator columns in flight at once: one TMEM latency on the ring slot's way back
          uint32_t a0[32], a1[32];
          tmem_ld_32x32(tm_dq + (g & 1) * 64 + lane_addr, a0);
          tmem_ld_32x32(tm_dq + (g & 1) * 64 + lane_addr + 32, a1);
          tmem_ld_wait();
          tc_fence_before();
          __syncwarp();
          if (lane == 0) mbar_arrive(&sm.dq_free[g & 1]);
#pragma unroll
          for (int j = 0; j < 8; ++j)
            asm volatile("st.shared.v4.b32 [%0], {%1, %2, %3, %4};" ::"r"(st_row + j * (BQ * 16)), "r"(a0[4 * j]),
                         "r"(a0[4 * j + 1]), "r"(a0[4 * j + 2]), "r"(a0[4 * j + 3]) : "memory");
#pragma unroll
          for (int j = 0; j < 8; ++j)
            asm volatile("st.shared.v4.b32 [%0], {%1, %2, %3, %4};" ::"r"(st_row + (8 + j) * (BQ * 16)), "r"(a1[4 * j]),
                         "r"(a1[4 * j + 1]), "r"(a1[4 * j + 2]), "r"(a1[4 * j + 3]) : "memory");
        }
        fence_proxy_async_smem();
        asm volatile("bar.sync 2, 128;" ::: "memory");
        if (issuer) {      // TMA engine adds the 32 KB tile into the fp32 accumulator; no LSU / register traffic
          bulk_reduce_add_f32(gq + size_t(i) * BQ * HD, sm.qdo[qs], BQ * HD * 4);
          bulk_commit_group();
          bulk_wait_group_read0();                          // staging tile read => the stage returns to the TMA producer
          mbar_arrive(&sm.q_empty[qs]);
        }
        if (++qs == QSTAGES) qs = 0;
      }
      // ---- dV then dK of this item (thread = key row).  (Tried: draining them before the item's last dQ tile, and on the
      // 16 compute warps -- both slower: the Q/dO ring slot that the dQ tile is staged in must go back to the producer first.)
      mbar_wait(&sm.dkv_full, n & 1);
      tc_fence_after();
      const int key = cur.k0 + r;
      const size_t off = (size_t(cur.b) * Lk + key) * lddkv + cur.h * HD;
#pragma unroll
      for (int t = 0; t < 2; ++t) {
        __nv_bfloat16* dst = (t ? dk : dv) + off;
        const float scale = t ? 0.125f : 1.0f;
#pragma unroll
        for (int c = 0; c < HD; c += 32) {
          uint32_t a[32];
          tmem_ld_32x32((t ? tm_dk : tm_dv) + lane_addr + c, a);
          tmem_ld_wait();
          if (t == 1 && c == 32) {        // every TMEM read of this item's accumulators is complete
            tc_fence_before();
            __syncwarp();
            if (lane == 0) mbar_arrive(&sm.dkv_free);
          }
          if (key < Lk) {
#pragma unroll
            for (int j = 0; j < 4; ++j)
              reinterpret_cast<uint4*>(dst + c)[j] =
                  make_uint4(pk2(__uint_as_float(a[8 * j]) * scale, __uint_as_float(a[8 * j + 1]) * scale),
                             pk2(__uint_as_float(a[8 * j + 2]) * scale, __uint_as_float(a[8 * j + 3]) * scale),
                             pk2(__uint_as_float(a[8 * j + 4]) * scale, __uint_as_float(a[8 * j + 5]) * scale),
                             pk2(__uint_as_float(a[8 * j + 6]) * scale, __uint_as_float(a[8 * j + 7]) * scale));
          }
          if (t == 0 && dv_colsum != nullptr) {
            // fused d(bv) = column sums of dV: the tile is in registers here (thread = key row, 32 columns), so the
            // separate 98 MB pass over d(k|v) per layer is not needed.  Transposing butterfly over the warp's 32 rows
            // (31 shuffles per 32 columns; lane l ends up with column l), one fp32 atomic per column and warp.  Rows of
            // padded keys hold exact zeros (P = 0 there), so every row may take part.
            float v[32];
#pragma unroll
            for (int i = 0; i < 32; ++i) v[i] = __uint_as_float(a[i]);
#pragma unroll
            for (int off = 16; off >= 1; off >>= 1) {
              const bool up = (lane & off) != 0;
#pragma unroll
              for (int i = 0; i < off; ++i) {
                const float send = up ? v[i] : v[i + off];
                const float keep = up ? v[i + off] : v[i];
                v[i] = keep + __shfl_xor_sync(0xffffffffu, send, off);
              }
            }
            atomicAdd(dv_colsum + cur.h * HD + c + lane, v[0]);
          }
        }
      }
      // ---- key visibility of item n+2 (its buffer was last read by item n, which is complete)
      if (w1 < total) {
        w2 = next_live(w1, nx2);
        if (w2 < total) fill_kbias(nx2, n + 2);
      }
      ++n;
      w = w1; cur = nx1;
      w1 = w2; nx1 = nx2;
      if (w >= total) break;
    }
    if (issuer) bulk_wait_group0();       // global reductions complete before the CTA retires its smem
  }
  tc_fence_before();
  __syncthreads();
  if (warp == COMPUTE_WARPS + DRAIN_WARPS + 1) {
    tc_fence_after();
    tmem_dealloc(tmem, 512);
  }
}

}  // namespace gxa

namespace gxa {

// dq(bf16)[b, q, h*64 + d] = 0.125 * acc[b, h, q-block, d/4, q%128, d%4]   (one block per 128-query tile; the
// transposition between the reduction-friendly accumulator layout and row-major dq goes through shared memory)
__global__ void __launch_bounds__(256)
attn_dq_finalize2_kernel(const float* __restrict__ acc, __nv_bfloat16* __restrict__ dq, int lddq, int H, int Lq, int nq,
                         float* __restrict__ dq_colsum) {
  __shared__ float4 tile[16][BQ + 1];
  pdl_launch_dependents();
  pdl_wait();
  const int blk = blockIdx.x;                       // (b*H + h)*nq + i
  const int i = blk % nq, bh = blk / nq, h = bh % H, b = bh / H;
  const float4* src = reinterpret_cast<const float4*>(acc) + size_t(blk) * 16 * BQ;
#pragma unroll
  for (int m = 0; m < 8; ++m) {
    const int idx = threadIdx.x + 256 * m;          // jj * 128 + row
    tile[idx >> 7][idx & 127] = src[idx];
  }
  __syncthreads();
#pragma unroll
  for (int m = 0; m < 8; ++m) {
    const int idx = threadIdx.x + 256 * m;          // row * 16 + jj
    const int row = idx >> 4, jj = idx & 15;
    const int qrow = i * BQ + row;
    if (qrow < Lq) {
      const float4 v = tile[jj][row];
      uint2 o;
      o.x = pk2(v.x * 0.125f, v.y * 0.125f);
      o.y = pk2(v.z * 0.125f, v.w * 0.125f);
      *reinterpret_cast<uint2*>(dq + (size_t(b) * Lq + qrow) * lddq + h * HD + jj * 4) = o;
    }
  }
  if (dq_colsum != nullptr) {      // fused d(bq): column sums of the bf16 dq rows this tile wrote
    const int c = threadIdx.x & 63, r0 = (threadIdx.x >> 6) * 32;
    float sum = 0.f;
    for (int row = r0; row < r0 + 32 && i * BQ + row < Lq; ++row)
      sum += __bfloat162float(__float2bfloat16_rn(reinterpret_cast<const float*>(&tile[c >> 2][row])[c & 3] * 0.125f));
    atomicAdd(dq_colsum + h * HD + c, sum);
  }
}

size_t attn_bwd_tc_workspace(int B, int H, int Lq);

// Persistent backward (see the header of this file).  Same contract as attn_bwd_tc.
int attn_bwd_tc2(const void* q, int ldq, const void* kv, int ldkv, const uint8_t* mask, const int* kv_end,
                 const void* dctx, int ldc, const float* lse, const float* delta, float* dq_acc, void* dq, int lddq,
                 void* dk, void* dv, int lddkv, int B, int H, int Lq, int Lk, float* dq_colsum, cudaStream_t s, DropCfg drop,
                 float* dv_colsum, bool acc_is_zero) {
  const int D = H * HD;
  const int LqPad = (Lq + BQ - 1) / BQ * BQ;
  CUtensorMap tmQ, tmKV, tmDO;
  if (int e = make_tmap_bf16_3d(&tmQ, q, D, Lq, B, ldq, size_t(Lq) * ldq, 64, BQ)) return e;
  if (int e = make_tmap_bf16_3d(&tmKV, kv, 2 * D, Lk, B, ldkv, size_t(Lk) * ldkv, 64, BK)) return e;
  if (int e = make_tmap_bf16_3d(&tmDO, dctx, D, Lq, B, ldc, size_t(Lq) * ldc, 64, BQ)) return e;
  if (int e = set_max_smem_once(attn_bwd_tc2_kernel, SMEM_BYTES)) return e;
  // the fp32 dQ accumulator must start from zero; a caller that zeroed it off the critical path says so
  if (!acc_is_zero && cudaMemsetAsync(dq_acc, 0, attn_bwd_tc_workspace(B, H, Lq), s) != cudaSuccess) return GXA_ERR_CUDA;
  const int total = B * H * ((Lk + BK - 1) / BK);
  int grid = total < num_sms() ? total : num_sms();
  if ((total + grid - 1) / grid > MAX_ITEMS) grid = (total + MAX_ITEMS - 1) / MAX_ITEMS;
#ifdef GXA_BWD_STAMPS
  long long* dbg = getenv("GXA_DBG_PTR") ? reinterpret_cast<long long*>(strtoull(getenv("GXA_DBG_PTR"), nullptr, 0)) : nullptr;
#else
  long long* dbg = nullptr;      // the stamp hook (tools/dbg_attn_bwd.py) exists in -DGXA_BWD_STAMPS builds only
#endif
  cudaError_t le = launch_pdl(attn_bwd_tc2_kernel, dim3(grid), dim3(THREADS), SMEM_BYTES, s, tmQ, tmKV, tmDO, mask, kv_end, lse,
                              delta, dq_acc, reinterpret_cast<__nv_bfloat16*>(dk), reinterpret_cast<__nv_bfloat16*>(dv), lddkv,
                              B, H, Lq, Lk, D, LqPad, drop, dv_colsum, dbg);
  count_launch();
  if (le != cudaSuccess || cudaGetLastError() != cudaSuccess) return GXA_ERR_CUDA;
  le = launch_pdl(attn_dq_finalize2_kernel, dim3(B * H * (LqPad / BQ)), dim3(256), 0, s, dq_acc,
                  reinterpret_cast<__nv_bfloat16*>(dq), lddq, H, Lq, LqPad / BQ, dq_colsum);
  count_launch();
  return le == cudaSuccess && cudaGetLastError() == cudaSuccess ? 0 : GXA_ERR_CUDA;
}

}  // namespace gxa
