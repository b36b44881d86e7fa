// 2-CTA (cta_group::2) tcgen05 GEMM: a CTA pair (cluster of 2 SMs of one TPC) computes one 256 x 256 tile.
//
// Why: with a 1-CTA 128x256 tile every 128-cycle MMA step reads 12 KB of smem while TMA writes another
// 12 KB -- more than the 128 B/clk shared-memory port (ncu: tensor pipe 60 % active, profiles/r01_a_*).
// In pair mode each CTA stages its own 128 rows of A and only HALF of B (128 of the 256 columns); the
// tensor cores of both SMs consume both halves, so per-SM smem traffic drops by a third and L2->SM traffic
// per FLOP drops from 1/85 to 1/128 B/FLOP.
//
//   warp 16 (both CTAs) : TMA producer for its CTA's A rows / B half; bytes complete on the LEADER's barrier
//   warp 17 (leader only): issues tcgen05.mma.cta_group::2, commits multicast to both CTAs' barriers
//   warps 0-15 (both)   : epilogue of the CTA's own 128 x 256 accumulator (TMEM -> smem transpose -> coalesced HBM)
#include "gxa_gemm.cuh"
#include "gxa_host.h"

namespace gxa {

constexpr int G2_BM = 256;        // per pair
constexpr int G2_BN = 256;
constexpr int G2_A_BYTES = 128 * GEMM_BK * 2;      // this CTA's 128 rows of A
constexpr int G2_B_BYTES = 128 * GEMM_BK * 2;      // this CTA's 128 columns of B
constexpr int G2_STAGE_BYTES = G2_A_BYTES + G2_B_BYTES;
// Epilogue width per fused epilogue: the GELU epilogues are latency-bound with 2 warps per scheduler (ncu: < 50 % issue
// utilisation), so they get 16 epilogue warps (4 per TMEM lane quarter, 64 accumulator columns each) and give one smem
// stage to the extra staging buffers; the others keep 8 warps and the 6-stage ring (long-K wgrad wants the depth).
//
// STG = 0 is the configuration above.  A non-zero STG overrides the ring depth: a shallower ring leaves shared memory on
// the SM for a co-resident CTA of an HBM-bound row-wise kernel -- experiment switch
// GXA_WGRAD_STAGES for the fp32-accumulating (weight-gradient / dxa) GEMMs (default 5, see launch2_layout).
template <int EPI, int STG = 0>
struct G2Cfg {
  static constexpr int EPI_WARPS = (EPI == EPI_BIAS_GELU || EPI == EPI_GELU_BWD) ? 16 : 8;
  static constexpr int THREADS = (EPI_WARPS + 2) * 32;
  static constexpr int COLS_PER_WARP = 256 / (EPI_WARPS / 4);
  static constexpr int STAGES = STG != 0 ? STG : (EPI_WARPS == 16 ? 5 : 6);
  static constexpr int EPI_STAGE_BYTES = EPI_WARPS * 4096;
  static constexpr int SMEM_BYTES = STAGES * G2_STAGE_BYTES + EPI_STAGE_BYTES + 1024 + 256;
};
constexpr int G2_TMEM_COLS = 512;

template <int EPI, bool A_MN, bool B_MN, int STG = 0>
__global__ void __cluster_dims__(2, 1, 1) __launch_bounds__(G2Cfg<EPI>::THREADS, 1)
gemm_tc2_kernel(const __grid_constant__ CUtensorMap tmA, const __grid_constant__ CUtensorMap tmB,
                const __grid_constant__ CUtensorMap tmC, const GemmParams p, const int tma_c) {
  constexpr int G2_EPI_WARPS = G2Cfg<EPI>::EPI_WARPS, G2_STAGES = G2Cfg<EPI, STG>::STAGES;
  constexpr int G2_COLS_PER_WARP = G2Cfg<EPI>::COLS_PER_WARP, G2_EPI_STAGE_BYTES = G2Cfg<EPI>::EPI_STAGE_BYTES;
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  uint8_t* epi_stage = smem + G2_STAGES * G2_STAGE_BYTES;
  uint64_t* bars = reinterpret_cast<uint64_t*>(epi_stage + G2_EPI_STAGE_BYTES);
  uint64_t* full = bars;                    // [STAGES] used in the leader CTA only
  uint64_t* empty = bars + G2_STAGES;       // [STAGES] one per CTA, signalled by the leader's multicast commit
  uint64_t* tfull = bars + 2 * G2_STAGES;   // [2] one per CTA (multicast commit)
  uint64_t* tempty = tfull + 2;             // [2] leader only: 16 arrivals (8 epilogue warps x 2 CTAs)
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(tempty + 2);

  const int warp = threadIdx.x >> 5;
  const int lane = threadIdx.x & 31;
  const uint32_t rank = cluster_ctarank();
  const bool leader = rank == 0;
  const int pair = blockIdx.x >> 1;
  const int num_pairs = gridDim.x >> 1;

  const int num_m = (p.M + G2_BM - 1) / G2_BM;
  const int num_n = (p.N + G2_BN - 1) / G2_BN;
  const int kb_total = (p.K + GEMM_BK - 1) / GEMM_BK;
  const int kb_per = (kb_total + p.k_splits - 1) / p.k_splits;
  const int total = num_m * num_n * p.k_splits;

  if (threadIdx.x == 0) {
    for (int i = 0; i < G2_STAGES; ++i) {
      mbar_init(&full[i], 1);
      mbar_init(&empty[i], 1);
    }
    for (int i = 0; i < 2; ++i) {
      mbar_init(&tfull[i], 1);
      mbar_init(&tempty[i], 2 * G2_EPI_WARPS);
    }
    mbar_fence_init();
  }
  if (warp == G2_EPI_WARPS && lane == 0) {
    tma_prefetch_desc(&tmA);
    tma_prefetch_desc(&tmB);
  }
  if (warp == G2_EPI_WARPS + 1) tmem_alloc_pair(tmem_slot, G2_TMEM_COLS);
  tc_fence_before();
  cluster_sync_all();                 // barrier inits + TMEM allocation visible in both CTAs
  tc_fence_after();
  pdl_wait();                         // prologue done; everything below reads / writes global memory
  const uint32_t tmem_base = *tmem_slot;

  if (warp == G2_EPI_WARPS) {
    // ------------------------------------------------------------ TMA producer (both CTAs)
    if (lane == 0) {
      int stage = 0;
      uint32_t phase = 0;
      for (int w = pair; w < total; w += num_pairs) {
        if (w + num_pairs >= total) pdl_launch_dependents();   // last tile of this CTA: the next kernel may be staged
        const int split = w % p.k_splits;
        const int tile = w / p.k_splits;
        const int m0 = (tile / num_n) * G2_BM + rank * 128;
        const int n0 = (tile % num_n) * G2_BN + rank * 128;
        const int kb0 = split * kb_per;
        const int kb1 = min(kb_total, kb0 + kb_per);
        for (int kb = kb0; kb < kb1; ++kb) {
          mbar_wait(&empty[stage], phase ^ 1);
          uint8_t* sA = smem + stage * G2_STAGE_BYTES;
          uint8_t* sB = sA + G2_A_BYTES;
          if (leader) mbar_arrive_expect_tx(&full[stage], 2 * G2_STAGE_BYTES);   // bytes of BOTH CTAs
          if constexpr (!A_MN) {
            tma_load_2d_pair(sA, &tmA, &full[stage], kb * GEMM_BK, m0);
          } else {
            tma_load_2d_pair(sA, &tmA, &full[stage], m0, kb * GEMM_BK);
            tma_load_2d_pair(sA + 8192, &tmA, &full[stage], m0 + 64, kb * GEMM_BK);
          }
          if constexpr (!B_MN) {
            tma_load_2d_pair(sB, &tmB, &full[stage], kb * GEMM_BK, n0);
          } else {
            tma_load_2d_pair(sB, &tmB, &full[stage], n0, kb * GEMM_BK);
            tma_load_2d_pair(sB + 8192, &tmB, &full[stage], n0 + 64, kb * GEMM_BK);
          }
          if (++stage == G2_STAGES) { stage = 0; phase ^= 1; }
        }
      }
    }
  } else if (warp == G2_EPI_WARPS + 1) {
    // ------------------------------------------------------------ MMA issuer (leader CTA only)
    if (leader) {
      constexpr uint32_t idesc = umma_idesc_bf16(G2_BM, G2_BN, A_MN, B_MN);
      // converged issue (gxa_ptx.cuh): descriptors are warp-uniform values computed by all lanes; only the tcgen05
      // instructions are predicated on the elected lane
      constexpr uint32_t HI = umma_desc_hi(1024);
      constexpr uint32_t A_LBO = A_MN ? 8192 : 16, B_LBO = B_MN ? 8192 : 16;
      constexpr uint32_t A_STEP = (A_MN ? 2048 : 32) >> 4, B_STEP = (B_MN ? 2048 : 32) >> 4;
      const uint32_t el = elect_one_pred();
      int stage = 0;
      uint32_t phase = 0;
      int as = 0;
      uint32_t aphase = 0;
      for (int w = pair; w < total; w += num_pairs) {
        const int split = w % p.k_splits;
        const int kb0 = split * kb_per;
        const int kb1 = min(kb_total, kb0 + kb_per);
        mbar_wait(&tempty[as], aphase ^ 1);
        tc_fence_after();
        const uint32_t tmem_d = tmem_base + as * G2_BN;
        for (int kb = kb0; kb < kb1; ++kb) {
          mbar_wait(&full[stage], phase);
          tc_fence_after();
          const uint32_t a_base = smem_u32(smem + stage * G2_STAGE_BYTES);
          const uint32_t da = umma_desc_lo(a_base, A_LBO), db = umma_desc_lo(a_base + G2_A_BYTES, B_LBO);
          const uint32_t acc0 = kb > kb0 ? 1u : 0u;
#pragma unroll
          for (int k = 0; k < GEMM_BK / 16; ++k)
            umma_bf16_pair_e(tmem_d, da + k * A_STEP, HI, db + k * B_STEP, HI, idesc, k > 0 ? 1u : acc0, el);
          umma_commit_pair_e(&empty[stage], 3, el);     // frees this smem slot in BOTH CTAs
          if (++stage == G2_STAGES) { stage = 0; phase ^= 1; }
        }
        umma_commit_pair_e(&tfull[as], 3, el);          // accumulator ready in both CTAs
        as ^= 1;
        if (as == 0) aphase ^= 1;
      }
    }
  } else {
    // ------------------------------------------------------------ epilogue warps 0..7 (both CTAs)
    const int quarter = warp & 3;
    const int half = warp >> 2;          // column group of this warp: G2_COLS_PER_WARP accumulator columns
    uint8_t* stage_buf = epi_stage + warp * 4096;
    float gate_t = 0.f;
    if constexpr (EPI == EPI_GATE_RES) gate_t = tanhf(__ldg(p.gate));
    const uint32_t tempty_leader0 = mapa_shared(smem_u32(&tempty[0]), 0);
    const uint32_t tempty_leader1 = mapa_shared(smem_u32(&tempty[1]), 0);
    int as = 0;
    uint32_t aphase = 0;
    constexpr int CHUNKS = G2_COLS_PER_WARP / 32;           // 32-column chunks per warp per tile
    AuxRegs<EPI> ax[2];                              // chunk operands from HBM, fetched one chunk ahead
    auto tile_origin = [&](int w, int& row0, int& n0) {
      const int tile = w / p.k_splits;
      row0 = (tile / num_n) * G2_BM + rank * 128 + quarter * 32;
      n0 = (tile % num_n) * G2_BN + half * G2_COLS_PER_WARP;
    };
    if (pair < total) {
      int r0, n0;
      tile_origin(pair, r0, n0);
      if (n0 < p.N) aux_prefetch<EPI>(p, r0, n0, lane, ax[0]);
    }
    for (int w = pair; w < total; w += num_pairs) {
      const int split = w % p.k_splits;
      int row0, n0;
      tile_origin(w, row0, n0);
      mbar_wait(&tfull[as], aphase);
      tc_fence_after();
#pragma unroll
      for (int ci = 0; ci < CHUNKS; ++ci) {
        const int col = n0 + ci * 32;
        if (ci + 1 < CHUNKS) {
          if (col + 32 < p.N) aux_prefetch<EPI>(p, row0, col + 32, lane, ax[(ci + 1) & 1]);
        } else if (w + num_pairs < total) {
          int r1, n1;
          tile_origin(w + num_pairs, r1, n1);
          if (n1 < p.N) aux_prefetch<EPI>(p, r1, n1, lane, ax[0]);
        }
        if (col < p.N) {
          uint32_t r[32];
          tmem_ld_32x32(tmem_base + (static_cast<uint32_t>(quarter * 32) << 16) + as * G2_BN + half * G2_COLS_PER_WARP + ci * 32, r);
          tmem_ld_wait();
          epilogue_chunk_v2<EPI>(p, stage_buf, row0, col, r, gate_t, split == 0, lane, ax[ci & 1], tma_c ? &tmC : nullptr);
        }
      }
      tc_fence_before();
      __syncwarp();
      if (lane == 0) mbar_arrive_cluster(as == 0 ? tempty_leader0 : tempty_leader1);
      as ^= 1;
      if (as == 0) aphase ^= 1;
    }
    if (tma_c && lane == 0) bulk_wait_group0();     // reductions complete before the CTA retires its shared memory
  }

  tc_fence_before();
  __syncwarp();
  cluster_sync_all();                 // nobody exits (or frees TMEM) while the peer may still signal / read
  if (warp == G2_EPI_WARPS + 1) {
    tc_fence_after();
    tmem_dealloc_pair(tmem_base, G2_TMEM_COLS);
  }
}

// ------------------------------------------------------------------ host side
int make_tmap_bf16(CUtensorMap* m, const void* base, uint64_t inner, uint64_t outer, uint64_t ld, uint32_t box_inner,
                   uint32_t box_outer);
int make_tmap_f32_2d(CUtensorMap* m, const void* base, uint64_t inner, uint64_t outer, uint64_t ld, uint32_t box_inner,
                     uint32_t box_outer);

template <int EPI, bool A_MN, bool B_MN, int STG = 0>
static int launch2_inst(const GemmOp& op, const GemmParams& p, cudaStream_t stream) {
  CUtensorMap tmA, tmB;
  int e;
  if (!A_MN) e = make_tmap_bf16(&tmA, op.A, op.K, op.M, op.lda, GEMM_BK, 128);
  else       e = make_tmap_bf16(&tmA, op.A, op.M, op.K, op.lda, 64, GEMM_BK);
  if (e) return e;
  if (!B_MN) e = make_tmap_bf16(&tmB, op.B, op.K, op.N, op.ldb, GEMM_BK, 128);
  else       e = make_tmap_bf16(&tmB, op.B, op.N, op.K, op.ldb, 64, GEMM_BK);
  if (e) return e;
  // fp32 accumulation (wgrad, split-K, dxa) without a bias goes out as TMA reduce-adds of 32 x 32 boxes
  CUtensorMap tmC = tmA;
  int tma_c = 0;
  if (EPI == EPI_F32 && (p.accumulate || p.k_splits > 1) && p.bias == nullptr &&
      make_tmap_f32_2d(&tmC, p.C, p.N, p.M, p.ldc, 32, 32) == 0)
    tma_c = 1;
  auto kern = gemm_tc2_kernel<EPI, A_MN, B_MN, STG>;
  if (int ae = set_max_smem_once(kern, G2Cfg<EPI, STG>::SMEM_BYTES)) return ae;
  const int num_m = (p.M + G2_BM - 1) / G2_BM, num_n = (p.N + G2_BN - 1) / G2_BN;
  const int total = num_m * num_n * p.k_splits;
  const int max_pairs = num_sms() / 2;
  const int pairs = total < max_pairs ? total : max_pairs;
  const cudaError_t le = launch_pdl(kern, dim3(2 * pairs), dim3(G2Cfg<EPI>::THREADS), G2Cfg<EPI, STG>::SMEM_BYTES, stream, tmA,
                                    tmB, tmC, p, tma_c);
  count_launch();
  return le == cudaSuccess && cudaGetLastError() == cudaSuccess ? 0 : GXA_ERR_CUDA;
}

template <int STG>
static int launch2_f32_shallow(const GemmOp& op, const GemmParams& p, cudaStream_t s) {
  if (!op.a_mn && !op.b_mn) return launch2_inst<EPI_F32, false, false, STG>(op, p, s);
  if (!op.a_mn && op.b_mn) return launch2_inst<EPI_F32, false, true, STG>(op, p, s);
  if (op.a_mn && op.b_mn) return launch2_inst<EPI_F32, true, true, STG>(op, p, s);
  return launch2_inst<EPI_F32, true, false, STG>(op, p, s);
}

template <int EPI>
static int launch2_layout(const GemmOp& op, const GemmParams& p, cudaStream_t s) {
  if constexpr (EPI == EPI_F32) {
    // default 5 (round 2, c3 step on one B200: 6 stages 86.9 ms, 5 stages 85.8 ms, 4 stages 86.5 ms): the shallower ring
    // leaves 32 KB of shared memory per SM, enough for a CTA of the HBM-bound row-wise kernels to run beside the
    // weight-gradient GEMM of the other stream instead of behind it.  GXA_WGRAD_STAGES=6 restores the deep ring.
    static const int shallow = [] { const char* e = getenv("GXA_WGRAD_STAGES"); return e ? atoi(e) : 5; }();
    if (shallow == 4) return launch2_f32_shallow<4>(op, p, s);
    if (shallow == 5) return launch2_f32_shallow<5>(op, p, s);
  }
  if (!op.a_mn && !op.b_mn) return launch2_inst<EPI, false, false>(op, p, s);
  if (!op.a_mn && op.b_mn) return launch2_inst<EPI, false, true>(op, p, s);
  if constexpr (EPI == EPI_F32) {
    if (op.a_mn && op.b_mn) return launch2_inst<EPI, true, true>(op, p, s);
    if (op.a_mn && !op.b_mn) return launch2_inst<EPI, true, false>(op, p, s);
  }
  return GXA_ERR_UNSUPPORTED;
}

int gemm2_launch(const GemmOp& op, const GemmParams& p, cudaStream_t s) {
  switch (op.epi) {
    case EPI_BF16: return launch2_layout<EPI_BF16>(op, p, s);
    case EPI_F32: return launch2_layout<EPI_F32>(op, p, s);
    case EPI_GATE_RES: return launch2_layout<EPI_GATE_RES>(op, p, s);
    case EPI_BIAS_GELU: return launch2_layout<EPI_BIAS_GELU>(op, p, s);
    case EPI_GELU_BWD: return launch2_layout<EPI_GELU_BWD>(op, p, s);
  }
  return GXA_ERR_UNSUPPORTED;
}

}  // namespace gxa
