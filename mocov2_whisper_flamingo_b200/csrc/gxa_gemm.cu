// tcgen05 GEMM kernel + host launcher (tensor-map construction, dispatch).
#include <cstdlib>
#include <mutex>

#include "gxa_gemm.cuh"
#include "gxa_host.h"

namespace gxa {

template <int BN, int EPI, bool A_MN, bool B_MN>
__global__ void __launch_bounds__(GEMM_THREADS, 1)
gemm_tc_kernel(const __grid_constant__ CUtensorMap tmA, const __grid_constant__ CUtensorMap tmB, const GemmParams p) {
  using Cfg = GemmCfg<BN>;
  constexpr int STAGES = Cfg::STAGES;
  extern __shared__ uint8_t smem_raw[];
  // 128B swizzle atoms need 1024-byte aligned tiles
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  uint64_t* bars = reinterpret_cast<uint64_t*>(smem + STAGES * Cfg::STAGE_BYTES);
  uint64_t* full = bars;                 // [STAGES]  TMA -> MMA
  uint64_t* empty = bars + STAGES;       // [STAGES]  MMA -> TMA
  uint64_t* tfull = bars + 2 * STAGES;   // [2]       MMA -> epilogue
  uint64_t* tempty = tfull + 2;          // [2]       epilogue -> MMA
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(tempty + 2);

  const int warp = threadIdx.x >> 5;
  const int lane = threadIdx.x & 31;

  const int num_m = (p.M + GEMM_BM - 1) / GEMM_BM;
  const int num_n = (p.N + BN - 1) / BN;
  const int kb_total = (p.K + GEMM_BK - 1) / GEMM_BK;
  const int kb_per = (kb_total + p.k_splits - 1) / p.k_splits;
  const int total = num_m * num_n * p.k_splits;

  if (threadIdx.x == 0) {
    for (int i = 0; i < STAGES; ++i) {
      mbar_init(&full[i], 1);
      mbar_init(&empty[i], 1);
    }
    for (int i = 0; i < 2; ++i) {
      mbar_init(&tfull[i], 1);
      mbar_init(&tempty[i], GEMM_EPI_WARPS * 32);
    }
    mbar_fence_init();
  }
  if (warp == 8 && lane == 0) {
    tma_prefetch_desc(&tmA);
    tma_prefetch_desc(&tmB);
  }
  if (warp == 9) tmem_alloc(tmem_slot, Cfg::TMEM_COLS);
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  pdl_wait();                           // prologue done; everything below reads / writes global memory
  const uint32_t tmem_base = *tmem_slot;

  if (warp == 8) {
    // ------------------------------------------------------------ TMA producer
    if (lane == 0) {
      int stage = 0;
      uint32_t phase = 0;
      for (int w = blockIdx.x; w < total; w += gridDim.x) {
        if (w + gridDim.x >= total) pdl_launch_dependents();   // last tile of this CTA: the next kernel may be staged
        const int split = w % p.k_splits;
        const int tile = w / p.k_splits;
        const int m0 = (tile / num_n) * GEMM_BM;
        const int n0 = (tile % num_n) * BN;
        const int kb0 = split * kb_per;
        const int kb1 = min(kb_total, kb0 + kb_per);
        for (int kb = kb0; kb < kb1; ++kb) {
          mbar_wait(&empty[stage], phase ^ 1);
          uint8_t* sA = smem + stage * Cfg::STAGE_BYTES;
          uint8_t* sB = sA + Cfg::A_BYTES;
          mbar_arrive_expect_tx(&full[stage], Cfg::STAGE_BYTES);
          if constexpr (!A_MN) {
            tma_load_2d(sA, &tmA, &full[stage], kb * GEMM_BK, m0);
          } else {
#pragma unroll
            for (int j = 0; j < GEMM_BM / 64; ++j) tma_load_2d(sA + j * 8192, &tmA, &full[stage], m0 + j * 64, kb * GEMM_BK);
          }
          if constexpr (!B_MN) {
            tma_load_2d(sB, &tmB, &full[stage], kb * GEMM_BK, n0);
          } else {
#pragma unroll
            for (int j = 0; j < BN / 64; ++j) tma_load_2d(sB + j * 8192, &tmB, &full[stage], n0 + j * 64, kb * GEMM_BK);
          }
          if (++stage == STAGES) { stage = 0; phase ^= 1; }
        }
      }
    }
  } else if (warp == 9) {
    // ------------------------------------------------------------ MMA issuer
    constexpr uint32_t idesc = umma_idesc_bf16(GEMM_BM, BN, A_MN, B_MN);
    int stage = 0;
    uint32_t phase = 0;
    int as = 0;
    uint32_t aphase = 0;
    for (int w = blockIdx.x; w < total; w += gridDim.x) {
      const int split = w % p.k_splits;
      const int kb0 = split * kb_per;
      const int kb1 = min(kb_total, kb0 + kb_per);
      mbar_wait(&tempty[as], aphase ^ 1);
      tc_fence_after();
      const uint32_t tmem_d = tmem_base + as * BN;
      for (int kb = kb0; kb < kb1; ++kb) {
        mbar_wait(&full[stage], phase);
        tc_fence_after();
        if (lane == 0) {
          const uint32_t a_base = smem_u32(smem + stage * Cfg::STAGE_BYTES);
          const uint32_t b_base = a_base + Cfg::A_BYTES;
#pragma unroll
          for (int k = 0; k < GEMM_BK / 16; ++k) {
            const uint64_t da = A_MN ? umma_desc_sw128(a_base + k * 2048, 8192, 1024)
                                     : umma_desc_sw128(a_base + k * 32, 16, 1024);
            const uint64_t db = B_MN ? umma_desc_sw128(b_base + k * 2048, 8192, 1024)
                                     : umma_desc_sw128(b_base + k * 32, 16, 1024);
            umma_bf16(tmem_d, da, db, idesc, (kb > kb0 || k > 0) ? 1u : 0u);
          }
          umma_commit(&empty[stage]);   // smem slot reusable once these MMAs retire
        }
        __syncwarp();
        if (++stage == STAGES) { stage = 0; phase ^= 1; }
      }
      if (lane == 0) umma_commit(&tfull[as]);   // accumulator complete
      __syncwarp();
      as ^= 1;
      if (as == 0) aphase ^= 1;
    }
  } else {
    // ------------------------------------------------------------ epilogue warps 0..7
    const int quarter = warp & 3;          // TMEM lane quarter this warp may access
    const int half = warp >> 2;            // which half of the BN columns
    float gate_t = 0.f;
    if constexpr (EPI == EPI_GATE_RES) gate_t = tanhf(__ldg(p.gate));
    if constexpr (EPI == EPI_F32) gate_t = p.gate != nullptr ? tanhf(__ldg(p.gate)) : 1.f;
    int as = 0;
    uint32_t aphase = 0;
    for (int w = blockIdx.x; w < total; w += gridDim.x) {
      const int split = w % p.k_splits;
      const int tile = w / p.k_splits;
      const int m0 = (tile / num_n) * GEMM_BM;
      const int n0 = (tile % num_n) * BN;
      mbar_wait(&tfull[as], aphase);
      tc_fence_after();
      const int row = m0 + quarter * 32 + lane;
#pragma unroll 1
      for (int c = half * (BN / 2); c < (half + 1) * (BN / 2); c += 32) {
        if (n0 + c >= p.N) break;
        uint32_t r[32];
        tmem_ld_32x32(tmem_base + (static_cast<uint32_t>(quarter * 32) << 16) + as * BN + c, r);
        tmem_ld_wait();
        epilogue_chunk<EPI>(p, row, n0 + c, r, gate_t, split == 0);
      }
      tc_fence_before();
      mbar_arrive(&tempty[as]);
      as ^= 1;
      if (as == 0) aphase ^= 1;
    }
  }

  tc_fence_before();
  __syncthreads();
  if (warp == 9) {
    tc_fence_after();
    tmem_dealloc(tmem_base, Cfg::TMEM_COLS);
  }
}

// ------------------------------------------------------------------ host side
static PFN_cuTensorMapEncodeTiled_t g_encode = nullptr;

static int get_encode() {
  if (g_encode) return 0;
  void* fn = nullptr;
  cudaDriverEntryPointQueryResult q;
  cudaError_t e = cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &fn, cudaEnableDefault, &q);
  if (e != cudaSuccess || q != cudaDriverEntryPointSuccess || fn == nullptr) return GXA_ERR_DRIVER;
  g_encode = reinterpret_cast<PFN_cuTensorMapEncodeTiled_t>(fn);
  return 0;
}

// Descriptor cache.  A training step encodes ~1000 tensor maps (2-3 per GEMM, 18 GEMMs per layer) and the caching allocator
// hands the same buffers back step after step, so nearly every encode repeats an earlier one; cuTensorMapEncodeTiled is a
// driver call of 1-2 us, which is host time the 8-samples-per-GPU step does not have.  Direct-mapped, keyed by every
// argument of the encode; a tensor map is a pure function of those arguments (it holds no reference to the memory).
struct TmapKey {
  const void* base;
  uint64_t inner, outer, ld;
  uint32_t box_inner, box_outer, kind, pad;
  uint64_t outer2, ld2;          // third dimension of the 3-D maps (0 otherwise)
  bool operator==(const TmapKey& o) const {
    return base == o.base && inner == o.inner && outer == o.outer && ld == o.ld && box_inner == o.box_inner &&
           box_outer == o.box_outer && kind == o.kind && outer2 == o.outer2 && ld2 == o.ld2;
  }
};
struct TmapSlot {
  TmapKey key;
  CUtensorMap map;
  bool valid;
};
constexpr int kTmapSlots = 4096;
static TmapSlot* g_tmap_cache = nullptr;
static std::mutex g_tmap_mu;
static const bool g_tmap_cache_on = [] { const char* e = getenv("GXA_TMAP_CACHE"); return !(e && e[0] == '0'); }();

static inline size_t tmap_hash(const TmapKey& k) {
  uint64_t h = reinterpret_cast<uint64_t>(k.base) * 0x9E3779B97F4A7C15ull;
  h ^= (k.inner * 0xC2B2AE3D27D4EB4Full) ^ (k.outer * 0x165667B19E3779F9ull) ^ (k.ld << 17) ^
       (uint64_t(k.box_inner) << 40) ^ (uint64_t(k.box_outer) << 48) ^ (uint64_t(k.kind) << 56) ^
       (k.outer2 * 0x27D4EB2F165667C5ull) ^ (k.ld2 << 7);
  return static_cast<size_t>((h ^ (h >> 29)) % kTmapSlots);
}
static bool tmap_lookup(const TmapKey& k, CUtensorMap* out) {
  if (!g_tmap_cache_on) return false;
  std::lock_guard<std::mutex> lk(g_tmap_mu);
  if (!g_tmap_cache) return false;
  const TmapSlot& s = g_tmap_cache[tmap_hash(k)];
  if (!s.valid || !(s.key == k)) return false;
  *out = s.map;
  return true;
}
static void tmap_store(const TmapKey& k, const CUtensorMap& m) {
  if (!g_tmap_cache_on) return;
  std::lock_guard<std::mutex> lk(g_tmap_mu);
  if (!g_tmap_cache) g_tmap_cache = new TmapSlot[kTmapSlots]();
  TmapSlot& s = g_tmap_cache[tmap_hash(k)];
  s.key = k;
  s.map = m;
  s.valid = true;
}

// 2-D bf16 tensor map: `inner` contiguous elements, `outer` rows `ld` elements apart.
int make_tmap_bf16(CUtensorMap* m, const void* base, uint64_t inner, uint64_t outer, uint64_t ld, uint32_t box_inner,
                   uint32_t box_outer) {
  if (int e = get_encode()) return e;
  if ((reinterpret_cast<uintptr_t>(base) & 15) || (ld * 2) % 16) return GXA_ERR_ALIGN;
  const TmapKey key{base, inner, outer, ld, box_inner, box_outer, 1u, 0u, 0, 0};
  if (tmap_lookup(key, m)) return 0;
  cuuint64_t dims[2] = {inner, outer};
  cuuint64_t strides[1] = {ld * 2};
  cuuint32_t box[2] = {box_inner, box_outer};
  cuuint32_t estr[2] = {1, 1};
  CUresult r = g_encode(m, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 2, const_cast<void*>(base), dims, strides, box, estr,
                        CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
                        CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  if (r != CUDA_SUCCESS) return GXA_ERR_TMAP;
  tmap_store(key, *m);
  return 0;
}


// 2-D fp32 tensor map (output side: TMA reduce-add of 128B-swizzled boxes)
int make_tmap_f32_2d(CUtensorMap* m, const void* base, uint64_t inner, uint64_t outer, uint64_t ld, uint32_t box_inner,
                     uint32_t box_outer) {
  if (int e = get_encode()) return e;
  if ((reinterpret_cast<uintptr_t>(base) & 15) || (ld * 4) % 16) return GXA_ERR_ALIGN;
  const TmapKey key{base, inner, outer, ld, box_inner, box_outer, 2u, 0u, 0, 0};
  if (tmap_lookup(key, m)) return 0;
  cuuint64_t dims[2] = {inner, outer};
  cuuint64_t strides[1] = {ld * 4};
  cuuint32_t box[2] = {box_inner, box_outer};
  cuuint32_t estr[2] = {1, 1};
  CUresult r = g_encode(m, CU_TENSOR_MAP_DATA_TYPE_FLOAT32, 2, const_cast<void*>(base), dims, strides, box, estr,
                        CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_NONE,
                        CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  if (r != CUDA_SUCCESS) return GXA_ERR_TMAP;
  tmap_store(key, *m);
  return 0;
}

// 3-D bf16 tensor map [d2][d1][d0] (d0 contiguous): box {b0, b1, 1}, 128B swizzle, OOB rows read as zero.
int make_tmap_bf16_3d(CUtensorMap* m, const void* base, uint64_t d0, uint64_t d1, uint64_t d2, uint64_t ld1,
                      uint64_t ld2, uint32_t b0, uint32_t b1) {
  if (int e = get_encode()) return e;
  if ((reinterpret_cast<uintptr_t>(base) & 15) || (ld1 * 2) % 16 || (ld2 * 2) % 16) return GXA_ERR_ALIGN;
  const TmapKey key{base, d0, d1, ld1, b0, b1, 3u, 0u, d2, ld2};
  if (tmap_lookup(key, m)) return 0;
  cuuint64_t dims[3] = {d0, d1, d2};
  cuuint64_t strides[2] = {ld1 * 2, ld2 * 2};
  cuuint32_t box[3] = {b0, b1, 1};
  cuuint32_t estr[3] = {1, 1, 1};
  CUresult r = g_encode(m, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 3, const_cast<void*>(base), dims, strides, box, estr,
                        CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
                        CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  if (r != CUDA_SUCCESS) return GXA_ERR_TMAP;
  tmap_store(key, *m);
  return 0;
}

template <int BN, int EPI, bool A_MN, bool B_MN>
static int launch_inst(const GemmOp& op, const GemmParams& p, cudaStream_t stream) {
  using Cfg = GemmCfg<BN>;
  CUtensorMap tmA, tmB;
  int e;
  if (!A_MN) e = make_tmap_bf16(&tmA, op.A, op.K, op.M, op.lda, GEMM_BK, GEMM_BM);
  else       e = make_tmap_bf16(&tmA, op.A, op.M, op.K, op.lda, 64, GEMM_BK);
  if (e) return e;
  if (!B_MN) e = make_tmap_bf16(&tmB, op.B, op.K, op.N, op.ldb, GEMM_BK, BN);
  else       e = make_tmap_bf16(&tmB, op.B, op.N, op.K, op.ldb, 64, GEMM_BK);
  if (e) return e;
  auto kern = gemm_tc_kernel<BN, EPI, A_MN, B_MN>;
  if (int ae = set_max_smem_once(kern, Cfg::SMEM_BYTES)) return ae;
  const int num_m = (p.M + GEMM_BM - 1) / GEMM_BM, num_n = (p.N + BN - 1) / BN;
  const int total = num_m * num_n * p.k_splits;
  const int grid = total < num_sms() ? total : num_sms();
  const cudaError_t le = launch_pdl(kern, dim3(grid), dim3(GEMM_THREADS), Cfg::SMEM_BYTES, stream, tmA, tmB, p);
  count_launch();
  return le == cudaSuccess && cudaGetLastError() == cudaSuccess ? 0 : GXA_ERR_CUDA;
}

template <int BN, int EPI>
static int launch_layout(const GemmOp& op, const GemmParams& p, cudaStream_t s) {
  if (!op.a_mn && !op.b_mn) return launch_inst<BN, EPI, false, false>(op, p, s);
  if (!op.a_mn && op.b_mn) return launch_inst<BN, EPI, false, true>(op, p, s);
  if constexpr (EPI == EPI_F32) {
    if (op.a_mn && op.b_mn) return launch_inst<BN, EPI, true, true>(op, p, s);
    if (op.a_mn && !op.b_mn) return launch_inst<BN, EPI, true, false>(op, p, s);
  }
  return GXA_ERR_UNSUPPORTED;
}

template <int BN>
static int launch_epi(const GemmOp& op, const GemmParams& p, cudaStream_t s) {
  switch (op.epi) {
    case EPI_BF16: return launch_layout<BN, EPI_BF16>(op, p, s);
    case EPI_F32: return launch_layout<BN, EPI_F32>(op, p, s);
    case EPI_GATE_RES: return launch_layout<BN, EPI_GATE_RES>(op, p, s);
    case EPI_BIAS_GELU: return launch_layout<BN, EPI_BIAS_GELU>(op, p, s);
    case EPI_GELU_BWD: return launch_layout<BN, EPI_GELU_BWD>(op, p, s);
  }
  return GXA_ERR_UNSUPPORTED;
}

int gemm2_launch(const GemmOp& op, const GemmParams& p, cudaStream_t s);

// split-K factor for the atomically-accumulating fp32 epilogue: fill every SM (or SM pair) in whole waves
static int pick_splits(int tiles, int workers, int kb_total) {
  int best = 1;
  double best_eff = 0.0;
  for (int ks = 1; ks <= 16; ++ks) {
    if (ks > 1 && kb_total / ks < 8) break;
    const int units = tiles * ks;
    const int waves = (units + workers - 1) / workers;
    const double eff = double(units) / (double(waves) * workers);
    if (eff > best_eff + 0.03) { best_eff = eff; best = ks; }
  }
  return best;
}

int gemm_launch(const GemmOp& op, cudaStream_t stream) {
  if (op.M <= 0 || op.N <= 0 || op.K <= 0) return GXA_ERR_SHAPE;
  if (op.N % 64 != 0) return GXA_ERR_SHAPE;
  GemmParams p;
  p.M = op.M; p.N = op.N; p.K = op.K;
  p.C = op.C; p.ldc = op.ldc; p.C2 = op.C2; p.ldc2 = op.ldc2;
  p.bias = op.bias; p.aux = op.aux; p.ld_aux = op.ld_aux; p.gate = op.gate; p.accumulate = op.accumulate; p.colsum = op.colsum; p.drop = op.drop;
  const int kb_total = (op.K + GEMM_BK - 1) / GEMM_BK;
  int bn = op.bn;
  if (bn == 0) {
    // CTA-pair 256x256 tiles for anything big; 1-CTA 128-wide tiles for small / narrow problems
    if (op.M >= 512 && op.N >= 512 && (op.N % 256 == 0 || op.N >= 1024)) bn = 512;
    else {
      const int num_m = (op.M + GEMM_BM - 1) / GEMM_BM;
      bn = (op.N % 256 == 0 && num_m * (op.N / 256) >= num_sms()) ? 256 : 128;
    }
  }
  if (op.epi == EPI_F32 && op.gate != nullptr && bn == 512) bn = 128;   // gated accumulation lives in the 1-CTA epilogue only
  const bool pair = bn == 512;
  int ks = op.k_splits;
  if (ks == 0) {
    ks = 1;
    if (op.epi == EPI_F32) {   // wgrad: few output tiles, very long K
      const int tiles = pair ? ((op.M + 255) / 256) * ((op.N + 255) / 256)
                             : ((op.M + GEMM_BM - 1) / GEMM_BM) * ((op.N + bn - 1) / bn);
      ks = pick_splits(tiles, pair ? num_sms() / 2 : num_sms(), kb_total);
    }
  }
  while (ks > 1 && (ks - 1) * ((kb_total + ks - 1) / ks) >= kb_total) --ks;   // no empty split
  p.k_splits = ks;
  if (ks > 1 && op.epi != EPI_F32) return GXA_ERR_UNSUPPORTED;
  if (ks > 1 && !op.accumulate) {
    // split-K partials are combined with fp32 atomics: the destination must start from zero
    if (op.ldc == op.N) {
      if (cudaMemsetAsync(op.C, 0, sizeof(float) * size_t(op.M) * op.N, stream) != cudaSuccess) return GXA_ERR_CUDA;
    } else {
      if (cudaMemset2DAsync(op.C, sizeof(float) * op.ldc, 0, sizeof(float) * op.N, op.M, stream) != cudaSuccess)
        return GXA_ERR_CUDA;
    }
  }
  if (op.colsum != nullptr && op.epi != EPI_GELU_BWD) return GXA_ERR_UNSUPPORTED;
  if (op.drop.thresh != 0 && op.epi != EPI_GATE_RES) return GXA_ERR_UNSUPPORTED;
  if (pair) return gemm2_launch(op, p, stream);
  int e = GXA_ERR_UNSUPPORTED;
  if (bn == 256) e = launch_epi<256>(op, p, stream);
  else if (bn == 128) e = launch_epi<128>(op, p, stream);
  // the small-problem kernel has no fused column sum: one extra pass over its (bf16) output
  if (e == 0 && op.colsum != nullptr) e = colsum(op.C, 0, op.ldc, op.colsum, op.M, op.N, stream);
  return e;
}

}  // namespace gxa
