// Few-row ("skinny") linear layers for the decode step (BASELINE config 5): Y[R, N] = X[R, K] W[N, K]^T with
// R = batch x beams <= 256 query rows (160 at config 5) and the epilogues of the block's four projections.
//
// Why not the 128-row-tile GEMMs: with 160 rows a layer's GEMMs have 12-48 output tiles, each CTA walks the whole K
// through a 4-6 stage ring (~85 GB/s per SM: bytes in flight / TMA latency), and the step is HBM-bound -- what matters is
// how many SMs pull weights at once.  Measured before this kernel (profiles/r02_gemm_experiments.txt item 5): Q 9.4 us,
// out-proj 8.3 + 1.8 (copy), FFN1 15.5, FFN2 8.3 + 1.8 us per layer for 11.8 MB of weights (2 us at HBM speed).
//
// Design.  The operands are swapped: the WEIGHT tile is the M = 128 side of the tcgen05 MMA and all R rows of X are its
// N side (R_pad <= 256 = one MMA instruction wide), so no tensor-core row is wasted on padding and X is never re-read
// per row tile.  K is split over a thread-block CLUSTER of `ks` CTAs (<= 8): CTA `s` of the cluster multiplies k-blocks
// [s*kbp, (s+1)*kbp) -- two to six 64-wide blocks, all in flight at once -- into a [128 x R_pad] fp32 accumulator in
// TMEM.  Row r of X is owned by CTA r % ks: after a first cluster barrier (every CTA's operand ring is dead) each CTA
// PUSHES its partial of every row into the owner's shared memory with st.shared::cluster (fire-and-forget, one 128-byte
// line per warp store; the first version PULLED with ld.shared::cluster and spent ~5 us of its 9 us waiting on
// dependent remote loads), a second cluster barrier publishes the stores, and the owner sums its ks slots from LOCAL
// shared memory and applies the fused epilogue.  Nothing goes through global atomics, the output needs no zero-fill,
// and non-linear epilogues (GELU) see the complete sum.
//   grid = (N / 128) * ks CTAs,  192 threads: warps 0-3 accumulator spill + reduction + epilogue, warp 4 TMA, warp 5 MMA
#include "gxa_gemm.cuh"
#include "gxa_host.h"

namespace gxa {

enum SkinnyEpi : int {
  SK_BF16 = 0,   // out(bf16) = acc + bias                                   (Q projection)
  SK_GELU = 1,   // out(bf16) = gelu_erf(bf16(acc + bias))                  (ff.0 + GELU; rounding point of EPI_BIAS_GELU)
  SK_GATE = 2,   // out(f32)  = res(f32) + tanh(gate) * (acc + bias)        (out-proj / ff.2 + gated residual)
};

struct SkinnyParams {
  int R, Rpad, N, K;
  int ks, kb_per;
  void* out;
  int ldo;
  const float* bias;
  const float* res;
  int ldres;
  const float* gate;
};

constexpr int SK_THREADS = 192;
constexpr int SK_STAGES = 4;
constexpr int SK_W_BYTES = 128 * GEMM_BK * 2;

__device__ __forceinline__ void st_dsmem_f32(uint32_t cluster_addr, float v) {
  asm volatile("st.shared::cluster.f32 [%0], %1;" ::"r"(cluster_addr), "f"(v) : "memory");
}

template <int EPI>
__global__ void __launch_bounds__(SK_THREADS, 1)
skinny_gemm_kernel(const __grid_constant__ CUtensorMap tmW, const __grid_constant__ CUtensorMap tmX, const SkinnyParams p) {
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  const int stage_bytes = SK_W_BYTES + p.Rpad * 128;                 // multiple of 1024 (Rpad % 32 == 0)
  const int ring_bytes = SK_STAGES * stage_bytes;
  const int rows_per = (p.R + p.ks - 1) / p.ks;                       // rows of X this CTA owns (r % ks == rank)
  const int part_bytes = p.ks * rows_per * 512;                       // fp32 slots [ks][rows_per][128], re-use the ring
  uint64_t* bars = reinterpret_cast<uint64_t*>(smem + (ring_bytes > part_bytes ? ring_bytes : part_bytes));
  uint64_t* full = bars;                    // [SK_STAGES]
  uint64_t* empty = bars + SK_STAGES;       // [SK_STAGES]
  uint64_t* tfull = bars + 2 * SK_STAGES;   // [1]
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(tfull + 1);

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const uint32_t split = cluster_ctarank();
  const int tile = blockIdx.x / p.ks;
  const int n0 = tile * 128;
  const int kb_total = (p.K + GEMM_BK - 1) / GEMM_BK;
  const int kb0 = split * p.kb_per;
  const int kb1 = min(kb_total, kb0 + p.kb_per);

  if (threadIdx.x == 0) {
    for (int i = 0; i < SK_STAGES; ++i) {
      mbar_init(&full[i], 1);
      mbar_init(&empty[i], 1);
    }
    mbar_init(tfull, 1);
    mbar_fence_init();
  }
  if (warp == 4 && lane == 0) {
    tma_prefetch_desc(&tmW);
    tma_prefetch_desc(&tmX);
  }
  if (warp == 5) tmem_alloc(tmem_slot, 256);
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  pdl_wait();                           // everything below reads / writes global memory
  const uint32_t tmem_base = *tmem_slot;

  if (warp == 4) {
    if (lane == 0) {
      pdl_launch_dependents();
      int stage = 0;
      uint32_t phase = 0;
      for (int kb = kb0; kb < kb1; ++kb) {
        mbar_wait(&empty[stage], phase ^ 1);
        uint8_t* sW = smem + stage * stage_bytes;
        mbar_arrive_expect_tx(&full[stage], stage_bytes);
        tma_load_2d(sW, &tmW, &full[stage], kb * GEMM_BK, n0);                 // rows past N read as zero
        tma_load_2d(sW + SK_W_BYTES, &tmX, &full[stage], kb * GEMM_BK, 0);     // rows past R read as zero
        if (++stage == SK_STAGES) { stage = 0; phase ^= 1; }
      }
    }
  } else if (warp == 5) {
    const uint32_t idesc = umma_idesc_bf16(128, p.Rpad, false, false);
    int stage = 0;
    uint32_t phase = 0;
    for (int kb = kb0; kb < kb1; ++kb) {
      mbar_wait(&full[stage], phase);
      tc_fence_after();
      if (lane == 0) {
        const uint32_t w_base = smem_u32(smem + stage * stage_bytes);
        const uint32_t x_base = w_base + SK_W_BYTES;
#pragma unroll
        for (int k = 0; k < GEMM_BK / 16; ++k)
          umma_bf16(tmem_base, umma_desc_sw128(w_base + k * 32, 16, 1024), umma_desc_sw128(x_base + k * 32, 16, 1024), idesc,
                    (kb > kb0 || k > 0) ? 1u : 0u);
        umma_commit(&empty[stage]);
      }
      __syncwarp();
      if (++stage == SK_STAGES) { stage = 0; phase ^= 1; }
    }
    if (lane == 0) umma_commit(tfull);
    __syncwarp();
  }
  if (warp < 4) {
    mbar_wait(tfull, 0);                // this CTA's MMAs have completed: its accumulator is final, its ring is dead
    tc_fence_after();
  }
  __syncwarp();
  cluster_sync_all();                   // ... and so is every peer's (their epilogue threads waited before arriving)

  if (warp < 4) {
    // accumulator -> the owners' shared memory: thread = TMEM lane = weight row n of the tile; slot [split][r / ks][n]
    const int n = warp * 32 + lane;
    const uint32_t slot = smem_u32(smem) + static_cast<uint32_t>((split * rows_per) * 128 + n) * 4u;
    for (int r0 = 0; r0 < p.Rpad; r0 += 32) {
      uint32_t v[32];
      tmem_ld_32x32(tmem_base + (static_cast<uint32_t>(warp * 32) << 16) + r0, v);
      tmem_ld_wait();
      int owner = r0 % p.ks, j = r0 / p.ks;
#pragma unroll
      for (int i = 0; i < 32; ++i) {
        if (r0 + i < p.R)
          st_dsmem_f32(mapa_shared(slot + static_cast<uint32_t>(j) * 512u, owner), __uint_as_float(v[i]));
        if (++owner == p.ks) { owner = 0; ++j; }
      }
    }
    tc_fence_before();
  }
  __syncwarp();
  cluster_sync_all();                   // all partials have landed in their owners' shared memory

  if (warp < 4) {
    // reduction + epilogue from LOCAL shared memory: a warp takes an owned row, a lane four columns
    const float* part = reinterpret_cast<const float*>(smem);
    const int c = n0 + lane * 4;
    float gate_t = 0.f;
    if constexpr (EPI == SK_GATE) gate_t = tanhf(__ldg(p.gate));
    float4 b = make_float4(0.f, 0.f, 0.f, 0.f);
    if (p.bias != nullptr && c < p.N) b = __ldg(reinterpret_cast<const float4*>(p.bias + c));
    for (int j = warp; j * p.ks + static_cast<int>(split) < p.R; j += 4) {
      const int r = j * p.ks + split;
      float4 acc = b;
      for (int s = 0; s < p.ks; ++s) {
        const float4 v = *reinterpret_cast<const float4*>(part + (s * rows_per + j) * 128 + lane * 4);
        acc.x += v.x; acc.y += v.y; acc.z += v.z; acc.w += v.w;
      }
      if (c >= p.N) continue;
      if constexpr (EPI == SK_GATE) {
        const float4 x = __ldg(reinterpret_cast<const float4*>(p.res + size_t(r) * p.ldres + c));
        float4 y = make_float4(x.x + gate_t * acc.x, x.y + gate_t * acc.y, x.z + gate_t * acc.z, x.w + gate_t * acc.w);
        *reinterpret_cast<float4*>(reinterpret_cast<float*>(p.out) + size_t(r) * p.ldo + c) = y;
      } else {
        if constexpr (EPI == SK_GELU) {
          acc.x = gelu_erf(bf16r(acc.x)); acc.y = gelu_erf(bf16r(acc.y));
          acc.z = gelu_erf(bf16r(acc.z)); acc.w = gelu_erf(bf16r(acc.w));
        }
        uint2 o;
        o.x = pack_bf16(acc.x, acc.y);
        o.y = pack_bf16(acc.z, acc.w);
        *reinterpret_cast<uint2*>(reinterpret_cast<__nv_bfloat16*>(p.out) + size_t(r) * p.ldo + c) = o;
      }
    }
  }
  __syncwarp();
  cluster_sync_all();                   // no CTA retires its shared memory while a peer may still read it
  if (warp == 5) {
    tc_fence_after();
    tmem_dealloc(tmem_base, 256);
  }
}

// ------------------------------------------------------------------ host side
template <int EPI>
static int skinny_launch(const CUtensorMap& tmW, const CUtensorMap& tmX, const SkinnyParams& p, int tiles, cudaStream_t s) {
  auto kern = skinny_gemm_kernel<EPI>;
  const int stage_bytes = SK_W_BYTES + p.Rpad * 128;
  const int part = p.ks * ((p.R + p.ks - 1) / p.ks) * 512;
  const int body = SK_STAGES * stage_bytes > part ? SK_STAGES * stage_bytes : part;
  const int smem = body + 1024 + 128;
  if (int e = set_max_smem_once(kern, 227 * 1024)) return e;
  cudaLaunchConfig_t cfg = {};
  cfg.gridDim = dim3(tiles * p.ks);
  cfg.blockDim = dim3(SK_THREADS);
  cfg.dynamicSmemBytes = smem;
  cfg.stream = s;
  cudaLaunchAttribute at[2];
  at[0].id = cudaLaunchAttributeClusterDimension;
  at[0].val.clusterDim.x = p.ks;
  at[0].val.clusterDim.y = 1;
  at[0].val.clusterDim.z = 1;
  at[1].id = cudaLaunchAttributeProgrammaticStreamSerialization;
  at[1].val.programmaticStreamSerializationAllowed = pdl_enabled() ? 1 : 0;
  cfg.attrs = at;
  cfg.numAttrs = 2;
  const cudaError_t le = cudaLaunchKernelEx(&cfg, kern, tmW, tmX, p);
  count_launch();
  return le == cudaSuccess && cudaGetLastError() == cudaSuccess ? 0 : GXA_ERR_CUDA;
}

bool skinny_supported(int R, int N, int K) {
  return R >= 1 && R <= 256 && N >= 64 && N % 4 == 0 && K >= 128 && K % 64 == 0;
}

// X: bf16 [R, K] (ldx), W: bf16 [N, K] (ldw); epi: SkinnyEpi.  out: bf16 [R, N] (SK_BF16 / SK_GELU) or f32 (SK_GATE).
int skinny_linear(const void* X, int ldx, const void* W, int ldw, int R, int N, int K, int epi, void* out, int ldo,
                  const float* bias, const float* res, int ldres, const float* gate, cudaStream_t s) {
  if (!skinny_supported(R, N, K)) return GXA_ERR_SHAPE;
  if (epi == SK_GATE && (!res || !gate)) return GXA_ERR_ARG;
  SkinnyParams p;
  p.R = R; p.Rpad = (R + 31) / 32 * 32; p.N = N; p.K = K;
  const int kb_total = K / GEMM_BK;
  int ks = kb_total / 2;
  if (ks > 8) ks = 8;
  if (ks < 1) ks = 1;
  const int tiles_n = (N + 127) / 128;
  while (ks > 1 && tiles_n * ks > 120) --ks;        // one wave: clusters cannot use every SM of a GPC, keep some slack
  p.kb_per = (kb_total + ks - 1) / ks;
  p.ks = (kb_total + p.kb_per - 1) / p.kb_per;                        // no empty split
  p.out = out; p.ldo = ldo; p.bias = bias; p.res = res; p.ldres = ldres; p.gate = gate;
  CUtensorMap tmW, tmX;
  if (int e = make_tmap_bf16(&tmW, W, K, N, ldw, GEMM_BK, 128)) return e;
  if (int e = make_tmap_bf16(&tmX, X, K, R, ldx, GEMM_BK, p.Rpad)) return e;
  const int tiles = (N + 127) / 128;
  switch (epi) {
    case SK_BF16: return skinny_launch<SK_BF16>(tmW, tmX, p, tiles, s);
    case SK_GELU: return skinny_launch<SK_GELU>(tmW, tmX, p, tiles, s);
    case SK_GATE: return skinny_launch<SK_GATE>(tmW, tmX, p, tiles, s);
  }
  return GXA_ERR_UNSUPPORTED;
}

}  // namespace gxa
