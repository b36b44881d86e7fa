// Microbenchmark: issue rate of tcgen05.mma (cta_group::1, bf16, M = 128, K = 16 per instruction) for the operand
// sources / majors the attention kernels use.  One CTA per SM, one issuing warp, operands resident in shared memory,
// 512 back-to-back MMAs per variant; prints cycles per MMA and the shared-memory operand bytes per cycle.
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -std=c++17 -I mocov2_whisper_flamingo_b200/csrc tools/mma_rate.cu -o gpurun_out/mma_rate -lcuda
#include <cstdio>
#include <cstdint>
#include <cuda.h>
#include <cuda_runtime.h>
#include "gxa_ptx.cuh"
using namespace gxa;

constexpr int TILE = 128 * 64 * 2;
struct Variant { const char* name; int N; int a_src; /*0 K-major smem, 1 MN-major smem, 2 TMEM*/ int b_mn; };
__constant__ int dummy;

template <int N, int A_SRC, int B_MN>
__global__ void __launch_bounds__(64, 1) rate_kernel(long long* out, int reps) {
  extern __shared__ __align__(1024) uint8_t smem[];
  __shared__ uint64_t bar;
  __shared__ uint32_t slot;
  const int warp = threadIdx.x >> 5;
  for (int i = threadIdx.x; i < 6 * TILE / 4; i += 64) reinterpret_cast<uint32_t*>(smem)[i] = 0x3c003c00u;
  if (threadIdx.x == 0) { mbar_init(&bar, 1); mbar_fence_init(); }
  if (warp == 1) tmem_alloc(&slot, 512);
  fence_proxy_async_smem();
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tm = slot;
  if (warp == 0) {
    constexpr uint32_t HI = umma_desc_hi(1024);
    constexpr uint32_t idesc = umma_idesc_bf16(128, N, A_SRC == 1, B_MN != 0);
    const uint32_t el = elect_one_pred();
    const uint32_t a_k = umma_desc_lo(smem_u32(smem), 16), a_mn = umma_desc_lo(smem_u32(smem), TILE);
    const uint32_t b_k = umma_desc_lo(smem_u32(smem) + 2 * TILE, 16), b_mn = umma_desc_lo(smem_u32(smem) + 2 * TILE, 8192);
    const long long t0 = clock64();
    for (int r = 0; r < reps; ++r) {
#pragma unroll
      for (int k = 0; k < 8; ++k) {
        // 8 k-steps over a 128-deep contraction, stepping the descriptors the way the attention kernels do
        const uint32_t bd = B_MN ? b_mn + k * 128 : b_k + (k >> 2) * (TILE >> 4) + (k & 3) * 2;
        if (A_SRC == 2) umma_bf16_ts_e(tm, tm + 256 + k * 8, bd, HI, idesc, 1u, el);
        else if (A_SRC == 1) umma_bf16_e(tm, a_mn + k * 128, HI, bd, HI, idesc, 1u, el);
        else umma_bf16_e(tm, a_k + (k >> 2) * (TILE >> 4) + (k & 3) * 2, HI, bd, HI, idesc, 1u, el);
      }
    }
    umma_commit_e(&bar, el);
    const long long t1 = clock64();
    mbar_wait(&bar, 0);
    const long long t2 = clock64();
    if (threadIdx.x == 0 && blockIdx.x == 0) { out[0] = t1 - t0; out[1] = t2 - t0; }
  }
  tc_fence_before();
  __syncthreads();
  if (warp == 1) { tc_fence_after(); tmem_dealloc(tm, 512); }
}

template <int N, int A_SRC, int B_MN>
void run(const char* name, long long* d_out) {
  const int reps = 64, smem = 6 * TILE + 1024;
  cudaFuncSetAttribute(rate_kernel<N, A_SRC, B_MN>, cudaFuncAttributeMaxDynamicSharedMemorySize, smem);
  long long h[2];
  for (int it = 0; it < 2; ++it) {
    rate_kernel<N, A_SRC, B_MN><<<148, 64, smem>>>(d_out, reps);
    if (cudaDeviceSynchronize() != cudaSuccess) { printf("%s: launch failed: %s\n", name, cudaGetErrorString(cudaGetLastError())); return; }
  }
  cudaMemcpy(h, d_out, sizeof(h), cudaMemcpyDeviceToHost);
  const double n = reps * 8.0, cyc = double(h[1]) / n;
  const double bytes = (A_SRC == 2 ? 0 : 128 * 16 * 2) + N * 16 * 2;
  printf("%-34s N=%3d  issue %6.1f clk/MMA  complete %6.1f clk/MMA  flops/clk %7.0f  smem operand B/clk %6.1f\n", name, N,
         double(h[0]) / n, cyc, 2.0 * 128 * N * 16 / cyc, bytes / cyc);
}

int main() {
  long long* d_out;
  cudaMalloc(&d_out, 16);
  run<128, 0, 0>("A K-major,  B K-major  (S, dP)", d_out);
  run<256, 0, 0>("A K-major,  B K-major", d_out);
  run<64, 0, 0>("A K-major,  B K-major", d_out);
  run<64, 0, 1>("A K-major,  B MN-major (dQ, PV v1)", d_out);
  run<64, 1, 1>("A MN-major, B MN-major (dV, dK)", d_out);
  run<64, 1, 0>("A MN-major, B K-major", d_out);
  run<64, 2, 1>("A TMEM,     B MN-major (PV v2)", d_out);
  run<64, 2, 0>("A TMEM,     B K-major", d_out);
  run<128, 2, 0>("A TMEM,     B K-major", d_out);
  run<128, 1, 1>("A MN-major, B MN-major", d_out);
  run<128, 0, 1>("A K-major,  B MN-major", d_out);
  return 0;
}
