// Host-side declarations shared by the .cu translation units (internal; the public C ABI is include/gxa.h).
#pragma once
#include <cuda.h>
#include <cuda_bf16.h>
#include <cuda_runtime.h>
#include <stdint.h>

#include "../../include/gxa.h"
#include "gxa_rng.cuh"

namespace gxa {

typedef CUresult (*PFN_cuTensorMapEncodeTiled_t)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*,
                                                 const cuuint64_t*, const cuuint32_t*, const cuuint32_t*,
                                                 CUtensorMapInterleave, CUtensorMapSwizzle, CUtensorMapL2promotion,
                                                 CUtensorMapFloatOOBfill);

int num_sms();
void count_launch();
void set_error_text(const char* msg);      // text returned by gxa_last_cuda_error() for the next failure
// cudaFuncAttributeMaxDynamicSharedMemorySize is per (kernel, device): set once for each pair (thread-safe)
int set_max_smem_once_impl(const void* kernel, int bytes);
template <typename K>
inline int set_max_smem_once(K kernel, int bytes) { return set_max_smem_once_impl(reinterpret_cast<const void*>(kernel), bytes); }

// Programmatic dependent launch: the kernel may be scheduled while its predecessor on the stream is still running, so
// its prologue (barrier init, TMEM allocation, tensor-map prefetch) and the launch latency overlap the predecessor's
// tail.  Every kernel launched through here executes pdl_wait() (gxa_ptx.cuh) before its first global-memory access.
// GXA_PDL=0 in the environment turns the attribute off (plain stream order) for A/B runs.
bool pdl_enabled();
#ifdef __CUDACC__
// device side: let the next kernel on the stream start its prologue / wait for every earlier kernel's memory
__device__ __forceinline__ void pdl_launch_dependents() { asm volatile("griddepcontrol.launch_dependents;" ::: "memory"); }
__device__ __forceinline__ void pdl_wait() { asm volatile("griddepcontrol.wait;" ::: "memory"); }
#endif
template <typename... KArgs, typename... Args>
inline cudaError_t launch_pdl(void (*kernel)(KArgs...), dim3 grid, dim3 block, size_t smem, cudaStream_t s, Args&&... args) {
  cudaLaunchConfig_t cfg = {};
  cfg.gridDim = grid;
  cfg.blockDim = block;
  cfg.dynamicSmemBytes = smem;
  cfg.stream = s;
  cudaLaunchAttribute at[1];
  at[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
  at[0].val.programmaticStreamSerializationAllowed = pdl_enabled() ? 1 : 0;
  cfg.attrs = at;
  cfg.numAttrs = 1;
  return cudaLaunchKernelEx(&cfg, kernel, static_cast<KArgs>(args)...);
}

// C[M,N] = A * B^T with the fused epilogue `epi` (see gxa_gemm.cuh).
struct GemmOp {
  const void* A = nullptr; int lda = 0; bool a_mn = false;   // a_mn: A stored [K, M] (M contiguous)
  const void* B = nullptr; int ldb = 0; bool b_mn = false;   // b_mn: B stored [K, N] (N contiguous)
  int M = 0, N = 0, K = 0;
  int epi = 0;
  void* C = nullptr; int ldc = 0;
  void* C2 = nullptr; int ldc2 = 0;
  const float* bias = nullptr;
  const void* aux = nullptr; int ld_aux = 0;
  const float* gate = nullptr;
  int accumulate = 0;
  DropCfg drop = DropCfg{0, 1.f, 0, 0};   // EPI_GATE_RES only: dropout on the linear output
  float* colsum = nullptr;   // EPI_GELU_BWD only: += column sums of the output (fused bias gradient)
  int bn = 0;        // 0 = choose
  int k_splits = 0;  // 0 = choose
};
int gemm_launch(const GemmOp& op, cudaStream_t stream);
int make_tmap_bf16(CUtensorMap* m, const void* base, uint64_t inner, uint64_t outer, uint64_t ld, uint32_t box_inner,
                   uint32_t box_outer);
int make_tmap_bf16_3d(CUtensorMap* m, const void* base, uint64_t d0, uint64_t d1, uint64_t d2, uint64_t ld1,
                      uint64_t ld2, uint32_t b0, uint32_t b1);

// ---- row-wise / elementwise kernels (gxa_rowwise.cu)
int ln_fwd(const float* x, const float* w, const float* b, void* out, int out_is_f32, float* mean, float* rstd, int rows,
           int D, cudaStream_t s, const float* add = nullptr, int add_period = 0);   // optional fused + add[row % period]
int sum3_f32(const float* a, const float* b, const float* c, float* out, size_t n, cudaStream_t s);
// dx = dres + LNbwd(g; x, w);  dw += sum g*xhat;  db += sum g;  optional gated output:
//   dgate_accum += sum(dx * o),  dout(bf16 or f32) = tanh(gate) * dx,  dbias_out += colsum(dout)
int ln_bwd(const void* g, int g_is_f32, const float* x, const float* w, const float* mean, const float* rstd,
           const float* dres, float* dx, float* dw, float* db, const void* o, int o_is_f32, const float* gate,
           float* dgate_accum, void* dout, float* dbias_out, int rows, int D, cudaStream_t s,
           float* dgate_out = nullptr, unsigned* done_counter = nullptr);   // fused gate_finish (last block; counter zeroed by the caller)
// dout = tanh(gate) * dy ; dgate_accum += sum(dy * f) ; dbias += colsum(dout)
int gate_bwd(const float* dy, const void* f, int f_is_f32, const float* gate, float* dgate_accum, void* dout,
             int out_is_f32, float* dbias, int rows, int D, cudaStream_t s, DropCfg drop = DropCfg{0, 1.f, 0, 0},
             float* dgate_out = nullptr, unsigned* done_counter = nullptr);   // fused gate_finish, as in ln_bwd
int gate_finish(const float* gate, const float* accum, float* dgate, cudaStream_t s);   // dgate = (1-tanh^2) * accum
int colsum(const void* src, int is_f32, int ld, float* dst, int rows, int cols, cudaStream_t s);   // dst += colsum
int cast_f32_bf16(const float* src, void* dst, size_t n, cudaStream_t s);
int cast_f32_bf16_multi(const float* const* src, void* const* dst, const size_t* n, int count, cudaStream_t s);   // <= 4
int cast_bf16_f32(const void* src, float* dst, size_t n, cudaStream_t s);
// fp32 -> bf16 hi/lo split for the 3xbf16 fp32-accurate GEMM mode.  pattern 0: [hi|lo|hi], 1: [hi|hi|lo].
// along_rows = 0: dst is [rows, 3*cols] (segments side by side); 1: dst is [3*rows, cols] (segments stacked).
int split3(const float* src, int ld, void* dst, int rows, int cols, int pattern, int along_rows, cudaStream_t s);
int mask_kv_end(const uint8_t* mask, int* kv_end, int B, int Lk, cudaStream_t s);
int gelu_f32(const float* u, float* a, size_t n, cudaStream_t s);
int gelu_bwd_f32(const float* da, const float* u, float* du, size_t n, cudaStream_t s);
// y = x + tanh(gate) * o;  with dropout (nn.Dropout after ff.2, fp32 mode): o <- dropout(o) in place first (cols = row width)
int gate_res_f32(const float* x, float* o, const float* gate, float* y, size_t n, cudaStream_t s, int cols = 0,
                 DropCfg drop = DropCfg{0, 1.f, 0, 0});
int add_inplace_f32(float* dst, const float* src, size_t n, cudaStream_t s);

// ---- attention (gxa_attn.cu: bf16 tensor-core path; gxa_attn_f32.cu: fp32 accuracy path)
int attn_fwd_bf16(const void* q, int ldq, const void* k, const void* v, int ldkv, const uint8_t* mask, const int* kv_end,
                  void* ctx, int ldc, float* lse, int B, int H, int Lq, int Lk, cudaStream_t s,
                  DropCfg drop = DropCfg{0, 1.f, 0, 0});
int attn_bwd_bf16(const void* q, int ldq, const void* k, const void* v, int ldkv, const uint8_t* mask, const int* kv_end,
                  const void* ctx, const void* dctx, int ldc, const float* lse, float* delta, void* dq, int lddq,
                  void* dk, void* dv, int lddkv, void* workspace, int B, int H, int Lq, int Lk, cudaStream_t s,
                  float* dq_colsum = nullptr,    // optional: += column sums of dq (fused bias gradient d bq)
                  DropCfg drop = DropCfg{0, 1.f, 0, 0},
                  float* dv_colsum = nullptr,    // optional: += column sums of dv (fused bias gradient d bv)
                  bool acc_is_zero = false);     // the fp32 dQ accumulator in `workspace` was zeroed by the caller
size_t attn_bwd_workspace_bytes(int B, int H, int Lq);
int attn_fwd_f32(const float* q, int ldq, const float* k, const float* v, int ldkv, const uint8_t* mask, float* ctx,
                 int ldc, float* P, int B, int H, int Lq, int Lk, cudaStream_t s, DropCfg drop = DropCfg{0, 1.f, 0, 0});
int attn_bwd_f32(const float* q, int ldq, const float* k, const float* v, int ldkv, const float* dctx, int ldc,
                 const float* P, float* dS, float* dq, int lddq, float* dk, float* dv, int lddkv, int B, int H, int Lq,
                 int Lk, cudaStream_t s, DropCfg drop = DropCfg{0, 1.f, 0, 0});

}  // namespace gxa
