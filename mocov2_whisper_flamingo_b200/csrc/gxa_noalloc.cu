// gxa_linear_fwd_static: y = x W^T + b (bf16 contraction, fp32 in / out) WITHOUT a caller-provided workspace and
// without any allocation.  No counterpart in the reference; added for callers
// that cannot call cudaMalloc (code under CUDA-graph capture, tight serving loops, harnesses that forbid it).
//
// gxa_linear_fwd needs scratch for the bf16 copies of its two operands (the tensor cores read bf16; the caller's
// tensors are fp32).  Here that scratch is a file-scope device array of a fixed, modest size: every GPU that loads the
// library holds kStaticScratchBytes (8 MiB) until the process exits.  Consequences, all documented in include/gxa.h:
//   * shapes whose operand copies do not fit are rejected (GXA_ERR_SHAPE + message), never silently truncated;
//   * every call on a device shares the one array: calls that overlap on the GPU (different streams, no ordering
//     between them) race on it.  Calls on ONE stream are ordered and safe;
//   * the array keeps the previous call's contents.  Nothing in it has to start from zero: both operand regions are
//     fully overwritten by the casts before the GEMM's tensor maps (which cover exactly rows x in_f and out_f x in_f
//     elements) read them, so there is nothing to clear.
//
// The FIRST call on a device is the set-up: it looks the array's device address up (and, with CUDA's default lazy module
// loading, that is also the moment the driver materialises the 8 MiB array on that device).  Callers that must not
// touch the allocator at all inside a region (capture, a timed loop) make one warm call before it.
//
// In host code `g_static_scratch` is only a host-side stand-in for the symbol, not a device address: the device
// address is taken once per device, with that device current, by cudaGetSymbolAddress below, and that pointer is what
// the kernels are given.
#include <cstdio>
#include <mutex>

#include "gxa_gemm.cuh"
#include "gxa_host.h"

namespace gxa {

constexpr size_t kStaticScratchBytes = size_t(8) << 20;
static __device__ __align__(256) unsigned char g_static_scratch[kStaticScratchBytes];

constexpr int kMaxDevices = 64;
static std::mutex g_scratch_mu;
static void* g_scratch_ptr[kMaxDevices] = {};

// device address of the scratch array on the CURRENT device (looked up once per device)
static int static_scratch(void** out) {
  int dev = -1;
  if (cudaGetDevice(&dev) != cudaSuccess || dev < 0 || dev >= kMaxDevices) {
    set_error_text("gxa_linear_fwd_static: cudaGetDevice failed or device ordinal >= 64");
    return GXA_ERR_CUDA;
  }
  std::lock_guard<std::mutex> lk(g_scratch_mu);
  if (g_scratch_ptr[dev] == nullptr) {
    void* p = nullptr;
    const cudaError_t e = cudaGetSymbolAddress(&p, g_static_scratch);
    if (e != cudaSuccess || p == nullptr) {
      set_error_text(e != cudaSuccess ? cudaGetErrorString(e) : "cudaGetSymbolAddress returned a null address");
      cudaGetLastError();
      return GXA_ERR_CUDA;
    }
    g_scratch_ptr[dev] = p;
  }
  *out = g_scratch_ptr[dev];
  return 0;
}

static size_t operand_a_elems(int rows, int in_f) { return (size_t(rows) * in_f + 127) & ~size_t(127); }
static size_t need_bytes(int rows, int in_f, int out_f) {
  return (operand_a_elems(rows, in_f) + size_t(out_f) * in_f) * sizeof(__nv_bfloat16);
}
static_assert(kStaticScratchBytes == GXA_LINEAR_FWD_STATIC_MAX_BYTES, "include/gxa.h states the limit");

}  // namespace gxa

using namespace gxa;

extern "C" {


int gxa_linear_fwd_static(const float* x, const float* w, const float* b, float* y, int rows, int in_f, int out_f,
                          void* stream) {
  if (!x || !w || !y) return GXA_ERR_ARG;
  if (rows <= 0 || in_f <= 0 || out_f <= 0 || in_f % 64 || out_f % 64) return GXA_ERR_SHAPE;
  const size_t need = need_bytes(rows, in_f, out_f);
  if (need > kStaticScratchBytes) {
    char msg[200];
    snprintf(msg, sizeof(msg),
             "gxa_linear_fwd_static: rows=%d in=%d out=%d needs %zu bytes of scratch, the built-in array holds %zu "
             "(use gxa_linear_fwd with a caller workspace)", rows, in_f, out_f, need, size_t(kStaticScratchBytes));
    set_error_text(msg);
    return GXA_ERR_SHAPE;
  }
  void* base = nullptr;
  if (int e = static_scratch(&base)) return e;
  cudaStream_t s = reinterpret_cast<cudaStream_t>(stream);
  __nv_bfloat16* A = reinterpret_cast<__nv_bfloat16*>(base);
  __nv_bfloat16* Bm = A + operand_a_elems(rows, in_f);
  if (int e = cast_f32_bf16(x, A, size_t(rows) * in_f, s)) return e;
  if (int e = cast_f32_bf16(w, Bm, size_t(out_f) * in_f, s)) return e;
  GemmOp op;
  op.A = A; op.lda = in_f; op.B = Bm; op.ldb = in_f;
  op.M = rows; op.N = out_f; op.K = in_f;
  op.epi = EPI_F32; op.C = y; op.ldc = out_f; op.bias = b;
  return gemm_launch(op, s);
}

}  // extern "C"
