// Masked-softmax attention over variable-length keys, head_dim 64, bf16 in / fp32 softmax: host-side dispatch to the
// tcgen05 / TMEM kernels (gxa_attn_tc.cu forward, gxa_attn_bwd_tc.cu persistent backward) and the small row-wise
// delta = rowsum(dO * O) kernel of the backward.
//
// There is ONE bf16 attention path.  Layouts it does not cover (K and V not packed side by side in one [B*Lk, 2D]
// matrix -- the K|V projection's natural output -- or more than 1024 keys in the forward) return GXA_ERR_UNSUPPORTED
// with a message in gxa_last_cuda_error(); there is no slower fallback to drop into silently.  (The round-1 mma.sync
// kernels that used to sit behind those cases are gone.)
#include <cstdio>
#include <cstdlib>

#include "gxa_host.h"

namespace gxa {

constexpr int AT = 64;            // head_dim

// ================================================================== backward: delta = rowsum(dctx * ctx) per head
// warp per query row; a lane owns 16 bytes (8 bf16) of a head, 8 lanes per head, 4 heads per pass: every load of the
// row is issued before the first reduction (H / 4 passes x 2 tensors in flight)
template <int HB>   // H <= 4 * HB heads
__global__ void __launch_bounds__(256)
attn_delta_kernel(const __nv_bfloat16* __restrict__ ctx, const __nv_bfloat16* __restrict__ dctx, int ldc,
                  float* __restrict__ delta, int B, int H, int Lq) {
  pdl_launch_dependents();
  pdl_wait();
  const int row = blockIdx.x * 8 + (threadIdx.x >> 5);
  const int l = threadIdx.x & 31;
  if (row >= B * Lq) return;
  const int b = row / Lq, qi = row % Lq;
  uint4 a[HB], d[HB];
#pragma unroll
  for (int i = 0; i < HB; ++i) {
    const int h = i * 4 + (l >> 3);
    a[i] = d[i] = make_uint4(0, 0, 0, 0);
    if (h < H) {
      const size_t off = size_t(row) * ldc + h * AT + (l & 7) * 8;
      a[i] = __ldg(reinterpret_cast<const uint4*>(ctx + off));
      d[i] = __ldg(reinterpret_cast<const uint4*>(dctx + off));
    }
  }
#pragma unroll
  for (int i = 0; i < HB; ++i) {
    const uint32_t* aw = reinterpret_cast<const uint32_t*>(&a[i]);
    const uint32_t* dw = reinterpret_cast<const uint32_t*>(&d[i]);
    float s = 0.f;
#pragma unroll
    for (int j = 0; j < 4; ++j) {
      s += __uint_as_float(aw[j] << 16) * __uint_as_float(dw[j] << 16);
      s += __uint_as_float(aw[j] & 0xffff0000u) * __uint_as_float(dw[j] & 0xffff0000u);
    }
    s += __shfl_xor_sync(0xffffffffu, s, 1);
    s += __shfl_xor_sync(0xffffffffu, s, 2);
    s += __shfl_xor_sync(0xffffffffu, s, 4);
    const int h = i * 4 + (l >> 3);
    if ((l & 7) == 0 && h < H) delta[(size_t(b) * H + h) * Lq + qi] = s;
  }
}

// ================================================================== host launchers
static inline int chk() { count_launch(); return cudaGetLastError() == cudaSuccess ? 0 : GXA_ERR_CUDA; }

int attn_fwd_tc(const void* q, int ldq, const void* kv, int ldkv, const uint8_t* mask, const int* kv_end, void* ctx,
                int ldc, float* lse, int B, int H, int Lq, int Lk, cudaStream_t s, DropCfg drop);
int attn_bwd_tc2(const void* q, int ldq, const void* kv, int ldkv, const uint8_t* mask, const int* kv_end,
                 const void* dctx, int ldc, const float* lse, const float* delta, float* dq_acc, void* dq, int lddq,
                 void* dk, void* dv, int lddkv, int B, int H, int Lq, int Lk, float* dq_colsum, cudaStream_t s, DropCfg drop,
                 float* dv_colsum, bool acc_is_zero);
size_t attn_bwd_tc_workspace(int B, int H, int Lq);

size_t attn_bwd_workspace_bytes(int B, int H, int Lq) { return attn_bwd_tc_workspace(B, H, Lq); }

int attn_fwd_bf16(const void* q, int ldq, const void* k, const void* v, int ldkv, const uint8_t* mask, const int* kv_end,
                  void* ctx, int ldc, float* lse, int B, int H, int Lq, int Lk, cudaStream_t s, DropCfg drop) {
  if (B <= 0 || H <= 0 || Lq <= 0 || Lk <= 0) return GXA_ERR_SHAPE;
  if ((ldq % 8) || (ldkv % 8) || (ldc % 8)) return GXA_ERR_ALIGN;
  const bool packed = ldkv == 2 * H * 64 && v == reinterpret_cast<const __nv_bfloat16*>(k) + H * 64 &&
                      (reinterpret_cast<uintptr_t>(ctx) & 15) == 0;
  if (!packed) {
    set_error_text("attention forward: K and V must be the two halves of one [B*Lk, 2*D] matrix (ldkv = 2*D, v = k + D)");
    return GXA_ERR_UNSUPPORTED;
  }
  if (Lk > 1024) {
    set_error_text("attention forward: more than 1024 keys per sample is not supported (the reference caps the visual "
                   "memory at max_frames = 400, config.py:28)");
    return GXA_ERR_UNSUPPORTED;
  }
  return attn_fwd_tc(q, ldq, k, ldkv, mask, kv_end, ctx, ldc, lse, B, H, Lq, Lk, s, drop);
}

int attn_bwd_bf16(const void* q, int ldq, const void* k, const void* v, int ldkv, const uint8_t* mask, const int* kv_end,
                  const void* ctx, const void* dctx, int ldc, const float* lse, float* delta, void* dq, int lddq,
                  void* dk, void* dv, int lddkv, void* workspace, int B, int H, int Lq, int Lk, cudaStream_t s,
                  float* dq_colsum, DropCfg drop, float* dv_colsum, bool acc_is_zero) {
  if (B <= 0 || H <= 0 || Lq <= 0 || Lk <= 0) return GXA_ERR_SHAPE;
  if ((ldq % 8) || (ldkv % 8) || (ldc % 8) || (lddq % 8) || (lddkv % 8)) return GXA_ERR_ALIGN;
  auto K = reinterpret_cast<const __nv_bfloat16*>(k);
  auto V = reinterpret_cast<const __nv_bfloat16*>(v);
  auto dO = reinterpret_cast<const __nv_bfloat16*>(dctx);
  const bool packed = workspace != nullptr && ldkv == 2 * H * 64 && lddkv == ldkv && V == K + H * 64 &&
                      reinterpret_cast<__nv_bfloat16*>(dv) == reinterpret_cast<__nv_bfloat16*>(dk) + H * 64;
  if (!packed) {
    set_error_text("attention backward: K|V and dK|dV must each be the halves of one [B*Lk, 2*D] matrix and the fp32 dQ "
                   "workspace (gxa_attn_bwd_workspace_bytes) must be given");
    return GXA_ERR_UNSUPPORTED;
  }
  if (H > 32 || (reinterpret_cast<uintptr_t>(ctx) & 15) || (reinterpret_cast<uintptr_t>(dctx) & 15)) return GXA_ERR_ALIGN;
  const dim3 grid((B * Lq + 7) / 8), block(256);
  const __nv_bfloat16* O = reinterpret_cast<const __nv_bfloat16*>(ctx);
  if (H <= 8) launch_pdl(attn_delta_kernel<2>, grid, block, 0, s, O, dO, ldc, delta, B, H, Lq);
  else if (H <= 16) launch_pdl(attn_delta_kernel<4>, grid, block, 0, s, O, dO, ldc, delta, B, H, Lq);
  else launch_pdl(attn_delta_kernel<8>, grid, block, 0, s, O, dO, ldc, delta, B, H, Lq);
  if (int e = chk()) return e;
  return attn_bwd_tc2(q, ldq, k, ldkv, mask, kv_end, dctx, ldc, lse, delta, reinterpret_cast<float*>(workspace), dq,
                      lddq, dk, dv, lddkv, B, H, Lq, Lk, dq_colsum, s, drop, dv_colsum, acc_is_zero);
}

}  // namespace gxa
