// fp32 accuracy path of the masked-softmax attention (GXA_PREC_FP32 only; small parity shapes such as
// BASELINE config 1).  Plain FFMA kernels, one warp per query row (or per key row in backward), the
// full probability matrix P [B,H,Lq,Lk] is materialised.  Not a performance path.
#include "gxa_host.h"

namespace gxa {

constexpr int HD = 64;

__device__ __forceinline__ float wsum(float v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
  return v;
}
__device__ __forceinline__ float wmax(float v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v = fmaxf(v, __shfl_xor_sync(0xffffffffu, v, o));
  return v;
}
__device__ __forceinline__ float dot64(const float* a /*smem*/, const float* __restrict__ g) {
  float s = 0.f;
#pragma unroll
  for (int d = 0; d < HD; d += 4) {
    float4 kv = *reinterpret_cast<const float4*>(g + d);
    s += a[d] * kv.x + a[d + 1] * kv.y + a[d + 2] * kv.z + a[d + 3] * kv.w;
  }
  return s;
}

// dropout multiplier (0 or 1/(1-p)) of probability (row, key): the same function of (seed, element) as the bf16 kernels
__device__ __forceinline__ float drop_mult(const DropCfg& c, uint32_t drop_row, int key) {
  if (c.thresh == 0) return 1.f;
  return drop_keep_scalar(c, drop_row + (static_cast<uint32_t>(key) >> 2), key & 3) ? c.scale : 0.f;
}

__global__ void __launch_bounds__(128)
attn_fwd_f32_kernel(const float* __restrict__ q, int ldq, const float* __restrict__ k, const float* __restrict__ v,
                    int ldkv, const uint8_t* __restrict__ mask, float* __restrict__ ctx, int ldc, float* __restrict__ P,
                    int B, int H, int Lq, int Lk, const DropCfg drop) {
  __shared__ float sq[4][HD];
  const int w = threadIdx.x >> 5, l = threadIdx.x & 31;
  const long long row = (long long)blockIdx.x * 4 + w;   // (b, h, qi)
  if (row >= (long long)B * H * Lq) return;
  const int qi = int(row % Lq), h = int((row / Lq) % H), b = int(row / ((long long)Lq * H));
  const float* qr = q + (size_t(b) * Lq + qi) * ldq + h * HD;
  sq[w][2 * l] = qr[2 * l];
  sq[w][2 * l + 1] = qr[2 * l + 1];
  __syncwarp();
  const float* kb = k + size_t(b) * Lk * ldkv + h * HD;
  const float* vb = v + size_t(b) * Lk * ldkv + h * HD;
  float* pr = P + size_t(row) * Lk;
  float mx = -INFINITY;
  for (int j = l; j < Lk; j += 32) {
    float s = (mask && mask[size_t(b) * Lk + j]) ? -INFINITY : 0.125f * dot64(sq[w], kb + size_t(j) * ldkv);
    pr[j] = s;
    mx = fmaxf(mx, s);
  }
  mx = wmax(mx);
  const float mu = (mx == -INFINITY) ? 0.f : mx;
  float sum = 0.f;
  for (int j = l; j < Lk; j += 32) {
    float e = expf(pr[j] - mu);
    pr[j] = e;
    sum += e;
  }
  sum = wsum(sum);
  const float inv = sum > 0.f ? 1.f / sum : 0.f;
  for (int j = l; j < Lk; j += 32) pr[j] *= inv;
  __syncwarp();
  float a0 = 0.f, a1 = 0.f;
  const uint32_t drop_row = static_cast<uint32_t>(row) * static_cast<uint32_t>((Lk + 3) >> 2);   // gxa_rng.cuh counter
  for (int j = 0; j < Lk; ++j) {
    const float p = pr[j] * drop_mult(drop, drop_row, j);       // P stays undropped in memory (the backward needs both)
    float2 vv = *reinterpret_cast<const float2*>(vb + size_t(j) * ldkv + 2 * l);
    a0 += p * vv.x;
    a1 += p * vv.y;
  }
  *reinterpret_cast<float2*>(ctx + (size_t(b) * Lq + qi) * ldc + h * HD + 2 * l) = make_float2(a0, a1);
}

// per query row: dS = P * (dP - delta), dq = 0.125 * dS K
__global__ void __launch_bounds__(128)
attn_bwd_f32_q_kernel(const float* __restrict__ k, const float* __restrict__ v, int ldkv,
                      const float* __restrict__ dctx, int ldc, const float* __restrict__ P, float* __restrict__ dS,
                      float* __restrict__ dq, int lddq, int B, int H, int Lq, int Lk, const DropCfg drop) {
  __shared__ float sd[4][HD];
  const int w = threadIdx.x >> 5, l = threadIdx.x & 31;
  const long long row = (long long)blockIdx.x * 4 + w;
  if (row >= (long long)B * H * Lq) return;
  const int qi = int(row % Lq), h = int((row / Lq) % H), b = int(row / ((long long)Lq * H));
  const float* dr = dctx + (size_t(b) * Lq + qi) * ldc + h * HD;
  sd[w][2 * l] = dr[2 * l];
  sd[w][2 * l + 1] = dr[2 * l + 1];
  __syncwarp();
  const float* kb = k + size_t(b) * Lk * ldkv + h * HD;
  const float* vb = v + size_t(b) * Lk * ldkv + h * HD;
  const float* pr = P + size_t(row) * Lk;
  float* ds = dS + size_t(row) * Lk;
  float delta = 0.f;
  const uint32_t drop_row = static_cast<uint32_t>(row) * static_cast<uint32_t>((Lk + 3) >> 2);
  for (int j = l; j < Lk; j += 32) {
    const float p = pr[j];
    // forward: ctx = (m * P) V  =>  dP = m * (dO V^T);  delta = rowsum(P * dP) = rowsum(dO * ctx)
    const float dp = p != 0.f ? drop_mult(drop, drop_row, j) * dot64(sd[w], vb + size_t(j) * ldkv) : 0.f;
    ds[j] = dp;
    delta += p * dp;
  }
  delta = wsum(delta);
  for (int j = l; j < Lk; j += 32) ds[j] = pr[j] * (ds[j] - delta);
  __syncwarp();
  float a0 = 0.f, a1 = 0.f;
  for (int j = 0; j < Lk; ++j) {
    const float g = ds[j];
    float2 kk = *reinterpret_cast<const float2*>(kb + size_t(j) * ldkv + 2 * l);
    a0 += g * kk.x;
    a1 += g * kk.y;
  }
  *reinterpret_cast<float2*>(dq + (size_t(b) * Lq + qi) * lddq + h * HD + 2 * l) = make_float2(0.125f * a0, 0.125f * a1);
}

// per key row: dk = 0.125 * dS^T Q, dv = P^T dctx
__global__ void __launch_bounds__(128)
attn_bwd_f32_k_kernel(const float* __restrict__ q, int ldq, const float* __restrict__ dctx, int ldc,
                      const float* __restrict__ P, const float* __restrict__ dS, float* __restrict__ dk,
                      float* __restrict__ dv, int lddkv, int B, int H, int Lq, int Lk, const DropCfg drop) {
  const int w = threadIdx.x >> 5, l = threadIdx.x & 31;
  const long long row = (long long)blockIdx.x * 4 + w;   // (b, h, j)
  if (row >= (long long)B * H * Lk) return;
  const int j = int(row % Lk), h = int((row / Lk) % H), b = int(row / ((long long)Lk * H));
  const float* qb = q + size_t(b) * Lq * ldq + h * HD + 2 * l;
  const float* db = dctx + size_t(b) * Lq * ldc + h * HD + 2 * l;
  const float* pb = P + (size_t(b) * H + h) * Lq * Lk + j;
  const float* sb = dS + (size_t(b) * H + h) * Lq * Lk + j;
  float k0 = 0.f, k1 = 0.f, v0 = 0.f, v1 = 0.f;
  const uint32_t half = static_cast<uint32_t>((Lk + 3) >> 2);
  const uint32_t drop_row0 = static_cast<uint32_t>((size_t(b) * H + h) * Lq) * half;
  for (int i = 0; i < Lq; ++i) {
    const float p = pb[size_t(i) * Lk] * drop_mult(drop, drop_row0 + static_cast<uint32_t>(i) * half, j);   // dV sees m * P
    const float g = sb[size_t(i) * Lk];
    float2 qq = *reinterpret_cast<const float2*>(qb + size_t(i) * ldq);
    float2 dd = *reinterpret_cast<const float2*>(db + size_t(i) * ldc);
    k0 += g * qq.x; k1 += g * qq.y;
    v0 += p * dd.x; v1 += p * dd.y;
  }
  const size_t o = (size_t(b) * Lk + j) * lddkv + h * HD + 2 * l;
  *reinterpret_cast<float2*>(dk + o) = make_float2(0.125f * k0, 0.125f * k1);
  *reinterpret_cast<float2*>(dv + o) = make_float2(v0, v1);
}

static inline int chk32() { count_launch(); return cudaGetLastError() == cudaSuccess ? 0 : GXA_ERR_CUDA; }

int attn_fwd_f32(const float* q, int ldq, const float* k, const float* v, int ldkv, const uint8_t* mask, float* ctx,
                 int ldc, float* P, int B, int H, int Lq, int Lk, cudaStream_t s, DropCfg drop) {
  const long long rows = (long long)B * H * Lq;
  attn_fwd_f32_kernel<<<(unsigned)((rows + 3) / 4), 128, 0, s>>>(q, ldq, k, v, ldkv, mask, ctx, ldc, P, B, H, Lq, Lk, drop);
  return chk32();
}
int attn_bwd_f32(const float* q, int ldq, const float* k, const float* v, int ldkv, const float* dctx, int ldc,
                 const float* P, float* dS, float* dq, int lddq, float* dk, float* dv, int lddkv, int B, int H, int Lq,
                 int Lk, cudaStream_t s, DropCfg drop) {
  const long long rows = (long long)B * H * Lq;
  attn_bwd_f32_q_kernel<<<(unsigned)((rows + 3) / 4), 128, 0, s>>>(k, v, ldkv, dctx, ldc, P, dS, dq, lddq, B, H, Lq, Lk, drop);
  if (int e = chk32()) return e;
  const long long krows = (long long)B * H * Lk;
  attn_bwd_f32_k_kernel<<<(unsigned)((krows + 3) / 4), 128, 0, s>>>(q, ldq, dctx, ldc, P, dS, dk, dv, lddkv, B, H, Lq, Lk, drop);
  return chk32();
}

}  // namespace gxa
