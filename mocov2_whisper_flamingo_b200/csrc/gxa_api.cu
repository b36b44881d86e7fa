// C ABI (include/gxa.h): block forward/backward orchestration over the kernels in gxa_gemm.cu,
// gxa_rowwise.cu, gxa_attn.cu and gxa_attn_f32.cu.  One call per layer and direction keeps the host
// cost per step at a handful of microseconds per launch (no per-kernel Python/ctypes round trips).
#include <cstdlib>
#include <atomic>
#include <cstring>
#include <map>
#include <mutex>
#include <utility>

#include "gxa_gemm.cuh"
#include "gxa_host.h"

namespace gxa {

static std::atomic<long long> g_launches{0};
static std::atomic<int> g_sms_dev[16];     // SM count per device ordinal (0 = not queried yet)
static char g_err[256] = "";
static std::mutex g_init_mu;               // guards the lazily created per-device state below

int set_max_smem_once_impl(const void* kernel, int bytes) {
  static std::map<std::pair<const void*, int>, int> done;
  int dev = 0;
  if (cudaGetDevice(&dev) != cudaSuccess) return GXA_ERR_CUDA;
  std::lock_guard<std::mutex> lk(g_init_mu);
  auto key = std::make_pair(kernel, dev);
  auto it = done.find(key);
  if (it != done.end() && it->second >= bytes) return 0;
  if (cudaFuncSetAttribute(kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, bytes) != cudaSuccess) return GXA_ERR_CUDA;
  done[key] = bytes;
  return 0;
}

void count_launch() { g_launches.fetch_add(1, std::memory_order_relaxed); }
void set_error_text(const char* msg) { strncpy(g_err, msg, sizeof(g_err) - 1); g_err[sizeof(g_err) - 1] = 0; }
// Measured at c3 (D = 1024, 24 layers, one B200): staging the next kernel early hides ~1.2 ms of launch gaps per step
// whatever the batch, but costs ~0.7% of the kernel time -> a gain below ~40 samples x 448 rows per GPU (-7% at 8, -4.7%
// at 16, -0.6% at 32) and a loss above (+0.7% at 64).  GXA_PDL=0 / 1 forces it off / on.
static thread_local bool g_pdl_small = true;
static void pdl_hint_rows(long long rows) { g_pdl_small = rows <= 16384; }
bool pdl_enabled() {
  static const int force = [] { const char* e = getenv("GXA_PDL"); return e ? (e[0] == '0' ? 0 : 1) : -1; }();
  return force >= 0 ? force == 1 : g_pdl_small;
}
static std::atomic<int> g_sm_limit{0};   // gxa_set_sm_limit: SMs the persistent grids may assume (0 = all)
int num_sms() {
  int dev = 0;
  if (cudaGetDevice(&dev) != cudaSuccess || dev < 0 || dev >= 16) dev = 0;
  int sms = g_sms_dev[dev].load(std::memory_order_relaxed);
  if (sms == 0) {
    int n = 0;
    sms = (cudaDeviceGetAttribute(&n, cudaDevAttrMultiProcessorCount, dev) == cudaSuccess && n > 0) ? n : 148;
    g_sms_dev[dev].store(sms, std::memory_order_relaxed);
  }
  const int lim = g_sm_limit.load(std::memory_order_relaxed);
  return lim > 0 && lim < sms ? lim : sms;
}
static int fail(int code) {
  cudaError_t e = cudaPeekAtLastError();
  if (e != cudaSuccess) {
    strncpy(g_err, cudaGetErrorString(e), sizeof(g_err) - 1);
    cudaGetLastError();
  }
  return code;
}
#define GXA_TRY(expr)                 \
  do {                                \
    int _e = (expr);                  \
    if (_e != 0) return fail(_e);     \
  } while (0)

// ---------------------------------------------------------------- buffer carving
struct Carver {
  uint8_t* base;
  size_t off = 0;
  explicit Carver(void* b) : base(reinterpret_cast<uint8_t*>(b)) {}
  template <typename T>
  T* take(size_t n) {
    T* p = base ? reinterpret_cast<T*>(base + off) : nullptr;
    off += (n * sizeof(T) + 255) & ~size_t(255);
    return p;
  }
};
typedef __nv_bfloat16 bf16;

struct SavedBf16 {
  float *mean1, *rstd1, *mean2, *rstd2, *lse, *x1;
  bf16 *h, *q, *kv, *ctx, *o, *h2, *u, *a, *f;
  size_t bytes;
  SavedBf16(const gxa_dims& d, void* base) {
    const size_t M = size_t(d.B) * d.Lq, Mk = size_t(d.B) * d.Lk, D = d.D;
    Carver c(base);
    mean1 = c.take<float>(M); rstd1 = c.take<float>(M); mean2 = c.take<float>(M); rstd2 = c.take<float>(M);
    lse = c.take<float>(size_t(d.B) * d.H * d.Lq);
    x1 = c.take<float>(M * D);
    h = c.take<bf16>(M * D); q = c.take<bf16>(M * D); kv = c.take<bf16>(Mk * 2 * D); ctx = c.take<bf16>(M * D);
    o = c.take<bf16>(M * D); h2 = c.take<bf16>(M * D); u = c.take<bf16>(M * 4 * D); a = c.take<bf16>(M * 4 * D);
    f = c.take<bf16>(M * D);
    bytes = c.off;
  }
};
struct WorkBf16 {
  float *gacc, *dx1, *delta;
  bf16 *df, *du, *dh2, *dob, *dctx, *dq, *dkv, *dh;
  float* dqacc;
  size_t bytes;
  WorkBf16(const gxa_dims& d, void* base) {
    const size_t M = size_t(d.B) * d.Lq, Mk = size_t(d.B) * d.Lk, D = d.D;
    Carver c(base);
    gacc = c.take<float>(64);
    dx1 = c.take<float>(M * D);
    delta = c.take<float>(size_t(d.B) * d.H * d.Lq);
    df = c.take<bf16>(M * D); du = c.take<bf16>(M * 4 * D); dh2 = c.take<bf16>(M * D); dob = c.take<bf16>(M * D);
    dctx = c.take<bf16>(M * D); dq = c.take<bf16>(M * D); dkv = c.take<bf16>(Mk * 2 * D); dh = c.take<bf16>(M * D);
    dqacc = c.take<float>(attn_bwd_workspace_bytes(d.B, d.H, d.Lq) / sizeof(float));
    bytes = c.off;
  }
};
struct SavedF32 {
  float *mean1, *rstd1, *mean2, *rstd2, *x1, *h, *q, *kv, *ctx, *P, *o, *h2, *u, *a, *f;
  size_t bytes;
  SavedF32(const gxa_dims& d, void* base) {
    const size_t M = size_t(d.B) * d.Lq, Mk = size_t(d.B) * d.Lk, D = d.D;
    Carver c(base);
    mean1 = c.take<float>(M); rstd1 = c.take<float>(M); mean2 = c.take<float>(M); rstd2 = c.take<float>(M);
    x1 = c.take<float>(M * D); h = c.take<float>(M * D); q = c.take<float>(M * D); kv = c.take<float>(Mk * 2 * D);
    ctx = c.take<float>(M * D); P = c.take<float>(size_t(d.B) * d.H * d.Lq * d.Lk); o = c.take<float>(M * D);
    h2 = c.take<float>(M * D); u = c.take<float>(M * 4 * D); a = c.take<float>(M * 4 * D); f = c.take<float>(M * D);
    bytes = c.off;
  }
};
struct WorkF32 {
  float *gacc, *dx1, *dS, *df, *da, *du, *dh2, *dob, *dctx, *dq, *dkv, *dh;
  bf16 *A3, *B3;
  size_t bytes;
  WorkF32(const gxa_dims& d, void* base) {
    const size_t M = size_t(d.B) * d.Lq, Mk = size_t(d.B) * d.Lk, D = d.D;
    const size_t Mx = M > Mk ? M : Mk;
    Carver c(base);
    gacc = c.take<float>(64);
    dx1 = c.take<float>(M * D); dS = c.take<float>(size_t(d.B) * d.H * d.Lq * d.Lk);
    df = c.take<float>(M * D); da = c.take<float>(M * 4 * D); du = c.take<float>(M * 4 * D); dh2 = c.take<float>(M * D);
    dob = c.take<float>(M * D); dctx = c.take<float>(M * D); dq = c.take<float>(M * D); dkv = c.take<float>(Mk * 2 * D);
    dh = c.take<float>(M * D);
    A3 = c.take<bf16>(3 * Mx * 4 * D);
    size_t b3 = 3 * Mx * 4 * D;                       // wgrad: B operand is an activation
    if (b3 < 3 * 4 * D * D) b3 = 3 * 4 * D * D;
    B3 = c.take<bf16>(b3);
    bytes = c.off;
  }
};

static int check_dims(const gxa_dims* d) {
  if (!d) return GXA_ERR_ARG;
  if (d->B <= 0 || d->Lq <= 0 || d->D <= 0 || d->H <= 0) return GXA_ERR_SHAPE;
  if (d->has_xa && d->Lk <= 0) return GXA_ERR_SHAPE;
  if (d->D != d->H * 64 || d->D % 128 != 0) return GXA_ERR_SHAPE;
  if (d->precision != GXA_PREC_BF16 && d->precision != GXA_PREC_FP32) return GXA_ERR_UNSUPPORTED;
  if (!(d->p_drop >= 0.f && d->p_drop < 1.f)) return GXA_ERR_ARG;
  return 0;
}

// fp32-accurate GEMM on the bf16 tensor cores: x = hi + lo, three bf16 products concatenated along K.
//   C[M,N] (+)= A * B^T + bias.   a_mn / b_mn as in GemmOp.
static int gemm_f32(const float* A, int lda, bool a_mn, const float* B, int ldb, bool b_mn, int M, int N, int K, float* C,
                    int ldc, const float* bias, int accumulate, bf16* A3, bf16* B3, cudaStream_t s) {
  GemmOp op;
  if (!a_mn) { GXA_TRY(split3(A, lda, A3, M, K, 0, 0, s)); op.A = A3; op.lda = 3 * K; }
  else       { GXA_TRY(split3(A, lda, A3, K, M, 0, 1, s)); op.A = A3; op.lda = M; }
  if (!b_mn) { GXA_TRY(split3(B, ldb, B3, N, K, 1, 0, s)); op.B = B3; op.ldb = 3 * K; }
  else       { GXA_TRY(split3(B, ldb, B3, K, N, 1, 1, s)); op.B = B3; op.ldb = N; }
  op.a_mn = a_mn; op.b_mn = b_mn;
  op.M = M; op.N = N; op.K = 3 * K;
  op.epi = EPI_F32; op.C = C; op.ldc = ldc; op.bias = bias; op.accumulate = accumulate;
  op.k_splits = 1;   // keep the fp32 path bit-reproducible
  return gemm_launch(op, s);
}

static int gemm_bf16(const void* A, int lda, bool a_mn, const void* B, int ldb, bool b_mn, int M, int N, int K, int epi,
                     void* C, int ldc, void* C2, int ldc2, const float* bias, const void* aux, int ld_aux,
                     const float* gate, int accumulate, cudaStream_t s, float* colsum_out = nullptr,
                     DropCfg drop = DropCfg{0, 1.f, 0, 0}, int k_splits = 0) {
  GemmOp op;
  op.k_splits = k_splits;
  op.colsum = colsum_out;
  op.drop = drop;
  op.A = A; op.lda = lda; op.a_mn = a_mn; op.B = B; op.ldb = ldb; op.b_mn = b_mn;
  op.M = M; op.N = N; op.K = K; op.epi = epi; op.C = C; op.ldc = ldc; op.C2 = C2; op.ldc2 = ldc2;
  op.bias = bias; op.aux = aux; op.ld_aux = ld_aux; op.gate = gate; op.accumulate = accumulate;
  return gemm_launch(op, s);
}

// bf16 weight cache offsets (elements)
struct WCache {
  const bf16 *in_proj, *out, *ff0, *ff2;
  WCache(const gxa_dims& d, const void* base) {
    const size_t D = d.D;
    const bf16* b = reinterpret_cast<const bf16*>(base);
    in_proj = b; out = b + 3 * D * D; ff0 = b + 4 * D * D; ff2 = b + 8 * D * D;
  }
};

// ================================================================== forward
// ------------------------------------------------------------------ side stream
// Work that is off the layer's dependency chain (the K|V projection in forward, the five weight-gradient GEMMs and the
// d(xa) GEMM in backward) is enqueued on a second stream, forked from and joined to the caller's stream by events inside
// the call.  Every GEMM is a persistent grid with a static tile split, so its last wave leaves SMs idle (448 tiles on 74
// CTA pairs = 6.05 waves for the N = 1024 GEMMs at c3); a kernel from the other stream fills them.  Nothing outlives the
// call: the join is enqueued before returning, so workspace reuse by the next layer stays ordered.  GXA_SIDE_STREAM=0
// puts everything back on one stream.
struct SideStream {
  cudaStream_t s = nullptr;
  cudaEvent_t fork[8] = {};
  cudaEvent_t join = nullptr;
  cudaEvent_t early = nullptr;      // a point in the middle of the side stream's work the caller's stream waits for
  bool ok = false;
};
static SideStream* side_stream() {
  static const bool enabled = [] { const char* e = getenv("GXA_SIDE_STREAM"); return !(e && e[0] == '0'); }();
  if (!enabled) return nullptr;
  static SideStream per_dev[16];
  int dev = 0;
  if (cudaGetDevice(&dev) != cudaSuccess || dev < 0 || dev >= 16) return nullptr;
  SideStream& sd = per_dev[dev];
  std::lock_guard<std::mutex> lk(g_init_mu);
  if (!sd.ok) {
    if (cudaStreamCreateWithFlags(&sd.s, cudaStreamNonBlocking) != cudaSuccess) return nullptr;
    for (auto& e : sd.fork)
      if (cudaEventCreateWithFlags(&e, cudaEventDisableTiming) != cudaSuccess) return nullptr;
    if (cudaEventCreateWithFlags(&sd.join, cudaEventDisableTiming) != cudaSuccess) return nullptr;
    if (cudaEventCreateWithFlags(&sd.early, cudaEventDisableTiming) != cudaSuccess) return nullptr;
    sd.ok = true;
  }
  return &sd;
}
// the side stream, ordered after everything enqueued on `s` so far (or `s` itself when there is no side stream)
static cudaStream_t side_fork(SideStream* sd, int slot, cudaStream_t s) {
  if (!sd) return s;
  if (cudaEventRecord(sd->fork[slot], s) != cudaSuccess || cudaStreamWaitEvent(sd->s, sd->fork[slot], 0) != cudaSuccess) return s;
  return sd->s;
}
static int side_join(SideStream* sd, cudaStream_t s) {
  if (!sd) return 0;
  if (cudaEventRecord(sd->join, sd->s) != cudaSuccess || cudaStreamWaitEvent(s, sd->join, 0) != cudaSuccess) return GXA_ERR_CUDA;
  return 0;
}

// y = x1 + tanh(gate) * (A W^T + bias) for a handful of rows (decode: M = batch x beams).  A 128-row-tile GEMM has only
// N/128 x ceil(M/128) CTAs there and each walks the whole K (33 us for ff.2 at Whisper-small), so the K range is split
// over the SMs instead: y <- x1, then every split adds its gated partial with fp32 atomics.  o / f are not rounded to
// bf16 on this route (nothing is saved for a backward), i.e. it is slightly MORE precise than the training forward.
static int gated_residual_small_m(const bf16* A, int K, const bf16* Wt, int M, int N, const float* bias, const float* x1,
                                  const float* gate, float* y, cudaStream_t s) {
  if (cudaMemcpyAsync(y, x1, sizeof(float) * size_t(M) * N, cudaMemcpyDeviceToDevice, s) != cudaSuccess)
    return fail(GXA_ERR_CUDA);
  const int kb = (K + 63) / 64, tiles = ((M + 127) / 128) * ((N + 127) / 128);
  int ks = num_sms() / (tiles > 0 ? tiles : 1);
  if (ks > kb / 4) ks = kb / 4;          // at least four 64-wide k-steps per split
  if (ks > 16) ks = 16;
  if (ks < 1) ks = 1;
  return gemm_bf16(A, K, false, Wt, K, false, M, N, K, EPI_F32, y, N, nullptr, 0, bias, nullptr, 0, gate, 1, s, nullptr,
                   DropCfg{0, 1.f, 0, 0}, ks);
}

// few-row linear layers of the decode step (gxa_skinny.cu): epi 0 bf16(+bias), 1 bias + GELU -> bf16, 2 gated residual f32
bool skinny_supported(int R, int N, int K);
int skinny_linear(const void* X, int ldx, const void* W, int ldw, int R, int N, int K, int epi, void* out, int ldo,
                  const float* bias, const float* res, int ldres, const float* gate, cudaStream_t s);

static int block_fwd_bf16(const gxa_dims& d, const gxa_block_params& p, const void* wcache, const float* x,
                          const void* xa_op, const uint8_t* mask, const int32_t* kv_end, float* y, void* saved,
                          cudaStream_t s, const bf16* kv_cached = nullptr) {
  pdl_hint_rows(static_cast<long long>(d.B) * d.Lq);
  const int M = d.B * d.Lq, Mk = d.B * d.Lk, D = d.D;
  gxa_dims dc = d;
  if (kv_cached != nullptr) dc.Lk = 1;     // decode: the activation carve has no K|V slot of its own ...
  SavedBf16 S(dc, saved);
  WCache W(d, wcache);
  if (kv_cached != nullptr) S.kv = const_cast<bf16*>(kv_cached);   // ... K|V were projected once (gxa_block_kv_prefill)
  const bool small_m = kv_cached != nullptr && M <= 256 && d.p_drop == 0.f;   // decode step: split-K gated residuals
  // decode step with <= 256 query rows: all four projections on the cluster split-K kernel (GXA_DECODE_SKINNY=0: A/B)
  static const bool skinny_on = [] { const char* e = getenv("GXA_DECODE_SKINNY"); return !(e && e[0] == '0'); }();
  const bool skinny = small_m && skinny_on && skinny_supported(M, D, D) && skinny_supported(M, 4 * D, D) &&
                      skinny_supported(M, D, 4 * D);
  const float* x1 = x;
  if (d.has_xa) {
    SideStream* sd = kv_cached == nullptr ? side_stream() : nullptr;
    if (kv_cached == nullptr)       // K|V depend on xa only: side stream, joined before the attention kernel
      GXA_TRY(gemm_bf16(xa_op, D, false, W.in_proj + size_t(D) * D, D, false, Mk, 2 * D, D, EPI_BF16, S.kv, 2 * D,
                        nullptr, 0, p.in_proj_b + D, nullptr, 0, nullptr, 0, side_fork(sd, 0, s)));
    GXA_TRY(ln_fwd(x, p.attn_ln_w, p.attn_ln_b, S.h, 0, S.mean1, S.rstd1, M, D, s));
    if (skinny)
      GXA_TRY(skinny_linear(S.h, D, W.in_proj, D, M, D, D, 0, S.q, D, p.in_proj_b, nullptr, 0, nullptr, s));
    else
      GXA_TRY(gemm_bf16(S.h, D, false, W.in_proj, D, false, M, D, D, EPI_BF16, S.q, D, nullptr, 0, p.in_proj_b, nullptr, 0,
                        nullptr, 0, s));
    GXA_TRY(side_join(sd, s));
    GXA_TRY(attn_fwd_bf16(S.q, D, S.kv, S.kv + D, 2 * D, mask, kv_end, S.ctx, D, S.lse, d.B, d.H, d.Lq, d.Lk, s,
                          make_drop_cfg(d.p_drop, d.seed, kDropStreamAttn)));
    if (skinny)
      GXA_TRY(skinny_linear(S.ctx, D, W.out, D, M, D, D, 2, S.x1, D, p.out_proj_b, x, D, p.attn_gate, s));
    else if (small_m)
      GXA_TRY(gated_residual_small_m(S.ctx, D, W.out, M, D, p.out_proj_b, x, p.attn_gate, S.x1, s));
    else
      GXA_TRY(gemm_bf16(S.ctx, D, false, W.out, D, false, M, D, D, EPI_GATE_RES, S.x1, D, S.o, D, p.out_proj_b, x, D,
                        p.attn_gate, 0, s));
    x1 = S.x1;
  }
  GXA_TRY(ln_fwd(x1, p.ff_ln_w, p.ff_ln_b, S.h2, 0, S.mean2, S.rstd2, M, D, s));
  if (skinny) {
    GXA_TRY(skinny_linear(S.h2, D, W.ff0, D, M, 4 * D, D, 1, S.a, 4 * D, p.ff0_b, nullptr, 0, nullptr, s));
    return skinny_linear(S.a, 4 * D, W.ff2, 4 * D, M, D, 4 * D, 2, y, D, p.ff2_b, x1, D, p.ff_gate, s);
  }
  GXA_TRY(gemm_bf16(S.h2, D, false, W.ff0, D, false, M, 4 * D, D, EPI_BIAS_GELU, S.u, 4 * D, S.a, 4 * D, p.ff0_b, nullptr,
                    0, nullptr, 0, s));
  if (small_m)
    return gated_residual_small_m(S.a, 4 * D, W.ff2, M, D, p.ff2_b, x1, p.ff_gate, y, s);
  GXA_TRY(gemm_bf16(S.a, 4 * D, false, W.ff2, 4 * D, false, M, D, 4 * D, EPI_GATE_RES, y, D, S.f, D, p.ff2_b, x1, D,
                    p.ff_gate, 0, s, nullptr, make_drop_cfg(d.p_drop, d.seed, kDropStreamFfn)));   // S.f = dropout(f)
  return 0;
}

static int block_fwd_f32(const gxa_dims& d, const gxa_block_params& p, const float* x, const void* xa_op,
                         const uint8_t* mask, float* y, void* saved, void* workspace, cudaStream_t s) {
  const int M = d.B * d.Lq, Mk = d.B * d.Lk, D = d.D;
  SavedF32 S(d, saved);
  WorkF32 Wk(d, workspace);
  const float* xa = reinterpret_cast<const float*>(xa_op);
  const float* x1 = x;
  if (d.has_xa) {
    GXA_TRY(ln_fwd(x, p.attn_ln_w, p.attn_ln_b, S.h, 1, S.mean1, S.rstd1, M, D, s));
    GXA_TRY(gemm_f32(S.h, D, false, p.in_proj_w, D, false, M, D, D, S.q, D, p.in_proj_b, 0, Wk.A3, Wk.B3, s));
    GXA_TRY(gemm_f32(xa, D, false, p.in_proj_w + size_t(D) * D, D, false, Mk, 2 * D, D, S.kv, 2 * D, p.in_proj_b + D, 0,
                     Wk.A3, Wk.B3, s));
    GXA_TRY(attn_fwd_f32(S.q, D, S.kv, S.kv + D, 2 * D, mask, S.ctx, D, S.P, d.B, d.H, d.Lq, d.Lk, s,
                         make_drop_cfg(d.p_drop, d.seed, kDropStreamAttn)));
    GXA_TRY(gemm_f32(S.ctx, D, false, p.out_proj_w, D, false, M, D, D, S.o, D, p.out_proj_b, 0, Wk.A3, Wk.B3, s));
    GXA_TRY(gate_res_f32(x, S.o, p.attn_gate, S.x1, size_t(M) * D, s));
    x1 = S.x1;
  }
  GXA_TRY(ln_fwd(x1, p.ff_ln_w, p.ff_ln_b, S.h2, 1, S.mean2, S.rstd2, M, D, s));
  GXA_TRY(gemm_f32(S.h2, D, false, p.ff0_w, D, false, M, 4 * D, D, S.u, 4 * D, p.ff0_b, 0, Wk.A3, Wk.B3, s));
  GXA_TRY(gelu_f32(S.u, S.a, size_t(M) * 4 * D, s));
  GXA_TRY(gemm_f32(S.a, 4 * D, false, p.ff2_w, 4 * D, false, M, D, 4 * D, S.f, D, p.ff2_b, 0, Wk.A3, Wk.B3, s));
  GXA_TRY(gate_res_f32(x1, S.f, p.ff_gate, y, size_t(M) * D, s, D, make_drop_cfg(d.p_drop, d.seed, kDropStreamFfn)));   // S.f = dropout(f)
  return 0;
}

// ================================================================== backward
static int block_bwd_bf16(const gxa_dims& d, const gxa_block_params& p, const void* wcache, const float* x,
                          const void* xa_op, const uint8_t* mask, const int32_t* kv_end, const void* saved,
                          const float* dy, float* dx, float* dxa, const gxa_block_grads& g, void* workspace,
                          cudaStream_t s) {
  pdl_hint_rows(static_cast<long long>(d.B) * d.Lq);
  const int M = d.B * d.Lq, Mk = d.B * d.Lk, D = d.D;
  SavedBf16 S(d, const_cast<void*>(saved));
  WorkBf16 Wk(d, workspace);
  WCache W(d, wcache);
  if (cudaMemsetAsync(Wk.gacc, 0, 64 * sizeof(float), s) != cudaSuccess) return fail(GXA_ERR_CUDA);
  SideStream* sd = side_stream();    // weight gradients leave the dgrad chain (joined before returning)
  // the fp32 dQ accumulator of the attention backward (134 MB at c3) is zeroed on the side stream, under the FFN
  // backward GEMMs (tensor-bound: HBM is idle there), instead of in front of the attention kernel
  bool dq_zeroed = false;
  if (sd && d.has_xa) {
    const cudaStream_t zs = side_fork(sd, 4, s);
    if (zs != s && cudaMemsetAsync(Wk.dqacc, 0, attn_bwd_workspace_bytes(d.B, d.H, d.Lq), zs) == cudaSuccess &&
        cudaEventRecord(sd->early, zs) == cudaSuccess)
      dq_zeroed = true;
  }
  const float* x1 = d.has_xa ? S.x1 : x;
  // ---- FFN branch
  unsigned* fin = reinterpret_cast<unsigned*>(Wk.gacc + 8);      // last-block-done counters of the fused gate finish
  GXA_TRY(gate_bwd(dy, S.f, 0, p.ff_gate, Wk.gacc + 1, Wk.df, 0, g.ff2_b, M, D, s,
                   make_drop_cfg(d.p_drop, d.seed, kDropStreamFfn), g.ff_gate, fin + 1));
  GXA_TRY(gemm_bf16(Wk.df, D, false, W.ff2, 4 * D, true, M, 4 * D, D, EPI_GELU_BWD, Wk.du, 4 * D, nullptr, 0, nullptr,
                    S.u, 4 * D, nullptr, 0, s, g.ff0_b));     // d(ff.0.bias) = colsum(du) fused in the epilogue
  GXA_TRY(gemm_bf16(Wk.df, D, true, S.a, 4 * D, true, D, 4 * D, M, EPI_F32, g.ff2_w, 4 * D, nullptr, 0, nullptr, nullptr,
                    0, nullptr, 1, side_fork(sd, 0, s)));
  GXA_TRY(gemm_bf16(Wk.du, 4 * D, false, W.ff0, D, true, M, D, 4 * D, EPI_BF16, Wk.dh2, D, nullptr, 0, nullptr, nullptr, 0,
                    nullptr, 0, s));
  GXA_TRY(gemm_bf16(Wk.du, 4 * D, true, S.h2, D, true, 4 * D, D, M, EPI_F32, g.ff0_w, D, nullptr, 0, nullptr, nullptr, 0,
                    nullptr, 1, side_fork(sd, 1, s)));
  if (!d.has_xa) {
    GXA_TRY(ln_bwd(Wk.dh2, 0, x1, p.ff_ln_w, S.mean2, S.rstd2, dy, dx, g.ff_ln_w, g.ff_ln_b, nullptr, 0, nullptr, nullptr,
                   nullptr, nullptr, M, D, s));
    return side_join(sd, s);
  }
  // ---- LN2 backward fused with the attention-gate backward: dx1, do = tanh(a)*dx1, d(attn_gate) (finished by the
  // kernel's last block), d(bo)
  GXA_TRY(ln_bwd(Wk.dh2, 0, x1, p.ff_ln_w, S.mean2, S.rstd2, dy, Wk.dx1, g.ff_ln_w, g.ff_ln_b, S.o, 0, p.attn_gate,
                 Wk.gacc, Wk.dob, g.out_proj_b, M, D, s, g.attn_gate, fin));
  GXA_TRY(gemm_bf16(Wk.dob, D, false, W.out, D, true, M, D, D, EPI_BF16, Wk.dctx, D, nullptr, 0, nullptr, nullptr, 0,
                    nullptr, 0, s));
  GXA_TRY(gemm_bf16(Wk.dob, D, true, S.ctx, D, true, D, D, M, EPI_F32, g.out_proj_w, D, nullptr, 0, nullptr, nullptr, 0,
                    nullptr, 1, side_fork(sd, 2, s)));
  // ---- attention core
  if (dq_zeroed && cudaStreamWaitEvent(s, sd->early, 0) != cudaSuccess) return fail(GXA_ERR_CUDA);
  // fused into the attention backward: d(bq) (column sums of dq, in the finalize kernel) and d(bv) (column sums of dV, in
  // the drain warps).  d(bk) is identically zero (every row of dS sums to zero -- softmax is invariant to a per-query
  // shift of the logits; the reference's autograd returns rounding noise there) and is left at zero.
  GXA_TRY(attn_bwd_bf16(S.q, D, S.kv, S.kv + D, 2 * D, mask, kv_end, S.ctx, Wk.dctx, D, S.lse, Wk.delta, Wk.dq, D,
                        Wk.dkv, Wk.dkv + D, 2 * D, Wk.dqacc, d.B, d.H, d.Lq, d.Lk, s, g.in_proj_b,
                        make_drop_cfg(d.p_drop, d.seed, kDropStreamAttn), g.in_proj_b + 2 * D, dq_zeroed));
  {   // everything that only consumes dq / dkv and is not on the way to dx: side stream
    const cudaStream_t ws = side_fork(sd, 3, s);
    GXA_TRY(gemm_bf16(Wk.dq, D, true, S.h, D, true, D, D, M, EPI_F32, g.in_proj_w, D, nullptr, 0, nullptr, nullptr, 0,
                      nullptr, 1, ws));
    GXA_TRY(gemm_bf16(Wk.dkv, 2 * D, true, xa_op, D, true, 2 * D, D, Mk, EPI_F32, g.in_proj_w + size_t(D) * D, D, nullptr,
                      0, nullptr, nullptr, 0, nullptr, 1, ws));
    if (dxa)
      GXA_TRY(gemm_bf16(Wk.dkv, 2 * D, false, W.in_proj + size_t(D) * D, D, true, Mk, D, 2 * D, EPI_F32, dxa, D, nullptr, 0,
                        nullptr, nullptr, 0, nullptr, 1, ws));
  }
  GXA_TRY(gemm_bf16(Wk.dq, D, false, W.in_proj, D, true, M, D, D, EPI_BF16, Wk.dh, D, nullptr, 0, nullptr, nullptr, 0,
                    nullptr, 0, s));
  GXA_TRY(ln_bwd(Wk.dh, 0, x, p.attn_ln_w, S.mean1, S.rstd1, Wk.dx1, dx, g.attn_ln_w, g.attn_ln_b, nullptr, 0, nullptr,
                 nullptr, nullptr, nullptr, M, D, s));
  return side_join(sd, s);
}

static int block_bwd_f32(const gxa_dims& d, const gxa_block_params& p, const float* x, const void* xa_op,
                         const void* saved, const float* dy, float* dx, float* dxa, const gxa_block_grads& g,
                         void* workspace, cudaStream_t s) {
  const int M = d.B * d.Lq, Mk = d.B * d.Lk, D = d.D;
  SavedF32 S(d, const_cast<void*>(saved));
  WorkF32 Wk(d, workspace);
  const float* xa = reinterpret_cast<const float*>(xa_op);
  if (cudaMemsetAsync(Wk.gacc, 0, 64 * sizeof(float), s) != cudaSuccess) return fail(GXA_ERR_CUDA);
  const float* x1 = d.has_xa ? S.x1 : x;
  GXA_TRY(gate_bwd(dy, S.f, 1, p.ff_gate, Wk.gacc + 1, Wk.df, 1, g.ff2_b, M, D, s,
                   make_drop_cfg(d.p_drop, d.seed, kDropStreamFfn)));
  GXA_TRY(gemm_f32(Wk.df, D, false, p.ff2_w, 4 * D, true, M, 4 * D, D, Wk.da, 4 * D, nullptr, 0, Wk.A3, Wk.B3, s));
  GXA_TRY(gelu_bwd_f32(Wk.da, S.u, Wk.du, size_t(M) * 4 * D, s));
  GXA_TRY(gemm_f32(Wk.df, D, true, S.a, 4 * D, true, D, 4 * D, M, g.ff2_w, 4 * D, nullptr, 1, Wk.A3, Wk.B3, s));
  GXA_TRY(colsum(Wk.du, 1, 4 * D, g.ff0_b, M, 4 * D, s));
  GXA_TRY(gemm_f32(Wk.du, 4 * D, false, p.ff0_w, D, true, M, D, 4 * D, Wk.dh2, D, nullptr, 0, Wk.A3, Wk.B3, s));
  GXA_TRY(gemm_f32(Wk.du, 4 * D, true, S.h2, D, true, 4 * D, D, M, g.ff0_w, D, nullptr, 1, Wk.A3, Wk.B3, s));
  GXA_TRY(gate_finish(p.ff_gate, Wk.gacc + 1, g.ff_gate, s));
  if (!d.has_xa) {
    GXA_TRY(ln_bwd(Wk.dh2, 1, x1, p.ff_ln_w, S.mean2, S.rstd2, dy, dx, g.ff_ln_w, g.ff_ln_b, nullptr, 0, nullptr, nullptr,
                   nullptr, nullptr, M, D, s));
    return 0;
  }
  GXA_TRY(ln_bwd(Wk.dh2, 1, x1, p.ff_ln_w, S.mean2, S.rstd2, dy, Wk.dx1, g.ff_ln_w, g.ff_ln_b, S.o, 1, p.attn_gate,
                 Wk.gacc, Wk.dob, g.out_proj_b, M, D, s));
  GXA_TRY(gate_finish(p.attn_gate, Wk.gacc, g.attn_gate, s));
  GXA_TRY(gemm_f32(Wk.dob, D, false, p.out_proj_w, D, true, M, D, D, Wk.dctx, D, nullptr, 0, Wk.A3, Wk.B3, s));
  GXA_TRY(gemm_f32(Wk.dob, D, true, S.ctx, D, true, D, D, M, g.out_proj_w, D, nullptr, 1, Wk.A3, Wk.B3, s));
  GXA_TRY(attn_bwd_f32(S.q, D, S.kv, S.kv + D, 2 * D, Wk.dctx, D, S.P, Wk.dS, Wk.dq, D, Wk.dkv, Wk.dkv + D, 2 * D, d.B,
                       d.H, d.Lq, d.Lk, s, make_drop_cfg(d.p_drop, d.seed, kDropStreamAttn)));
  GXA_TRY(colsum(Wk.dq, 1, D, g.in_proj_b, M, D, s));
  GXA_TRY(colsum(Wk.dkv + D, 1, 2 * D, g.in_proj_b + 2 * D, Mk, D, s));      // d(bv); d(bk) == 0 (see the bf16 path)
  GXA_TRY(gemm_f32(Wk.dq, D, true, S.h, D, true, D, D, M, g.in_proj_w, D, nullptr, 1, Wk.A3, Wk.B3, s));
  GXA_TRY(gemm_f32(Wk.dkv, 2 * D, true, xa, D, true, 2 * D, D, Mk, g.in_proj_w + size_t(D) * D, D, nullptr, 1, Wk.A3,
                   Wk.B3, s));
  GXA_TRY(gemm_f32(Wk.dq, D, false, p.in_proj_w, D, true, M, D, D, Wk.dh, D, nullptr, 0, Wk.A3, Wk.B3, s));
  if (dxa)
    GXA_TRY(gemm_f32(Wk.dkv, 2 * D, false, p.in_proj_w + size_t(D) * D, D, true, Mk, D, 2 * D, dxa, D, nullptr, 1, Wk.A3,
                     Wk.B3, s));
  GXA_TRY(ln_bwd(Wk.dh, 1, x, p.attn_ln_w, S.mean1, S.rstd1, Wk.dx1, dx, g.attn_ln_w, g.attn_ln_b, nullptr, 0, nullptr,
                 nullptr, nullptr, nullptr, M, D, s));
  return 0;
}

}  // namespace gxa

// ================================================================== extern "C"
using namespace gxa;

extern "C" {

int gxa_query_caps(int* sm_count, int* cc_major, int* cc_minor) {
  int dev = 0;
  if (cudaGetDevice(&dev) != cudaSuccess) return fail(GXA_ERR_CUDA);
  cudaDeviceProp pr;
  if (cudaGetDeviceProperties(&pr, dev) != cudaSuccess) return fail(GXA_ERR_CUDA);
  if (sm_count) *sm_count = pr.multiProcessorCount;
  if (cc_major) *cc_major = pr.major;
  if (cc_minor) *cc_minor = pr.minor;
  return 0;
}
const char* gxa_version(void) { return "gxa-b200 0.1 (sm_100a; tcgen05+TMA GEMMs)"; }
const char* gxa_last_cuda_error(void) { return g_err; }
long long gxa_launch_count(void) { return g_launches.load(); }
int gxa_dropout_multipliers_host(float p_drop, uint64_t seed, int site, int rows, int cols, float* out) {
  if (!out || rows < 0 || cols < 0) return GXA_ERR_ARG;
  if (!(p_drop >= 0.f && p_drop < 1.f) || (site != int(kDropStreamAttn) && site != int(kDropStreamFfn))) return GXA_ERR_ARG;
  const DropCfg c = make_drop_cfg(p_drop, seed, uint32_t(site));
  const uint32_t quarter = uint32_t(cols + 3) >> 2;
  for (int r = 0; r < rows; ++r)
    for (int col = 0; col < cols; ++col) {
      const bool keep = c.thresh == 0 || drop_keep_scalar(c, uint32_t(r) * quarter + (uint32_t(col) >> 2), col & 3);   // gxa_rng.cuh
      out[size_t(r) * cols + col] = keep ? c.scale : 0.f;
    }
  return 0;
}
int gxa_set_sm_limit(int n_sms) { return g_sm_limit.exchange(n_sms < 0 ? 0 : n_sms); }

size_t gxa_block_wcache_bytes(const gxa_dims* d) {
  if (check_dims(d)) return 0;
  return d->precision == GXA_PREC_BF16 ? size_t(12) * d->D * d->D * sizeof(bf16) : 256;
}
size_t gxa_block_saved_bytes(const gxa_dims* d) {
  if (check_dims(d)) return 0;
  return d->precision == GXA_PREC_BF16 ? SavedBf16(*d, nullptr).bytes : SavedF32(*d, nullptr).bytes;
}
size_t gxa_block_workspace_bytes(const gxa_dims* d) {
  if (check_dims(d)) return 0;
  return d->precision == GXA_PREC_BF16 ? WorkBf16(*d, nullptr).bytes : WorkF32(*d, nullptr).bytes;
}
size_t gxa_xa_operand_bytes(const gxa_dims* d) {
  if (check_dims(d)) return 0;
  return size_t(d->B) * d->Lk * d->D * (d->precision == GXA_PREC_BF16 ? sizeof(bf16) : sizeof(float));
}

int gxa_block_prepare_weights(const gxa_dims* d, const gxa_block_params* p, void* wcache, void* stream) {
  GXA_TRY(check_dims(d));
  if (!p || !wcache) return GXA_ERR_ARG;
  if (d->precision != GXA_PREC_BF16) return 0;   // fp32 mode splits operands on the fly
  cudaStream_t s = reinterpret_cast<cudaStream_t>(stream);
  const size_t DD = size_t(d->D) * d->D;
  bf16* w = reinterpret_cast<bf16*>(wcache);
  const float* src[4] = {p->in_proj_w, p->out_proj_w, p->ff0_w, p->ff2_w};
  void* dst[4] = {w, w + 3 * DD, w + 4 * DD, w + 8 * DD};
  const size_t n[4] = {3 * DD, DD, 4 * DD, 4 * DD};
  GXA_TRY(cast_f32_bf16_multi(src, dst, n, 4, s));      // one launch for the layer's four weight matrices
  return 0;
}

int gxa_mask_prepare(const uint8_t* mask, int32_t* kv_end, int B, int Lk, void* stream) {
  if (!mask || !kv_end || B <= 0 || Lk <= 0) return GXA_ERR_ARG;
  GXA_TRY(mask_kv_end(mask, kv_end, B, Lk, reinterpret_cast<cudaStream_t>(stream)));
  return 0;
}

int gxa_xa_prepare(const gxa_dims* d, const float* xa, void* xa_op, void* stream) {
  GXA_TRY(check_dims(d));
  if (!xa || !xa_op) return GXA_ERR_ARG;
  cudaStream_t s = reinterpret_cast<cudaStream_t>(stream);
  const size_t n = size_t(d->B) * d->Lk * d->D;
  if (d->precision == GXA_PREC_BF16) {
    GXA_TRY(cast_f32_bf16(xa, xa_op, n, s));
  } else if (cudaMemcpyAsync(xa_op, xa, n * sizeof(float), cudaMemcpyDeviceToDevice, s) != cudaSuccess) {
    return fail(GXA_ERR_CUDA);
  }
  return 0;
}

int gxa_block_fwd(const gxa_dims* d, const gxa_block_params* p, const void* wcache, const float* x, const void* xa_op,
                  const uint8_t* mask, const int32_t* kv_end, float* y, void* saved, void* workspace, void* stream) {
  GXA_TRY(check_dims(d));
  if (!p || !x || !y || x == y || !saved) return GXA_ERR_ARG;
  if (d->has_xa && !xa_op) return GXA_ERR_ARG;
  cudaStream_t s = reinterpret_cast<cudaStream_t>(stream);
  if (d->precision == GXA_PREC_BF16) {
    if (!wcache) return GXA_ERR_ARG;
    return block_fwd_bf16(*d, *p, wcache, x, xa_op, mask, kv_end, y, saved, s);
  }
  if (!workspace) return GXA_ERR_ARG;
  return block_fwd_f32(*d, *p, x, xa_op, mask, y, saved, workspace, s);
}

// ---- decode-time path: K|V of the (fixed) visual memory are projected once per layer and cached;
// every decoding step then runs the block on the new query rows only.  Beams of one sample share its cache: pass them
// as the Lq dimension (the block has no interaction between query positions).
size_t gxa_block_kv_cache_bytes(const gxa_dims* d) {
  if (check_dims(d) || !d->has_xa) return 0;
  return size_t(d->B) * d->Lk * 2 * d->D * sizeof(bf16);
}
int gxa_block_kv_prefill(const gxa_dims* d, const gxa_block_params* p, const void* wcache, const void* xa_op,
                         void* kv_cache, void* stream) {
  GXA_TRY(check_dims(d));
  if (!p || !wcache || !xa_op || !kv_cache || !d->has_xa) return GXA_ERR_ARG;
  if (d->precision != GXA_PREC_BF16) return GXA_ERR_UNSUPPORTED;
  const int Mk = d->B * d->Lk, D = d->D;
  WCache W(*d, wcache);
  GXA_TRY(gemm_bf16(xa_op, D, false, W.in_proj + size_t(D) * D, D, false, Mk, 2 * D, D, EPI_BF16, kv_cache, 2 * D, nullptr,
                    0, p->in_proj_b + D, nullptr, 0, nullptr, 0, reinterpret_cast<cudaStream_t>(stream)));
  return 0;
}
size_t gxa_block_decode_workspace_bytes(const gxa_dims* d) {
  if (check_dims(d)) return 0;
  gxa_dims q = *d;
  q.Lk = 1;                               // the K|V slot of the activation carve is not used (external cache)
  return SavedBf16(q, nullptr).bytes;
}
int gxa_block_decode(const gxa_dims* d, const gxa_block_params* p, const void* wcache, const float* x,
                     const void* kv_cache, const uint8_t* mask, const int32_t* kv_end, float* y, void* workspace,
                     void* stream) {
  GXA_TRY(check_dims(d));
  if (!p || !wcache || !x || !y || x == y || !workspace || !kv_cache || !d->has_xa) return GXA_ERR_ARG;
  if (d->precision != GXA_PREC_BF16 || d->p_drop != 0.f) return GXA_ERR_UNSUPPORTED;   // inference: bf16, no dropout
  return block_fwd_bf16(*d, *p, wcache, x, nullptr, mask, kv_end, y, workspace, reinterpret_cast<cudaStream_t>(stream),
                        reinterpret_cast<const bf16*>(kv_cache));
}

int gxa_block_bwd(const gxa_dims* d, const gxa_block_params* p, const void* wcache, const float* x, const void* xa_op,
                  const uint8_t* mask, const int32_t* kv_end, const void* saved, const float* dy, float* dx, float* dxa,
                  const gxa_block_grads* g, void* workspace, void* stream) {
  GXA_TRY(check_dims(d));
  if (!p || !x || !saved || !dy || !dx || !g || !workspace) return GXA_ERR_ARG;
  cudaStream_t s = reinterpret_cast<cudaStream_t>(stream);
  if (d->precision == GXA_PREC_BF16) {
    if (!wcache) return GXA_ERR_ARG;
    return block_bwd_bf16(*d, *p, wcache, x, xa_op, mask, kv_end, saved, dy, dx, dxa, *g, workspace, s);
  }
  return block_bwd_f32(*d, *p, x, xa_op, saved, dy, dx, dxa, *g, workspace, s);
}

int gxa_layernorm_fwd(const float* x, const float* w, const float* b, void* out, int out_is_f32, float* mean,
                      float* rstd, int rows, int D, void* stream) {
  if (!x || !w || !b || !out) return GXA_ERR_ARG;
  GXA_TRY(ln_fwd(x, w, b, out, out_is_f32, mean, rstd, rows, D, reinterpret_cast<cudaStream_t>(stream)));
  return 0;
}
int gxa_layernorm_pe_fwd(const float* x, const float* w, const float* b, const float* pe, int T, float* out, float* mean,
                         float* rstd, int rows, int D, void* stream) {
  if (!x || !w || !b || !pe || !out || T <= 0) return GXA_ERR_ARG;
  GXA_TRY(ln_fwd(x, w, b, out, 1, mean, rstd, rows, D, reinterpret_cast<cudaStream_t>(stream), pe, T));
  return 0;
}
int gxa_sum3_f32(const float* a, const float* b, const float* c, float* out, size_t n, void* stream) {
  if (!a || !b || !c || !out) return GXA_ERR_ARG;
  GXA_TRY(sum3_f32(a, b, c, out, n, reinterpret_cast<cudaStream_t>(stream)));
  return 0;
}
int gxa_layernorm_bwd(const float* g, const float* x, const float* w, const float* mean, const float* rstd,
                      const float* dres, float* dx, float* dw, float* db, int rows, int D, void* stream) {
  if (!g || !x || !w || !mean || !rstd || !dx || !dw || !db) return GXA_ERR_ARG;
  GXA_TRY(ln_bwd(g, 1, x, w, mean, rstd, dres, dx, dw, db, nullptr, 0, nullptr, nullptr, nullptr, nullptr, rows, D,
                 reinterpret_cast<cudaStream_t>(stream)));
  return 0;
}

size_t gxa_linear_workspace_bytes(int rows, int in_f, int out_f, int precision) {
  const size_t mx = size_t(in_f) > size_t(out_f) ? in_f : out_f;
  const size_t mult = precision == GXA_PREC_BF16 ? 1 : 3;
  // A operand (activation or dy) + B operand (weight or activation), bf16
  return (mult * size_t(rows) * mx * 2 + mult * (size_t(rows) > size_t(in_f) ? size_t(rows) : size_t(in_f)) * mx * 2 +
          1024) & ~size_t(255);
}
int gxa_linear_fwd(const float* x, const float* w, const float* b, float* y, int rows, int in_f, int out_f,
                   int precision, void* workspace, void* stream) {
  if (!x || !w || !y || !workspace) return GXA_ERR_ARG;
  if (in_f % 64 || out_f % 64) return GXA_ERR_SHAPE;
  cudaStream_t s = reinterpret_cast<cudaStream_t>(stream);
  const size_t mx = size_t(in_f) > size_t(out_f) ? in_f : out_f;
  const size_t mult = precision == GXA_PREC_BF16 ? 1 : 3;
  bf16* A = reinterpret_cast<bf16*>(workspace);
  bf16* Bm = A + ((mult * size_t(rows) * mx + 127) & ~size_t(127));
  if (precision == GXA_PREC_BF16) {
    GXA_TRY(cast_f32_bf16(x, A, size_t(rows) * in_f, s));
    GXA_TRY(cast_f32_bf16(w, Bm, size_t(out_f) * in_f, s));
    GXA_TRY(gemm_bf16(A, in_f, false, Bm, in_f, false, rows, out_f, in_f, EPI_F32, y, out_f, nullptr, 0, b, nullptr, 0,
                      nullptr, 0, s));
    return 0;
  }
  GXA_TRY(gemm_f32(x, in_f, false, w, in_f, false, rows, out_f, in_f, y, out_f, b, 0, A, Bm, s));
  return 0;
}
int gxa_linear_bwd(const float* dy, const float* x, const float* w, float* dx, float* dw, float* db, int rows, int in_f,
                   int out_f, int precision, void* workspace, void* stream) {
  if (!dy || !x || !w || !workspace) return GXA_ERR_ARG;
  if (in_f % 64 || out_f % 64) return GXA_ERR_SHAPE;
  cudaStream_t s = reinterpret_cast<cudaStream_t>(stream);
  const size_t mx = size_t(in_f) > size_t(out_f) ? in_f : out_f;
  const size_t mult = precision == GXA_PREC_BF16 ? 1 : 3;
  bf16* A = reinterpret_cast<bf16*>(workspace);
  bf16* Bm = A + ((mult * size_t(rows) * mx + 127) & ~size_t(127));
  if (db) GXA_TRY(colsum(dy, 1, out_f, db, rows, out_f, s));
  if (precision == GXA_PREC_BF16) {
    GXA_TRY(cast_f32_bf16(dy, A, size_t(rows) * out_f, s));
    if (dx) {
      GXA_TRY(cast_f32_bf16(w, Bm, size_t(out_f) * in_f, s));
      GXA_TRY(gemm_bf16(A, out_f, false, Bm, in_f, true, rows, in_f, out_f, EPI_F32, dx, in_f, nullptr, 0, nullptr,
                        nullptr, 0, nullptr, 0, s));
    }
    if (dw) {
      GXA_TRY(cast_f32_bf16(x, Bm, size_t(rows) * in_f, s));
      GXA_TRY(gemm_bf16(A, out_f, true, Bm, in_f, true, out_f, in_f, rows, EPI_F32, dw, in_f, nullptr, 0, nullptr,
                        nullptr, 0, nullptr, 1, s));
    }
    return 0;
  }
  if (dx) GXA_TRY(gemm_f32(dy, out_f, false, w, in_f, true, rows, in_f, out_f, dx, in_f, nullptr, 0, A, Bm, s));
  if (dw) GXA_TRY(gemm_f32(dy, out_f, true, x, in_f, true, out_f, in_f, rows, dw, in_f, nullptr, 1, A, Bm, s));
  return 0;
}

int gxa_gemm_bf16(const void* A, int lda, int a_mn, const void* B, int ldb, int b_mn, int M, int N, int K, int epi,
                  void* C, int ldc, void* C2, int ldc2, const float* bias, const void* aux, int ld_aux,
                  const float* gate, int accumulate, int bn, int k_splits, void* stream) {
  if (!A || !B || !C) return GXA_ERR_ARG;
  GemmOp op;
  op.A = A; op.lda = lda; op.a_mn = a_mn != 0; op.B = B; op.ldb = ldb; op.b_mn = b_mn != 0;
  op.M = M; op.N = N; op.K = K; op.epi = epi; op.C = C; op.ldc = ldc; op.C2 = C2; op.ldc2 = ldc2;
  op.bias = bias; op.aux = aux; op.ld_aux = ld_aux; op.gate = gate; op.accumulate = accumulate;
  op.bn = bn; op.k_splits = k_splits;
  GXA_TRY(gemm_launch(op, reinterpret_cast<cudaStream_t>(stream)));
  return 0;
}

int gxa_linear_few_rows_bf16(const void* X, int ldx, const void* W, int ldw, int R, int N, int K, int epi, void* out,
                             int ldo, const float* bias, const float* res, int ldres, const float* gate, void* stream) {
  if (!X || !W || !out) return GXA_ERR_ARG;
  if (epi < 0 || epi > 2) return GXA_ERR_UNSUPPORTED;
  GXA_TRY(skinny_linear(X, ldx, W, ldw, R, N, K, epi, out, ldo, bias, res, ldres, gate,
                        reinterpret_cast<cudaStream_t>(stream)));
  return 0;
}

int gxa_attn_fwd(const void* q, int ldq, const void* k, const void* v, int ldkv, const uint8_t* mask,
                 const int32_t* kv_end, void* ctx, int ldc, float* lse, int B, int H, int Lq, int Lk, void* stream) {
  if (!q || !k || !v || !ctx) return GXA_ERR_ARG;
  GXA_TRY(attn_fwd_bf16(q, ldq, k, v, ldkv, mask, kv_end, ctx, ldc, lse, B, H, Lq, Lk,
                        reinterpret_cast<cudaStream_t>(stream)));
  return 0;
}
int gxa_attn_fwd_dropout(const void* q, int ldq, const void* k, const void* v, int ldkv, const uint8_t* mask,
                         const int32_t* kv_end, void* ctx, int ldc, float* lse, int B, int H, int Lq, int Lk,
                         float p_drop, uint64_t seed, void* stream) {
  if (!q || !k || !v || !ctx || !(p_drop >= 0.f && p_drop < 1.f)) return GXA_ERR_ARG;
  GXA_TRY(attn_fwd_bf16(q, ldq, k, v, ldkv, mask, kv_end, ctx, ldc, lse, B, H, Lq, Lk,
                        reinterpret_cast<cudaStream_t>(stream), make_drop_cfg(p_drop, seed, kDropStreamAttn)));
  return 0;
}
size_t gxa_attn_bwd_workspace_bytes(int B, int H, int Lq) { return attn_bwd_workspace_bytes(B, H, Lq); }

int gxa_attn_bwd(const void* q, int ldq, const void* k, const void* v, int ldkv, const uint8_t* mask,
                 const int32_t* kv_end, const void* ctx, const void* dctx, int ldc, const float* lse, float* delta,
                 void* dq, int lddq, void* dk, void* dv, int lddkv, void* workspace, int B, int H, int Lq, int Lk,
                 void* stream) {
  if (!q || !k || !v || !ctx || !dctx || !lse || !delta || !dq || !dk || !dv) return GXA_ERR_ARG;
  GXA_TRY(attn_bwd_bf16(q, ldq, k, v, ldkv, mask, kv_end, ctx, dctx, ldc, lse, delta, dq, lddq, dk, dv, lddkv, workspace,
                        B, H, Lq, Lk, reinterpret_cast<cudaStream_t>(stream)));
  return 0;
}

int gxa_attn_bwd_dropout(const void* q, int ldq, const void* k, const void* v, int ldkv, const uint8_t* mask,
                         const int32_t* kv_end, const void* ctx, const void* dctx, int ldc, const float* lse,
                         float* delta, void* dq, int lddq, void* dk, void* dv, int lddkv, void* workspace, int B, int H,
                         int Lq, int Lk, float p_drop, uint64_t seed, void* stream) {
  if (!q || !k || !v || !ctx || !dctx || !lse || !delta || !dq || !dk || !dv) return GXA_ERR_ARG;
  if (!(p_drop >= 0.f && p_drop < 1.f)) return GXA_ERR_ARG;
  GXA_TRY(attn_bwd_bf16(q, ldq, k, v, ldkv, mask, kv_end, ctx, dctx, ldc, lse, delta, dq, lddq, dk, dv, lddkv, workspace,
                        B, H, Lq, Lk, reinterpret_cast<cudaStream_t>(stream), nullptr,
                        make_drop_cfg(p_drop, seed, kDropStreamAttn)));
  return 0;
}

}  // extern "C"
