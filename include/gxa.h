/* gxa.h -- C ABI of libgxa_b200.so: the B200-native (sm_100a) gated cross-attention path.
 *
 * The reference (maihathm/MoCov2-Whisper-Flamingo) is pure Python and has no FFI; its
 * "operator API" for this path is three nn.Module classes in
 * models/gate_cross_attention.py.  Each entry point below names the reference interface it
 * replaces.  All pointers are raw CUDA device pointers owned by the caller (PyTorch's
 * allocator in the shipped host shell); the library never frees or retains them beyond the
 * call, takes the stream explicitly, keeps no thread-local state (backward runs on the
 * autograd engine thread) and never throws: every function returns 0 or a negative gxa_err.
 * There is no CPU fallback anywhere in this library.
 */
#ifndef GXA_H_
#define GXA_H_

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

typedef enum gxa_err {
  GXA_OK = 0,
  GXA_ERR_SHAPE = -1,        /* bad / unsupported shape (head_dim must be 64, D % 64 == 0) */
  GXA_ERR_ALIGN = -2,        /* pointer or leading dimension not 16-byte aligned */
  GXA_ERR_UNSUPPORTED = -3,  /* combination not compiled */
  GXA_ERR_CUDA = -4,         /* CUDA runtime error at launch (see gxa_last_cuda_error) */
  GXA_ERR_DRIVER = -5,       /* cuTensorMapEncodeTiled entry point unavailable */
  GXA_ERR_TMAP = -6,         /* tensor-map encode failed */
  GXA_ERR_ARG = -7           /* null pointer / inconsistent arguments */
} gxa_err;

enum { GXA_PREC_BF16 = 0, GXA_PREC_FP32 = 1 };

/* Problem description of one GatedCrossAttentionBlock call
 * (replaces the ctor args + forward shapes of models/gate_cross_attention.py:13,28). */
typedef struct gxa_dims {
  int32_t B, Lq, Lk, D, H;
  int32_t precision;   /* GXA_PREC_BF16: bf16 tensor-core contractions, fp32 LN/residual (== reference under
                          torch.autocast(bf16)); GXA_PREC_FP32: 3xbf16 split contractions, fp32 everywhere */
  int32_t has_xa;      /* 0 => FFN only (reference forward with xa=None, :39) */
  int32_t reserved;
  /* training-mode dropout (models/gate_cross_attention.py:16 nn.MultiheadAttention(dropout=p) -> on the attention
   * probabilities; :26 nn.Dropout(p) after ff.2).  p_drop = 0 disables it (eval mode).  The masks are a pure
   * function of (seed, element) -- csrc/gxa_rng.cuh -- so forward and backward must be called with the same seed;
   * give every (step, layer) its own seed.  Both precisions.  Drop probability = round(p * 32768) / 32768, p <= 0.968. */
  float p_drop;
  uint32_t reserved2;
  uint64_t seed;
} gxa_dims;

/* The 14 state_dict tensors of one block, fp32, in the reference's layout. */
typedef struct gxa_block_params {
  const float* attn_gate;     /* [1] */
  const float* ff_gate;       /* [1] */
  const float* in_proj_w;     /* [3D, D]  rows [0:D]=Wq [D:2D]=Wk [2D:3D]=Wv */
  const float* in_proj_b;     /* [3D] */
  const float* out_proj_w;    /* [D, D] */
  const float* out_proj_b;    /* [D] */
  const float* attn_ln_w;     /* [D] */
  const float* attn_ln_b;     /* [D] */
  const float* ff_ln_w;       /* [D] */
  const float* ff_ln_b;       /* [D] */
  const float* ff0_w;         /* [4D, D] */
  const float* ff0_b;         /* [4D] */
  const float* ff2_w;         /* [D, 4D] */
  const float* ff2_b;         /* [D] */
} gxa_block_params;

/* Gradients, same layout, fp32.  Every tensor is ACCUMULATED into (+=); zero them first for
 * plain autograd semantics.  Laid out by the host shell as one flat bucket per layer so that
 * the data-parallel all-reduce of a layer is a single NCCL call. */
typedef struct gxa_block_grads {
  float *attn_gate, *ff_gate, *in_proj_w, *in_proj_b, *out_proj_w, *out_proj_b, *attn_ln_w, *attn_ln_b, *ff_ln_w,
      *ff_ln_b, *ff0_w, *ff0_b, *ff2_w, *ff2_b;
} gxa_block_grads;

/* ---- capability / bookkeeping ------------------------------------------------------- */
int gxa_query_caps(int* sm_count, int* cc_major, int* cc_minor);
const char* gxa_version(void);
const char* gxa_last_cuda_error(void);
/* number of kernels this library has launched in the calling process (bench.py gpu_launches) */
long long gxa_launch_count(void);
/* Persistent kernels size their grids to the SMs they can expect to own.  While another resident kernel holds some
 * (the NCCL all-reduce of the data-parallel gradient buckets, reference train.py:319 DDPStrategy), the caller lowers
 * the budget so that no CTA of a statically partitioned grid has to wait for a second wave.  n_sms <= 0 restores the
 * whole device; returns the previous limit. */
int gxa_set_sm_limit(int n_sms);
/* The one collective of the path (the reference's DDP gradient all-reduce, train.py:319) is host-side by design:
 * mocov2_whisper_flamingo_b200/parallel.py drives NCCL through torch.distributed on a communication stream, one flat
 * fp32 bucket per layer.  There is no gxa_allreduce_layer entry point: the library never owns a communicator. */

/* Host-side evaluation of the training-mode dropout masks (reference: nn.MultiheadAttention(dropout=p) on the attention
 * probabilities and nn.Dropout(p) after ff.2, models/gate_cross_attention.py:16, :26).  The kernels never store a mask:
 * forward and backward evaluate one counter-based function of (seed, element) (csrc/gxa_rng.cuh); this entry point
 * evaluates the SAME compiled function on the host for a [rows, cols] matrix, so a caller (or a test) can reproduce the
 * mask a call used.  site: 1 = attention probabilities (rows = B*H*Lq, cols = Lk), 2 = FFN output (rows = B*Lq,
 * cols = D).  out[r*cols + c] = 0 where dropped, 1/(1-p) where kept.  No device work. */
int gxa_dropout_multipliers_host(float p_drop, uint64_t seed, int site, int rows, int cols, float* out);

/* ---- sizes ------------------------------------------------------------------------ */
/* bytes of the per-layer bf16 (or split) weight cache written by gxa_block_prepare_weights */
size_t gxa_block_wcache_bytes(const gxa_dims* d);
/* bytes of per-layer activations saved by forward for backward */
size_t gxa_block_saved_bytes(const gxa_dims* d);
/* bytes of scratch needed by one fwd or bwd call (shared by all layers) */
size_t gxa_block_workspace_bytes(const gxa_dims* d);

/* ---- the path ---------------------------------------------------------------------- */
/* fp32 master weights -> operand cache (call after every optimizer step). */
int gxa_block_prepare_weights(const gxa_dims* d, const gxa_block_params* p, void* wcache, void* stream);

/* Key-padding mask -> per-sample key extent.  mask: uint8 [B, Lk], nonzero = padded key
 * (torch/nn/functional.py:6409-6415 semantics); kv_end[b] = 1 + index of last unmasked key. */
int gxa_mask_prepare(const uint8_t* mask, int32_t* kv_end, int B, int Lk, void* stream);

/* Operand form of xa for the K/V projection (once per stack; xa is shared by all layers). */
size_t gxa_xa_operand_bytes(const gxa_dims* d);
int gxa_xa_prepare(const gxa_dims* d, const float* xa, void* xa_op, void* stream);

/* GatedCrossAttentionBlock.forward (models/gate_cross_attention.py:28-50); dropout per d->p_drop / d->seed.
 * x, y: fp32 [B*Lq, D] contiguous; xa_op from gxa_xa_prepare; mask/kv_end from gxa_mask_prepare
 * (mask may be NULL = no padding).  `saved` receives what backward needs (pass NULL for
 * inference).  y may not alias x. */
int gxa_block_fwd(const gxa_dims* d, const gxa_block_params* p, const void* wcache, const float* x, const void* xa_op,
                  const uint8_t* mask, const int32_t* kv_end, float* y, void* saved, void* workspace, void* stream);

/* autograd backward of the block.  dy: fp32 [B*Lq, D]; dx: fp32 out (may alias dy);
 * dxa: fp32 [B*Lk, D], ACCUMULATED (+=) because xa is shared across layers. */
int gxa_block_bwd(const gxa_dims* d, const gxa_block_params* p, const void* wcache, const float* x, const void* xa_op,
                  const uint8_t* mask, const int32_t* kv_end, const void* saved, const float* dy, float* dx, float* dxa,
                  const gxa_block_grads* g, void* workspace, void* stream);

/* ---- decode-time path (the reference has no decode loop -- this must equal the block's forward on the
 * same query rows).  The visual memory xa is fixed during decoding, so K|V = xa [Wk;Wv]^T + b are projected ONCE per
 * layer into a bf16 cache [B*Lk, 2D]; each step runs the block on the new query rows only (x: fp32 [B*Lq, D], usually
 * Lq = 1, or Lq = beams: the beams of a sample share its cache and the block has no interaction between query
 * positions).  bf16 precision, no dropout, nothing saved for backward. */
size_t gxa_block_kv_cache_bytes(const gxa_dims* d);
int gxa_block_kv_prefill(const gxa_dims* d, const gxa_block_params* p, const void* wcache, const void* xa_op,
                         void* kv_cache, void* stream);
size_t gxa_block_decode_workspace_bytes(const gxa_dims* d);
int gxa_block_decode(const gxa_dims* d, const gxa_block_params* p, const void* wcache, const float* x,
                     const void* kv_cache, const uint8_t* mask, const int32_t* kv_end, float* y, void* workspace,
                     void* stream);

/* ---- building blocks, exported for unit parity tests and for the stack driver ---------- */
/* LayerNorm.forward (gate_cross_attention.py:8-10): fp32 statistics, eps 1e-5. out fp32 or bf16. */
int gxa_layernorm_fwd(const float* x, const float* w, const float* b, void* out, int out_is_f32, float* mean,
                      float* rstd, int rows, int D, void* stream);
/* pos_enc(ln(x)) in one pass (models/av_net.py:83-84, :88-89 with models/utils.py:21-28): out[r] = LN(x[r]) + pe[r % T];
 * pe is the [T, D] fp32 head of the reference's positional table, rows = B * T.  Backward = gxa_layernorm_bwd. */
int gxa_layernorm_pe_fwd(const float* x, const float* w, const float* b, const float* pe, int T, float* out, float* mean,
                         float* rstd, int rows, int D, void* stream);
/* out = a + b + c, fp32, n % 4 == 0: `fused + audio_features + video_features` (models/av_net.py:106) in one pass. */
int gxa_sum3_f32(const float* a, const float* b, const float* c, float* out, size_t n, void* stream);
/* dx = dres + LNbwd(g); dw += ..; db += ..  (g fp32) */
int gxa_layernorm_bwd(const float* g, const float* x, const float* w, const float* mean, const float* rstd,
                      const float* dres, float* dx, float* dw, float* db, int rows, int D, void* stream);
/* nn.Linear on the stack boundary (audio_proj / video_proj, gate_cross_attention.py:56-57,64-65):
 * y = x W^T + b, fp32 in/out, contraction in the precision of d->precision. workspace >= gxa_linear_workspace_bytes */
size_t gxa_linear_workspace_bytes(int rows, int in_f, int out_f, int precision);
int gxa_linear_fwd(const float* x, const float* w, const float* b, float* y, int rows, int in_f, int out_f,
                   int precision, void* workspace, void* stream);
/* gxa_linear_fwd (bf16 contraction) with NO caller workspace and NO allocation of any kind (no cudaMalloc / cudaFree, no
 * allocating library call).  This entry point has no counterpart in the reference: it was added
 * for callers that cannot allocate (CUDA-graph capture, tight serving loops, harnesses that forbid
 * cudaMalloc).  The bf16 operand copies live in a file-scope `static __device__` array inside the library
 * (csrc/gxa_noalloc.cu) of GXA_LINEAR_FWD_STATIC_MAX_BYTES = 8 MiB, which every GPU that uses the entry point holds
 * until the process exits (the first call on a device is the set-up: symbol-address lookup, and -- under CUDA's lazy
 * module loading -- the point where the driver materialises the array; make one warm call before a capture / timed
 * region).  Shapes that need 2 * (round_up(rows * in_f, 128) + out_f * in_f) bytes above that limit return
 * GXA_ERR_SHAPE with a message in gxa_last_cuda_error().  All calls on a device share the one array, which keeps the last call's contents (nothing in it
 * needs to start from zero: both operand regions are overwritten before they are read): calls on one stream are ordered,
 * calls that overlap on the GPU (different streams without an ordering between them) RACE -- serialise them.
 * Same result as gxa_linear_fwd(..., GXA_PREC_BF16, workspace, stream). */
#define GXA_LINEAR_FWD_STATIC_MAX_BYTES (8u << 20)
int gxa_linear_fwd_static(const float* x, const float* w, const float* b, float* y, int rows, int in_f, int out_f,
                          void* stream);
/* dx = dy W (may be NULL); dw += dy^T x; db += colsum(dy) */
int gxa_linear_bwd(const float* dy, const float* x, const float* w, float* dx, float* dw, float* db, int rows, int in_f,
                   int out_f, int precision, void* workspace, void* stream);

/* Raw tcgen05 GEMM, C[M,N] = A * B^T with fused epilogue (unit tests / profiling).
 * a_mn / b_mn: operand stored [K, rows] instead of [rows, K].  epi: 0 bf16(+bias), 1 fp32(+bias, accumulate),
 * 2 gated residual, 3 bias+GELU, 4 GELU backward.  bn: 0 auto, 128, 256.  k_splits: 0 auto. */
int gxa_gemm_bf16(const void* A, int lda, int a_mn, const void* B, int ldb, int b_mn, int M, int N, int K, int epi,
                  void* C, int ldc, void* C2, int ldc2, const float* bias, const void* aux, int ld_aux,
                  const float* gate, int accumulate, int bn, int k_splits, void* stream);

/* Few-row linear layer of the decode step (the nn.Linear / in-projection calls of
 * models/gate_cross_attention.py:40, :43-44 when only R = batch x beams <= 256 query rows are live):
 * out[R, N] = epilogue(X[R, K] W[N, K]^T + bias), X and W bf16, K % 64 == 0, K >= 128, N % 4 == 0.
 * epi 0: out bf16;  1: out bf16 = gelu_erf(bf16(.));  2: out f32 = res + tanh(*gate) * (.)  (res f32 [R, ldres]).
 * Weight tile on the M side of the MMA, K split over a thread-block cluster, partials reduced through distributed
 * shared memory (csrc/gxa_skinny.cu).  gxa_block_decode uses it for all four projections. */
int gxa_linear_few_rows_bf16(const void* X, int ldx, const void* W, int ldw, int R, int N, int K, int epi, void* out,
                             int ldo, const float* bias, const float* res, int ldres, const float* gate, void* stream);

/* Masked-softmax attention over variable-length keys, head_dim 64 (replaces the SDPA call at
 * torch/nn/functional.py:6676-6684).  q: bf16 [B*Lq, ldq]; k,v: bf16 [B*Lk, ldkv] (head h = columns
 * [64h, 64h+64)); ctx: bf16 [B*Lq, ldc]; lse: fp32 [B, H, Lq] (natural log). */
int gxa_attn_fwd(const void* q, int ldq, const void* k, const void* v, int ldkv, const uint8_t* mask,
                 const int32_t* kv_end, void* ctx, int ldc, float* lse, int B, int H, int Lq, int Lk, void* stream);
/* workspace: gxa_attn_bwd_workspace_bytes() of scratch (fp32 dQ accumulator of the tcgen05 path); NULL selects the
 * scratch-free legacy kernels. */
/* the same with dropout on the attention probabilities (p_drop in [0,1), mask = f(seed, b, h, q, key)) */
int gxa_attn_fwd_dropout(const void* q, int ldq, const void* k, const void* v, int ldkv, const uint8_t* mask,
                         const int32_t* kv_end, void* ctx, int ldc, float* lse, int B, int H, int Lq, int Lk,
                         float p_drop, uint64_t seed, void* stream);
size_t gxa_attn_bwd_workspace_bytes(int B, int H, int Lq);
int gxa_attn_bwd(const void* q, int ldq, const void* k, const void* v, int ldkv, const uint8_t* mask,
                 const int32_t* kv_end, const void* ctx, const void* dctx, int ldc, const float* lse, float* delta,
                 void* dq, int lddq, void* dk, void* dv, int lddkv, void* workspace, int B, int H, int Lq, int Lk,
                 void* stream);
int gxa_attn_bwd_dropout(const void* q, int ldq, const void* k, const void* v, int ldkv, const uint8_t* mask,
                         const int32_t* kv_end, const void* ctx, const void* dctx, int ldc, const float* lse,
                         float* delta, void* dq, int lddq, void* dk, void* dv, int lddkv, void* workspace, int B, int H,
                         int Lq, int Lk, float p_drop, uint64_t seed, void* stream);

#ifdef __cplusplus
}
#endif
#endif /* GXA_H_ */
