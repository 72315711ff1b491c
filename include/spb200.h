/*
 * spb200.h - C ABI of libspb200.so: the B200 (sm_100a) attention-block hot path.
 *
 * The reference (prashantpandeygit/solvingpapers) has NO native code and NO plugin
 * registry; its "operator interface" for this path is a handful of Python function
 * signatures.  Each entry point below cites the reference function it replaces
 * (file:line = raw line of the .ipynb JSON under the reference checkout).  A host
 * binding (ctypes today, jax.ffi where JAX exists - see INTEGRATION.md) maps the
 * notebook's functions 1:1 onto these symbols.
 *
 * Conventions
 *  - extern "C", plain pointers and sizes; no torch / XLA types.
 *  - every pointer is a DEVICE pointer unless said otherwise; the caller owns all
 *    buffers (inputs, outputs and the workspace).  Nothing here allocates.
 *  - work is enqueued on the cudaStream_t passed as `stream`; no call synchronises.
 *  - return 0 = ok; < 0 = invalid argument / unsupported shape (see SPB_E_*);
 *    > 0 = cudaError_t (or 1000 + ncclResult_t).  spb_last_error() has the text.
 *    Unsupported shapes are an ERROR, never a silent slow path; there is no CPU
 *    fallback anywhere in the library.
 *  - activations are fp32 or bf16 (`dtype`); weight GRADIENTS are always fp32.
 *  - N = B*T tokens; matrices are row-major with an explicit leading dimension.
 */
#ifndef SPB200_H
#define SPB200_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

typedef void* spb_stream_t; /* cudaStream_t */

enum { SPB_F32 = 0, SPB_BF16 = 1 };
enum { SPB_NORM_NONE = 0, SPB_NORM_RMS = 1, SPB_NORM_LAYER = 2 };
enum { SPB_ROPE_NONE = 0, SPB_ROPE_TABLE = 1 };
enum { SPB_W_IN_OUT = 0 /* jax/flax: y = x @ W[in,out] */, SPB_W_OUT_IN = 1 /* torch: y = x @ W[out,in]^T */ };
enum { SPB_IMPL_AUTO = 0, SPB_IMPL_SIMT = 1, SPB_IMPL_TCGEN05 = 2 };

enum {
  SPB_OK = 0,
  SPB_E_INVALID = -1,     /* null pointer, negative size, misaligned pointer */
  SPB_E_UNSUPPORTED = -2, /* shape/dtype this build has no kernel for */
  SPB_E_STATE = -3,       /* e.g. communicator not initialised, driver entry point missing */
  SPB_E_LIMIT = -4        /* size beyond a documented fixed limit */
};

/* One attention block.  Mirrors the static arguments of
 *   llama3/LLaMA-jax.ipynb:809  attention(params,x,mask,freqs_cis,n_heads,n_kv_heads,cache=None,position=0)
 *   + :536 rms_norm(x, weight, eps=1e-6) as called at :881,
 *   gpt/gpt-jax.ipynb:321-357 CausalSelfAttention + :414 nn.LayerNorm,
 *   gemma/gemma.ipynb:218-259 MQA + :139-150 RMSNorm,
 *   vision transformer/ViT.ipynb:205-208 LayerNorm + nn.MultiheadAttention.          */
typedef struct spb_block_desc {
  int32_t B, T, C;        /* x is [B,T,C] */
  int32_t Hq, Hkv, hd;    /* q heads, kv heads (Hq % Hkv == 0), head dim; q width Hq*hd need not equal C (Gemma) */
  int32_t dtype;          /* SPB_F32 | SPB_BF16: activations AND weights */
  int32_t norm_kind;      /* SPB_NORM_* */
  float   norm_eps;       /* inside the sqrt (llama3:538) */
  int32_t norm_upcast;    /* 1: normalise in fp32, round to dtype, THEN multiply weight (gemma:148-150) */
  int32_t rope_kind;      /* SPB_ROPE_TABLE: q,k pairs (2j,2j+1) <- 2x2 matrix from `rope` table */
  int32_t causal;         /* mask is not None => causal; -1e9 add / -1e4 replace / -inf are all
                             exp()==0 exactly, so masked tiles are skipped */
  float   scale;          /* scores multiplied by this (1/sqrt(hd); Gemma 1/sqrt(768); ViT hd^-0.5) */
  int32_t weight_layout;  /* SPB_W_IN_OUT | SPB_W_OUT_IN */
  int32_t qkv_bias;       /* 1: bq,bk,bv present (ViT in_proj_bias) */
  int32_t out_bias;       /* 1: bo present (GPT proj bias, ViT out_proj.bias) */
  int32_t attn_impl;      /* SPB_IMPL_*: AUTO = tcgen05 flash kernels when dtype==bf16 and hd in {64,128,256}, the tcgen05 GEMM
                             pipeline of spb_attn_wide_* when dtype==bf16 and hd > 256, else SIMT */
  int32_t gemm_impl;      /* SPB_IMPL_*: AUTO = tcgen05 GEMM; for dtype==f32 on two-term bf16 splits of both operands (three
                             products in one GEMM, fp32 accumulation: within the 1e-4 tolerance), FFMA GEMM where a shape or a
                             bias rules that out.  SIMT = FFMA GEMM always */
  int32_t residual;       /* 0: y = block(x) (what `attention` returns, llama3:809-829).  1: y = x + block(x), the caller's
                             `h = x + attention(...)` (llama3:885, gpt:420, gemma:335, ViT:222) fused into the output
                             projection's epilogue; spb_block_bwd then returns dx = d block/dx + dy (fused into the norm
                             VJP's store).  SURVEY section 8 row f1 */
} spb_block_desc;

/* Weights.  W_IN_OUT: wq is [C, Hq*hd] with row stride ldw_qkv, wk/wv are [C, Hkv*hd] with the same stride
 * (separate arrays: ldw = their own width is NOT assumed - pass three pointers into one packed [C, Ntot] array,
 * e.g. flax qkv/kernel [C,3C]: wk = wq + C, ldw_qkv = 3C; or three separate arrays via ldw_q/ldw_k/ldw_v).
 * W_OUT_IN: wq is [Hq*hd, C], wk/wv [Hkv*hd, C], row stride C (torch in_proj_weight [3E,E]: wk = wq + E*E).
 * wo: W_IN_OUT [Hq*hd, C]; W_OUT_IN [C, Hq*hd].                                                              */
typedef struct spb_block_params {
  const void* norm_w;  /* [C] dtype; RMS weight / LayerNorm scale */
  const void* norm_b;  /* [C] dtype; LayerNorm bias or NULL */
  const void* wq; const void* wk; const void* wv;
  int64_t ldw_q, ldw_k, ldw_v;
  const void* bq; const void* bk; const void* bv; /* [Hq*hd], [Hkv*hd], [Hkv*hd] dtype or NULL */
  const void* wo; int64_t ldw_o;
  const void* bo;      /* [C] dtype or NULL */
  const float* rope;   /* [T, hd/2, 4] fp32 (a,b,c,d): o_even = a*x_even + b*x_odd ; o_odd = c*x_even + d*x_odd.
                          Llama (llama3:563-567,592-601): (cos,-sin,sin,cos) with angle t*theta^(-j/hd).
                          Gemma (gemma:182-200): (cos,cos,-sin,sin), one angle per position.  NULL if ROPE_NONE */
} spb_block_params;

/* Gradients (all fp32 except dx which has the activation dtype).  Same shapes/strides as the weights.
 * Any pointer may be NULL to skip that gradient.  Buffers are OVERWRITTEN (not accumulated).           */
typedef struct spb_block_grads {
  void*  dx;            /* [B,T,C] dtype */
  float* dnorm_w; float* dnorm_b;
  float* dwq; float* dwk; float* dwv;
  int64_t ldw_q, ldw_k, ldw_v;
  float* dbq; float* dbk; float* dbv;
  float* dwo; int64_t ldw_o;
  float* dbo;
} spb_block_grads;

/* Byte offsets inside the caller-owned workspace (residuals of the forward that the backward reads, and scratch). */
typedef struct spb_ws_layout {
  size_t total;
  size_t xhat;     /* [N, C] dtype: normalised input = GEMM operand                                  */
  size_t stats;    /* [N, 2] f32: (mean, rstd)                                                       */
  size_t qkv;      /* [N, ld_qkv] dtype, post-RoPE: q cols [0,Hq*hd), k [Hq*hd, +Hkv*hd), v after    */
  size_t attn_o;   /* [N, Hq*hd] dtype: attention output before the out projection                   */
  size_t lse;      /* [B, Hq, T] f32 log-sum-exp of the scaled scores                                */
  size_t delta;    /* [B, Hq, T] f32 rowsum(dO*O)                 (backward)                         */
  size_t dattn_o;  /* [N, Hq*hd] dtype                            (backward)                         */
  size_t dqkv;     /* [N, ld_qkv] dtype                           (backward)                         */
  size_t dq_acc;   /* [N, Hq*hd] f32 dQ accumulator of the tcgen05 flash backward                    */
  size_t dxhat;    /* [N, C] dtype                                (backward)                         */
  size_t partial;  /* f32 partial sums (norm-weight / bias column reductions)                         */
  size_t partial_bytes;
  int64_t ld_qkv;  /* (Hq + 2*Hkv) * hd */
  size_t attn_scratch;        /* scratch of the wide-head attention path (bf16, hd > 256); 0 bytes otherwise   */
  size_t attn_scratch_bytes;
  size_t gemm_scratch;        /* fp32 mode: bf16 hi/lo splits of the operands of the block's GEMMs (tensor-core fp32 path)     */
  size_t gemm_scratch_bytes;
} spb_ws_layout;

/* ---- library ------------------------------------------------------------------------------------ */
const char* spb_version(void);
const char* spb_last_error(void);                       /* thread-local text of the last non-zero return */
int spb_device_info(int32_t* sm_count, int32_t* cc_major, int32_t* cc_minor, size_t* smem_optin);
/* number of kernels this library has launched in this process (bench.py's gpu_launches) */
int64_t spb_launch_count(void);

/* Optional stage timing for bench.py's per-kernel roofline (no counterpart in the reference, which has no profiler
 * hooks: SURVEY section 5).  spb_prof_enable(max_steps) creates CUDA events once (max_steps <= 0 turns it off);
 * afterwards every spb_block_fwd starts a new step and both block calls record an event on the caller's stream at each
 * stage boundary.  After synchronising, spb_prof_read(step, ...) returns the number of stages and fills names[]/ms[]
 * with the device time between consecutive boundaries.  Not thread-safe; measurement aid only.                      */
int spb_prof_enable(int max_steps);
int spb_prof_steps(void);
int spb_prof_read(int step, int max_n, const char** names, float* ms);

/* ---- per-stage entry points (per-kernel parity tests, ncu) -------------------------------------- */

/* llama3:536-538 rms_norm / gpt:414 nn.LayerNorm / gemma:139-150 RMSNorm / ViT:205 nn.LayerNorm.
 * y[N,C] = norm(x[N,C]); stats[N,2] = (mean, rstd) saved for the backward (may be NULL).            */
int spb_norm_fwd(int dtype, int norm_kind, int upcast, float eps, const void* x, const void* w, const void* b,
                 void* y, float* stats, int64_t N, int32_t C, spb_stream_t stream);
/* VJP of the above.  dw/db are fp32 [C]; `partial` is scratch of spb_norm_bwd_partial_bytes(N,C).    */
size_t spb_norm_bwd_partial_bytes(int64_t N, int32_t C);
int spb_norm_bwd(int dtype, int norm_kind, int upcast, const void* x, const void* w, const float* stats,
                 const void* dy, void* dx, float* dw, float* db, void* partial, int64_t N, int32_t C,
                 spb_stream_t stream);

/* llama3:592-601 apply_rotary_emb (and gemma:241-242 M_t @ q_t) IN PLACE on x[N, heads*hd] (row stride ld):
 * token n has position n % T; pair j of every head uses table[t][j].  transpose=1 applies the VJP (M^T).   */
int spb_rope_apply(int dtype, void* x, int64_t ld, int64_t N, int32_t T, int32_t heads, int32_t hd,
                   const float* table, int transpose, spb_stream_t stream);

/* The jnp.dot / nn.Dense / nn.Linear of llama3:812-814,829, gpt:329-330, gemma:232-236, and their VJPs.
 * D[M,N] (ldd) = opA(A) . opB(B) (+ bias[N]) (+ D if accumulate).
 *   transA=0: A stored [M,K] (lda).   transA=1: A stored [K,M] (lda).
 *   transB=0: B stored [K,N] (ldb).   transB=1: B stored [N,K] (ldb).
 * in_dtype: A,B,bias element type; out_dtype: D element type (SPB_F32 for weight gradients).
 * impl: AUTO = tcgen05 for bf16 inputs, SIMT for fp32 inputs.  tcgen05 needs 16-byte aligned pointers and
 * lda/ldb multiples of 8 elements; M,N,K are arbitrary (TMA zero-fills the tails).                       */
int spb_gemm(int in_dtype, int out_dtype, int impl, int64_t M, int64_t N, int64_t K, const void* A, int64_t lda,
             int transA, const void* Bm, int64_t ldb, int transB, void* D, int64_t ldd, const void* bias,
             int accumulate, spb_stream_t stream);

/* bias gradient: out[c] = sum_n x[n, c], x is [N, cols] (ld), out fp32; partial = scratch (same size rule). */
int spb_colsum(int dtype, const void* x, int64_t ld, int64_t N, int32_t cols, float* out, void* partial,
               spb_stream_t stream);

/* spb_colsum without caller-provided scratch.  No counterpart in the reference (it has no native code); it exists for
 * hosts that have no way to hand this library scratch memory: an XLA/jax.ffi handler that was registered without a
 * scratch result (XLA owns every buffer and the handler may not cudaMalloc), or a process whose own pool has already
 * claimed all of HBM so that any further driver allocation fails or stalls the device.  The two-stage reduction's
 * partial sums live in a fixed 4 MiB array inside the library image (every GPU that loads the library carries it until
 * the process exits).  Limits: cols <= 8192 (wider matrices return SPB_E_LIMIT - use spb_colsum); any N.
 * This call performs no cudaMalloc/cudaFree and calls no allocating routine.  The first call on each device looks up
 * the array's device address (cudaGetSymbolAddress) and caches it - make that first call before starting any stream
 * capture.  Every call on a device shares the one array: calls that can overlap on the GPU (different streams, or
 * different host threads) race on it and must be ordered by the caller.  Nothing in the array has to start at zero
 * (stage one overwrites every partial that stage two reads), so stale contents from the previous call are harmless.   */
int spb_colsum_static(int dtype, const void* x, int64_t ld, int64_t N, int32_t cols, float* out, spb_stream_t stream);

/* llama3:820-828 (repeat_kv + transpose + QK^T/sqrt(hd) + mask + softmax + PV + merge heads), never
 * materialising scores or repeated K/V.  q[N,Hq*hd] (ldq), k,v[N,Hkv*hd] (ldk,ldv), o[N,Hq*hd] (ldo),
 * lse[B,Hq,T] fp32.  kv head of q head h is h / (Hq/Hkv)  (llama3:626-627 consecutive grouping).        */
int spb_flash_fwd(int dtype, int impl, int32_t B, int32_t T, int32_t Hq, int32_t Hkv, int32_t hd, int causal,
                  float scale, const void* q, int64_t ldq, const void* k, int64_t ldk, const void* v, int64_t ldv,
                  void* o, int64_t ldo, float* lse, spb_stream_t stream);
/* VJP of the above (jax.value_and_grad at llama3:995 has no hand-written counterpart).
 * delta[B,Hq,T] fp32 scratch; dq_acc: fp32 [N, Hq*hd] scratch (tcgen05 path only, may be NULL for SIMT). */
int spb_flash_bwd(int dtype, int impl, int32_t B, int32_t T, int32_t Hq, int32_t Hkv, int32_t hd, int causal,
                  float scale, const void* q, int64_t ldq, const void* k, int64_t ldk, const void* v, int64_t ldv,
                  const void* o, int64_t ldo, const void* d_o, int64_t lddo, const float* lse, float* delta,
                  void* dq, int64_t lddq, void* dk, int64_t lddk, void* dv, int64_t lddv, float* dq_acc,
                  spb_stream_t stream);

/* Attention for head dims the fused kernels cannot hold on chip (bf16, hd > 256, hd % 8 == 0; gemma/gemma.ipynb:218-259
 * literal: 2 q heads and one shared k/v of width 768).  Same arguments and results as spb_flash_fwd / spb_flash_bwd plus
 * `scratch` (spb_attn_wide_scratch_bytes, 256-byte aligned): per (batch, q head) and block of 2048 query rows the scores
 * of the keys the block can see go through HBM once - S = Q K^T, softmax, O = P V, and in the backward dP = dO V^T,
 * dV += P^T dO, dK += dS^T Q, dQ = dS K - with every contraction on the tcgen05 GEMM.  Scratch is O(2048 * T + T * hd),
 * independent of B and the head counts; nothing O(T^2) is kept between forward and backward.                         */
size_t spb_attn_wide_scratch_bytes(int32_t B, int32_t T, int32_t Hq, int32_t Hkv, int32_t hd);
int spb_attn_wide_fwd(int32_t B, int32_t T, int32_t Hq, int32_t Hkv, int32_t hd, int causal, float scale, const void* q, int64_t ldq,
                      const void* k, int64_t ldk, const void* v, int64_t ldv, void* o, int64_t ldo, float* lse, void* scratch,
                      spb_stream_t stream);
int spb_attn_wide_bwd(int32_t B, int32_t T, int32_t Hq, int32_t Hkv, int32_t hd, int causal, float scale, const void* q, int64_t ldq,
                      const void* k, int64_t ldk, const void* v, int64_t ldv, const void* o, int64_t ldo, const void* d_o, int64_t lddo,
                      const float* lse, float* delta, void* dq, int64_t lddq, void* dk, int64_t lddk, void* dv, int64_t lddv,
                      void* scratch, spb_stream_t stream);

/* ---- the block ---------------------------------------------------------------------------------- */
int spb_block_ws_layout(const spb_block_desc* d, spb_ws_layout* out);
/* llama3:881  attention(params['attention'], rms_norm(x, params['attention_norm']), mask, freqs_cis, ...)
 * y[B,T,C]; residuals (x-hat, stats, post-RoPE q|k|v, attention output, lse) are left in `ws`.
 * The (k, v) "new_cache" of llama3:819 is the k / v column range of ws+layout.qkv.                    */
int spb_block_fwd(const spb_block_desc* d, const spb_block_params* p, const void* x, void* y, void* ws,
                  spb_stream_t stream);
/* VJP of spb_block_fwd given dy[B,T,C]; reads the residuals spb_block_fwd left in `ws`.
 * If `comm` != NULL the weight gradients are summed over ranks (ncclAllReduce, fp32) before returning
 * control of the stream: dwo's all-reduce is launched on a side stream as soon as its GEMM is done and
 * overlaps the attention backward.                                                        */
int spb_block_bwd(const spb_block_desc* d, const spb_block_params* p, const void* x, const void* dy, void* ws,
                  const spb_block_grads* g, void* comm, spb_stream_t stream);

/* ---- the gated feed-forward sub-block ---------------------------------------- */
/* llama3:886,890  ffn_output = feed_forward(params['ffn'], rms_norm(h, params['ffn_norm']));  out = h + ffn_output
 * llama3:854-855  feed_forward(p, x) = dot(silu(dot(x, p['w3'])) * dot(x, p['w1']), p['w2'])
 * Weights as init_ffn_weights stores them (llama3:713-718): w1, w3 [C, H], w2 [H, C] for SPB_W_IN_OUT (H = 4*C in the
 * notebook); SPB_W_OUT_IN: w1, w3 [H, C], w2 [C, H].                                                              */
#define SPB_ACT_SWIGLU    0  /* silu(x.w3) * (x.w1), w2                                   llama3:854-855                   */
#define SPB_ACT_GEGLU     1  /* gelu_erf(x.w3) * (x.w1), w2: w3 = linear_layer1 (gate), w1 = linear_layer2 (value),
                                w2 = linear_layer3, torch [out,in] weights                  gemma:277-285                    */
#define SPB_ACT_GELU_TANH 2  /* plain MLP gelu_tanh(x.w1 + b1).w2 + b2 (flax nn.gelu default approximate=True); w3 unused
                                                                                           gpt:381-388                      */

typedef struct spb_ffn_desc {
  int32_t N, C, H;        /* tokens (B*T), model width, hidden width */
  int32_t dtype;          /* SPB_F32 | SPB_BF16: activations AND weights */
  int32_t norm_kind;      /* SPB_NORM_*: the ffn_norm in front (SPB_NORM_NONE = plain feed_forward(x)) */
  float   norm_eps;
  int32_t norm_upcast;
  int32_t act;            /* SPB_ACT_* */
  int32_t weight_layout;  /* SPB_W_IN_OUT | SPB_W_OUT_IN */
  int32_t residual;       /* 1: y = h + ffn(norm(h)) (llama3:890) in the w2 GEMM's epilogue; backward dx += dy */
  int32_t gemm_impl;      /* SPB_IMPL_* */
} spb_ffn_desc;

typedef struct spb_ffn_params {
  const void* norm_w; const void* norm_b;     /* [C] dtype */
  const void* w1; const void* w3; int64_t ldw_13;  /* same row stride for both; packed [C, 2H] (w3 = w1 + H) is one GEMM */
  const void* w2; int64_t ldw_2;
  const void* b1;     /* [H] dtype or NULL: bias of the up projection (SPB_ACT_GELU_TANH: flax nn.Dense default use_bias) */
  const void* b2;     /* [C] dtype or NULL: bias of the down projection                                                  */
} spb_ffn_params;

typedef struct spb_ffn_grads {
  void*  dx;                                  /* [N, C] dtype */
  float* dnorm_w; float* dnorm_b;             /* [C] */
  float* dw1; float* dw3; int64_t ldw_13;     /* fp32, same layout as the weights */
  float* dw2; int64_t ldw_2;
  float* db1; float* db2;                     /* [H], [C] fp32 or NULL */
} spb_ffn_grads;

/* Byte offsets inside the caller-owned workspace */
typedef struct spb_ffn_ws {
  size_t total;
  size_t xhat;     /* [N, C] dtype normalised input (unused when norm_kind == NONE)        */
  size_t stats;    /* [N, 2] f32                                                            */
  size_t g13;      /* [N, 2H] dtype: x.w1 in columns [0,H), x.w3 in [H,2H)  (residual); [N, H] for the plain MLP */
  size_t act;      /* [N, H] dtype: silu(g3)*g1 = the w2 GEMM's operand     (residual)      */
  size_t dact;     /* [N, H] dtype                                           (backward)     */
  size_t dg13;     /* [N, 2H] dtype                                          (backward)     */
  size_t dxhat;    /* [N, C] dtype                                           (backward)     */
  size_t partial;  /* f32 partial sums of the norm-weight gradient                          */
  size_t partial_bytes;
} spb_ffn_ws;

int spb_ffn_ws_layout(const spb_ffn_desc* d, spb_ffn_ws* out);
int spb_ffn_fwd(const spb_ffn_desc* d, const spb_ffn_params* p, const void* x, void* y, void* ws, spb_stream_t stream);
/* VJP given dy [N, C]; with `comm` the weight gradients are summed over ranks like spb_block_bwd's. */
int spb_ffn_bwd(const spb_ffn_desc* d, const spb_ffn_params* p, const void* x, const void* dy, void* ws,
                const spb_ffn_grads* g, void* comm, spb_stream_t stream);

/* ---- the loss head's memory-bound half --------------------------------------- */
/* llama3:959-967  loss = -mean(take_along_axis(log_softmax(logits), targets[:, None], axis=1)) over logits [N, V] (row pitch ld
 * >= V elements; V = 50257 in the notebook).  lse[N] and row_nll[N] are fp32 outputs (lse is the backward's residual);
 * `loss` (one fp32, may be NULL) receives the mean.  The logits GEMM of llama3:930 and its backward GEMMs are spb_gemm
 * (the tcgen05 path needs leading dimensions that are multiples of 8: pad the vocabulary axis of the weight and of logits). */
int spb_nll_fwd(int dtype, const void* logits, int64_t ld, const int64_t* targets, int64_t N, int32_t V, float* lse,
                float* row_nll, float* loss, spb_stream_t stream);
/* dlogits[n, j] = (softmax(logits)[n, j] - [j == targets[n]]) * dloss / N; may run in place (dlogits == logits, ldd == ld). */
int spb_nll_bwd(int dtype, const void* logits, int64_t ld, const int64_t* targets, int64_t N, int32_t V, const float* lse,
                float dloss, void* dlogits, int64_t ldd, spb_stream_t stream);

/* The whole loss head, llama3:930 + :959-967: loss = nll(dot(h, w), targets) with h [N, C] (row stride ldh) and the output weight
 * w [C, V] as the notebook stores it (row stride ldw; V = 50257 there: NOT a multiple of 8 - the library pads the vocabulary axis of
 * a private copy in `ws` when the pitch is not TMA-addressable, the caller pads nothing).  bf16.  The [N, V] logits are never
 * written: the forward GEMM's epilogue keeps the online (max, sum exp, target logit) of llama3:961-965 per row and 256-column tile
 * and a small kernel merges them (lse, row_nll [N] fp32, loss = mean); the backward recomputes each logits tile from h, w and lse
 * and emits dlogits = (softmax - onehot) * dloss / N in bf16 straight from the epilogue for the two gradient GEMMs
 * (dh [N, C] dtype or NULL; dw [C, V] fp32, row stride lddw, or NULL).  ws: spb_loss_head_ws_bytes(N, C, V), 256-byte aligned. */
size_t spb_loss_head_ws_bytes(int64_t N, int32_t C, int32_t V);
int spb_loss_head_fwd(int dtype, const void* h, int64_t ldh, const void* w, int64_t ldw, const int64_t* targets, int64_t N, int32_t C,
                      int32_t V, float* lse, float* row_nll, float* loss, void* ws, spb_stream_t stream);
int spb_loss_head_bwd(int dtype, const void* h, int64_t ldh, const void* w, int64_t ldw, const int64_t* targets, int64_t N, int32_t C,
                      int32_t V, const float* lse, float dloss, void* dh, int64_t lddh, float* dw, int64_t lddw, void* ws,
                      spb_stream_t stream);

/* ---- KV-cache decode ----------------------------------------------------------- */
/* llama3:815-829 with cache != None: Tq new query rows per sequence (already rotated) against Tk cached keys/values
 * (the new tokens' own k, v included: the caller - or spb_block_decode - appended them first).  No mask: every query row
 * sees all Tk keys, which is what the notebook computes for the shapes its cache branch accepts (Tq = 1, or mask=None).
 * q [B*Tq, ldq], head h at column h*hd; caches: token j of batch b at b*batch_stride + j*ld (elements), kv head g at
 * column g*hd; o [B*Tq, ldo]; lse [B, Hq, Tq] fp32 or NULL; scratch = spb_attn_decode_scratch_bytes(...) bytes.
 * fp32 or bf16, hd in {64, 128, 256}.  Bound by HBM bandwidth: reads the K and V caches once.                          */
size_t spb_attn_decode_scratch_bytes(int32_t B, int32_t Tq, int32_t Hq, int32_t Hkv, int32_t hd, int32_t Tk);
int spb_attn_decode(int dtype, int32_t B, int32_t Tq, int32_t Hq, int32_t Hkv, int32_t hd, int32_t Tk, float scale, const void* q,
                    int64_t ldq, const void* k_cache, int64_t ldk, int64_t k_batch_stride, const void* v_cache, int64_t ldv,
                    int64_t v_batch_stride, void* o, int64_t ldo, float* lse, void* scratch, spb_stream_t stream);

/* One decode step of the block: attention(params, norm(x), mask, freqs_cis, n_heads, n_kv_heads, cache=(k_cache, v_cache),
 * position=cache_len) of llama3:809-829 for x [B, Tq, C] with d->T = Tq.  The new tokens' k (rotated at positions
 * cache_len .. cache_len+Tq-1) and v are APPENDED IN PLACE into the caches [B, Tmax, Hkv, hd] (the notebook's
 * jnp.concatenate, :817-818, without copying the old rows); afterwards the caches hold cache_len + Tq valid rows.
 * p->rope must cover positions [0, cache_len + Tq).  d->causal with Tq > 1 is SPB_E_UNSUPPORTED (the notebook's own code
 * cannot express it: its mask slice does not broadcast against [.., Tq, Tk]).  SPB_W_IN_OUT weights; M = B*Tq <= 64.
 * Workspace offsets: spb_block_decode_ws_layout (Tmax bounds the attention scratch).                                   */
typedef struct spb_decode_ws {
  size_t total;
  size_t xhat;       /* [M, C] dtype                                */
  size_t qkv;        /* [M, (Hq+2Hkv)*hd] dtype, q rotated          */
  size_t attn_o;     /* [M, Hq*hd] dtype                            */
  size_t gemm_part;  /* fp32 K-split partial sums of the projections */
  size_t attn_part;  /* fp32 partial (O, m, l) of the key splits    */
} spb_decode_ws;
int spb_block_decode_ws_layout(const spb_block_desc* d, int32_t Tmax, spb_decode_ws* out);
int spb_block_decode(const spb_block_desc* d, const spb_block_params* p, const void* x, void* k_cache, void* v_cache,
                     int32_t cache_len, int32_t Tmax, void* y, void* ws, spb_stream_t stream);
/* The same step with the cache length in DEVICE memory (one int32 the caller owns and initialises): the kernels read it,
 * and the step's last kernel advances it by Tq (never past Tmax).  No launch parameter depends on the length - the grids
 * are sized for Tmax and share the valid key tiles evenly - so a step recorded once into a CUDA graph
 * (llama3:499-511 `generate` loops over it) is replayed for every token: one driver call instead of seven launches, which
 * is what bounds the eager step.  p->rope must cover positions [0, Tmax).  A step that would pass Tmax writes no cache row
 * and attends to Tmax keys: keep the host's own count and do not replay past the capacity.                                */
int spb_block_decode_dev(const spb_block_desc* d, const spb_block_params* p, const void* x, void* k_cache, void* v_cache,
                         int32_t* cache_len_dev, int32_t Tmax, void* y, void* ws, spb_stream_t stream);

/* ---- data parallelism: batch-sharded, weight gradients all-reduced (no reference precedent) ---- */
int spb_comm_unique_id(void* out128);                        /* HOST buffer, 128 bytes (ncclUniqueId)      */
int spb_comm_init_rank(const void* id128, int nranks, int rank, void** comm_out);
int spb_comm_allreduce_f32(void* comm, float* const* ptrs, const size_t* counts, int n, spb_stream_t stream);
int spb_comm_destroy(void* comm);
/* one communicator per device 0..ndev-1 of THIS process (ncclCommInitAll): the single-process harness of SURVEY 8b/8e.
 * A communicator handle is used by one host thread at a time (XLA calls from one thread per device: one handle each).  */
int spb_comm_init_all(int ndev, void** comms_out);
/* Gradient buffers the all-reduce can work on IN PLACE: memory from NCCL's allocator (cuMem, multicast-capable) registered
 * with the communicator (symmetric window where this NCCL offers it, plain registration otherwise), so the NVLink-switch
 * reduction reads and writes the gradients directly instead of staging them through NCCL's own buffers - fewer CTAs taken
 * from the GEMM the all-reduce overlaps, less HBM traffic.  Collective: every rank calls spb_comm_alloc in the same order with
 * the same size.  Any device pointer still works with spb_block_bwd; registered ones are faster.  Freed by spb_comm_free or
 * spb_comm_destroy.  spb_comm_buffer_kind: 0 not registered, 1 registered, 2 symmetric window.                          */
int spb_comm_alloc(void* comm, size_t bytes, void** ptr_out);
int spb_comm_free(void* comm, void* ptr);
int spb_comm_buffer_kind(void* comm, const void* ptr);

#ifdef __cplusplus
}
#endif
#endif /* SPB200_H */
