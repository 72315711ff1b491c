// kernels.h - internal launcher declarations (one per .cu); the C ABI in api.cu dispatches to these.
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

namespace spb {

int gemm_simt(int in_dtype, int out_dtype, int64_t M, int64_t N, int64_t K, const void* A, int64_t lda, int transA, const void* B,
              int64_t ldb, int transB, void* D, int64_t ldd, const void* bias, int accumulate, cudaStream_t st,
              const void* residual = nullptr, int64_t ldr = 0);
// Attention-backward preparation fused into the epilogue of the d(attn_o) = dy . wo^T GEMM (D = dO [B*T, H*hd] bf16):
// delta[b,h,t] = sum_c dO[n, h*hd + c] * o[n, h*hd + c] and zero[n, :] = 0 (the fp32 dq accumulator) - what
// flash_bwd_tc's FB_PHASE_PRE (a memset + a kernel over o and dO) otherwise does in two extra passes over HBM.
struct GemmTcFuse {
  const void* o;   // attention output [B*T, H*hd] bf16, row stride ldo
  int64_t ldo;
  float* delta;    // [B, H, T] fp32
  float* zero;     // [B*T, H*hd] fp32, row stride ldz
  int64_t ldz;
  int hd, T, H;
};
// Loss-head epilogues of the logits GEMM (llama3:930 + :959-967): the [N, V] logits never reach memory.
//  mode 1 (forward): no D store; per (row, 256-column tile) the online (max, sum exp) of the fp32 logits -> part[row][tile] (float2,
//          ceil(N/256) tiles per row) and the logit of targets[row] -> tgt_logit[row]
//  mode 2 (backward): D = (exp(logit - lse[row]) - [col == targets[row]]) * gscale in bf16 (dlogits for the dh / dW GEMMs)
struct GemmTcHead {
  int mode;
  void* part;
  float* tgt_logit;
  const int64_t* targets;
  const float* lse;
  float gscale;
};
// rope != nullptr: fused rotary epilogue on output columns < rope_cols (row n has position n % rope_T)
int gemm_tc(int out_dtype, int64_t M, int64_t N, int64_t K, const void* A, int64_t lda, int transA, const void* B, int64_t ldb,
            int transB, void* D, int64_t ldd, const void* bias, int accumulate, cudaStream_t st, const float* rope = nullptr,
            int rope_cols = 0, int rope_T = 0, int rope_hd = 0, const void* residual = nullptr, int64_t ldr = 0,
            const GemmTcFuse* fuse = nullptr, const GemmTcHead* head = nullptr);
// > 0: the CTA-pair GEMMs launched by this host thread use one cluster per `tiles_per_cluster` tiles (SMs shared with a collective)
void gemm_tc_hint_shared_sms(int tiles_per_cluster);
void gemm_tc_hint_bias_f32(int on);  // the next gemm_tc calls of this host thread read `bias` as fp32
// residual != nullptr (both GEMMs): D = A.B (+bias) + residual, residual [M,N] of D's type with row stride ldr

int attn_simt_fwd(int dtype, int B, int Tn, int Hq, int Hkv, int hd, int causal, float scale, const void* q, int64_t ldq,
                  const void* k, int64_t ldk, const void* v, int64_t ldv, void* o, int64_t ldo, float* lse, cudaStream_t st);
int attn_simt_bwd(int dtype, int B, int Tn, int Hq, int Hkv, int hd, int causal, float scale, const void* q, int64_t ldq,
                  const void* k, int64_t ldk, const void* v, int64_t ldv, const void* o, int64_t ldo, const void* d_o,
                  int64_t lddo, const float* lse, float* delta, void* dq, int64_t lddq, void* dk, int64_t lddk, void* dv,
                  int64_t lddv, cudaStream_t st);

// gemm_split.cu: fp32 GEMM on the tcgen05 tensor cores through a two-term bf16 split of both operands (three products in one
// GEMM over a 3x longer contraction axis); scratch = gemm_f32_split_scratch_bytes(M, N, K)
size_t gemm_f32_split_scratch_bytes(int64_t M, int64_t N, int64_t K);
bool gemm_f32_split_ok(int64_t M, int64_t N, int64_t K, const void* A, int64_t lda, int transA, const void* B, int64_t ldb, int transB);
int gemm_f32_split(int64_t M, int64_t N, int64_t K, const void* A, int64_t lda, int transA, const void* B, int64_t ldb, int transB, void* D,
                   int64_t ldd, const void* bias, int accumulate, const void* residual, int64_t ldr, void* scratch, size_t scratch_bytes, cudaStream_t st);

// attn_ffma.cu: register-tiled fp32 attention (hd in {64, 128, 256}, 16-byte aligned operands), used by attn_simt_fwd / bwd
bool attn_ffma_supported(int dtype, int hd, const void* const* ptrs, const int64_t* lds, int n);
int attn_ffma_fwd(int B, int Tn, int Hq, int Hkv, int hd, int causal, float scale, const void* q, int64_t ldq, const void* k, int64_t ldk,
                  const void* v, int64_t ldv, void* o, int64_t ldo, float* lse, cudaStream_t st);
int attn_ffma_bwd(int B, int Tn, int Hq, int Hkv, int hd, int causal, float scale, const void* q, int64_t ldq, const void* k, int64_t ldk,
                  const void* v, int64_t ldv, const void* o, int64_t ldo, const void* d_o, int64_t lddo, const float* lse, float* delta, void* dq,
                  int64_t lddq, void* dk, int64_t lddk, void* dv, int64_t lddv, cudaStream_t st);

// tcgen05 flash attention (bf16, hd in {64,128})
bool flash_tc_supported(int dtype, int hd);
bool flash_bwd_tc_supported(int dtype, int hd);
int flash_fwd_tc(int B, int Tn, int Hq, int Hkv, int hd, int causal, float scale, const void* q, int64_t ldq, const void* k,
                 int64_t ldk, const void* v, int64_t ldv, void* o, int64_t ldo, float* lse, cudaStream_t st);
constexpr int FB_PHASE_PRE = 1, FB_PHASE_MAIN = 2;  // flash_bwd_tc phases: scratch preparation / main kernel + dq conversion
int flash_bwd_tc(int B, int Tn, int Hq, int Hkv, int hd, int causal, float scale, const void* q, int64_t ldq, const void* k,
                 int64_t ldk, const void* v, int64_t ldv, const void* o, int64_t ldo, const void* d_o, int64_t lddo,
                 const float* lse, float* delta, void* dq, int64_t lddq, void* dk, int64_t lddk, void* dv, int64_t lddv,
                 float* dq_acc, cudaStream_t st, const float* rope = nullptr, int phase = 3);  // rope: fused rotary VJP on dq and dk

// comm.cu
// all-reduce on the communicator's side stream, ordered after everything already enqueued on `main`
int ffn_act_fwd(int dtype, int act, const void* g, int64_t ldg, void* a, int64_t lda, int64_t N, int H, cudaStream_t st);
int ffn_act_bwd(int dtype, int act, const void* g, int64_t ldg, const void* da, int64_t lda, void* dg, int64_t ldd, int64_t N, int H,
                cudaStream_t st);
int norm_bwd_add(int dtype, int norm_kind, int upcast, const void* x, const void* w, const float* stats, const void* dy, void* dx,
                 float* dw, float* db, void* partial, int64_t N, int32_t C, const void* add, cudaStream_t st);
int comm_allreduce_side(void* comm, float* const* ptrs, const size_t* counts, int n, cudaStream_t main);
// all-reduce on `main`, then make `main` wait for the side stream's outstanding all-reduce
int comm_allreduce_join(void* comm, float* const* ptrs, const size_t* counts, int n, cudaStream_t main);
// error path: if a side-stream all-reduce is outstanding, make `main` wait for it and forget it
void comm_side_cleanup(void* comm, cudaStream_t main);

// attn_wide.cu: bf16 attention for head dims > 256 (Gemma literal, hd 768) as per-(batch, head) tcgen05 GEMM pipelines
bool attn_wide_supported(int dtype, int hd);
size_t attn_wide_scratch_bytes(int B, int T, int Hq, int Hkv, int hd);
int attn_wide_fwd(int B, int T, int Hq, int Hkv, int hd, int causal, float scale, const void* q, int64_t ldq, const void* k, int64_t ldk,
                  const void* v, int64_t ldv, void* o, int64_t ldo, float* lse, void* scratch, cudaStream_t st);
int attn_wide_bwd(int B, int T, int Hq, int Hkv, int hd, int causal, float scale, const void* q, int64_t ldq, const void* k, int64_t ldk,
                  const void* v, int64_t ldv, const void* o, int64_t ldo, const void* d_o, int64_t lddo, const float* lse, float* delta, void* dq,
                  int64_t lddq, void* dk, int64_t lddk, void* dv, int64_t lddv, void* scratch, cudaStream_t st);

// decode.cu
size_t small_m_gemm_scratch_bytes(int M, int N, int K, int dtype);
// advance != nullptr: the finishing kernel also does *advance = min(*advance + by, cap) (the device-side cache length of a recorded step)
int small_m_gemm(int dtype, int M, int N, int K, const void* A, int64_t lda, const void* W, int64_t ldw, void* D, int64_t ldd, const void* bias,
                 const void* res, int64_t ldr, float* part, cudaStream_t st, int* advance = nullptr, int by = 0, int cap = 0);
bool attn_decode_supported(int dtype, int hd);
size_t attn_decode_scratch_bytes(int B, int Tq, int Hq, int Hkv, int hd, int Tk);
// tk_dev != nullptr: the number of valid keys is min(*tk_dev + tk_add, Tk), read on the device (a step recorded in a CUDA graph);
// Tk is then the cache capacity the grid and the scratch are sized for
int attn_decode(int dtype, int B, int Tq, int Hq, int Hkv, int hd, int Tk, float scale, const void* q, int64_t ldq, const void* k, int64_t ldk,
                int64_t kbs, const void* v, int64_t ldv, int64_t vbs, void* o, int64_t ldo, float* lse, void* scratch, cudaStream_t st,
                const int* tk_dev = nullptr, int tk_add = 0);
// decode_gemm_mma.cu: bf16 D[M <= 64, N] = A . W[K, N] as a weight stream on mma.sync; fp32 partials per K slice
bool small_m_mma_ok(int dtype, int M, int N, int K, const void* A, int64_t lda, const void* W, int64_t ldw);
size_t small_m_mma_scratch_bytes(int M, int N, int K);
int small_m_mma(int M, int N, int K, const void* A, int64_t lda, const void* W, int64_t ldw, float* part, int* nsplit_out, cudaStream_t st);
// attn_decode_mma.cu: bf16 decode attention on mma.sync (16 query rows per kv head and CTA); partials merged by attn_decode_combine
bool attn_decode_mma_supported(int dtype, int hd);
size_t attn_decode_mma_scratch_bytes(int B, int Tq, int Hq, int Hkv, int hd, int Tk);
int attn_decode_mma(int B, int Tq, int Hq, int Hkv, int hd, int Tk, float scale, const void* q, int64_t ldq, const void* k, int64_t ldk, int64_t kbs,
                    const void* v, int64_t ldv, int64_t vbs, void* o, int64_t ldo, float* lse, void* scratch, cudaStream_t st,
                    const int* tk_dev = nullptr, int tk_add = 0);
int attn_decode_combine(int dtype, const float* po, const float* pml, void* o, int64_t ldo, float* lse, int B, int Tq, int Hq, int hd, int splits,
                        cudaStream_t st);
// part != nullptr: also sums the projection's K-slice partials [nsplit][B*Tq][ld] (+ bias) - the reduce pass and the rotary/append in one launch
int decode_rope_append(int dtype, void* qkv, int64_t ld, int B, int Tq, int Hq, int Hkv, int hd, const float* rope, int pos0, void* kc, void* vc,
                       int64_t ldc, int64_t cbs, cudaStream_t st, const float* part = nullptr, int nsplit = 0, const void* bias = nullptr,
                       const int* pos0_dev = nullptr, int cap = 0);

}  // namespace spb
