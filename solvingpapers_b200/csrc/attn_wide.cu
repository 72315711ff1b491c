// attn_wide.cu - attention for head dims the flash kernels cannot hold on chip (hd > 256; Gemma literal: 2 q heads and one
// shared k/v of width 768, gemma/gemma.ipynb:218-259), bf16, every contraction on the tcgen05 GEMM (gemm_tc.cu).
//
// Why not a flash-style kernel: tensor memory is 128 lanes x 512 fp32 columns.  A 128-row output accumulator of width 768 is
// 768 columns by itself, and the backward's dK and dV accumulators are 1536 - so a fused kernel has to give each CTA a
// 128-256 column slice of the outputs and RECOMPUTE the scores (contraction over all 768 dims) in every slice: 3x the
// score work forward, 6x backward (the hd-256 backward already pays 2x), with K, V, Q and dO streamed again for each slice
// (a [128, 768] bf16 tile is 192 KB - none of them can stay resident in the 227 KB of shared memory).  At hd = 768 the
// arithmetic per score element is high enough (2 * 768 FLOPs per contraction) that writing the scores of ONE (batch, head)
// row block to HBM is cheap next to the GEMMs that make and use them: 12 bytes per score (S and dP in fp32, P and dS in
// bf16) against 7680 FLOPs.  So, per (batch, q head) and per block of WIDE_ROWS query rows (causal: only the keys the
// block can see):
//   forward   S = Q_blk K^T (fp32)  ->  row softmax (scale, causal mask, lse; P in bf16)  ->  O_blk = P V
//   backward  S = Q_blk K^T, dP = dO_blk V^T (fp32)  ->  P = exp(S*scale - lse), dS = P o (dP - delta) * scale (bf16)
//             ->  dV += P^T dO_blk,  dK += dS^T Q_blk  (fp32 accumulators over the row blocks and the q heads of the kv group:
//             repeat_kv's VJP),  dQ_blk = dS K
// The scratch is O(WIDE_ROWS * T) whatever B and H are; nothing O(T^2) is saved between forward and backward (the backward
// recomputes S like the flash kernels do).  Algorithmic FLOPs as for the flash path: 4*B*Hq*T^2*hd*m forward, x2.5 backward.
#include "common.cuh"
#include "kernels.h"

namespace spb {

constexpr int WIDE_ROWS = 2048;  // query rows per block: the [rows, T] GEMMs then fill the machine (>= 128 tiles of 128 x 256 at T = 4096)

static inline int64_t wide_ld(int64_t kvlen) { return (kvlen + 7) / 8 * 8; }

bool attn_wide_supported(int dtype, int hd) { return dtype == SPB_BF16 && hd > 256 && hd % 8 == 0; }

// scratch layout: S fp32 [rows, ld] | dP fp32 [rows, ld] | P bf16 [rows, ld] | dS bf16 [rows, ld] | dK_acc, dV_acc fp32 [T, hd] each
size_t attn_wide_scratch_bytes(int B, int T, int Hq, int Hkv, int hd) {
  (void)B; (void)Hq; (void)Hkv;
  const size_t rows = T < WIDE_ROWS ? T : WIDE_ROWS, ld = (size_t)wide_ld(T);
  return rows * ld * (4 + 4 + 2 + 2) + 2 * (size_t)T * hd * 4 + 1024;
}

// One CTA per query row: P[i, j] = exp(S[i, j]*scale - m) / l for the visible j < n_vis, 0 for n_vis <= j < kvlen; lse = m + log l
__global__ void __launch_bounds__(256) wide_softmax_kernel(const float* __restrict__ S, int64_t lds, __nv_bfloat16* __restrict__ P, int64_t ldp,
                                                           float* __restrict__ lse, int r0, int rows, int kvlen, int T, int causal, float scale) {
  __shared__ float red[8];
  const int i = blockIdx.x;
  const int t = r0 + i;
  const int n_vis = causal ? min(t + 1, kvlen) : kvlen;
  const float* s = S + (int64_t)i * lds;
  float m = -INFINITY;
  for (int j = threadIdx.x; j < n_vis; j += 256) m = fmaxf(m, s[j]);
  m = warp_max(m);
  if ((threadIdx.x & 31) == 0) red[threadIdx.x >> 5] = m;
  __syncthreads();
  m = red[0];
#pragma unroll
  for (int w = 1; w < 8; ++w) m = fmaxf(m, red[w]);
  m *= scale;  // scale > 0: scaling commutes with max
  __syncthreads();
  float l = 0.f;
  for (int j = threadIdx.x; j < n_vis; j += 256) l += __expf(s[j] * scale - m);
  l = warp_sum(l);
  if ((threadIdx.x & 31) == 0) red[threadIdx.x >> 5] = l;
  __syncthreads();
  l = 0.f;
#pragma unroll
  for (int w = 0; w < 8; ++w) l += red[w];
  const float inv = 1.0f / l;
  __nv_bfloat16* p = P + (int64_t)i * ldp;
  for (int j = threadIdx.x; j < kvlen; j += 256) p[j] = __float2bfloat16_rn(j < n_vis ? __expf(s[j] * scale - m) * inv : 0.f);
  if (threadIdx.x == 0) lse[t] = m + logf(l);
  (void)rows; (void)T;
}

// P = exp(S*scale - lse), dS = P * (dP - delta) * scale, both bf16, masked entries 0
__global__ void __launch_bounds__(256) wide_ds_kernel(const float* __restrict__ S, const float* __restrict__ dP, int64_t lds, __nv_bfloat16* __restrict__ P,
                                                      __nv_bfloat16* __restrict__ dS, int64_t ldp, const float* __restrict__ lse,
                                                      const float* __restrict__ delta, int r0, int kvlen, int causal, float scale) {
  const int i = blockIdx.x;
  const int t = r0 + i;
  const int n_vis = causal ? min(t + 1, kvlen) : kvlen;
  const float l = lse[t], dl = delta[t];
  const float* s = S + (int64_t)i * lds;
  const float* dp = dP + (int64_t)i * lds;
  __nv_bfloat16* p = P + (int64_t)i * ldp;
  __nv_bfloat16* ds = dS + (int64_t)i * ldp;
  for (int j = threadIdx.x; j < kvlen; j += 256) {
    float pv = 0.f, dv = 0.f;
    if (j < n_vis) {
      pv = __expf(s[j] * scale - l);
      dv = pv * (dp[j] - dl) * scale;
    }
    p[j] = __float2bfloat16_rn(pv);
    ds[j] = __float2bfloat16_rn(dv);
  }
}

// delta[b, h, t] = sum_c dO[n, h*hd + c] * O[n, h*hd + c]: one warp per (token, head)
__global__ void __launch_bounds__(256) wide_delta_kernel(const __nv_bfloat16* __restrict__ o, int64_t ldo, const __nv_bfloat16* __restrict__ d_o,
                                                         int64_t lddo, float* __restrict__ delta, int B, int T, int Hq, int hd) {
  const int64_t w = ((int64_t)blockIdx.x * blockDim.x + threadIdx.x) >> 5;
  const int lane = threadIdx.x & 31;
  if (w >= (int64_t)B * T * Hq) return;
  const int h = (int)(w % Hq);
  const int64_t tok = w / Hq;
  const __nv_bfloat16* a = o + tok * ldo + (int64_t)h * hd;
  const __nv_bfloat16* g = d_o + tok * lddo + (int64_t)h * hd;
  float s = 0.f;
  for (int c = lane * 8; c < hd; c += 256) {
    const uint4 av = *reinterpret_cast<const uint4*>(a + c), gv = *reinterpret_cast<const uint4*>(g + c);
    const uint32_t aw[4] = {av.x, av.y, av.z, av.w}, gw[4] = {gv.x, gv.y, gv.z, gv.w};
#pragma unroll
    for (int i = 0; i < 4; ++i) {
      const float2 x = unpack_bf16(aw[i]), y = unpack_bf16(gw[i]);
      s = fmaf(x.x, y.x, fmaf(x.y, y.y, s));
    }
  }
  s = warp_sum(s);
  if (lane == 0) delta[((tok / T) * Hq + h) * T + tok % T] = s;
}

// out[n, c] (bf16, row stride ldo) = acc[n, c] (fp32 [rows, cols])
__global__ void __launch_bounds__(256) wide_store_bf16_kernel(const float* __restrict__ acc, __nv_bfloat16* __restrict__ out, int64_t ldo, int64_t rows,
                                                              int cols) {
  const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= rows * (cols / 4)) return;
  const int64_t r = i / (cols / 4);
  const int c = (int)(i % (cols / 4)) * 4;
  const float4 v = *reinterpret_cast<const float4*>(acc + r * cols + c);
  uint2 o;
  o.x = pack_bf16(v.x, v.y);
  o.y = pack_bf16(v.z, v.w);
  *reinterpret_cast<uint2*>(out + r * ldo + c) = o;
}

static int wide_check(int hd, const void* const* ptrs, const int64_t* lds, int n, const void* scratch) {
  SPB_CHECK_ARG(hd % 8 == 0, SPB_E_UNSUPPORTED, "wide attention: head dim %d must be a multiple of 8", hd);
  for (int i = 0; i < n; ++i)
    SPB_CHECK_ARG(((uintptr_t)ptrs[i] % 16 == 0) && (lds[i] % 8 == 0), SPB_E_UNSUPPORTED,
                  "wide attention: operand %d must be 16-byte aligned with a leading dimension multiple of 8", i);
  SPB_CHECK_ARG(scratch != nullptr && (uintptr_t)scratch % 256 == 0, SPB_E_INVALID, "wide attention: scratch must be 256-byte aligned");
  return 0;
}

int attn_wide_fwd(int B, int T, int Hq, int Hkv, int hd, int causal, float scale, const void* q, int64_t ldq, const void* k, int64_t ldk,
                  const void* v, int64_t ldv, void* o, int64_t ldo, float* lse, void* scratch, cudaStream_t st) {
  const void* ptrs[4] = {q, k, v, o};
  const int64_t lds[4] = {ldq, ldk, ldv, ldo};
  int rc = wide_check(hd, ptrs, lds, 4, scratch);
  if (rc) return rc;
  SPB_CHECK_ARG(scale > 0.f, SPB_E_INVALID, "wide attention: scale must be positive");
  const int n_rep = Hq / Hkv;
  const int rows_max = T < WIDE_ROWS ? T : WIDE_ROWS;
  const int64_t ld = wide_ld(T);
  float* S = (float*)scratch;
  __nv_bfloat16* P = (__nv_bfloat16*)((char*)scratch + (size_t)rows_max * ld * 8);
  const __nv_bfloat16 *qb = (const __nv_bfloat16*)q, *kb = (const __nv_bfloat16*)k, *vb = (const __nv_bfloat16*)v;
  __nv_bfloat16* ob = (__nv_bfloat16*)o;
  for (int b = 0; b < B; ++b) {
    for (int h = 0; h < Hq; ++h) {
      const int g = h / n_rep;  // repeat_kv's consecutive grouping (llama3:626-627); Gemma: every q head uses the one k/v
      for (int r0 = 0; r0 < T; r0 += rows_max) {
        const int rows = min(rows_max, T - r0);
        const int kvlen = causal ? r0 + rows : T;
        const int64_t ldk_s = wide_ld(kvlen);
        // S = Q_blk K^T
        rc = gemm_tc(SPB_F32, rows, kvlen, hd, qb + ((int64_t)b * T + r0) * ldq + (int64_t)h * hd, ldq, 0, kb + (int64_t)b * T * ldk + (int64_t)g * hd, ldk,
                     1, S, ldk_s, nullptr, 0, st);
        if (rc) return rc;
        wide_softmax_kernel<<<rows, 256, 0, st>>>(S, ldk_s, P, ldk_s, lse + ((int64_t)b * Hq + h) * T, r0, rows, kvlen, T, causal, scale);
        SPB_LAUNCH_CHECK();
        // O_blk = P V
        rc = gemm_tc(SPB_BF16, rows, hd, kvlen, P, ldk_s, 0, vb + (int64_t)b * T * ldv + (int64_t)g * hd, ldv, 0,
                     ob + ((int64_t)b * T + r0) * ldo + (int64_t)h * hd, ldo, nullptr, 0, st);
        if (rc) return rc;
      }
    }
  }
  return 0;
}

int attn_wide_bwd(int B, int T, int Hq, int Hkv, int hd, int causal, float scale, const void* q, int64_t ldq, const void* k, int64_t ldk,
                  const void* v, int64_t ldv, const void* o, int64_t ldo, const void* d_o, int64_t lddo, const float* lse, float* delta, void* dq,
                  int64_t lddq, void* dk, int64_t lddk, void* dv, int64_t lddv, void* scratch, cudaStream_t st) {
  const void* ptrs[8] = {q, k, v, o, d_o, dq, dk, dv};
  const int64_t lds[8] = {ldq, ldk, ldv, ldo, lddo, lddq, lddk, lddv};
  int rc = wide_check(hd, ptrs, lds, 8, scratch);
  if (rc) return rc;
  const int n_rep = Hq / Hkv;
  const int rows_max = T < WIDE_ROWS ? T : WIDE_ROWS;
  const int64_t ld = wide_ld(T);
  char* sc = (char*)scratch;
  float* S = (float*)sc;
  float* dP = (float*)(sc + (size_t)rows_max * ld * 4);
  __nv_bfloat16* P = (__nv_bfloat16*)(sc + (size_t)rows_max * ld * 8);
  __nv_bfloat16* dS = (__nv_bfloat16*)(sc + (size_t)rows_max * ld * 10);
  float* dk_acc = (float*)(sc + (size_t)rows_max * ld * 12);
  float* dv_acc = dk_acc + (size_t)T * hd;
  const __nv_bfloat16 *qb = (const __nv_bfloat16*)q, *kb = (const __nv_bfloat16*)k, *vb = (const __nv_bfloat16*)v, *dob = (const __nv_bfloat16*)d_o;
  {
    const int64_t warps = (int64_t)B * T * Hq;
    wide_delta_kernel<<<(unsigned)((warps * 32 + 255) / 256), 256, 0, st>>>((const __nv_bfloat16*)o, ldo, dob, lddo, delta, B, T, Hq, hd);
    SPB_LAUNCH_CHECK();
  }
  for (int b = 0; b < B; ++b) {
    for (int g = 0; g < Hkv; ++g) {
      SPB_CUDA(cudaMemsetAsync(dk_acc, 0, 2 * (size_t)T * hd * sizeof(float), st));
      const __nv_bfloat16* kg = kb + (int64_t)b * T * ldk + (int64_t)g * hd;
      const __nv_bfloat16* vg = vb + (int64_t)b * T * ldv + (int64_t)g * hd;
      for (int hh = 0; hh < n_rep; ++hh) {
        const int h = g * n_rep + hh;
        for (int r0 = 0; r0 < T; r0 += rows_max) {
          const int rows = min(rows_max, T - r0);
          const int kvlen = causal ? r0 + rows : T;
          const int64_t ldk_s = wide_ld(kvlen);
          const __nv_bfloat16* q_blk = qb + ((int64_t)b * T + r0) * ldq + (int64_t)h * hd;
          const __nv_bfloat16* do_blk = dob + ((int64_t)b * T + r0) * lddo + (int64_t)h * hd;
          rc = gemm_tc(SPB_F32, rows, kvlen, hd, q_blk, ldq, 0, kg, ldk, 1, S, ldk_s, nullptr, 0, st);       // S = Q_blk K^T
          if (rc) return rc;
          rc = gemm_tc(SPB_F32, rows, kvlen, hd, do_blk, lddo, 0, vg, ldv, 1, dP, ldk_s, nullptr, 0, st);    // dP = dO_blk V^T
          if (rc) return rc;
          wide_ds_kernel<<<rows, 256, 0, st>>>(S, dP, ldk_s, P, dS, ldk_s, lse + ((int64_t)b * Hq + h) * T, delta + ((int64_t)b * Hq + h) * T, r0, kvlen,
                                              causal, scale);
          SPB_LAUNCH_CHECK();
          rc = gemm_tc(SPB_F32, kvlen, hd, rows, P, ldk_s, 1, do_blk, lddo, 0, dv_acc, hd, nullptr, 1, st);  // dV[0:kvlen] += P^T dO_blk
          if (rc) return rc;
          rc = gemm_tc(SPB_F32, kvlen, hd, rows, dS, ldk_s, 1, q_blk, ldq, 0, dk_acc, hd, nullptr, 1, st);   // dK[0:kvlen] += dS^T Q_blk
          if (rc) return rc;
          rc = gemm_tc(SPB_BF16, rows, hd, kvlen, dS, ldk_s, 0, kg, ldk, 0,                                   // dQ_blk = dS K
                       (__nv_bfloat16*)dq + ((int64_t)b * T + r0) * lddq + (int64_t)h * hd, lddq, nullptr, 0, st);
          if (rc) return rc;
        }
      }
      const int64_t n4 = (int64_t)T * (hd / 4);
      wide_store_bf16_kernel<<<(unsigned)((n4 + 255) / 256), 256, 0, st>>>(dk_acc, (__nv_bfloat16*)dk + (int64_t)b * T * lddk + (int64_t)g * hd, lddk, T, hd);
      SPB_LAUNCH_CHECK();
      wide_store_bf16_kernel<<<(unsigned)((n4 + 255) / 256), 256, 0, st>>>(dv_acc, (__nv_bfloat16*)dv + (int64_t)b * T * lddv + (int64_t)g * hd, lddv, T, hd);
      SPB_LAUNCH_CHECK();
    }
  }
  return 0;
}

}  // namespace spb
