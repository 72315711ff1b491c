// attn_simt.cu - fp32-accurate softmax attention forward/backward on CUDA cores (the fp32 parity mode, and head dims
// the tcgen05 kernels do not tile: hd = 16 (ViT notebook), 256 (GPT), 768 (Gemma literal)).
// Never materialises scores or repeated K/V in HBM (llama3:820-828, :626-627): one warp owns one query row (forward and
// dQ) or one key row (dK/dV); the row of scores lives in shared memory; GQA is an index (kv head = q head / n_rep) and the
// dK/dV group sum happens in registers by looping over the group's q heads.  Deterministic (no atomics).
#include <cstdlib>

#include "common.cuh"
#include "kernels.h"

namespace spb {

constexpr int AS_WARPS = 4;
constexpr int AS_MAXC = 24;  // hd <= 32*24 = 768

template <typename T, int NC>
__device__ __forceinline__ void load_row(const T* p, int hd, int lane, float* r) {
#pragma unroll
  for (int c = 0; c < NC; ++c) {
    const int d = lane + 32 * c;
    r[c] = (d < hd) ? to_f32<T>(p[d]) : 0.f;
  }
}

// dot(reg row, global row) reduced over the warp; NC >= ceil(hd/32) register chunks per lane.
template <typename T, int NC>
__device__ __forceinline__ float dot_row(const float* r, const T* p, int hd, int lane) {
  float s = 0.f;
#pragma unroll
  for (int c = 0; c < NC; ++c) {
    const int d = lane + 32 * c;
    if (d < hd) s = fmaf(r[c], to_f32<T>(p[d]), s);
  }
  return warp_sum(s);
}

// ---------------------------------------------------------------------------------------------- forward
template <typename T, int NC>
__global__ void __launch_bounds__(AS_WARPS * 32) attn_simt_fwd_kernel(int Tn, int Hq, int Hkv, int hd, int causal, float scale,
                                                                      const T* __restrict__ q, int64_t ldq,
                                                                      const T* __restrict__ k, int64_t ldk,
                                                                      const T* __restrict__ v, int64_t ldv, T* __restrict__ o,
                                                                      int64_t ldo, float* __restrict__ lse) {
  extern __shared__ float smem[];
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int i = blockIdx.x * AS_WARPS + warp;
  const int h = blockIdx.y, b = blockIdx.z;
  if (i >= Tn) return;
  float* sc = smem + (size_t)warp * Tn;
  const int kvh = h / (Hq / Hkv);
  const int64_t tok0 = (int64_t)b * Tn;
  float qr[NC];
  load_row<T, NC>(q + (tok0 + i) * ldq + (int64_t)h * hd, hd, lane, qr);
  const int L = causal ? i + 1 : Tn;
  float m = -INFINITY;
  for (int j = 0; j < L; ++j) {
    const float s = dot_row<T, NC>(qr, k + (tok0 + j) * ldk + (int64_t)kvh * hd, hd, lane) * scale;
    if (lane == 0) sc[j] = s;
    m = fmaxf(m, s);
  }
  __syncwarp();
  float sum = 0.f;
  for (int j = lane; j < L; j += 32) {
    const float e = expf(sc[j] - m);
    sc[j] = e;
    sum += e;
  }
  sum = warp_sum(sum);
  __syncwarp();
  const float inv = 1.0f / sum;
  if (lane == 0) lse[((int64_t)b * Hq + h) * Tn + i] = m + logf(sum);
  float acc[NC];
#pragma unroll
  for (int c = 0; c < NC; ++c) acc[c] = 0.f;
  for (int j = 0; j < L; ++j) {
    const float p = sc[j];
    const T* vr = v + (tok0 + j) * ldv + (int64_t)kvh * hd;
#pragma unroll
    for (int c = 0; c < NC; ++c) {
      {
        const int d = lane + 32 * c;
        if (d < hd) acc[c] = fmaf(p, to_f32<T>(vr[d]), acc[c]);
      }
    }
  }
  T* orow = o + (tok0 + i) * ldo + (int64_t)h * hd;
#pragma unroll
  for (int c = 0; c < NC; ++c) {
    {
      const int d = lane + 32 * c;
      if (d < hd) orow[d] = from_f32<T>(acc[c] * inv);
    }
  }
}

// ---------------------------------------------------------------------------------------------- backward: dQ (+delta)
// Per query row i:  P_j = exp(s_j - lse_i);  dP_j = dO_i . V_j;  D_i = dO_i . O_i;  dS_j = P_j (dP_j - D_i);
//                   dQ_i = scale * sum_j dS_j K_j
template <typename T, int NC>
__global__ void __launch_bounds__(AS_WARPS * 32) attn_simt_bwd_dq_kernel(int Tn, int Hq, int Hkv, int hd, int causal, float scale,
                                                                         const T* __restrict__ q, int64_t ldq,
                                                                         const T* __restrict__ k, int64_t ldk,
                                                                         const T* __restrict__ v, int64_t ldv,
                                                                         const T* __restrict__ o, int64_t ldo,
                                                                         const T* __restrict__ d_o, int64_t lddo,
                                                                         const float* __restrict__ lse, float* __restrict__ delta,
                                                                         T* __restrict__ dq, int64_t lddq) {
  extern __shared__ float smem[];
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int i = blockIdx.x * AS_WARPS + warp;
  const int h = blockIdx.y, b = blockIdx.z;
  if (i >= Tn) return;
  float* sc = smem + (size_t)warp * Tn;
  const int kvh = h / (Hq / Hkv);
  const int64_t tok0 = (int64_t)b * Tn;
  float qr[NC], dor[NC];
  load_row<T, NC>(q + (tok0 + i) * ldq + (int64_t)h * hd, hd, lane, qr);
  load_row<T, NC>(d_o + (tok0 + i) * lddo + (int64_t)h * hd, hd, lane, dor);
  const float Di = dot_row<T, NC>(dor, o + (tok0 + i) * ldo + (int64_t)h * hd, hd, lane);
  const float l = lse[((int64_t)b * Hq + h) * Tn + i];
  if (lane == 0) delta[((int64_t)b * Hq + h) * Tn + i] = Di;
  const int L = causal ? i + 1 : Tn;
  for (int j = 0; j < L; ++j) {
    const float s = dot_row<T, NC>(qr, k + (tok0 + j) * ldk + (int64_t)kvh * hd, hd, lane) * scale;
    const float dp = dot_row<T, NC>(dor, v + (tok0 + j) * ldv + (int64_t)kvh * hd, hd, lane);
    if (lane == 0) sc[j] = expf(s - l) * (dp - Di) * scale;
  }
  __syncwarp();
  float acc[NC];
#pragma unroll
  for (int c = 0; c < NC; ++c) acc[c] = 0.f;
  for (int j = 0; j < L; ++j) {
    const float ds = sc[j];
    const T* kr = k + (tok0 + j) * ldk + (int64_t)kvh * hd;
#pragma unroll
    for (int c = 0; c < NC; ++c) {
      {
        const int d = lane + 32 * c;
        if (d < hd) acc[c] = fmaf(ds, to_f32<T>(kr[d]), acc[c]);
      }
    }
  }
  T* drow = dq + (tok0 + i) * lddq + (int64_t)h * hd;
#pragma unroll
  for (int c = 0; c < NC; ++c) {
    {
      const int d = lane + 32 * c;
      if (d < hd) drow[d] = from_f32<T>(acc[c]);
    }
  }
}

// ---------------------------------------------------------------------------------------------- backward: dK, dV
// Per key row j of kv head g:  for every q head h of the group and every visible query i:
//   P = exp(scale q_i.k_j - lse_i);  dV_j += P dO_i;  dS = P (dO_i.v_j - D_i);  dK_j += scale dS q_i
template <typename T, int NC>
__global__ void __launch_bounds__(AS_WARPS * 32) attn_simt_bwd_dkv_kernel(int Tn, int Hq, int Hkv, int hd, int causal, float scale,
                                                                          const T* __restrict__ q, int64_t ldq,
                                                                          const T* __restrict__ k, int64_t ldk,
                                                                          const T* __restrict__ v, int64_t ldv,
                                                                          const T* __restrict__ d_o, int64_t lddo,
                                                                          const float* __restrict__ lse,
                                                                          const float* __restrict__ delta, T* __restrict__ dk,
                                                                          int64_t lddk, T* __restrict__ dv, int64_t lddv) {
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int j = blockIdx.x * AS_WARPS + warp;
  const int g = blockIdx.y, b = blockIdx.z;
  if (j >= Tn) return;
  const int nrep = Hq / Hkv;
  const int64_t tok0 = (int64_t)b * Tn;
  float kr[NC], vr[NC], dkr[NC], dvr[NC];
  load_row<T, NC>(k + (tok0 + j) * ldk + (int64_t)g * hd, hd, lane, kr);
  load_row<T, NC>(v + (tok0 + j) * ldv + (int64_t)g * hd, hd, lane, vr);
#pragma unroll
  for (int c = 0; c < NC; ++c) dkr[c] = dvr[c] = 0.f;
  const int i0 = causal ? j : 0;
  for (int r = 0; r < nrep; ++r) {
    const int h = g * nrep + r;
    const float* lrow = lse + ((int64_t)b * Hq + h) * Tn;
    const float* drow = delta + ((int64_t)b * Hq + h) * Tn;
    for (int i = i0; i < Tn; ++i) {
      const T* qrow = q + (tok0 + i) * ldq + (int64_t)h * hd;
      const T* dorow = d_o + (tok0 + i) * lddo + (int64_t)h * hd;
      const float s = dot_row<T, NC>(kr, qrow, hd, lane) * scale;
      const float dp = dot_row<T, NC>(vr, dorow, hd, lane);
      const float p = expf(s - lrow[i]);
      const float ds = p * (dp - drow[i]) * scale;
#pragma unroll
      for (int c = 0; c < NC; ++c) {
        {
          const int d = lane + 32 * c;
          if (d < hd) {
            dvr[c] = fmaf(p, to_f32<T>(dorow[d]), dvr[c]);
            dkr[c] = fmaf(ds, to_f32<T>(qrow[d]), dkr[c]);
          }
        }
      }
    }
  }
  T* dkrow = dk + (tok0 + j) * lddk + (int64_t)g * hd;
  T* dvrow = dv + (tok0 + j) * lddv + (int64_t)g * hd;
#pragma unroll
  for (int c = 0; c < NC; ++c) {
    {
      const int d = lane + 32 * c;
      if (d < hd) {
        dkrow[d] = from_f32<T>(dkr[c]);
        dvrow[d] = from_f32<T>(dvr[c]);
      }
    }
  }
}

// ---------------------------------------------------------------------------------------------- host
template <typename T, int NC>
static int attn_simt_fwd_t(int B, int Tn, int Hq, int Hkv, int hd, int causal, float scale, const void* q, int64_t ldq,
                           const void* k, int64_t ldk, const void* v, int64_t ldv, void* o, int64_t ldo, float* lse,
                           cudaStream_t st) {
  const size_t smem = (size_t)AS_WARPS * Tn * sizeof(float);
  SPB_CUDA(cudaFuncSetAttribute(attn_simt_fwd_kernel<T, NC>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
  dim3 grid((Tn + AS_WARPS - 1) / AS_WARPS, Hq, B);
  attn_simt_fwd_kernel<T, NC><<<grid, AS_WARPS * 32, smem, st>>>(Tn, Hq, Hkv, hd, causal, scale, (const T*)q, ldq, (const T*)k, ldk,
                                                             (const T*)v, ldv, (T*)o, ldo, lse);
  SPB_LAUNCH_CHECK();
  return 0;
}

template <typename T, int NC>
static int attn_simt_bwd_t(int B, int Tn, int Hq, int Hkv, int hd, int causal, float scale, const void* q, int64_t ldq,
                           const void* k, int64_t ldk, const void* v, int64_t ldv, const void* o, int64_t ldo, const void* d_o,
                           int64_t lddo, const float* lse, float* delta, void* dq, int64_t lddq, void* dk, int64_t lddk, void* dv,
                           int64_t lddv, cudaStream_t st) {
  const size_t smem = (size_t)AS_WARPS * Tn * sizeof(float);
  SPB_CUDA(cudaFuncSetAttribute(attn_simt_bwd_dq_kernel<T, NC>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
  dim3 grid((Tn + AS_WARPS - 1) / AS_WARPS, Hq, B);
  attn_simt_bwd_dq_kernel<T, NC><<<grid, AS_WARPS * 32, smem, st>>>(Tn, Hq, Hkv, hd, causal, scale, (const T*)q, ldq, (const T*)k,
                                                                ldk, (const T*)v, ldv, (const T*)o, ldo, (const T*)d_o, lddo, lse,
                                                                delta, (T*)dq, lddq);
  SPB_LAUNCH_CHECK();
  dim3 grid2((Tn + AS_WARPS - 1) / AS_WARPS, Hkv, B);
  attn_simt_bwd_dkv_kernel<T, NC><<<grid2, AS_WARPS * 32, 0, st>>>(Tn, Hq, Hkv, hd, causal, scale, (const T*)q, ldq, (const T*)k, ldk,
                                                               (const T*)v, ldv, (const T*)d_o, lddo, lse, delta, (T*)dk, lddk,
                                                               (T*)dv, lddv);
  SPB_LAUNCH_CHECK();
  return 0;
}

int attn_simt_fwd(int dtype, int B, int Tn, int Hq, int Hkv, int hd, int causal, float scale, const void* q, int64_t ldq,
                  const void* k, int64_t ldk, const void* v, int64_t ldv, void* o, int64_t ldo, float* lse, cudaStream_t st) {
  {  // fp32 with a tile-able head dim: the register-tiled kernels of attn_ffma.cu (SPB_ATTN_FFMA=0: the warp-per-row kernels below)
    static const bool use_ffma = [] { const char* e = getenv("SPB_ATTN_FFMA"); return e == nullptr || atoi(e) != 0; }();
    const void* ptrs[4] = {q, k, v, o};
    const int64_t lds[4] = {ldq, ldk, ldv, ldo};
    if (use_ffma && attn_ffma_supported(dtype, hd, ptrs, lds, 4)) return attn_ffma_fwd(B, Tn, Hq, Hkv, hd, causal, scale, q, ldq, k, ldk, v, ldv, o, ldo, lse, st);
  }
  SPB_CHECK_ARG(hd <= 32 * AS_MAXC, SPB_E_UNSUPPORTED, "attention (SIMT): hd %d > %d", hd, 32 * AS_MAXC);
  SPB_CHECK_ARG((size_t)AS_WARPS * Tn * sizeof(float) <= 200 * 1024, SPB_E_UNSUPPORTED, "attention (SIMT): T %d > 12800", Tn);
#define SPB_AF(TT, NCC) return attn_simt_fwd_t<TT, NCC>(B, Tn, Hq, Hkv, hd, causal, scale, q, ldq, k, ldk, v, ldv, o, ldo, lse, st)
#define SPB_AF_NC(TT)            \
  do {                           \
    if (hd <= 32) SPB_AF(TT, 1); \
    if (hd <= 64) SPB_AF(TT, 2); \
    if (hd <= 128) SPB_AF(TT, 4);\
    if (hd <= 256) SPB_AF(TT, 8);\
    SPB_AF(TT, AS_MAXC);         \
  } while (0)
  if (dtype == SPB_F32) SPB_AF_NC(float);
  if (dtype == SPB_BF16) SPB_AF_NC(__nv_bfloat16);
#undef SPB_AF_NC
#undef SPB_AF
  SPB_CHECK_ARG(false, SPB_E_UNSUPPORTED, "attention (SIMT): dtype %d", dtype);
}

int attn_simt_bwd(int dtype, int B, int Tn, int Hq, int Hkv, int hd, int causal, float scale, const void* q, int64_t ldq,
                  const void* k, int64_t ldk, const void* v, int64_t ldv, const void* o, int64_t ldo, const void* d_o,
                  int64_t lddo, const float* lse, float* delta, void* dq, int64_t lddq, void* dk, int64_t lddk, void* dv,
                  int64_t lddv, cudaStream_t st) {
  {
    static const bool use_ffma = [] { const char* e = getenv("SPB_ATTN_FFMA"); return e == nullptr || atoi(e) != 0; }();
    const void* ptrs[8] = {q, k, v, o, d_o, dq, dk, dv};
    const int64_t lds[8] = {ldq, ldk, ldv, ldo, lddo, lddq, lddk, lddv};
    if (use_ffma && attn_ffma_supported(dtype, hd, ptrs, lds, 8))
      return attn_ffma_bwd(B, Tn, Hq, Hkv, hd, causal, scale, q, ldq, k, ldk, v, ldv, o, ldo, d_o, lddo, lse, delta, dq, lddq, dk, lddk, dv, lddv, st);
  }
  SPB_CHECK_ARG(hd <= 32 * AS_MAXC, SPB_E_UNSUPPORTED, "attention (SIMT): hd %d > %d", hd, 32 * AS_MAXC);
  SPB_CHECK_ARG((size_t)AS_WARPS * Tn * sizeof(float) <= 200 * 1024, SPB_E_UNSUPPORTED, "attention (SIMT): T %d > 12800", Tn);
#define SPB_AB(TT, NCC)                                                                                                   \
  return attn_simt_bwd_t<TT, NCC>(B, Tn, Hq, Hkv, hd, causal, scale, q, ldq, k, ldk, v, ldv, o, ldo, d_o, lddo, lse, delta, dq, \
                                  lddq, dk, lddk, dv, lddv, st)
#define SPB_AB_NC(TT)            \
  do {                           \
    if (hd <= 32) SPB_AB(TT, 1); \
    if (hd <= 64) SPB_AB(TT, 2); \
    if (hd <= 128) SPB_AB(TT, 4);\
    if (hd <= 256) SPB_AB(TT, 8);\
    SPB_AB(TT, AS_MAXC);         \
  } while (0)
  if (dtype == SPB_F32) SPB_AB_NC(float);
  if (dtype == SPB_BF16) SPB_AB_NC(__nv_bfloat16);
#undef SPB_AB_NC
#undef SPB_AB
  SPB_CHECK_ARG(false, SPB_E_UNSUPPORTED, "attention (SIMT): dtype %d", dtype);
}

}  // namespace spb
