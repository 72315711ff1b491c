// nll.cu - the loss head's memory-bound half: llama3:959-967
//   loss = -mean(take_along_axis(log_softmax(logits), targets[:, None], axis=1))
// over logits [N, V] (V = 50257 in the notebook).  The logits GEMM (llama3:930) and its two backward GEMMs are spb_gemm.
//   forward : one CTA per row, one pass: online (max, sum exp) per thread, block-combined; row_nll = lse - logit[target];
//             lse is kept for the backward; a second tiny kernel averages row_nll deterministically.       reads N*V*e
//   backward: dlogits[n, j] = (exp(logit - lse) - [j == target]) * scale, scale = dloss / N                reads + writes N*V*e
// Arithmetic in fp32; columns are read with 128-bit loads when the row pitch allows it.
#include "common.cuh"
#include "kernels.h"

namespace spb {

constexpr int NLL_THREADS = 256;

template <typename T>
struct NllVec;
template <>
struct NllVec<float> {
  static constexpr int N = 4;
  static __device__ __forceinline__ void load(const float* p, float* f) {
    const float4 v = *reinterpret_cast<const float4*>(p);
    f[0] = v.x; f[1] = v.y; f[2] = v.z; f[3] = v.w;
  }
  static __device__ __forceinline__ void store(float* p, const float* f) { *reinterpret_cast<float4*>(p) = make_float4(f[0], f[1], f[2], f[3]); }
};
template <>
struct NllVec<__nv_bfloat16> {
  static constexpr int N = 8;
  static __device__ __forceinline__ void load(const __nv_bfloat16* p, float* f) {
    const uint4 v = *reinterpret_cast<const uint4*>(p);
    const float2 a = unpack_bf16(v.x), b = unpack_bf16(v.y), c = unpack_bf16(v.z), d = unpack_bf16(v.w);
    f[0] = a.x; f[1] = a.y; f[2] = b.x; f[3] = b.y; f[4] = c.x; f[5] = c.y; f[6] = d.x; f[7] = d.y;
  }
  static __device__ __forceinline__ void store(__nv_bfloat16* p, const float* f) {
    uint4 v;
    v.x = pack_bf16(f[0], f[1]); v.y = pack_bf16(f[2], f[3]); v.z = pack_bf16(f[4], f[5]); v.w = pack_bf16(f[6], f[7]);
    *reinterpret_cast<uint4*>(p) = v;
  }
};

// combine two online-softmax states (m, s): s counts exp(x - m)
__device__ __forceinline__ void nll_merge(float& m, float& s, float m2, float s2) {
  const float mm = fmaxf(m, m2);
  s = (m == -INFINITY ? 0.f : s * __expf(m - mm)) + (m2 == -INFINITY ? 0.f : s2 * __expf(m2 - mm));
  m = mm;
}

template <typename T, bool VEC>
__global__ void __launch_bounds__(NLL_THREADS) nll_fwd_kernel(const T* __restrict__ logits, int64_t ld, const int64_t* __restrict__ targets,
                                                              int V, float* __restrict__ lse, float* __restrict__ row_nll) {
  constexpr int VN = VEC ? NllVec<T>::N : 1;
  __shared__ float sm[NLL_THREADS / 32], ss[NLL_THREADS / 32];
  const int64_t n = blockIdx.x;
  const T* row = logits + n * ld;
  float m = -INFINITY, s = 0.f;
  // two vectors per trip: twice the loads in flight per thread (a bf16 row of 50257 logits is only 25 trips of 4 KB per CTA)
  for (int j0 = threadIdx.x * VN; j0 < V; j0 += 2 * NLL_THREADS * VN) {
    float f[2][VN];
#pragma unroll
    for (int u = 0; u < 2; ++u) {
      const int j = j0 + u * NLL_THREADS * VN;
      if (VEC && j + VN <= V) {
        NllVec<T>::load(row + j, f[u]);
      } else {
#pragma unroll
        for (int e = 0; e < VN; ++e) f[u][e] = (j + e < V) ? to_f32<T>(row[j + e]) : -INFINITY;
      }
    }
    float cm = f[0][0];
#pragma unroll
    for (int u = 0; u < 2; ++u)
#pragma unroll
      for (int e = 0; e < VN; ++e) cm = fmaxf(cm, f[u][e]);
    if (cm > m) {
      s = (m == -INFINITY) ? 0.f : s * __expf(m - cm);
      m = cm;
    }
    // everything seen so far is -inf (vocabulary-masked logits): exp(-inf - (-inf)) would be NaN, and the sum stays 0
    if (m == -INFINITY) continue;
#pragma unroll
    for (int u = 0; u < 2; ++u)
#pragma unroll
      for (int e = 0; e < VN; ++e) s += __expf(f[u][e] - m);
  }
#pragma unroll
  for (int off = 16; off > 0; off >>= 1) nll_merge(m, s, __shfl_xor_sync(0xffffffffu, m, off), __shfl_xor_sync(0xffffffffu, s, off));
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  if (lane == 0) {
    sm[warp] = m;
    ss[warp] = s;
  }
  __syncthreads();
  if (threadIdx.x == 0) {
    for (int w = 1; w < NLL_THREADS / 32; ++w) nll_merge(m, s, sm[w], ss[w]);
    const float l = m + logf(s);
    lse[n] = l;
    const int64_t t = targets[n];
    row_nll[n] = (t >= 0 && t < V) ? l - to_f32<T>(row[t]) : 0.f;
  }
}

// loss = sum(row_nll) / N: one block, fixed summation order (deterministic)
__global__ void __launch_bounds__(1024) nll_mean_kernel(const float* __restrict__ row_nll, int64_t N, float* __restrict__ loss) {
  __shared__ float red[32];
  float s = 0.f;
  for (int64_t i = threadIdx.x; i < N; i += 1024) s += row_nll[i];
  s = warp_sum(s);
  if ((threadIdx.x & 31) == 0) red[threadIdx.x >> 5] = s;
  __syncthreads();
  if (threadIdx.x < 32) {
    s = warp_sum(red[threadIdx.x]);
    if (threadIdx.x == 0) *loss = s / (float)N;
  }
}

// dlogits = (softmax(logits) - onehot(target)) * scale; may run in place (dlogits == logits)
template <typename T, bool VEC>
__global__ void __launch_bounds__(NLL_THREADS) nll_bwd_kernel(const T* logits, int64_t ld, const int64_t* __restrict__ targets, int V,
                                                              const float* __restrict__ lse, float scale, T* dlogits, int64_t ldd) {
  constexpr int VN = VEC ? NllVec<T>::N : 1;
  const int64_t n = blockIdx.x;
  const T* row = logits + n * ld;
  T* out = dlogits + n * ldd;
  const float l = lse[n];
  const int t = (int)targets[n];
  for (int j = threadIdx.x * VN; j < V; j += NLL_THREADS * VN) {
    float f[VN];
    const bool full = VEC && j + VN <= V;
    if (full) {
      NllVec<T>::load(row + j, f);
    } else {
#pragma unroll
      for (int e = 0; e < VN; ++e) f[e] = (j + e < V) ? to_f32<T>(row[j + e]) : -INFINITY;
    }
#pragma unroll
    for (int e = 0; e < VN; ++e) f[e] = (__expf(f[e] - l) - ((j + e == t) ? 1.0f : 0.0f)) * scale;
    if (full) {
      NllVec<T>::store(out + j, f);
    } else {
#pragma unroll
      for (int e = 0; e < VN; ++e)
        if (j + e < V) out[j + e] = from_f32<T>(f[e]);
    }
  }
}

template <typename T>
static int nll_launch(bool bwd, const void* logits, int64_t ld, const int64_t* targets, int64_t N, int V, float* lse, float* row_nll,
                      float* loss, float scale, void* dlogits, int64_t ldd, cudaStream_t st) {
  constexpr int VN = NllVec<T>::N;
  const bool vec = (ld % VN == 0) && ((uintptr_t)logits % 16 == 0) && (!bwd || ((ldd % VN == 0) && ((uintptr_t)dlogits % 16 == 0)));
  if (!bwd) {
    if (vec) nll_fwd_kernel<T, true><<<(unsigned)N, NLL_THREADS, 0, st>>>((const T*)logits, ld, targets, V, lse, row_nll);
    else nll_fwd_kernel<T, false><<<(unsigned)N, NLL_THREADS, 0, st>>>((const T*)logits, ld, targets, V, lse, row_nll);
    SPB_LAUNCH_CHECK();
    if (loss != nullptr) {
      nll_mean_kernel<<<1, 1024, 0, st>>>(row_nll, N, loss);
      SPB_LAUNCH_CHECK();
    }
  } else {
    if (vec) nll_bwd_kernel<T, true><<<(unsigned)N, NLL_THREADS, 0, st>>>((const T*)logits, ld, targets, V, lse, scale, (T*)dlogits, ldd);
    else nll_bwd_kernel<T, false><<<(unsigned)N, NLL_THREADS, 0, st>>>((const T*)logits, ld, targets, V, lse, scale, (T*)dlogits, ldd);
    SPB_LAUNCH_CHECK();
  }
  return 0;
}

// ---- fused loss head (logits GEMM + NLL): merge the GEMM epilogue's per-tile partials ------------------------------
// lse[n] = log sum_j exp(logit[n, j]) from part[n][tile] = (max, sum exp(x - max)); row_nll[n] = lse[n] - logit[n, target[n]]
__global__ void __launch_bounds__(256) head_merge_kernel(const float2* __restrict__ part, int tiles, const float* __restrict__ tgt_logit,
                                                         const int64_t* __restrict__ targets, int V, int64_t N, float* __restrict__ lse,
                                                         float* __restrict__ row_nll) {
  const int64_t n = (int64_t)blockIdx.x * (blockDim.x / 32) + (threadIdx.x >> 5);
  const int lane = threadIdx.x & 31;
  if (n >= N) return;
  float m = -INFINITY, s = 0.f;
  for (int t = lane; t < tiles; t += 32) {
    const float2 p = part[n * tiles + t];
    nll_merge(m, s, p.x, p.y);
  }
#pragma unroll
  for (int off = 16; off > 0; off >>= 1) nll_merge(m, s, __shfl_xor_sync(0xffffffffu, m, off), __shfl_xor_sync(0xffffffffu, s, off));
  if (lane == 0) {
    const float l = m + logf(s);
    lse[n] = l;
    const int64_t t = targets[n];
    row_nll[n] = (t >= 0 && t < V) ? l - tgt_logit[n] : 0.f;
  }
}

static inline size_t head_align(size_t v) { return (v + 255) / 256 * 256; }
static inline int64_t head_vp(int V) { return ((int64_t)V + 7) / 8 * 8; }
struct HeadWs {
  size_t w_pad, part, tgt, dlogits, total;
};
static HeadWs head_ws(int64_t N, int C, int V) {
  HeadWs w;
  size_t off = 0;
  auto take = [&](size_t b) { size_t o = off; off = head_align(off + b); return o; };
  w.w_pad = take((size_t)C * head_vp(V) * 2);                       // [C, Vp] bf16 copy of the weight when its pitch is not TMA-addressable
  w.part = take((size_t)N * ((V + 255) / 256) * sizeof(float2));    // (max, sum) per (row, 256-column tile)
  w.tgt = take((size_t)N * sizeof(float));
  w.dlogits = take((size_t)N * head_vp(V) * 2);                     // backward only: dlogits [N, Vp] bf16
  w.total = off;
  return w;
}

// the weight as the tcgen05 GEMM can address it: itself, or a copy with the vocabulary axis padded to a multiple of 8 elements
static int head_weight(const void* w, int64_t ldw, int C, int V, char* ws, const HeadWs& L, const void** w_out, int64_t* ld_out, cudaStream_t st) {
  if (ldw % 8 == 0 && (uintptr_t)w % 16 == 0) {
    *w_out = w;
    *ld_out = ldw;
    return 0;
  }
  const int64_t vp = head_vp(V);
  SPB_CUDA(cudaMemcpy2DAsync(ws + L.w_pad, (size_t)vp * 2, w, (size_t)ldw * 2, (size_t)V * 2, (size_t)C, cudaMemcpyDeviceToDevice, st));
  *w_out = ws + L.w_pad;
  *ld_out = vp;
  return 0;
}

}  // namespace spb

extern "C" {

size_t spb_loss_head_ws_bytes(int64_t N, int32_t C, int32_t V) {
  if (N <= 0 || C <= 0 || V <= 0) return 0;
  return spb::head_ws(N, C, V).total;
}

int spb_loss_head_fwd(int dtype, const void* h, int64_t ldh, const void* w, int64_t ldw, const int64_t* targets, int64_t N, int32_t C, int32_t V,
                      float* lse, float* row_nll, float* loss, void* ws, spb_stream_t stream) {
  using namespace spb;
  SPB_CHECK_ARG(h && w && targets && lse && row_nll && ws, SPB_E_INVALID, "spb_loss_head_fwd: null pointer");
  SPB_CHECK_ARG(N > 0 && N < (1LL << 31) && C > 0 && V > 128 && ldh >= C && ldw >= V, SPB_E_INVALID, "spb_loss_head_fwd: bad shape (N=%lld C=%d V=%d)",
                (long long)N, C, V);
  SPB_CHECK_ARG(dtype == SPB_BF16, SPB_E_UNSUPPORTED, "spb_loss_head_fwd: the fused head runs on the bf16 tcgen05 GEMM (fp32: spb_gemm + spb_nll_fwd)");
  SPB_CHECK_ARG((uintptr_t)ws % 256 == 0, SPB_E_INVALID, "spb_loss_head_fwd: workspace must be 256-byte aligned");
  cudaStream_t st = (cudaStream_t)stream;
  char* w8 = (char*)ws;
  const HeadWs L = head_ws(N, C, V);
  const void* wp;
  int64_t ldp;
  int rc = head_weight(w, ldw, C, V, w8, L, &wp, &ldp, st);
  if (rc) return rc;
  GemmTcHead hd;
  hd.mode = 1;
  hd.part = w8 + L.part;
  hd.tgt_logit = (float*)(w8 + L.tgt);
  hd.targets = targets;
  hd.lse = nullptr;
  hd.gscale = 0.f;
  // logits = h . w (llama3:930) with the reduction of llama3:961-965 in the epilogue: nothing [N, V]-sized is written
  rc = gemm_tc(SPB_BF16, N, V, C, h, ldh, 0, wp, ldp, 0, w8 + L.part, head_vp(V), nullptr, 0, st, nullptr, 0, 0, 0, nullptr, 0, nullptr, &hd);
  if (rc) return rc;
  head_merge_kernel<<<(unsigned)((N + 7) / 8), 256, 0, st>>>((const float2*)(w8 + L.part), (V + 255) / 256, (const float*)(w8 + L.tgt), targets, V, N, lse,
                                                             row_nll);
  SPB_LAUNCH_CHECK();
  if (loss != nullptr) {
    nll_mean_kernel<<<1, 1024, 0, st>>>(row_nll, N, loss);
    SPB_LAUNCH_CHECK();
  }
  return 0;
}

int spb_loss_head_bwd(int dtype, const void* h, int64_t ldh, const void* w, int64_t ldw, const int64_t* targets, int64_t N, int32_t C, int32_t V,
                      const float* lse, float dloss, void* dh, int64_t lddh, float* dw, int64_t lddw, void* ws, spb_stream_t stream) {
  using namespace spb;
  SPB_CHECK_ARG(h && w && targets && lse && ws && (dh || dw), SPB_E_INVALID, "spb_loss_head_bwd: null pointer");
  SPB_CHECK_ARG(N > 0 && N < (1LL << 31) && C > 0 && V > 128 && ldh >= C && ldw >= V, SPB_E_INVALID, "spb_loss_head_bwd: bad shape");
  SPB_CHECK_ARG(dtype == SPB_BF16, SPB_E_UNSUPPORTED, "spb_loss_head_bwd: the fused head runs on the bf16 tcgen05 GEMM");
  SPB_CHECK_ARG((uintptr_t)ws % 256 == 0, SPB_E_INVALID, "spb_loss_head_bwd: workspace must be 256-byte aligned");
  cudaStream_t st = (cudaStream_t)stream;
  char* w8 = (char*)ws;
  const HeadWs L = head_ws(N, C, V);
  const int64_t vp = head_vp(V);
  const void* wp;
  int64_t ldp;
  int rc = head_weight(w, ldw, C, V, w8, L, &wp, &ldp, st);
  if (rc) return rc;
  GemmTcHead hd;
  hd.mode = 2;
  hd.part = nullptr;
  hd.tgt_logit = nullptr;
  hd.targets = targets;
  hd.lse = lse;
  hd.gscale = dloss / (float)N;
  // dlogits = (softmax(h . w) - onehot) * dloss / N, recomputed tile by tile from h, w and lse: the logits are never stored
  rc = gemm_tc(SPB_BF16, N, V, C, h, ldh, 0, wp, ldp, 0, w8 + L.dlogits, vp, nullptr, 0, st, nullptr, 0, 0, 0, nullptr, 0, nullptr, &hd);
  if (rc) return rc;
  if (dh != nullptr) {  // dh = dlogits . w^T
    rc = gemm_tc(SPB_BF16, N, C, V, w8 + L.dlogits, vp, 0, wp, ldp, 1, dh, lddh, nullptr, 0, st);
    if (rc) return rc;
  }
  if (dw != nullptr) {  // dw = h^T . dlogits (fp32, the weight's own [C, V] layout)
    rc = gemm_tc(SPB_F32, C, V, N, h, ldh, 1, w8 + L.dlogits, vp, 0, dw, lddw, nullptr, 0, st);
    if (rc) return rc;
  }
  return 0;
}

int spb_nll_fwd(int dtype, const void* logits, int64_t ld, const int64_t* targets, int64_t N, int32_t V, float* lse, float* row_nll,
                float* loss, spb_stream_t stream) {
  SPB_CHECK_ARG(logits && targets && lse && row_nll, SPB_E_INVALID, "spb_nll_fwd: null pointer");
  SPB_CHECK_ARG(N > 0 && N < (1LL << 31) && V > 0 && ld >= V, SPB_E_INVALID, "spb_nll_fwd: bad shape (N=%lld V=%d ld=%lld)", (long long)N, V,
                (long long)ld);
  cudaStream_t st = (cudaStream_t)stream;
  if (dtype == SPB_F32) return spb::nll_launch<float>(false, logits, ld, targets, N, V, lse, row_nll, loss, 0.f, nullptr, 0, st);
  if (dtype == SPB_BF16) return spb::nll_launch<__nv_bfloat16>(false, logits, ld, targets, N, V, lse, row_nll, loss, 0.f, nullptr, 0, st);
  SPB_CHECK_ARG(false, SPB_E_UNSUPPORTED, "spb_nll_fwd: dtype %d", dtype);
}

int spb_nll_bwd(int dtype, const void* logits, int64_t ld, const int64_t* targets, int64_t N, int32_t V, const float* lse, float dloss,
                void* dlogits, int64_t ldd, spb_stream_t stream) {
  SPB_CHECK_ARG(logits && targets && lse && dlogits, SPB_E_INVALID, "spb_nll_bwd: null pointer");
  SPB_CHECK_ARG(N > 0 && N < (1LL << 31) && V > 0 && ld >= V && ldd >= V, SPB_E_INVALID, "spb_nll_bwd: bad shape");
  cudaStream_t st = (cudaStream_t)stream;
  const float scale = dloss / (float)N;
  if (dtype == SPB_F32) return spb::nll_launch<float>(true, logits, ld, targets, N, V, const_cast<float*>(lse), nullptr, nullptr, scale, dlogits, ldd, st);
  if (dtype == SPB_BF16)
    return spb::nll_launch<__nv_bfloat16>(true, logits, ld, targets, N, V, const_cast<float*>(lse), nullptr, nullptr, scale, dlogits, ldd, st);
  SPB_CHECK_ARG(false, SPB_E_UNSUPPORTED, "spb_nll_bwd: dtype %d", dtype);
}

}  // extern "C"
