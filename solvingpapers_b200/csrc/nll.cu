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

}  // namespace spb

extern "C" {

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
