// ffn.cu - element-wise kernels of the gated feed-forward sub-block (llama3:854-855
// `dot(silu(dot(x, w3)) * dot(x, w1), w2)`).  The three contractions are the tcgen05 / FFMA GEMMs of gemm_tc.cu / gemm_simt.cu;
// what is left is HBM-bound: 128-bit accesses, one thread per 128-bit vector of the [N, H] activation.
//   forward :  a[n,j]   = silu(g3[n,j]) * g1[n,j]                               reads 2*N*H*e, writes N*H*e
//   backward:  dg1[n,j] = da * silu(g3)        dg3[n,j] = da * g1 * silu'(g3)   reads 3*N*H*e, writes 2*N*H*e
// with g13 = [g1 | g3] stored as one [N, 2H] array (the output of the packed x.[w1|w3] GEMM), silu(z) = z*sigmoid(z),
// silu'(z) = sigmoid(z) * (1 + z * (1 - sigmoid(z))).  Arithmetic in fp32.
#include "common.cuh"
#include "kernels.h"

namespace spb {

template <typename T>
struct FVec;  // 128-bit vector of T as floats
template <>
struct FVec<float> {
  static constexpr int N = 4;
  static __device__ __forceinline__ void load(const float* p, float* f) {
    const float4 v = *reinterpret_cast<const float4*>(p);
    f[0] = v.x; f[1] = v.y; f[2] = v.z; f[3] = v.w;
  }
  static __device__ __forceinline__ void store(float* p, const float* f) { *reinterpret_cast<float4*>(p) = make_float4(f[0], f[1], f[2], f[3]); }
};
template <>
struct FVec<__nv_bfloat16> {
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

__device__ __forceinline__ float sigmoidf_(float z) { return 1.0f / (1.0f + __expf(-z)); }

template <typename T, bool VEC>
__global__ void __launch_bounds__(256) swiglu_fwd_kernel(const T* __restrict__ g13, int64_t ld13, T* __restrict__ a, int64_t lda,
                                                         int64_t N, int H) {
  constexpr int VN = VEC ? FVec<T>::N : 1;
  const int vpr = H / VN;
  const int64_t idx = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (idx >= N * vpr) return;
  const int64_t n = idx / vpr;
  const int j = (int)(idx % vpr) * VN;
  float g1[VN], g3[VN], o[VN];
  if (VEC) {
    FVec<T>::load(g13 + n * ld13 + j, g1);
    FVec<T>::load(g13 + n * ld13 + H + j, g3);
  } else {
    g1[0] = to_f32<T>(g13[n * ld13 + j]);
    g3[0] = to_f32<T>(g13[n * ld13 + H + j]);
  }
#pragma unroll
  for (int e = 0; e < VN; ++e) o[e] = g3[e] * sigmoidf_(g3[e]) * g1[e];
  if (VEC) FVec<T>::store(a + n * lda + j, o);
  else a[n * lda + j] = from_f32<T>(o[0]);
}

template <typename T, bool VEC>
__global__ void __launch_bounds__(256) swiglu_bwd_kernel(const T* __restrict__ g13, int64_t ld13, const T* __restrict__ da, int64_t lda,
                                                         T* __restrict__ dg13, int64_t ldd, int64_t N, int H) {
  constexpr int VN = VEC ? FVec<T>::N : 1;
  const int vpr = H / VN;
  const int64_t idx = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (idx >= N * vpr) return;
  const int64_t n = idx / vpr;
  const int j = (int)(idx % vpr) * VN;
  float g1[VN], g3[VN], d[VN], o1[VN], o3[VN];
  if (VEC) {
    FVec<T>::load(g13 + n * ld13 + j, g1);
    FVec<T>::load(g13 + n * ld13 + H + j, g3);
    FVec<T>::load(da + n * lda + j, d);
  } else {
    g1[0] = to_f32<T>(g13[n * ld13 + j]);
    g3[0] = to_f32<T>(g13[n * ld13 + H + j]);
    d[0] = to_f32<T>(da[n * lda + j]);
  }
#pragma unroll
  for (int e = 0; e < VN; ++e) {
    const float s = sigmoidf_(g3[e]);
    o1[e] = d[e] * g3[e] * s;
    o3[e] = d[e] * g1[e] * s * (1.0f + g3[e] * (1.0f - s));
  }
  if (VEC) {
    FVec<T>::store(dg13 + n * ldd + j, o1);
    FVec<T>::store(dg13 + n * ldd + H + j, o3);
  } else {
    dg13[n * ldd + j] = from_f32<T>(o1[0]);
    dg13[n * ldd + H + j] = from_f32<T>(o3[0]);
  }
}

template <typename T>
static int swiglu_launch(bool bwd, const void* g13, int64_t ld13, const void* da, void* out, int64_t ldo, int64_t lda, int64_t N, int H,
                         cudaStream_t st) {
  constexpr int VN = FVec<T>::N;
  const bool vec = (H % VN == 0) && (ld13 % VN == 0) && (ldo % VN == 0) && (lda % VN == 0) && ((uintptr_t)g13 % 16 == 0) &&
                   ((uintptr_t)out % 16 == 0) && ((uintptr_t)da % 16 == 0);
  const int64_t total = N * (H / (vec ? VN : 1));
  const unsigned grid = (unsigned)((total + 255) / 256);
  if (!bwd) {
    if (vec) swiglu_fwd_kernel<T, true><<<grid, 256, 0, st>>>((const T*)g13, ld13, (T*)out, ldo, N, H);
    else swiglu_fwd_kernel<T, false><<<grid, 256, 0, st>>>((const T*)g13, ld13, (T*)out, ldo, N, H);
  } else {
    if (vec) swiglu_bwd_kernel<T, true><<<grid, 256, 0, st>>>((const T*)g13, ld13, (const T*)da, lda, (T*)out, ldo, N, H);
    else swiglu_bwd_kernel<T, false><<<grid, 256, 0, st>>>((const T*)g13, ld13, (const T*)da, lda, (T*)out, ldo, N, H);
  }
  SPB_LAUNCH_CHECK();
  return 0;
}

// a[N,H] = silu(g13[:, H:2H]) * g13[:, 0:H]
int swiglu_fwd(int dtype, const void* g13, int64_t ld13, void* a, int64_t lda, int64_t N, int H, cudaStream_t st) {
  SPB_CHECK_ARG(g13 && a && N > 0 && H > 0, SPB_E_INVALID, "swiglu_fwd: bad argument");
  if (dtype == SPB_F32) return swiglu_launch<float>(false, g13, ld13, nullptr, a, lda, 0, N, H, st);
  if (dtype == SPB_BF16) return swiglu_launch<__nv_bfloat16>(false, g13, ld13, nullptr, a, lda, 0, N, H, st);
  SPB_CHECK_ARG(false, SPB_E_UNSUPPORTED, "swiglu_fwd: dtype %d", dtype);
}
// dg13[N,2H] from da[N,H] and the saved pre-activations
int swiglu_bwd(int dtype, const void* g13, int64_t ld13, const void* da, int64_t lda, void* dg13, int64_t ldd, int64_t N, int H,
               cudaStream_t st) {
  SPB_CHECK_ARG(g13 && da && dg13 && N > 0 && H > 0, SPB_E_INVALID, "swiglu_bwd: bad argument");
  if (dtype == SPB_F32) return swiglu_launch<float>(true, g13, ld13, da, dg13, ldd, lda, N, H, st);
  if (dtype == SPB_BF16) return swiglu_launch<__nv_bfloat16>(true, g13, ld13, da, dg13, ldd, lda, N, H, st);
  SPB_CHECK_ARG(false, SPB_E_UNSUPPORTED, "swiglu_bwd: dtype %d", dtype);
}

}  // namespace spb
