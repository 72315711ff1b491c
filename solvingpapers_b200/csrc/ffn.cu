// ffn.cu - element-wise kernels of the feed-forward sub-block: SwiGLU (llama3:854-855
// `dot(silu(dot(x, w3)) * dot(x, w1), w2)`), GeGLU (gemma:281-285) and the plain GELU MLP (gpt:385-389).  The three contractions are the tcgen05 / FFMA GEMMs of gemm_tc.cu / gemm_simt.cu;
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

// activations and their derivatives (fp32).  ACT: 0 silu (llama3:855), 1 gelu/erf (torch default: gemma:282), 2 gelu/tanh (flax
// nn.gelu default approximate=True: gpt:386)
template <int ACT>
__device__ __forceinline__ void act_eval(float z, float& f, float& df) {
  if (ACT == 0) {
    const float s = 1.0f / (1.0f + __expf(-z));
    f = z * s;
    df = s * (1.0f + z * (1.0f - s));
  } else if (ACT == 1) {
    const float cdf = 0.5f * (1.0f + erff(z * 0.70710678118654752f));
    f = z * cdf;
    df = cdf + z * 0.3989422804014327f * __expf(-0.5f * z * z);
  } else {
    const float k = 0.7978845608028654f, c = 0.044715f;
    const float t = tanhf(k * (z + c * z * z * z));
    f = 0.5f * z * (1.0f + t);
    df = 0.5f * (1.0f + t) + 0.5f * z * (1.0f - t * t) * k * (1.0f + 3.0f * c * z * z);
  }
}

// forward: gated   a = act(g[:, H:2H]) * g[:, 0:H]      (g = [value | gate], e.g. [x.w1 | x.w3])
//          plain   a = act(g[:, 0:H])
template <typename T, bool VEC, int ACT, bool GATED>
__global__ void __launch_bounds__(256) ffn_act_fwd_kernel(const T* __restrict__ g, int64_t ldg, T* __restrict__ a, int64_t lda, int64_t N,
                                                          int H) {
  constexpr int VN = VEC ? FVec<T>::N : 1;
  const int vpr = H / VN;
  const int64_t idx = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (idx >= N * vpr) return;
  const int64_t n = idx / vpr;
  const int j = (int)(idx % vpr) * VN;
  float v[VN], z[VN], o[VN];
  if (VEC) {
    FVec<T>::load(g + n * ldg + j, v);
    if (GATED) FVec<T>::load(g + n * ldg + H + j, z);
  } else {
    v[0] = to_f32<T>(g[n * ldg + j]);
    if (GATED) z[0] = to_f32<T>(g[n * ldg + H + j]);
  }
#pragma unroll
  for (int e = 0; e < VN; ++e) {
    float f, df;
    act_eval<ACT>(GATED ? z[e] : v[e], f, df);
    o[e] = GATED ? f * v[e] : f;
  }
  if (VEC) FVec<T>::store(a + n * lda + j, o);
  else a[n * lda + j] = from_f32<T>(o[0]);
}

// backward: gated  dvalue = da * act(gate), dgate = da * value * act'(gate);  plain  dg = da * act'(g)
template <typename T, bool VEC, int ACT, bool GATED>
__global__ void __launch_bounds__(256) ffn_act_bwd_kernel(const T* __restrict__ g, int64_t ldg, const T* __restrict__ da, int64_t lda,
                                                          T* __restrict__ dg, int64_t ldd, int64_t N, int H) {
  constexpr int VN = VEC ? FVec<T>::N : 1;
  const int vpr = H / VN;
  const int64_t idx = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (idx >= N * vpr) return;
  const int64_t n = idx / vpr;
  const int j = (int)(idx % vpr) * VN;
  float v[VN], z[VN], d[VN], o1[VN], o3[VN];
  if (VEC) {
    FVec<T>::load(g + n * ldg + j, v);
    if (GATED) FVec<T>::load(g + n * ldg + H + j, z);
    FVec<T>::load(da + n * lda + j, d);
  } else {
    v[0] = to_f32<T>(g[n * ldg + j]);
    if (GATED) z[0] = to_f32<T>(g[n * ldg + H + j]);
    d[0] = to_f32<T>(da[n * lda + j]);
  }
#pragma unroll
  for (int e = 0; e < VN; ++e) {
    float f, df;
    act_eval<ACT>(GATED ? z[e] : v[e], f, df);
    if (GATED) {
      o1[e] = d[e] * f;
      o3[e] = d[e] * v[e] * df;
    } else {
      o1[e] = d[e] * df;
    }
  }
  if (VEC) {
    FVec<T>::store(dg + n * ldd + j, o1);
    if (GATED) FVec<T>::store(dg + n * ldd + H + j, o3);
  } else {
    dg[n * ldd + j] = from_f32<T>(o1[0]);
    if (GATED) dg[n * ldd + H + j] = from_f32<T>(o3[0]);
  }
}

template <typename T, int ACT, bool GATED>
static int ffn_act_launch(bool bwd, const void* g, int64_t ldg, const void* da, void* out, int64_t ldo, int64_t lda, int64_t N, int H,
                          cudaStream_t st) {
  constexpr int VN = FVec<T>::N;
  const bool vec = (H % VN == 0) && (ldg % VN == 0) && (ldo % VN == 0) && (lda % VN == 0) && ((uintptr_t)g % 16 == 0) &&
                   ((uintptr_t)out % 16 == 0) && ((uintptr_t)da % 16 == 0);
  const int64_t total = N * (H / (vec ? VN : 1));
  const unsigned grid = (unsigned)((total + 255) / 256);
  if (!bwd) {
    if (vec) ffn_act_fwd_kernel<T, true, ACT, GATED><<<grid, 256, 0, st>>>((const T*)g, ldg, (T*)out, ldo, N, H);
    else ffn_act_fwd_kernel<T, false, ACT, GATED><<<grid, 256, 0, st>>>((const T*)g, ldg, (T*)out, ldo, N, H);
  } else {
    if (vec) ffn_act_bwd_kernel<T, true, ACT, GATED><<<grid, 256, 0, st>>>((const T*)g, ldg, (const T*)da, lda, (T*)out, ldo, N, H);
    else ffn_act_bwd_kernel<T, false, ACT, GATED><<<grid, 256, 0, st>>>((const T*)g, ldg, (const T*)da, lda, (T*)out, ldo, N, H);
  }
  SPB_LAUNCH_CHECK();
  return 0;
}

template <typename T>
static int ffn_act_dispatch(int act, bool bwd, const void* g, int64_t ldg, const void* da, void* out, int64_t ldo, int64_t lda, int64_t N,
                            int H, cudaStream_t st) {
  switch (act) {
    case SPB_ACT_SWIGLU: return ffn_act_launch<T, 0, true>(bwd, g, ldg, da, out, ldo, lda, N, H, st);
    case SPB_ACT_GEGLU: return ffn_act_launch<T, 1, true>(bwd, g, ldg, da, out, ldo, lda, N, H, st);
    case SPB_ACT_GELU_TANH: return ffn_act_launch<T, 2, false>(bwd, g, ldg, da, out, ldo, lda, N, H, st);
    default: break;
  }
  SPB_CHECK_ARG(false, SPB_E_UNSUPPORTED, "ffn activation %d", act);
}

// a[N,H] from the pre-activations g (gated kinds: [N,2H] = [value | gate]; SPB_ACT_GELU_TANH: [N,H])
int ffn_act_fwd(int dtype, int act, const void* g, int64_t ldg, void* a, int64_t lda, int64_t N, int H, cudaStream_t st) {
  SPB_CHECK_ARG(g && a && N > 0 && H > 0, SPB_E_INVALID, "ffn_act_fwd: bad argument");
  if (dtype == SPB_F32) return ffn_act_dispatch<float>(act, false, g, ldg, nullptr, a, lda, 0, N, H, st);
  if (dtype == SPB_BF16) return ffn_act_dispatch<__nv_bfloat16>(act, false, g, ldg, nullptr, a, lda, 0, N, H, st);
  SPB_CHECK_ARG(false, SPB_E_UNSUPPORTED, "ffn_act_fwd: dtype %d", dtype);
}
// dg (same shape as g) from da[N,H] and the saved pre-activations
int ffn_act_bwd(int dtype, int act, const void* g, int64_t ldg, const void* da, int64_t lda, void* dg, int64_t ldd, int64_t N, int H,
                cudaStream_t st) {
  SPB_CHECK_ARG(g && da && dg && N > 0 && H > 0, SPB_E_INVALID, "ffn_act_bwd: bad argument");
  if (dtype == SPB_F32) return ffn_act_dispatch<float>(act, true, g, ldg, da, dg, ldd, lda, N, H, st);
  if (dtype == SPB_BF16) return ffn_act_dispatch<__nv_bfloat16>(act, true, g, ldg, da, dg, ldd, lda, N, H, st);
  SPB_CHECK_ARG(false, SPB_E_UNSUPPORTED, "ffn_act_bwd: dtype %d", dtype);
}

}  // namespace spb
