// gemm_split.cu - fp32 GEMMs of the fp32 parity mode on the tcgen05 tensor cores (the notebooks' own default dtype, llama3:349-358;
// BASELINE configs C1 and C2-literal, whose projections ran on the FFMA GEMM at 4-10 TFLOP/s).
// One bf16 pass cannot meet the mode's 1e-4 tolerance (8 mantissa bits), so every fp32 operand is split into two bf16 terms,
//   a = a_hi + a_lo,  a_hi = bf16(a),  a_lo = bf16(a - a_hi)           (16 mantissa bits: |a - a_hi - a_lo| <= 2^-17 |a|)
// and the product keeps the three leading terms  a_hi b_hi + a_lo b_hi + a_hi b_lo  (the dropped a_lo b_lo is 2^-16 relative).
// The three partial GEMMs are ONE tcgen05 GEMM over a three times longer contraction axis,
//   A' = [A_hi | A_lo | A_hi]   B' = [B_hi ; B_hi ; B_lo]   =>   A' B' = A_hi B_hi + A_lo B_hi + A_hi B_lo,
// accumulated in fp32 in tensor memory without intermediate rounding.  A' and B' are written by one element-wise pass per
// operand into caller scratch (3 bf16 per fp32 element), in the operand's own orientation - concatenated along the contiguous
// axis for a K-contiguous operand, stacked along the rows for a K-strided one - so no transposition is involved.
#include "common.cuh"
#include "kernels.h"

namespace spb {

// in [R, C] fp32 (row stride ld).  CONCAT: out [R, 3C] (row stride ldo), segments (s0, s1, s2) of hi/lo.  STACK: out [3R, C].
// pattern_b = 0: (hi, lo, hi) for the A operand; 1: (hi, hi, lo) for the B operand
template <bool CONCAT>
__global__ void __launch_bounds__(256) split3_kernel(const float* __restrict__ in, int64_t ld, __nv_bfloat16* __restrict__ out, int64_t ldo, int64_t R,
                                                     int C, int pattern_b) {
  const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;  // one thread per 4 elements (C % 4 == 0)
  const int c4 = C / 4;
  if (i >= R * c4) return;
  const int64_t r = i / c4;
  const int c = (int)(i % c4) * 4;
  const float4 v = __ldg(reinterpret_cast<const float4*>(in + r * ld + c));
  const float x[4] = {v.x, v.y, v.z, v.w};
  __nv_bfloat16 hi[4], lo[4];
#pragma unroll
  for (int e = 0; e < 4; ++e) {
    hi[e] = __float2bfloat16_rn(x[e]);
    lo[e] = __float2bfloat16_rn(x[e] - __bfloat162float(hi[e]));
  }
  const uint2 h2 = make_uint2(((uint32_t)__bfloat16_as_ushort(hi[1]) << 16) | __bfloat16_as_ushort(hi[0]),
                              ((uint32_t)__bfloat16_as_ushort(hi[3]) << 16) | __bfloat16_as_ushort(hi[2]));
  const uint2 l2 = make_uint2(((uint32_t)__bfloat16_as_ushort(lo[1]) << 16) | __bfloat16_as_ushort(lo[0]),
                              ((uint32_t)__bfloat16_as_ushort(lo[3]) << 16) | __bfloat16_as_ushort(lo[2]));
  const uint2 seg[3] = {h2, pattern_b ? h2 : l2, pattern_b ? l2 : h2};
#pragma unroll
  for (int s = 0; s < 3; ++s) {
    __nv_bfloat16* dst = CONCAT ? out + r * ldo + (int64_t)s * C + c : out + ((int64_t)s * R + r) * ldo + c;
    *reinterpret_cast<uint2*>(dst) = seg[s];
  }
}

static inline int64_t round8(int64_t v) { return (v + 7) / 8 * 8; }

// bytes of scratch gemm_f32_split needs for D[M,N] = opA(A) opB(B) with contraction K
size_t gemm_f32_split_scratch_bytes(int64_t M, int64_t N, int64_t K) {
  // worst case over the orientations: concatenated [rows, 3K] or stacked [3K, round8(cols)]
  const size_t a = (size_t)(3 * K) * (size_t)round8(M) * 2, b = (size_t)(3 * K) * (size_t)round8(N) * 2;
  return a + b + 512;
}

bool gemm_f32_split_ok(int64_t M, int64_t N, int64_t K, const void* A, int64_t lda, int transA, const void* B, int64_t ldb, int transB) {
  if (K % 8 != 0 || M < 1 || N < 1) return false;          // concatenated operands: segments of K elements, TMA rows of 3K
  if ((uintptr_t)A % 16 != 0 || (uintptr_t)B % 16 != 0 || lda % 4 != 0 || ldb % 4 != 0) return false;
  if (transA && M % 4 != 0) return false;                  // 128-bit loads along the contiguous axis
  if (!transB && N % 4 != 0) return false;
  return true;
}

// D[M,N] (fp32, ldd) = opA(A) opB(B) (+ residual | + D): A, B fp32.  transA/transB as in spb_gemm.
int gemm_f32_split(int64_t M, int64_t N, int64_t K, const void* A, int64_t lda, int transA, const void* B, int64_t ldb, int transB, void* D,
                   int64_t ldd, const void* bias, int accumulate, const void* residual, int64_t ldr, void* scratch, size_t scratch_bytes, cudaStream_t st) {
  SPB_CHECK_ARG(scratch != nullptr && (uintptr_t)scratch % 256 == 0 && scratch_bytes >= gemm_f32_split_scratch_bytes(M, N, K), SPB_E_INVALID,
                "fp32 split GEMM: scratch too small or misaligned");
  __nv_bfloat16* As = (__nv_bfloat16*)scratch;
  int64_t lda_s, ldb_s;
  // A operand: stored [M, K] (transA = 0: concatenate the three segments along K) or [K, M] (stack them along the rows)
  if (!transA) {
    lda_s = 3 * K;
    const int64_t n4 = M * (K / 4);
    split3_kernel<true><<<(unsigned)((n4 + 255) / 256), 256, 0, st>>>((const float*)A, lda, As, lda_s, M, (int)K, 0);
  } else {
    lda_s = round8(M);
    const int64_t n4 = K * (M / 4);
    split3_kernel<false><<<(unsigned)((n4 + 255) / 256), 256, 0, st>>>((const float*)A, lda, As, lda_s, K, (int)M, 0);
  }
  SPB_LAUNCH_CHECK();
  __nv_bfloat16* Bs = As + (size_t)(3 * K) * (size_t)round8(M);
  Bs = (__nv_bfloat16*)(((uintptr_t)Bs + 255) / 256 * 256);
  if (transB) {  // stored [N, K]
    ldb_s = 3 * K;
    const int64_t n4 = N * (K / 4);
    split3_kernel<true><<<(unsigned)((n4 + 255) / 256), 256, 0, st>>>((const float*)B, ldb, Bs, ldb_s, N, (int)K, 1);
  } else {       // stored [K, N]
    ldb_s = round8(N);
    const int64_t n4 = K * (N / 4);
    split3_kernel<false><<<(unsigned)((n4 + 255) / 256), 256, 0, st>>>((const float*)B, ldb, Bs, ldb_s, K, (int)N, 1);
  }
  SPB_LAUNCH_CHECK();
  gemm_tc_hint_bias_f32(1);  // fp32 mode: the bias (if any) is fp32
  const int rc = gemm_tc(SPB_F32, M, N, 3 * K, As, lda_s, transA, Bs, ldb_s, transB, D, ldd, bias, accumulate, st, nullptr, 0, 0, 0, residual, ldr);
  gemm_tc_hint_bias_f32(0);
  return rc;
}

}  // namespace spb
