// decode.cu - SURVEY 8 row f3: the KV-cache branch of `attention` (llama3/LLaMA-jax.ipynb:815-819, caller `generate`
// :499-511): a few new tokens per sequence (Tq, usually 1) attend to everything already in the cache plus themselves.
//   k = concatenate([cache[0], k], axis=1); v = concatenate([cache[1], v], axis=1)           (:817-818)
//   scores = q . k^T / sqrt(hd)  [B, H, Tq, Tk]; softmax; . v; out-projection                  (:820-829)
// Everything here is bound by HBM bandwidth, not by the tensor cores: one decode step reads every weight once
// (2 * C * (Hq + 2 Hkv + Hq) * hd bytes in bf16) and the whole K/V cache once (2 * B * Tk * Hkv * hd elements), for a
// handful of query rows.  So the kernels are SIMT with 128-bit coalesced loads and fp32 accumulation:
//   K5  small_m_gemm_kernel   D[M<=8 rows, N] += A[M, k-slice] . W[k-slice, N]: a thread owns 8 (bf16) / 4 (fp32) adjacent
//                             output columns, streams its k-slice of the [in,out] weight with 16-byte loads (a warp reads
//                             512 contiguous bytes of a weight row), the M activations rows sit in shared memory; the K
//                             dimension is split over CTAs (>= 2 CTAs per SM) into fp32 partials, summed in a fixed order by
//   K5b small_m_reduce_kernel (+ bias, + residual, -> dtype): deterministic, no atomics.
//   K6  decode_rope_append_kernel  rotary at positions Tc .. Tc+Tq-1 on q (in place) and k, and the new k, v rows written
//                             into the caches in place (the jnp.concatenate of :817-818 without copying the old rows).
//   K7  attn_decode_kernel    flash-decoding: grid (key splits, kv head x row groups, batch); a CTA streams its slice of K
//                             (thread = key: dot products against the group's <= 8 query rows, broadcast from shared
//                             memory) and of V (thread = 16-byte column piece, 128/TPR keys in flight), online softmax per
//                             128-key tile, GQA by index (the group's n_rep q heads share one pass over the K/V head);
//   K7b attn_decode_combine_kernel merges the splits' (max, sum, partial O) in a fixed order -> o, lse.
// Algorithmic bytes: weights once + K,V cache once + O(M*C) activations.
#include "common.cuh"
#include "kernels.h"

namespace spb {

constexpr int DEC_THREADS = 128;

template <typename T>
struct Vec16;  // one 16-byte load of T -> fp32 lanes
template <>
struct Vec16<float> {
  static constexpr int N = 4;
  static __device__ __forceinline__ void load(const float* p, float* f) {
    const float4 v = __ldg(reinterpret_cast<const float4*>(p));
    f[0] = v.x; f[1] = v.y; f[2] = v.z; f[3] = v.w;
  }
};
template <>
struct Vec16<__nv_bfloat16> {
  static constexpr int N = 8;
  static __device__ __forceinline__ void load(const __nv_bfloat16* p, float* f) {
    const uint4 v = __ldg(reinterpret_cast<const uint4*>(p));
    const float2 a = unpack_bf16(v.x), b = unpack_bf16(v.y), c = unpack_bf16(v.z), d = unpack_bf16(v.w);
    f[0] = a.x; f[1] = a.y; f[2] = b.x; f[3] = b.y; f[4] = c.x; f[5] = c.y; f[6] = d.x; f[7] = d.y;
  }
};

// ------------------------------------------------------------------------------------------------ K5: small-M GEMM
// grid (column panels, k splits, row groups of MB); W is [K, N] row-major (row stride ldw): SPB_W_IN_OUT
template <typename T, int MB>
__global__ void __launch_bounds__(DEC_THREADS) small_m_gemm_kernel(const T* __restrict__ A, int64_t lda, const T* __restrict__ W, int64_t ldw,
                                                                  float* __restrict__ part, int M, int N, int K, int k_per_split) {
  constexpr int VE = Vec16<T>::N;
  constexpr int KC = 64;  // k rows staged per shared-memory refill
  __shared__ float xs[MB][KC];
  const int m0 = blockIdx.z * MB;
  const int col = (blockIdx.x * DEC_THREADS + threadIdx.x) * VE;
  const int kb = blockIdx.y * k_per_split, ke = min(K, kb + k_per_split);
  const bool live = col < N;  // N % VE == 0 is checked by the launcher
  float acc[MB][VE];
#pragma unroll
  for (int m = 0; m < MB; ++m)
#pragma unroll
    for (int e = 0; e < VE; ++e) acc[m][e] = 0.f;
  for (int k0 = kb; k0 < ke; k0 += KC) {
    const int kn = min(KC, ke - k0);
    __syncthreads();
    for (int i = threadIdx.x; i < MB * KC; i += DEC_THREADS) {
      const int m = i / KC, kk = i % KC;
      xs[m][kk] = (m0 + m < M && kk < kn) ? to_f32<T>(A[(int64_t)(m0 + m) * lda + k0 + kk]) : 0.f;
    }
    __syncthreads();
    if (!live) continue;
    const T* wp = W + (int64_t)k0 * ldw + col;
    int kk = 0;
    for (; kk + 4 <= kn; kk += 4) {  // four weight rows in flight per thread
      float w[4][VE];
#pragma unroll
      for (int u = 0; u < 4; ++u) Vec16<T>::load(wp + (int64_t)(kk + u) * ldw, w[u]);
#pragma unroll
      for (int u = 0; u < 4; ++u)
#pragma unroll
        for (int m = 0; m < MB; ++m) {
          const float xv = xs[m][kk + u];
#pragma unroll
          for (int e = 0; e < VE; ++e) acc[m][e] = fmaf(xv, w[u][e], acc[m][e]);
        }
    }
    for (; kk < kn; ++kk) {
      float w[VE];
      Vec16<T>::load(wp + (int64_t)kk * ldw, w);
#pragma unroll
      for (int m = 0; m < MB; ++m) {
        const float xv = xs[m][kk];
#pragma unroll
        for (int e = 0; e < VE; ++e) acc[m][e] = fmaf(xv, w[e], acc[m][e]);
      }
    }
  }
  if (!live) return;
#pragma unroll
  for (int m = 0; m < MB; ++m) {
    if (m0 + m >= M) break;
    float* dst = part + ((int64_t)blockIdx.y * M + (m0 + m)) * N + col;
#pragma unroll
    for (int e = 0; e < VE; e += 4) *reinterpret_cast<float4*>(dst + e) = make_float4(acc[m][e], acc[m][e + 1], acc[m][e + 2], acc[m][e + 3]);
  }
}

// out[m, n] = sum_s part[s, m, n] (+ bias[n]) (+ res[m, n]); fixed summation order
template <typename T>
__global__ void __launch_bounds__(256) small_m_reduce_kernel(const float* __restrict__ part, int splits, T* __restrict__ out, int64_t ldo,
                                                            const T* __restrict__ bias, const T* __restrict__ res, int64_t ldr, int M, int N) {
  const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= (int64_t)M * N) return;
  const int m = (int)(i / N), n = (int)(i % N);
  float s = 0.f;
  for (int p = 0; p < splits; ++p) s += part[((int64_t)p * M + m) * N + n];
  if (bias != nullptr) s += to_f32<T>(bias[n]);
  if (res != nullptr) s += to_f32<T>(res[(int64_t)m * ldr + n]);
  out[(int64_t)m * ldo + n] = from_f32<T>(s);
}

static int small_m_splits(int M, int N, int K, int ve) {
  const int panels = (N + DEC_THREADS * ve - 1) / (DEC_THREADS * ve);
  int want = (2 * sm_count() + panels - 1) / panels;
  int max_splits = K / 32;
  if (max_splits < 1) max_splits = 1;
  if (want > max_splits) want = max_splits;
  if (want > 64) want = 64;
  if (want < 1) want = 1;
  return want;
}

size_t small_m_gemm_scratch_bytes(int M, int N, int K, int dtype) {
  return (size_t)small_m_splits(M, N, K, dtype == SPB_BF16 ? 8 : 4) * M * N * sizeof(float);
}

template <typename T>
static int small_m_gemm_t(int M, int N, int K, const T* A, int64_t lda, const T* W, int64_t ldw, T* D, int64_t ldd, const T* bias, const T* res,
                          int64_t ldr, float* part, cudaStream_t st) {
  constexpr int VE = Vec16<T>::N;
  SPB_CHECK_ARG(N % VE == 0 && ldw % VE == 0 && (uintptr_t)W % 16 == 0, SPB_E_UNSUPPORTED,
                "small-M GEMM: the weight needs 16-byte aligned rows (N=%d ldw=%lld)", N, (long long)ldw);
  const int splits = small_m_splits(M, N, K, VE);
  const int k_per = ((K + splits - 1) / splits + 3) / 4 * 4;
  const int nsplit = (K + k_per - 1) / k_per;
  const int MB = M >= 8 ? 8 : (M > 2 ? 4 : (M > 1 ? 2 : 1));
  dim3 grid((N + DEC_THREADS * VE - 1) / (DEC_THREADS * VE), nsplit, (M + MB - 1) / MB);
  if (MB == 8) small_m_gemm_kernel<T, 8><<<grid, DEC_THREADS, 0, st>>>(A, lda, W, ldw, part, M, N, K, k_per);
  else if (MB == 4) small_m_gemm_kernel<T, 4><<<grid, DEC_THREADS, 0, st>>>(A, lda, W, ldw, part, M, N, K, k_per);
  else if (MB == 2) small_m_gemm_kernel<T, 2><<<grid, DEC_THREADS, 0, st>>>(A, lda, W, ldw, part, M, N, K, k_per);
  else small_m_gemm_kernel<T, 1><<<grid, DEC_THREADS, 0, st>>>(A, lda, W, ldw, part, M, N, K, k_per);
  SPB_LAUNCH_CHECK();
  const int64_t total = (int64_t)M * N;
  small_m_reduce_kernel<T><<<(unsigned)((total + 255) / 256), 256, 0, st>>>(part, nsplit, D, ldd, bias, res, ldr, M, N);
  SPB_LAUNCH_CHECK();
  return 0;
}

// D[M,N] = A[M,K] . W[K,N] (+bias) (+res); M <= 64 rows; `part` = small_m_gemm_scratch_bytes(M, N, K, dtype)
int small_m_gemm(int dtype, int M, int N, int K, const void* A, int64_t lda, const void* W, int64_t ldw, void* D, int64_t ldd, const void* bias,
                 const void* res, int64_t ldr, float* part, cudaStream_t st) {
  SPB_CHECK_ARG(M >= 1 && M <= 64 && N >= 1 && K >= 1, SPB_E_LIMIT, "small-M GEMM: M=%d rows (1..64 supported; longer inputs go through the prefill path)", M);
  if (dtype == SPB_BF16)
    return small_m_gemm_t<__nv_bfloat16>(M, N, K, (const __nv_bfloat16*)A, lda, (const __nv_bfloat16*)W, ldw, (__nv_bfloat16*)D, ldd,
                                         (const __nv_bfloat16*)bias, (const __nv_bfloat16*)res, ldr, part, st);
  return small_m_gemm_t<float>(M, N, K, (const float*)A, lda, (const float*)W, ldw, (float*)D, ldd, (const float*)bias, (const float*)res, ldr, part, st);
}

// ------------------------------------------------------------------------------------------------ K6: rotary + cache append
// qkv [B*Tq, ld] (q | k | v columns).  q is rotated in place; k is rotated and written, v copied, into the caches at token
// rows pos0 .. pos0+Tq-1 of each batch entry.  One thread per adjacent pair (2j, 2j+1) of a (row, head).
template <typename T>
__global__ void __launch_bounds__(256) decode_rope_append_kernel(T* __restrict__ qkv, int64_t ld, int B, int Tq, int Hq, int Hkv, int hd,
                                                                const float4* __restrict__ rope, int pos0, T* __restrict__ kc,
                                                                T* __restrict__ vc, int64_t ldc, int64_t cache_batch_stride) {
  const int hp = hd / 2;
  const int heads = Hq + 2 * Hkv;
  const int64_t idx = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  const int64_t total = (int64_t)B * Tq * heads * hp;
  if (idx >= total) return;
  const int j = (int)(idx % hp);
  const int h = (int)((idx / hp) % heads);
  const int64_t row = idx / ((int64_t)hp * heads);  // b * Tq + t
  const int t = (int)(row % Tq), b = (int)(row / Tq);
  T* src = qkv + row * ld + (int64_t)h * hd + 2 * j;
  float xe = to_f32<T>(src[0]), xo = to_f32<T>(src[1]);
  if (h < Hq + Hkv && rope != nullptr) {  // q and k heads: pair <- 2x2 block of the table row of position pos0 + t
    const float4 m = rope[(int64_t)(pos0 + t) * hp + j];
    const float e2 = m.x * xe + m.y * xo, o2 = m.z * xe + m.w * xo;
    xe = e2;
    xo = o2;
  }
  if (h < Hq) {
    src[0] = from_f32<T>(xe);
    src[1] = from_f32<T>(xo);
    return;
  }
  const bool is_k = h < Hq + Hkv;
  const int g = is_k ? h - Hq : h - Hq - Hkv;
  T* dst = (is_k ? kc : vc) + (int64_t)b * cache_batch_stride + (int64_t)(pos0 + t) * ldc + (int64_t)g * hd + 2 * j;
  dst[0] = from_f32<T>(xe);
  dst[1] = from_f32<T>(xo);
}

// ------------------------------------------------------------------------------------------------ K7: flash-decoding attention
struct DecArgs {
  const void* q;      // [B*Tq, ldq]: head h of row (b, t) at (b*Tq + t) * ldq + h * hd
  int64_t ldq;
  const void* k;      // cache: token j of batch b, kv head g at b * kbs + j * ldk + g * hd
  const void* v;
  int64_t ldk, ldv, kbs, vbs;
  float* po;          // partial O   [B, Hq, Tq, splits, hd] fp32 (unnormalised)
  float* pml;         // partial (m, l) [B, Hq, Tq, splits, 2]
  int B, Tq, Hq, Hkv, Tk, splits, tiles_per_split, row_groups;
  float scale;
};

template <typename T, int HD, int RB>
__global__ void __launch_bounds__(DEC_THREADS) attn_decode_kernel(const DecArgs p) {
  constexpr int VE = Vec16<T>::N;
  constexpr int TPR = HD / VE;              // threads per V row
  constexpr int G = DEC_THREADS / TPR;      // V rows in flight
  constexpr int KT = DEC_THREADS;           // keys per tile
  static_assert(DEC_THREADS % TPR == 0 && TPR >= 1, "head dim does not tile the CTA");
  __shared__ float qs[RB][HD];
  __shared__ float ps[RB][KT];
  __shared__ float red[2][4][RB];
  extern __shared__ float osum[];  // [G][RB][HD] for the final cross-group reduction

  const int split = blockIdx.x;
  const int g = blockIdx.y / p.row_groups, rg = blockIdx.y % p.row_groups;
  const int b = blockIdx.z;
  const int n_rep = p.Hq / p.Hkv;
  const int rtot = n_rep * p.Tq;
  const int r0 = rg * RB;
  const int nr = min(RB, rtot - r0);
  const int t = threadIdx.x, lane = t & 31, warp = t >> 5;

  // the group's query rows -> shared memory, pre-multiplied by the score scale.  Row r = (local q head, tq).
  for (int i = t; i < RB * HD; i += DEC_THREADS) {
    const int r = i / HD, d = i % HD;
    float qv = 0.f;
    if (r < nr) {
      const int hq = g * n_rep + (r0 + r) / p.Tq, tq = (r0 + r) % p.Tq;
      qv = to_f32<T>(((const T*)p.q)[((int64_t)b * p.Tq + tq) * p.ldq + (int64_t)hq * HD + d]) * p.scale;
    }
    qs[r][d] = qv;
  }
  __syncthreads();

  const T* kbase = (const T*)p.k + (int64_t)b * p.kbs + (int64_t)g * HD;
  const T* vbase = (const T*)p.v + (int64_t)b * p.vbs + (int64_t)g * HD;
  const int vg = t / TPR, vc = (t % TPR) * VE;
  float m_run[RB], l_run[RB], acc[RB][VE];
#pragma unroll
  for (int r = 0; r < RB; ++r) {
    m_run[r] = -INFINITY;
    l_run[r] = 0.f;
#pragma unroll
    for (int e = 0; e < VE; ++e) acc[r][e] = 0.f;
  }
  const int tile0 = split * p.tiles_per_split;
  const int tile1 = min(tile0 + p.tiles_per_split, (p.Tk + KT - 1) / KT);
  for (int tile = tile0; tile < tile1; ++tile) {
    const int k0 = tile * KT;
    const int j = k0 + t;
    // ---- scores of key j against the group's rows
    float s[RB];
#pragma unroll
    for (int r = 0; r < RB; ++r) s[r] = 0.f;
    if (j < p.Tk) {
      const T* kr = kbase + (int64_t)j * p.ldk;
      constexpr int NV = HD / VE;
      constexpr int CH = NV < 8 ? NV : 8;  // 16-byte loads in flight per thread
#pragma unroll
      for (int c0 = 0; c0 < NV; c0 += CH) {
        float kv[CH][VE];
#pragma unroll
        for (int c = 0; c < CH; ++c) Vec16<T>::load(kr + (c0 + c) * VE, kv[c]);
#pragma unroll
        for (int c = 0; c < CH; ++c)
#pragma unroll
          for (int e = 0; e < VE; e += 4) {
#pragma unroll
            for (int r = 0; r < RB; ++r) {
              const float4 qv = *reinterpret_cast<const float4*>(&qs[r][(c0 + c) * VE + e]);
              s[r] = fmaf(qv.x, kv[c][e], fmaf(qv.y, kv[c][e + 1], fmaf(qv.z, kv[c][e + 2], fmaf(qv.w, kv[c][e + 3], s[r]))));
            }
          }
      }
    } else {
#pragma unroll
      for (int r = 0; r < RB; ++r) s[r] = -INFINITY;
    }
    // ---- tile max per row
#pragma unroll
    for (int r = 0; r < RB; ++r) {
      const float wm = warp_max(s[r]);
      if (lane == 0) red[0][warp][r] = wm;
    }
    __syncthreads();  // also: the previous tile's P*V readers are done with ps
    float alpha[RB];
#pragma unroll
    for (int r = 0; r < RB; ++r) {
      const float tm = fmaxf(fmaxf(red[0][0][r], red[0][1][r]), fmaxf(red[0][2][r], red[0][3][r]));
      const float mn = fmaxf(m_run[r], tm);
      alpha[r] = (m_run[r] == -INFINITY) ? 0.f : __expf(m_run[r] - mn);
      m_run[r] = mn;
      const float pv = (s[r] == -INFINITY) ? 0.f : __expf(s[r] - mn);
      ps[r][t] = pv;
      const float wsum = warp_sum(pv);
      if (lane == 0) red[1][warp][r] = wsum;
    }
    __syncthreads();
#pragma unroll
    for (int r = 0; r < RB; ++r) {
      l_run[r] = l_run[r] * alpha[r] + (red[1][0][r] + red[1][1][r] + red[1][2][r] + red[1][3][r]);
#pragma unroll
      for (int e = 0; e < VE; ++e) acc[r][e] *= alpha[r];
    }
    // ---- acc[r][.] += sum_j P[r][j] * V[j][vc ..]: thread owns a 16-byte column piece, G keys in flight per pass
    const int kn = min(KT, p.Tk - k0);
    constexpr int UN = 4;
    int jj = vg;
    for (; jj + (UN - 1) * G < kn; jj += UN * G) {
      float vv[UN][VE];
#pragma unroll
      for (int u = 0; u < UN; ++u) Vec16<T>::load(vbase + (int64_t)(k0 + jj + u * G) * p.ldv + vc, vv[u]);
#pragma unroll
      for (int u = 0; u < UN; ++u)
#pragma unroll
        for (int r = 0; r < RB; ++r) {
          const float pv = ps[r][jj + u * G];
#pragma unroll
          for (int e = 0; e < VE; ++e) acc[r][e] = fmaf(pv, vv[u][e], acc[r][e]);
        }
    }
    for (; jj < kn; jj += G) {
      float vv[VE];
      Vec16<T>::load(vbase + (int64_t)(k0 + jj) * p.ldv + vc, vv);
#pragma unroll
      for (int r = 0; r < RB; ++r) {
        const float pv = ps[r][jj];
#pragma unroll
        for (int e = 0; e < VE; ++e) acc[r][e] = fmaf(pv, vv[e], acc[r][e]);
      }
    }
  }
  // ---- sum the G key groups' accumulators, write this split's partial (unnormalised O, m, l)
  __syncthreads();
#pragma unroll
  for (int r = 0; r < RB; ++r)
#pragma unroll
    for (int e = 0; e < VE; ++e) osum[((int64_t)vg * RB + r) * HD + vc + e] = acc[r][e];
  __syncthreads();
  for (int i = t; i < nr * HD; i += DEC_THREADS) {
    const int r = i / HD, d = i % HD;
    float o = 0.f;
#pragma unroll
    for (int gg = 0; gg < G; ++gg) o += osum[((int64_t)gg * RB + r) * HD + d];
    const int hq = g * n_rep + (r0 + r) / p.Tq, tq = (r0 + r) % p.Tq;
    const int64_t row = (((int64_t)b * p.Hq + hq) * p.Tq + tq) * p.splits + split;
    p.po[row * HD + d] = o;
  }
  if (t == 0) {
#pragma unroll
    for (int r = 0; r < RB; ++r) {
      if (r >= nr) break;
      const int hq = g * n_rep + (r0 + r) / p.Tq, tq = (r0 + r) % p.Tq;
      const int64_t row = (((int64_t)b * p.Hq + hq) * p.Tq + tq) * p.splits + split;
      p.pml[row * 2] = m_run[r];
      p.pml[row * 2 + 1] = l_run[r];
    }
  }
}

// o[b, tq, hq, :] = sum_s w_s O_s / sum_s w_s l_s, w_s = exp(m_s - max m); lse = max m + log(sum): fixed order over splits
template <typename T>
__global__ void __launch_bounds__(128) attn_decode_combine_kernel(const float* __restrict__ po, const float* __restrict__ pml, T* __restrict__ o,
                                                                  int64_t ldo, float* __restrict__ lse, int B, int Tq, int Hq, int hd, int splits) {
  const int64_t row = blockIdx.x;  // (b * Hq + hq) * Tq + tq
  const int tq = (int)(row % Tq), hq = (int)((row / Tq) % Hq), b = (int)(row / ((int64_t)Tq * Hq));
  float M = -INFINITY;
  for (int s = 0; s < splits; ++s) M = fmaxf(M, pml[(row * splits + s) * 2]);
  float L = 0.f;
  for (int s = 0; s < splits; ++s) {
    const float m = pml[(row * splits + s) * 2];
    if (m != -INFINITY) L += __expf(m - M) * pml[(row * splits + s) * 2 + 1];
  }
  const float inv = 1.0f / L;
  for (int d = threadIdx.x; d < hd; d += blockDim.x) {
    float acc = 0.f;
    for (int s = 0; s < splits; ++s) {
      const float m = pml[(row * splits + s) * 2];
      if (m != -INFINITY) acc += __expf(m - M) * po[(row * splits + s) * hd + d];
    }
    o[((int64_t)b * Tq + tq) * ldo + (int64_t)hq * hd + d] = from_f32<T>(acc * inv);
  }
  if (lse != nullptr && threadIdx.x == 0) lse[row] = M + logf(L);
}

static void decode_plan(int B, int Tq, int Hq, int Hkv, int Tk, int& RB, int& row_groups, int& splits, int& tiles_per_split) {
  const int rtot = (Hq / Hkv) * Tq;
  RB = rtot >= 8 ? 8 : (rtot > 2 ? 4 : (rtot > 1 ? 2 : 1));
  row_groups = (rtot + RB - 1) / RB;
  const int tiles = (Tk + DEC_THREADS - 1) / DEC_THREADS;
  const int ctas = B * Hkv * row_groups;
  int want = (4 * sm_count() + ctas - 1) / ctas;  // ~4 CTAs per SM
  if (want > tiles) want = tiles;
  if (want > 64) want = 64;
  if (want < 1) want = 1;
  tiles_per_split = (tiles + want - 1) / want;
  splits = (tiles + tiles_per_split - 1) / tiles_per_split;
}

size_t attn_decode_scratch_bytes(int B, int Tq, int Hq, int Hkv, int hd, int Tk) {
  int RB, rgs, splits, tps;
  decode_plan(B, Tq, Hq, Hkv, Tk, RB, rgs, splits, tps);
  return (size_t)B * Hq * Tq * splits * ((size_t)hd + 2) * sizeof(float) + 256;
}

bool attn_decode_supported(int dtype, int hd) {
  return (dtype == SPB_BF16 || dtype == SPB_F32) && (hd == 64 || hd == 128 || hd == 256);
}

template <typename T, int HD, int RB>
static int attn_decode_launch_rb(const DecArgs& a, cudaStream_t st) {
  constexpr int G = DEC_THREADS / (HD / Vec16<T>::N);
  const size_t dyn = (size_t)G * RB * HD * sizeof(float);
  SPB_CUDA(cudaFuncSetAttribute(attn_decode_kernel<T, HD, RB>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)dyn));
  dim3 grid(a.splits, a.Hkv * a.row_groups, a.B);
  attn_decode_kernel<T, HD, RB><<<grid, DEC_THREADS, dyn, st>>>(a);
  SPB_LAUNCH_CHECK();
  return 0;
}

template <typename T, int HD>
static int attn_decode_launch(DecArgs a, int RB, void* o, int64_t ldo, float* lse, cudaStream_t st) {
  int rc;
  if (RB == 8) rc = attn_decode_launch_rb<T, HD, 8>(a, st);
  else if (RB == 4) rc = attn_decode_launch_rb<T, HD, 4>(a, st);
  else if (RB == 2) rc = attn_decode_launch_rb<T, HD, 2>(a, st);
  else rc = attn_decode_launch_rb<T, HD, 1>(a, st);
  if (rc) return rc;
  attn_decode_combine_kernel<T><<<(unsigned)((int64_t)a.B * a.Hq * a.Tq), 128, 0, st>>>(a.po, a.pml, (T*)o, ldo, lse, a.B, a.Tq, a.Hq, HD, a.splits);
  SPB_LAUNCH_CHECK();
  return 0;
}

int attn_decode(int dtype, int B, int Tq, int Hq, int Hkv, int hd, int Tk, float scale, const void* q, int64_t ldq, const void* k, int64_t ldk,
                int64_t kbs, const void* v, int64_t ldv, int64_t vbs, void* o, int64_t ldo, float* lse, void* scratch, cudaStream_t st) {
  SPB_CHECK_ARG(attn_decode_supported(dtype, hd), SPB_E_UNSUPPORTED, "attention decode: dtype %d / head dim %d (fp32 or bf16; 64, 128, 256)", dtype, hd);
  const int ve = dtype == SPB_BF16 ? 8 : 4;
  SPB_CHECK_ARG(ldk % ve == 0 && ldv % ve == 0 && kbs % ve == 0 && vbs % ve == 0 && (uintptr_t)k % 16 == 0 && (uintptr_t)v % 16 == 0, SPB_E_UNSUPPORTED,
                "attention decode: the caches need 16-byte aligned rows");
  SPB_CHECK_ARG((int64_t)B * Hq * Tq < (1LL << 31) && B <= 65535 && Tq >= 1 && Tk >= 1, SPB_E_LIMIT, "attention decode: size beyond the grid limits");
  DecArgs a;
  int RB;
  decode_plan(B, Tq, Hq, Hkv, Tk, RB, a.row_groups, a.splits, a.tiles_per_split);
  SPB_CHECK_ARG((int64_t)Hkv * a.row_groups <= 65535, SPB_E_LIMIT, "attention decode: too many query rows per kv head");
  a.q = q; a.ldq = ldq; a.k = k; a.v = v; a.ldk = ldk; a.ldv = ldv; a.kbs = kbs; a.vbs = vbs;
  a.B = B; a.Tq = Tq; a.Hq = Hq; a.Hkv = Hkv; a.Tk = Tk; a.scale = scale;
  a.po = (float*)scratch;
  a.pml = a.po + (size_t)B * Hq * Tq * a.splits * hd;
#define SPB_DEC_CASE(TY, HDV) return attn_decode_launch<TY, HDV>(a, RB, o, ldo, lse, st)
  if (dtype == SPB_BF16) {
    if (hd == 64) SPB_DEC_CASE(__nv_bfloat16, 64);
    if (hd == 128) SPB_DEC_CASE(__nv_bfloat16, 128);
    SPB_DEC_CASE(__nv_bfloat16, 256);
  }
  if (hd == 64) SPB_DEC_CASE(float, 64);
  if (hd == 128) SPB_DEC_CASE(float, 128);
  SPB_DEC_CASE(float, 256);
#undef SPB_DEC_CASE
}

template <typename T>
static int rope_append_t(void* qkv, int64_t ld, int B, int Tq, int Hq, int Hkv, int hd, const float* rope, int pos0, void* kc, void* vc,
                         int64_t ldc, int64_t cbs, cudaStream_t st) {
  const int64_t total = (int64_t)B * Tq * (Hq + 2 * Hkv) * (hd / 2);
  decode_rope_append_kernel<T><<<(unsigned)((total + 255) / 256), 256, 0, st>>>((T*)qkv, ld, B, Tq, Hq, Hkv, hd, (const float4*)rope, pos0, (T*)kc,
                                                                             (T*)vc, ldc, cbs);
  SPB_LAUNCH_CHECK();
  return 0;
}

int decode_rope_append(int dtype, void* qkv, int64_t ld, int B, int Tq, int Hq, int Hkv, int hd, const float* rope, int pos0, void* kc, void* vc,
                       int64_t ldc, int64_t cbs, cudaStream_t st) {
  SPB_CHECK_ARG(hd % 2 == 0, SPB_E_INVALID, "decode: odd head dim");
  if (dtype == SPB_BF16) return rope_append_t<__nv_bfloat16>(qkv, ld, B, Tq, Hq, Hkv, hd, rope, pos0, kc, vc, ldc, cbs, st);
  return rope_append_t<float>(qkv, ld, B, Tq, Hq, Hkv, hd, rope, pos0, kc, vc, ldc, cbs, st);
}

}  // namespace spb
