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

__device__ __forceinline__ void cp_async16(void* smem_dst, const void* gsrc) {
  asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"(smem_u32(smem_dst)), "l"(gsrc) : "memory");
}
__device__ __forceinline__ void cp_async_commit() { asm volatile("cp.async.commit_group;" ::: "memory"); }
template <int N>
__device__ __forceinline__ void cp_async_wait() {
  asm volatile("cp.async.wait_group %0;" ::"n"(N) : "memory");
}
template <typename T>
__device__ __forceinline__ void smem_vec(const uint8_t* p, float* f);
template <>
__device__ __forceinline__ void smem_vec<float>(const uint8_t* p, float* f) {
  const float4 v = *reinterpret_cast<const float4*>(p);
  f[0] = v.x; f[1] = v.y; f[2] = v.z; f[3] = v.w;
}
template <>
__device__ __forceinline__ void smem_vec<__nv_bfloat16>(const uint8_t* p, float* f) {
  const uint4 v = *reinterpret_cast<const uint4*>(p);
  const float2 a = unpack_bf16(v.x), b = unpack_bf16(v.y), c = unpack_bf16(v.z), d = unpack_bf16(v.w);
  f[0] = a.x; f[1] = a.y; f[2] = b.x; f[3] = b.y; f[4] = c.x; f[5] = c.y; f[6] = d.x; f[7] = d.y;
}

// ------------------------------------------------------------------------------------------------ K5: small-M GEMM
// 1-D TMA bulk copy (no tensor map): `bytes` contiguous bytes global -> shared, completion credited to an mbarrier
__device__ __forceinline__ void bulk_load(void* smem_dst, const void* gsrc, uint32_t bytes, uint64_t* bar) {
  asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(smem_u32(smem_dst)), "l"(gsrc),
               "r"(bytes), "r"(smem_u32(bar))
               : "memory");
}

constexpr int SMG_ROWS = 16;    // weight rows per pipeline stage
constexpr int SMG_STAGES = 4;   // stages in flight: 4 x 16 rows x 2 KB = 128 KB of weight bytes per SM
constexpr int SMG_PANEL_BYTES = DEC_THREADS * 16;

// grid (column panels, k splits, row groups of MB); W is [K, N] row-major (row stride ldw): SPB_W_IN_OUT.
// One CTA per (panel, k slice): thread 0 keeps SMG_STAGES stages of 16 weight-row segments (2 KB each, contiguous in
// HBM) in flight with bulk copies; all 128 threads consume a stage from shared memory (thread = one 16-byte column piece)
// against the MB activation rows held in shared memory as fp32.
template <typename T, int MB>
__global__ void __launch_bounds__(DEC_THREADS) small_m_gemm_kernel(const T* __restrict__ A, int64_t lda, const T* __restrict__ W, int64_t ldw,
                                                                  float* __restrict__ part, int M, int N, int K, int k_per_split) {
  constexpr int VE = Vec16<T>::N;
  extern __shared__ __align__(128) uint8_t smg[];
  uint8_t* stages = smg;                                                        // [SMG_STAGES][SMG_ROWS][SMG_PANEL_BYTES]
  float* xs = reinterpret_cast<float*>(smg + SMG_STAGES * SMG_ROWS * SMG_PANEL_BYTES);  // [MB][k_per_split]
  __shared__ uint64_t full[SMG_STAGES];
  const int m0 = blockIdx.z * MB;
  const int col0 = blockIdx.x * DEC_THREADS * VE;
  const int col = col0 + threadIdx.x * VE;
  const int kb = blockIdx.y * k_per_split, ke = min(K, kb + k_per_split);
  const int klen = ke - kb;
  const int ncols = min(DEC_THREADS * VE, N - col0);
  const uint32_t row_bytes = (uint32_t)ncols * sizeof(T);
  const bool live = col < N;  // N % VE == 0 is checked by the launcher
  const int nst = (klen + SMG_ROWS - 1) / SMG_ROWS;

  auto issue = [&](int st) {  // thread 0: rows [st*16, st*16+16) of the slice -> stage st % SMG_STAGES
    const int r0 = st * SMG_ROWS, rn = min(SMG_ROWS, klen - r0);
    uint64_t* bar = &full[st % SMG_STAGES];
    mbar_arrive_expect_tx(bar, rn * row_bytes);
    uint8_t* dst = stages + (size_t)(st % SMG_STAGES) * SMG_ROWS * SMG_PANEL_BYTES;
    for (int r = 0; r < rn; ++r) bulk_load(dst + r * SMG_PANEL_BYTES, W + (int64_t)(kb + r0 + r) * ldw + col0, row_bytes, bar);
  };
  if (threadIdx.x == 0) {
    for (int i = 0; i < SMG_STAGES; ++i) mbar_init(&full[i], 1);
    fence_barrier_init();
    for (int st = 0; st < min(nst, SMG_STAGES); ++st) issue(st);
  }
  for (int i = threadIdx.x; i < MB * klen; i += DEC_THREADS) {
    const int m = i / klen, kk = i % klen;
    xs[m * k_per_split + kk] = (m0 + m < M) ? to_f32<T>(A[(int64_t)(m0 + m) * lda + kb + kk]) : 0.f;
  }
  float acc[MB][VE];
#pragma unroll
  for (int m = 0; m < MB; ++m)
#pragma unroll
    for (int e = 0; e < VE; ++e) acc[m][e] = 0.f;
  __syncthreads();  // barriers initialised, xs complete
  for (int st = 0; st < nst; ++st) {
    const int r0 = st * SMG_ROWS, rn = min(SMG_ROWS, klen - r0);
    mbar_wait(&full[st % SMG_STAGES], (st / SMG_STAGES) & 1);
    if (live) {
      const uint8_t* src = stages + (size_t)(st % SMG_STAGES) * SMG_ROWS * SMG_PANEL_BYTES + threadIdx.x * 16;
      if (rn == SMG_ROWS) {
#pragma unroll
        for (int r = 0; r < SMG_ROWS; ++r) {
          float w[VE];
          smem_vec<T>(src + r * SMG_PANEL_BYTES, w);
#pragma unroll
          for (int m = 0; m < MB; ++m) {
            const float xv = xs[m * k_per_split + r0 + r];
#pragma unroll
            for (int e = 0; e < VE; ++e) acc[m][e] = fmaf(xv, w[e], acc[m][e]);
          }
        }
      } else {
        for (int r = 0; r < rn; ++r) {
          float w[VE];
          smem_vec<T>(src + r * SMG_PANEL_BYTES, w);
#pragma unroll
          for (int m = 0; m < MB; ++m) {
            const float xv = xs[m * k_per_split + r0 + r];
#pragma unroll
            for (int e = 0; e < VE; ++e) acc[m][e] = fmaf(xv, w[e], acc[m][e]);
          }
        }
      }
    }
    __syncthreads();  // everyone is done with this stage: refill it
    if (threadIdx.x == 0 && st + SMG_STAGES < nst) issue(st + SMG_STAGES);
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
                                                            const T* __restrict__ bias, const T* __restrict__ res, int64_t ldr, int M, int N,
                                                            int* __restrict__ advance, int by, int cap) {
  pdl_launch_dependents();
  pdl_wait();
  const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (advance != nullptr && i == 0) *advance = min(*advance + by, cap);  // every reader of the length in this step is complete (stream order)
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
  int want = sm_count() / panels;  // one CTA per SM (each keeps 128 KB of weight bytes in flight), the whole grid in ONE wave
  int max_splits = K / 32;
  if (max_splits < 1) max_splits = 1;
  if (want > max_splits) want = max_splits;
  if (want > 64) want = 64;
  if (want < 1) want = 1;
  return want;
}

size_t small_m_gemm_scratch_bytes(int M, int N, int K, int dtype) {
  size_t b = (size_t)small_m_splits(M, N, K, dtype == SPB_BF16 ? 8 : 4) * M * N * sizeof(float);
  if (dtype == SPB_BF16 && M <= 64) {
    const size_t m = small_m_mma_scratch_bytes(M, N, K);
    if (m > b) b = m;
  }
  return b;
}

template <typename T>
static int small_m_gemm_t(int M, int N, int K, const T* A, int64_t lda, const T* W, int64_t ldw, T* D, int64_t ldd, const T* bias, const T* res,
                          int64_t ldr, float* part, cudaStream_t st, int* advance, int by, int cap) {
  constexpr int VE = Vec16<T>::N;
  SPB_CHECK_ARG(N % VE == 0 && ldw % VE == 0 && (uintptr_t)W % 16 == 0, SPB_E_UNSUPPORTED,
                "small-M GEMM: the weight needs 16-byte aligned rows (N=%d ldw=%lld)", N, (long long)ldw);
  const int splits = small_m_splits(M, N, K, VE);
  const int k_per = ((K + splits - 1) / splits + 3) / 4 * 4;
  const int nsplit = (K + k_per - 1) / k_per;
  const int MB = M >= 8 ? 8 : (M > 2 ? 4 : (M > 1 ? 2 : 1));
  dim3 grid((N + DEC_THREADS * VE - 1) / (DEC_THREADS * VE), nsplit, (M + MB - 1) / MB);
  const size_t dyn = (size_t)SMG_STAGES * SMG_ROWS * SMG_PANEL_BYTES + (size_t)MB * k_per * sizeof(float);
  SPB_CHECK_ARG(dyn <= 200 * 1024, SPB_E_LIMIT, "small-M GEMM: K slice of %d rows does not fit shared memory", k_per);
#define SPB_SMG_LAUNCH(MBV)                                                                                                      \
  do {                                                                                                                           \
    SPB_CUDA(cudaFuncSetAttribute(small_m_gemm_kernel<T, MBV>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)dyn));          \
    small_m_gemm_kernel<T, MBV><<<grid, DEC_THREADS, dyn, st>>>(A, lda, W, ldw, part, M, N, K, k_per);                           \
  } while (0)
  if (MB == 8) SPB_SMG_LAUNCH(8);
  else if (MB == 4) SPB_SMG_LAUNCH(4);
  else if (MB == 2) SPB_SMG_LAUNCH(2);
  else SPB_SMG_LAUNCH(1);
#undef SPB_SMG_LAUNCH
  SPB_LAUNCH_CHECK();
  const int64_t total = (int64_t)M * N;
  small_m_reduce_kernel<T><<<(unsigned)((total + 255) / 256), 256, 0, st>>>(part, nsplit, D, ldd, bias, res, ldr, M, N, advance, by, cap);
  SPB_LAUNCH_CHECK();
  return 0;
}

// D[M,N] = A[M,K] . W[K,N] (+bias) (+res); M <= 64 rows; `part` = small_m_gemm_scratch_bytes(M, N, K, dtype)
int small_m_gemm(int dtype, int M, int N, int K, const void* A, int64_t lda, const void* W, int64_t ldw, void* D, int64_t ldd, const void* bias,
                 const void* res, int64_t ldr, float* part, cudaStream_t st, int* advance, int by, int cap) {
  SPB_CHECK_ARG(M >= 1 && M <= 64 && N >= 1 && K >= 1, SPB_E_LIMIT, "small-M GEMM: M=%d rows (1..64 supported; longer inputs go through the prefill path)", M);
  // bf16: the weight stream on mma.sync (decode_gemm_mma.cu): K slices -> fp32 partials, summed below
  if (small_m_mma_ok(dtype, M, N, K, A, lda, W, ldw)) {
    int nsplit = 1;
    int rc = small_m_mma(M, N, K, A, lda, W, ldw, part, &nsplit, st);
    if (rc) return rc;
    const int64_t total = (int64_t)M * N;
    SPB_CUDA(launch_pdl(small_m_reduce_kernel<__nv_bfloat16>, dim3((unsigned)((total + 255) / 256)), dim3(256), 0, st, decode_pdl_enabled(),
                        (const float*)part, nsplit, (__nv_bfloat16*)D, ldd, (const __nv_bfloat16*)bias, (const __nv_bfloat16*)res, ldr, M, N, advance, by, cap));
    SPB_LAUNCH_CHECK();
    return 0;
  }
  // SPB_DECODE_GEMM_MMA=0: the tcgen05 GEMM with its K-split for few-tile problems (one 128-row tile, N/256 column tiles, K cut into up to 8
  // parts that TMA-reduce into a zeroed fp32 [M, N]) streams the weight at HBM speed with no CUDA-core arithmetic at all -
  // the SIMT kernel below is bound by FMA issue on one warp per scheduler (ncu r02d: 28 % issue-active, 34 k cycles for 50 MB).
  // The fp32 result is finished (bias, residual, conversion) by the reduce kernel with one "split".
  if (dtype == SPB_BF16 && lda % 8 == 0 && ldw % 8 == 0 && (uintptr_t)A % 16 == 0 && (uintptr_t)W % 16 == 0 && N % 4 == 0 && K >= 1024) {
    int rc = gemm_tc(SPB_F32, M, N, K, A, lda, 0, W, ldw, 0, part, N, nullptr, 0, st);
    if (rc) return rc;
    const int64_t total = (int64_t)M * N;
    small_m_reduce_kernel<__nv_bfloat16><<<(unsigned)((total + 255) / 256), 256, 0, st>>>(part, 1, (__nv_bfloat16*)D, ldd, (const __nv_bfloat16*)bias,
                                                                                      (const __nv_bfloat16*)res, ldr, M, N, advance, by, cap);
    SPB_LAUNCH_CHECK();
    return 0;
  }
  if (dtype == SPB_BF16)
    return small_m_gemm_t<__nv_bfloat16>(M, N, K, (const __nv_bfloat16*)A, lda, (const __nv_bfloat16*)W, ldw, (__nv_bfloat16*)D, ldd,
                                         (const __nv_bfloat16*)bias, (const __nv_bfloat16*)res, ldr, part, st, advance, by, cap);
  return small_m_gemm_t<float>(M, N, K, (const float*)A, lda, (const float*)W, ldw, (float*)D, ldd, (const float*)bias, (const float*)res, ldr, part, st,
                               advance, by, cap);
}

// ------------------------------------------------------------------------------------------------ K6: rotary + cache append
// qkv [B*Tq, ld] (q | k | v columns).  q is rotated in place; k is rotated and written, v copied, into the caches at token
// rows pos0 .. pos0+Tq-1 of each batch entry.  One thread per adjacent pair (2j, 2j+1) of a (row, head).
// With `part` != nullptr the kernel is also the finish of the projection: the pair is first summed over the K slices'
// fp32 partials [nsplit][B*Tq][ld] (+ bias) and rounded to T - what small_m_reduce_kernel would have stored - so the
// projection's reduce pass and this one are a single launch.
template <typename T>
__global__ void __launch_bounds__(256) decode_rope_append_kernel(T* __restrict__ qkv, int64_t ld, int B, int Tq, int Hq, int Hkv, int hd,
                                                                const float4* __restrict__ rope, int pos0, T* __restrict__ kc,
                                                                T* __restrict__ vc, int64_t ldc, int64_t cache_batch_stride,
                                                                const float* __restrict__ part, int nsplit, const T* __restrict__ bias,
                                                                const int* __restrict__ pos0_dev, int cap) {
  pdl_launch_dependents();
  pdl_wait();
  if (pos0_dev != nullptr) pos0 = *pos0_dev;  // recorded step: the cache length is read here, not baked into the launch
  const int hp = hd / 2;
  const int heads = Hq + 2 * Hkv;
  const int64_t idx = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  const int64_t total = (int64_t)B * Tq * heads * hp;
  if (idx >= total) return;
  const int j = (int)(idx % hp);
  const int h = (int)((idx / hp) % heads);
  const int64_t row = idx / ((int64_t)hp * heads);  // b * Tq + t
  const int t = (int)(row % Tq), b = (int)(row / Tq);
  if (pos0_dev != nullptr && pos0 + t >= cap) return;  // a replay past the capacity writes nothing (the host wrapper refuses it first)
  T* src = qkv + row * ld + (int64_t)h * hd + 2 * j;
  float xe, xo;
  if (part != nullptr) {
    const int64_t col = (int64_t)h * hd + 2 * j;
    const int64_t slab = (int64_t)B * Tq * ld;
    float se = 0.f, so = 0.f;
    for (int p = 0; p < nsplit; ++p) {
      const float2 v = *reinterpret_cast<const float2*>(part + p * slab + row * ld + col);
      se += v.x;
      so += v.y;
    }
    if (bias != nullptr) {
      se += to_f32<T>(bias[col]);
      so += to_f32<T>(bias[col + 1]);
    }
    xe = to_f32<T>(from_f32<T>(se));
    xo = to_f32<T>(from_f32<T>(so));
  } else {
    xe = to_f32<T>(src[0]);
    xo = to_f32<T>(src[1]);
  }
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
  const int* tk_dev;  // != nullptr: valid keys = min(*tk_dev + tk_add, Tk), read on the device (a recorded step)
  int tk_add;
};

// A CTA of 256 threads streams its slice of the cache in tiles of KT keys through a two-stage shared-memory ring filled with
// 16-byte cp.async copies (coalesced: TPR consecutive threads cover one head row; a whole tile of K and of V - none of it
// through registers - is in flight while the previous tile is being used).  Scores: two threads per key (interleaved 16-byte
// pieces, one shuffle to combine); P.V: thread = 16-byte column piece of every G-th key.  Rows are padded by 32 bytes so
// that both access patterns are free of bank conflicts.  One CTA per SM (the ring takes most of the shared memory); the
// host picks the key split so that the grid is a single wave where it can.
constexpr int DA_THREADS = 256;
template <typename T, int HD>
struct DecTile {
  static constexpr int ROWB = HD * (int)sizeof(T) + 32;            // padded row pitch in bytes
  static constexpr int KT = (2 * 2 * 128 * ROWB <= 200 * 1024) ? 128 : ((2 * 2 * 64 * ROWB <= 200 * 1024) ? 64 : 32);  // two stages of K and V must fit
  static constexpr int STAGE_BYTES = 2 * KT * ROWB;
};

template <typename T, int HD, int RB>
__global__ void __launch_bounds__(DA_THREADS) attn_decode_kernel(const DecArgs p) {
  constexpr int VE = Vec16<T>::N;
  constexpr int TPR = HD / VE;              // 16-byte pieces (threads) per head row
  constexpr int G = DA_THREADS / TPR;       // rows covered per pass of the CTA
  constexpr int KT = DecTile<T, HD>::KT;
  constexpr int ROWB = DecTile<T, HD>::ROWB;
  constexpr int NW = DA_THREADS / 32;
  static_assert(DA_THREADS % TPR == 0 && TPR % 2 == 0, "head dim does not tile the CTA");
  __shared__ float qs[RB][HD];
  __shared__ float ps[RB][KT];
  __shared__ float red[2][NW][RB];
  extern __shared__ __align__(16) uint8_t dsm[];  // [2 stages][K tile [KT][ROWB] | V tile [KT][ROWB]]; reused for the final reduction

  const int split = blockIdx.x;
  const int g = blockIdx.y / p.row_groups, rg = blockIdx.y % p.row_groups;
  const int b = blockIdx.z;
  const int n_rep = p.Hq / p.Hkv;
  const int rtot = n_rep * p.Tq;
  const int r0 = rg * RB;
  const int nr = min(RB, rtot - r0);
  const int t = threadIdx.x, lane = t & 31, warp = t >> 5;
  const T* kbase = (const T*)p.k + (int64_t)b * p.kbs + (int64_t)g * HD;
  const T* vbase = (const T*)p.v + (int64_t)b * p.vbs + (int64_t)g * HD;
  const int vg = t / TPR, vc = (t % TPR) * VE;
  int Tk = p.Tk, tps = p.tiles_per_split;
  if (p.tk_dev != nullptr) {  // grid sized for the capacity: share the tiles that exist evenly over its splits
    Tk = min(*p.tk_dev + p.tk_add, p.Tk);
    tps = ((Tk + KT - 1) / KT + p.splits - 1) / p.splits;
  }
  const int tile0 = split * tps;
  const int tile1 = min(tile0 + tps, (Tk + KT - 1) / KT);

  auto stage_tile = [&](int tile) {  // one cp.async group per tile: K and V rows of the tile -> stage (tile - tile0) & 1
    const int k0 = tile * KT, kn = min(KT, Tk - k0);
    uint8_t* sK = dsm + ((tile - tile0) & 1) * DecTile<T, HD>::STAGE_BYTES;
    uint8_t* sV = sK + KT * ROWB;
    for (int jj = vg; jj < kn; jj += G) {
      cp_async16(sK + jj * ROWB + vc * (int)sizeof(T), kbase + (int64_t)(k0 + jj) * p.ldk + vc);
      cp_async16(sV + jj * ROWB + vc * (int)sizeof(T), vbase + (int64_t)(k0 + jj) * p.ldv + vc);
    }
    cp_async_commit();
  };
  if (tile0 < tile1) stage_tile(tile0);
  if (tile0 + 1 < tile1) stage_tile(tile0 + 1);

  // the group's query rows -> shared memory, pre-multiplied by the score scale.  Row r = (local q head, tq).
  for (int i = t; i < RB * HD; i += DA_THREADS) {
    const int r = i / HD, d = i % HD;
    float qv = 0.f;
    if (r < nr) {
      const int hq = g * n_rep + (r0 + r) / p.Tq, tq = (r0 + r) % p.Tq;
      qv = to_f32<T>(((const T*)p.q)[((int64_t)b * p.Tq + tq) * p.ldq + (int64_t)hq * HD + d]) * p.scale;
    }
    qs[r][d] = qv;
  }

  float m_run[RB], l_run[RB], acc[RB][VE];
#pragma unroll
  for (int r = 0; r < RB; ++r) {
    m_run[r] = -INFINITY;
    l_run[r] = 0.f;
#pragma unroll
    for (int e = 0; e < VE; ++e) acc[r][e] = 0.f;
  }
  const int key = t >> 1, half = t & 1;  // score phase: two threads per key, 16-byte pieces 2i + half
  for (int tile = tile0; tile < tile1; ++tile) {
    const int k0 = tile * KT;
    const int kn = min(KT, Tk - k0);
    const uint8_t* sK = dsm + ((tile - tile0) & 1) * DecTile<T, HD>::STAGE_BYTES;
    const uint8_t* sV = sK + KT * ROWB;
    if (tile + 1 < tile1) cp_async_wait<1>(); else cp_async_wait<0>();  // this thread's copies of this tile have landed
    __syncthreads();                                                    // everybody's (and qs)
    // ---- scores
    float s[RB];
#pragma unroll
    for (int r = 0; r < RB; ++r) s[r] = 0.f;
    if (key < kn) {
      const uint8_t* kr = sK + key * ROWB;
#pragma unroll
      for (int i = 0; i < TPR / 2; ++i) {
        const int c = 2 * i + half;
        float kv[VE];
        smem_vec<T>(kr + c * 16, kv);
#pragma unroll
        for (int e = 0; e < VE; e += 4) {
#pragma unroll
          for (int r = 0; r < RB; ++r) {
            const float4 qv = *reinterpret_cast<const float4*>(&qs[r][c * VE + e]);
            s[r] = fmaf(qv.x, kv[e], fmaf(qv.y, kv[e + 1], fmaf(qv.z, kv[e + 2], fmaf(qv.w, kv[e + 3], s[r]))));
          }
        }
      }
    }
#pragma unroll
    for (int r = 0; r < RB; ++r) {  // the two halves of a key (all lanes take part: no shuffle under divergence)
      s[r] += __shfl_xor_sync(0xffffffffu, s[r], 1);
      if (key >= kn) s[r] = -INFINITY;
    }
    // ---- tile max per row (both threads of a key hold the same score)
#pragma unroll
    for (int r = 0; r < RB; ++r) {
      const float wm = warp_max(s[r]);
      if (lane == 0) red[0][warp][r] = wm;
    }
    __syncthreads();
    float alpha[RB];
#pragma unroll
    for (int r = 0; r < RB; ++r) {
      float tm = red[0][0][r];
#pragma unroll
      for (int w = 1; w < NW; ++w) tm = fmaxf(tm, red[0][w][r]);
      const float mn = fmaxf(m_run[r], tm);
      alpha[r] = (m_run[r] == -INFINITY) ? 0.f : __expf(m_run[r] - mn);
      m_run[r] = mn;
      const float pv = (s[r] == -INFINITY || half) ? 0.f : __expf(s[r] - mn);  // one of the two threads of a key contributes
      if (!half && key < KT) ps[r][key] = pv;
      const float wsum = warp_sum(pv);
      if (lane == 0) red[1][warp][r] = wsum;
    }
    __syncthreads();
#pragma unroll
    for (int r = 0; r < RB; ++r) {
      float ts = red[1][0][r];
#pragma unroll
      for (int w = 1; w < NW; ++w) ts += red[1][w][r];
      l_run[r] = l_run[r] * alpha[r] + ts;
#pragma unroll
      for (int e = 0; e < VE; ++e) acc[r][e] *= alpha[r];
    }
    // ---- acc[r][.] += sum_j P[r][j] * V[j][vc ..]: thread owns a 16-byte column piece of every G-th key
    for (int jj = vg; jj < kn; jj += G) {
      float vv[VE];
      smem_vec<T>(sV + jj * ROWB + vc * (int)sizeof(T), vv);
#pragma unroll
      for (int r = 0; r < RB; ++r) {
        const float pv = ps[r][jj];
#pragma unroll
        for (int e = 0; e < VE; ++e) acc[r][e] = fmaf(pv, vv[e], acc[r][e]);
      }
    }
    __syncthreads();  // this stage (and ps) is free again
    if (tile + 2 < tile1) stage_tile(tile + 2);
  }
  // ---- sum the G key groups' accumulators (through the tile buffer), write this split's partial (unnormalised O, m, l)
  float* osum = reinterpret_cast<float*>(dsm);  // [G][RB][HD]
#pragma unroll
  for (int r = 0; r < RB; ++r)
#pragma unroll
    for (int e = 0; e < VE; ++e) osum[((int64_t)vg * RB + r) * HD + vc + e] = acc[r][e];
  __syncthreads();
  for (int i = t; i < nr * HD; i += DA_THREADS) {
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
  pdl_launch_dependents();
  pdl_wait();
  const int64_t row = blockIdx.x;  // (b * Hq + hq) * Tq + tq
  const int tq = (int)(row % Tq), hq = (int)((row / Tq) % Hq), b = (int)(row / ((int64_t)Tq * Hq));
  // the splits' (m, l) once into shared memory (splits <= 64 < blockDim), then w_s = exp(m_s - M) and w_s * l_s once per split
  __shared__ float sm_m[64], sm_w[64], sm_wl[64];
  const int t = threadIdx.x;
  float lv = 0.f;
  if (t < splits) {
    const float2 ml = *reinterpret_cast<const float2*>(pml + (row * splits + t) * 2);
    sm_m[t] = ml.x;
    lv = ml.y;
  }
  __syncthreads();
  float M = -INFINITY;
  for (int s = 0; s < splits; ++s) M = fmaxf(M, sm_m[s]);
  if (t < splits) {
    const float w = (sm_m[t] == -INFINITY) ? 0.f : __expf(sm_m[t] - M);
    sm_w[t] = w;
    sm_wl[t] = w * lv;
  }
  __syncthreads();
  float L = 0.f;
  for (int s = 0; s < splits; ++s) L += sm_wl[s];  // fixed order, every thread the same sum
  const float inv = 1.0f / L;
  for (int d = t; d < hd; d += blockDim.x) {
    float acc = 0.f;
#pragma unroll 4
    for (int s = 0; s < splits; ++s) {
      const float w = sm_w[s];
      if (w != 0.f) acc += w * po[(row * splits + s) * hd + d];
    }
    o[((int64_t)b * Tq + tq) * ldo + (int64_t)hq * hd + d] = from_f32<T>(acc * inv);
  }
  if (lse != nullptr && t == 0) lse[row] = M + logf(L);
}

int attn_decode_combine(int dtype, const float* po, const float* pml, void* o, int64_t ldo, float* lse, int B, int Tq, int Hq, int hd, int splits,
                        cudaStream_t st) {
  const unsigned rows = (unsigned)((int64_t)B * Hq * Tq);
  if (dtype == SPB_BF16)
    SPB_CUDA(launch_pdl(attn_decode_combine_kernel<__nv_bfloat16>, dim3(rows), dim3(128), 0, st, decode_pdl_enabled(), po, pml, (__nv_bfloat16*)o, ldo, lse, B,
                        Tq, Hq, hd, splits));
  else
    SPB_CUDA(launch_pdl(attn_decode_combine_kernel<float>, dim3(rows), dim3(128), 0, st, decode_pdl_enabled(), po, pml, (float*)o, ldo, lse, B, Tq, Hq, hd,
                        splits));
  SPB_LAUNCH_CHECK();
  return 0;
}

static int decode_kt(int dtype, int hd) {  // DecTile<T, HD>::KT on the host
  const int rowb = hd * (dtype == SPB_BF16 ? 2 : 4) + 32;
  return (2 * 2 * 128 * rowb <= 200 * 1024) ? 128 : ((2 * 2 * 64 * rowb <= 200 * 1024) ? 64 : 32);
}

// One CTA per SM: choose the key split so that the grid is one wave (or whole waves) of the machine where the tile count allows.
static void decode_plan(int dtype, int hd, int B, int Tq, int Hq, int Hkv, int Tk, int& RB, int& row_groups, int& splits, int& tiles_per_split) {
  const int rtot = (Hq / Hkv) * Tq;
  RB = rtot >= 8 ? 8 : (rtot > 2 ? 4 : (rtot > 1 ? 2 : 1));
  row_groups = (rtot + RB - 1) / RB;
  const int kt = decode_kt(dtype, hd);
  const int tiles = (Tk + kt - 1) / kt;
  const int ctas = B * Hkv * row_groups;
  const int sms = sm_count();
  int want = sms / ctas;               // splits that still fit one wave
  if (want < 1) want = 1;
  if (want > tiles) want = tiles;
  if (want > 64) want = 64;
  tiles_per_split = (tiles + want - 1) / want;
  splits = (tiles + tiles_per_split - 1) / tiles_per_split;
}

size_t attn_decode_scratch_bytes(int B, int Tq, int Hq, int Hkv, int hd, int Tk) {
  // worst case over both element types (the split count does not depend on the type except through the tile height)
  size_t worst = 0;
  for (int dtype = 0; dtype < 2; ++dtype) {
    int RB, rgs, splits, tps;
    decode_plan(dtype, hd, B, Tq, Hq, Hkv, Tk, RB, rgs, splits, tps);
    const size_t b = (size_t)B * Hq * Tq * splits * ((size_t)hd + 2) * sizeof(float) + 256;
    if (b > worst) worst = b;
  }
  if (attn_decode_mma_supported(SPB_BF16, hd)) {
    const size_t b = attn_decode_mma_scratch_bytes(B, Tq, Hq, Hkv, hd, Tk);
    if (b > worst) worst = b;
  }
  return worst;
}

bool attn_decode_supported(int dtype, int hd) {
  return (dtype == SPB_BF16 || dtype == SPB_F32) && (hd == 64 || hd == 128 || hd == 256);
}

template <typename T, int HD, int RB>
static int attn_decode_launch_rb(const DecArgs& a, cudaStream_t st) {
  constexpr int G = DA_THREADS / (HD / Vec16<T>::N);
  size_t dyn = 2 * (size_t)DecTile<T, HD>::STAGE_BYTES;
  if ((size_t)G * RB * HD * sizeof(float) > dyn) dyn = (size_t)G * RB * HD * sizeof(float);  // final reduction reuses the buffer
  SPB_CUDA(cudaFuncSetAttribute(attn_decode_kernel<T, HD, RB>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)dyn));
  dim3 grid(a.splits, a.Hkv * a.row_groups, a.B);
  attn_decode_kernel<T, HD, RB><<<grid, DA_THREADS, dyn, st>>>(a);
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
                int64_t kbs, const void* v, int64_t ldv, int64_t vbs, void* o, int64_t ldo, float* lse, void* scratch, cudaStream_t st,
                const int* tk_dev, int tk_add) {
  SPB_CHECK_ARG(attn_decode_supported(dtype, hd), SPB_E_UNSUPPORTED, "attention decode: dtype %d / head dim %d (fp32 or bf16; 64, 128, 256)", dtype, hd);
  const int ve = dtype == SPB_BF16 ? 8 : 4;
  SPB_CHECK_ARG(ldk % ve == 0 && ldv % ve == 0 && kbs % ve == 0 && vbs % ve == 0 && (uintptr_t)k % 16 == 0 && (uintptr_t)v % 16 == 0, SPB_E_UNSUPPORTED,
                "attention decode: the caches need 16-byte aligned rows");
  SPB_CHECK_ARG((int64_t)B * Hq * Tq < (1LL << 31) && B <= 65535 && Tq >= 1 && Tk >= 1, SPB_E_LIMIT, "attention decode: size beyond the grid limits");
  if (attn_decode_mma_supported(dtype, hd) && ldq % 8 == 0 && (uintptr_t)q % 16 == 0)  // bf16: the tensor-core kernel (attn_decode_mma.cu)
    return attn_decode_mma(B, Tq, Hq, Hkv, hd, Tk, scale, q, ldq, k, ldk, kbs, v, ldv, vbs, o, ldo, lse, scratch, st, tk_dev, tk_add);
  DecArgs a;
  int RB;
  decode_plan(dtype, hd, B, Tq, Hq, Hkv, Tk, RB, a.row_groups, a.splits, a.tiles_per_split);
  SPB_CHECK_ARG((int64_t)Hkv * a.row_groups <= 65535, SPB_E_LIMIT, "attention decode: too many query rows per kv head");
  a.q = q; a.ldq = ldq; a.k = k; a.v = v; a.ldk = ldk; a.ldv = ldv; a.kbs = kbs; a.vbs = vbs;
  a.B = B; a.Tq = Tq; a.Hq = Hq; a.Hkv = Hkv; a.Tk = Tk; a.scale = scale;
  a.tk_dev = tk_dev; a.tk_add = tk_add;
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
                         int64_t ldc, int64_t cbs, const float* part, int nsplit, const void* bias, const int* pos0_dev, int cap, cudaStream_t st) {
  const int64_t total = (int64_t)B * Tq * (Hq + 2 * Hkv) * (hd / 2);
  SPB_CUDA(launch_pdl(decode_rope_append_kernel<T>, dim3((unsigned)((total + 255) / 256)), dim3(256), 0, st, decode_pdl_enabled(), (T*)qkv, ld, B, Tq, Hq,
                      Hkv, hd, (const float4*)rope, pos0, (T*)kc, (T*)vc, ldc, cbs, part, nsplit, (const T*)bias, pos0_dev, cap));
  SPB_LAUNCH_CHECK();
  return 0;
}

int decode_rope_append(int dtype, void* qkv, int64_t ld, int B, int Tq, int Hq, int Hkv, int hd, const float* rope, int pos0, void* kc, void* vc,
                       int64_t ldc, int64_t cbs, cudaStream_t st, const float* part, int nsplit, const void* bias, const int* pos0_dev, int cap) {
  SPB_CHECK_ARG(hd % 2 == 0, SPB_E_INVALID, "decode: odd head dim");
  if (dtype == SPB_BF16)
    return rope_append_t<__nv_bfloat16>(qkv, ld, B, Tq, Hq, Hkv, hd, rope, pos0, kc, vc, ldc, cbs, part, nsplit, bias, pos0_dev, cap, st);
  return rope_append_t<float>(qkv, ld, B, Tq, Hq, Hkv, hd, rope, pos0, kc, vc, ldc, cbs, part, nsplit, bias, pos0_dev, cap, st);
}

}  // namespace spb
