// decode_gemm_mma.cu - SURVEY 8 row f3, bf16: the projections of the KV-cache branch (llama3/LLaMA-jax.ipynb:803-805, :829)
// for a handful of token rows: D[M <= 64, N] = A[M, K] . W[K, N], W in the notebook's [in, out] layout.
//
// One decode step reads every weight byte once and does ~M FMAs per weight element: the kernel must be a weight STREAM with
// next to no work per byte.  The SIMT version (decode.cu K5) converts each weight to fp32 and issues M FMAs per element - it
// is bound by instruction issue; the big tcgen05 GEMM with its K-split reads the weights well but pays a zeroing launch, a
// 128-row tile's prologue/epilogue and TMA reduce-adds for 4 live rows.  Here the token rows are the 16-row M of
// mma.sync.m16n8k16 (up to four such tiles), the weight goes HBM -> cp.async -> shared memory -> ldmatrix.trans -> tensor core
// as bf16, and a CTA is one (128-column panel, K slice): 64 weight rows x 256 bytes per stage, four stages in flight, two
// CTAs per SM.  The K slices write fp32 partials [slice, M, N]; decode.cu's small_m_reduce_kernel sums them in a fixed
// order (+ bias, + residual, -> bf16).  No atomics, no pre-zeroed buffer.
#include "common.cuh"
#include "kernels.h"

namespace spb {

namespace {

constexpr int SG_THREADS = 128;
constexpr int SG_PN = 128;                      // output columns per CTA (32 per warp)
constexpr int SG_KS = 64;                       // weight rows per stage
constexpr int SG_STAGES = 4;
constexpr int SG_WPITCH = SG_PN * 2 + 16;       // 272: 8 consecutive rows land in 8 different 16-byte bank groups
constexpr int SG_APITCH = SG_KS * 2 + 16;       // 144
constexpr int SG_WBYTES = SG_KS * SG_WPITCH;    // 17408

template <int MT>
struct SgCfg {
  static constexpr int ABYTES = 16 * MT * SG_APITCH;
  static constexpr int STAGE = SG_WBYTES + ABYTES;
  static constexpr int SMEM = SG_STAGES * STAGE;
};

__device__ __forceinline__ void sg_cp16(void* smem_dst, const void* gsrc, bool valid) {
  const int n = valid ? 16 : 0;  // src-size 0: zeros are written, nothing is read
  asm volatile("cp.async.cg.shared.global [%0], [%1], 16, %2;" ::"r"(smem_u32(smem_dst)), "l"(gsrc), "r"(n) : "memory");
}
__device__ __forceinline__ void sg_commit() { asm volatile("cp.async.commit_group;" ::: "memory"); }
template <int N>
__device__ __forceinline__ void sg_wait() {
  asm volatile("cp.async.wait_group %0;" ::"n"(N) : "memory");
}
__device__ __forceinline__ void sg_ldsm(uint32_t addr, uint32_t (&r)[4]) {
  asm volatile("ldmatrix.sync.aligned.m8n8.x4.shared.b16 {%0,%1,%2,%3}, [%4];" : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]) : "r"(addr));
}
__device__ __forceinline__ void sg_ldsm_t(uint32_t addr, uint32_t (&r)[4]) {
  asm volatile("ldmatrix.sync.aligned.m8n8.x4.trans.shared.b16 {%0,%1,%2,%3}, [%4];" : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]) : "r"(addr));
}
__device__ __forceinline__ void sg_mma(float (&d)[4], const uint32_t (&a)[4], uint32_t b0, uint32_t b1) {
  asm volatile("mma.sync.aligned.m16n8k16.row.col.f32.bf16.bf16.f32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, {%0,%1,%2,%3};"
               : "+f"(d[0]), "+f"(d[1]), "+f"(d[2]), "+f"(d[3])
               : "r"(a[0]), "r"(a[1]), "r"(a[2]), "r"(a[3]), "r"(b0), "r"(b1));
}

// grid (column panels, K slices).  part[slice][M][N] fp32.
template <int MT>
__global__ void __launch_bounds__(SG_THREADS) small_m_mma_kernel(const __nv_bfloat16* __restrict__ A, int64_t lda, const __nv_bfloat16* __restrict__ W,
                                                                 int64_t ldw, float* __restrict__ part, int M, int N, int K, int k_per_split) {
  using C = SgCfg<MT>;
  extern __shared__ __align__(128) uint8_t sg_smem[];
  const int t = threadIdx.x, lane = t & 31, warp = t >> 5;
  const int col0 = blockIdx.x * SG_PN;
  const int kb = blockIdx.y * k_per_split, ke = min(K, kb + k_per_split);
  const int nst = (ke - kb + SG_KS - 1) / SG_KS;

  // the weight does not depend on the previous kernel in the stream, the token rows do: under programmatic dependent
  // launch the first weight stages are already in flight while the producer of A is finishing
  auto stage_w = [&](int st) {
    if (st < nst) {
      const int k0 = kb + st * SG_KS;
      uint8_t* sW = sg_smem + (st % SG_STAGES) * C::STAGE;
#pragma unroll
      for (int i = t; i < SG_KS * (SG_PN / 8); i += SG_THREADS) {
        const int r = i / (SG_PN / 8), c = i % (SG_PN / 8);
        const bool ok = (k0 + r < ke) && (col0 + c * 8 < N);
        sg_cp16(sW + r * SG_WPITCH + c * 16, W + (ok ? ((int64_t)(k0 + r) * ldw + col0 + c * 8) : 0), ok);
      }
    }
  };
  auto stage_a = [&](int st) {  // closes the stage's commit group (empty past the end: the wait counts stay uniform)
    if (st < nst) {
      const int k0 = kb + st * SG_KS;
      uint8_t* sA = sg_smem + (st % SG_STAGES) * C::STAGE + SG_WBYTES;
#pragma unroll
      for (int i = t; i < 16 * MT * (SG_KS / 8); i += SG_THREADS) {
        const int r = i / (SG_KS / 8), c = i % (SG_KS / 8);
        const bool ok = (r < M) && (k0 + c * 8 < ke);
        sg_cp16(sA + r * SG_APITCH + c * 16, A + (ok ? ((int64_t)r * lda + k0 + c * 8) : 0), ok);
      }
    }
    sg_commit();
  };
  pdl_launch_dependents();
#pragma unroll
  for (int s = 0; s < SG_STAGES - 1; ++s) stage_w(s);
  pdl_wait();  // A (and the partial-sum scratch this kernel overwrites) belong to the kernels before
#pragma unroll
  for (int s = 0; s < SG_STAGES - 1; ++s) stage_a(s);  // group 0 also carries every weight copy issued above

  float acc[MT][4][4];
#pragma unroll
  for (int mt = 0; mt < MT; ++mt)
#pragma unroll
    for (int nb = 0; nb < 4; ++nb)
#pragma unroll
      for (int j = 0; j < 4; ++j) acc[mt][nb][j] = 0.f;

  const int mi = lane >> 3;
  const uint32_t a_off = (uint32_t)((lane & 15) * SG_APITCH + (lane >> 4) * 16);
  // W through ldmatrix.trans: matrices (k 0-7, n 0-7), (k 8-15, n 0-7), (k 0-7, n 8-15), (k 8-15, n 8-15) of the warp's 32 columns
  const uint32_t w_off = (uint32_t)(((mi & 1) * 8 + (lane & 7)) * SG_WPITCH + (warp * 32 + (mi >> 1) * 8) * 2);

  for (int st = 0; st < nst; ++st) {
    sg_wait<SG_STAGES - 2>();
    __syncthreads();
    stage_w(st + SG_STAGES - 1);
    stage_a(st + SG_STAGES - 1);
    const uint32_t sW = smem_u32(sg_smem + (st % SG_STAGES) * C::STAGE);
    const uint32_t sA = sW + SG_WBYTES;
#pragma unroll
    for (int ks = 0; ks < SG_KS / 16; ++ks) {
      uint32_t a[MT][4];
#pragma unroll
      for (int mt = 0; mt < MT; ++mt) sg_ldsm(sA + mt * 16 * SG_APITCH + a_off + ks * 32, a[mt]);
#pragma unroll
      for (int nb2 = 0; nb2 < 2; ++nb2) {
        uint32_t b[4];
        sg_ldsm_t(sW + ks * 16 * SG_WPITCH + w_off + nb2 * 32, b);
#pragma unroll
        for (int mt = 0; mt < MT; ++mt) {
          sg_mma(acc[mt][2 * nb2], a[mt], b[0], b[1]);
          sg_mma(acc[mt][2 * nb2 + 1], a[mt], b[2], b[3]);
        }
      }
    }
  }
  sg_wait<0>();
  // accumulator (mt, nb, j): row mt*16 + lane/4 (+8 for j >= 2), column warp*32 + nb*8 + 2*(lane%4) + (j & 1)
  float* dst = part + (int64_t)blockIdx.y * M * N;
#pragma unroll
  for (int mt = 0; mt < MT; ++mt)
#pragma unroll
    for (int nb = 0; nb < 4; ++nb) {
      const int col = col0 + warp * 32 + nb * 8 + 2 * (lane & 3);
      if (col >= N) continue;
      const int r_lo = mt * 16 + (lane >> 2);
      if (r_lo < M) *reinterpret_cast<float2*>(dst + (int64_t)r_lo * N + col) = make_float2(acc[mt][nb][0], acc[mt][nb][1]);
      if (r_lo + 8 < M) *reinterpret_cast<float2*>(dst + (int64_t)(r_lo + 8) * N + col) = make_float2(acc[mt][nb][2], acc[mt][nb][3]);
    }
}

void sg_plan(int N, int K, int& k_per, int& nsplit) {
  const int panels = (N + SG_PN - 1) / SG_PN;
  int want = sm_count() * 2 / panels;  // two CTAs per SM, the whole grid resident at once
  if (want < 1) want = 1;
  if (want > 32) want = 32;
  k_per = ((K + want - 1) / want + SG_KS - 1) / SG_KS * SG_KS;
  if (k_per < 2 * SG_KS) k_per = 2 * SG_KS;
  nsplit = (K + k_per - 1) / k_per;
}

}  // namespace

bool small_m_mma_ok(int dtype, int M, int N, int K, const void* A, int64_t lda, const void* W, int64_t ldw) {
  static const bool off = [] { const char* e = getenv("SPB_DECODE_GEMM_MMA"); return e && e[0] == '0'; }();
  return !off && dtype == SPB_BF16 && M >= 1 && M <= 64 && N % 8 == 0 && K % 8 == 0 && lda % 8 == 0 && ldw % 8 == 0 && (uintptr_t)A % 16 == 0 &&
         (uintptr_t)W % 16 == 0;
}

size_t small_m_mma_scratch_bytes(int M, int N, int K) {
  int k_per, nsplit;
  sg_plan(N, K, k_per, nsplit);
  return (size_t)nsplit * M * N * sizeof(float);
}

// the partial sums [nsplit][M][N] land in `part`; returns the number of slices through *nsplit_out
int small_m_mma(int M, int N, int K, const void* A, int64_t lda, const void* W, int64_t ldw, float* part, int* nsplit_out, cudaStream_t st) {
  int k_per, nsplit;
  sg_plan(N, K, k_per, nsplit);
  dim3 grid((N + SG_PN - 1) / SG_PN, nsplit);
  const int MT = (M + 15) / 16;
#define SPB_SG_LAUNCH(MTV)                                                                                                            \
  do {                                                                                                                                \
    SPB_CUDA(cudaFuncSetAttribute(small_m_mma_kernel<MTV>, cudaFuncAttributeMaxDynamicSharedMemorySize, SgCfg<MTV>::SMEM));           \
    SPB_CUDA(launch_pdl(small_m_mma_kernel<MTV>, grid, dim3(SG_THREADS), SgCfg<MTV>::SMEM, st, decode_pdl_enabled(), (const __nv_bfloat16*)A, lda, \
                        (const __nv_bfloat16*)W, ldw, part, M, N, K, k_per));                                                         \
  } while (0)
  if (MT == 1) SPB_SG_LAUNCH(1);
  else if (MT == 2) SPB_SG_LAUNCH(2);
  else if (MT == 3) SPB_SG_LAUNCH(3);
  else SPB_SG_LAUNCH(4);
#undef SPB_SG_LAUNCH
  SPB_LAUNCH_CHECK();
  *nsplit_out = nsplit;
  return 0;
}

}  // namespace spb
