// attn_decode_mma.cu - SURVEY 8 row f3, bf16: the attention of the KV-cache branch (llama3/LLaMA-jax.ipynb:815-829) for a
// few query rows per kv head, with the two small products on the warp-level tensor-core path (mma.sync m16n8k16).
//
// Why not the SIMT kernel of decode.cu: that one converts every cached element to fp32 and spends ~2 FMAs per element and
// query row on the CUDA cores; at a Llama-3-8B-shaped head (4 query heads per kv head, hd 128) it runs at a sixth of the
// HBM rate.  Here the group's query rows (n_rep * Tq <= 16, padded to the 16 rows of one MMA) multiply 16 cached keys
// per warp and step; K and V stay bf16 from HBM to the tensor core (cp.async -> shared memory -> ldmatrix), so the SM's
// work per cached byte is a few instructions and the kernel is bound by the cache read.  tcgen05 has no use here: its
// smallest tile is 64 rows and an accumulator in TMEM buys nothing for a 16-row product that is over in two steps.
//
// grid (key splits, kv heads x row groups, batch); CTA = 4 warps.  A CTA streams its slice of the cache in tiles of 64 keys
// through a three-stage cp.async ring (rows padded by 16 bytes: ldmatrix without bank conflicts); warp w owns keys
// 16w .. 16w+15 of every tile and keeps its own running (max, sum, O[16 x hd]) - flash-decoding inside the CTA - merged
// through shared memory at the end.  The CTA writes one partial (unnormalised O, max, sum) per query row; the splits are
// merged by decode.cu's attn_decode_combine_kernel in a fixed order.
#include "common.cuh"
#include "kernels.h"

namespace spb {

namespace {

constexpr int DM_THREADS = 128;
constexpr int DM_ROWS = 16;  // query rows per CTA = the M of mma.m16n8k16

template <int HD>
struct DecMma {
  static constexpr int KT = 64;                 // keys per stage (16 per warp)
  static constexpr int ROWB = HD * 2 + 16;      // padded row pitch: pitch mod 128 = 16 -> 8 consecutive rows hit 8 different 16-byte bank groups
  static constexpr int STAGES = 3;
  static constexpr int TILE_BYTES = KT * ROWB;  // K (or V) tile
  static constexpr int STAGE_BYTES = 2 * TILE_BYTES;
  static constexpr int Q_BYTES = DM_ROWS * ROWB;
  static constexpr int MERGE_BYTES = 4 * DM_ROWS * (HD + 2) * 4;  // per-warp (O, m, l) at the end, inside the ring
  static constexpr int SMEM = Q_BYTES + (STAGES * STAGE_BYTES > MERGE_BYTES ? STAGES * STAGE_BYTES : MERGE_BYTES);
};

struct DecMmaArgs {
  const __nv_bfloat16* q;
  int64_t ldq;
  const __nv_bfloat16* k;
  const __nv_bfloat16* v;
  int64_t ldk, ldv, kbs, vbs;
  float* po;   // [B, Hq, Tq, splits, hd]
  float* pml;  // [B, Hq, Tq, splits, 2]
  int B, Tq, Hq, Hkv, Tk, splits, tiles_per_split, row_groups;
  float scale_log2;  // score scale * log2(e)
  const int* tk_dev; // != nullptr: the cache length lives in device memory (graph replay): Tk = min(*tk_dev + tk_add, Tk)
  int tk_add;
};

__device__ __forceinline__ void cp_async16_zfill(void* smem_dst, const void* gsrc, bool valid) {
  const int n = valid ? 16 : 0;  // src-size 0: the 16 bytes are written as zeros, nothing is read
  asm volatile("cp.async.cg.shared.global [%0], [%1], 16, %2;" ::"r"(smem_u32(smem_dst)), "l"(gsrc), "r"(n) : "memory");
}
__device__ __forceinline__ void cpa_commit() { asm volatile("cp.async.commit_group;" ::: "memory"); }
template <int N>
__device__ __forceinline__ void cpa_wait() {
  asm volatile("cp.async.wait_group %0;" ::"n"(N) : "memory");
}
__device__ __forceinline__ void ldsm_x4(uint32_t addr, uint32_t& r0, uint32_t& r1, uint32_t& r2, uint32_t& r3) {
  asm volatile("ldmatrix.sync.aligned.m8n8.x4.shared.b16 {%0,%1,%2,%3}, [%4];" : "=r"(r0), "=r"(r1), "=r"(r2), "=r"(r3) : "r"(addr));
}
__device__ __forceinline__ void ldsm_x4_t(uint32_t addr, uint32_t& r0, uint32_t& r1, uint32_t& r2, uint32_t& r3) {
  asm volatile("ldmatrix.sync.aligned.m8n8.x4.trans.shared.b16 {%0,%1,%2,%3}, [%4];" : "=r"(r0), "=r"(r1), "=r"(r2), "=r"(r3) : "r"(addr));
}
__device__ __forceinline__ void mma_bf16(float (&d)[4], const uint32_t (&a)[4], uint32_t b0, uint32_t b1) {
  asm volatile("mma.sync.aligned.m16n8k16.row.col.f32.bf16.bf16.f32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, {%0,%1,%2,%3};"
               : "+f"(d[0]), "+f"(d[1]), "+f"(d[2]), "+f"(d[3])
               : "r"(a[0]), "r"(a[1]), "r"(a[2]), "r"(a[3]), "r"(b0), "r"(b1));
}
__device__ __forceinline__ uint32_t pack2_bf16(float lo, float hi) {
  const __nv_bfloat162 h = __floats2bfloat162_rn(lo, hi);
  return *reinterpret_cast<const uint32_t*>(&h);
}

template <int HD>
__global__ void __launch_bounds__(DM_THREADS) attn_decode_mma_kernel(const DecMmaArgs p) {
  using C = DecMma<HD>;
  constexpr int PIECES = HD / 8;  // 16-byte pieces per row
  extern __shared__ __align__(128) uint8_t dsm[];
  uint8_t* sQ = dsm;
  uint8_t* ring = dsm + C::Q_BYTES;

  const int split = blockIdx.x;
  const int g = blockIdx.y / p.row_groups, rg = blockIdx.y % p.row_groups;
  const int b = blockIdx.z;
  const int n_rep = p.Hq / p.Hkv;
  const int rtot = n_rep * p.Tq;
  const int r0 = rg * DM_ROWS;
  const int nr = min(DM_ROWS, rtot - r0);
  const int t = threadIdx.x, lane = t & 31, warp = t >> 5;
  const __nv_bfloat16* kbase = p.k + (int64_t)b * p.kbs + (int64_t)g * HD;
  const __nv_bfloat16* vbase = p.v + (int64_t)b * p.vbs + (int64_t)g * HD;
  pdl_launch_dependents();
  pdl_wait();  // q, the newest cache rows (and the device-side length) come from the kernels before
  int Tk = p.Tk, tps = p.tiles_per_split;
  if (p.tk_dev != nullptr) {  // the grid was sized for the cache capacity; share the tiles that exist evenly over its splits
    Tk = min(*p.tk_dev + p.tk_add, p.Tk);
    tps = ((Tk + C::KT - 1) / C::KT + p.splits - 1) / p.splits;
  }
  const int ntiles = (Tk + C::KT - 1) / C::KT;
  const int tile0 = split * tps;
  const int tile1 = min(tile0 + tps, ntiles);

  // one commit group per call, even when the tile does not exist: the wait counts below stay uniform
  auto stage_tile = [&](int tile) {
    if (tile < tile1) {
      const int k0 = tile * C::KT;
      uint8_t* sK = ring + ((tile - tile0) % C::STAGES) * C::STAGE_BYTES;
      uint8_t* sV = sK + C::TILE_BYTES;
#pragma unroll 4
      for (int i = t; i < C::KT * PIECES; i += DM_THREADS) {
        const int jj = i / PIECES, c = i % PIECES;
        const bool ok = k0 + jj < Tk;  // rows past the end of the cache are zero-filled: 0 * V must stay 0
        const int64_t row = ok ? (k0 + jj) : 0;
        cp_async16_zfill(sK + jj * C::ROWB + c * 16, kbase + row * p.ldk + c * 8, ok);
        cp_async16_zfill(sV + jj * C::ROWB + c * 16, vbase + row * p.ldv + c * 8, ok);
      }
    }
    cpa_commit();
  };
  stage_tile(tile0);
  stage_tile(tile0 + 1);

  // the group's query rows (bf16 as they are; the score scale is applied to the fp32 products), zero rows beyond nr
  for (int i = t; i < DM_ROWS * PIECES; i += DM_THREADS) {
    const int r = i / PIECES, c = i % PIECES;
    uint4 val = make_uint4(0u, 0u, 0u, 0u);
    if (r < nr) {
      const int hq = g * n_rep + (r0 + r) / p.Tq, tq = (r0 + r) % p.Tq;
      val = *reinterpret_cast<const uint4*>(p.q + ((int64_t)b * p.Tq + tq) * p.ldq + (int64_t)hq * HD + c * 8);
    }
    *reinterpret_cast<uint4*>(sQ + r * C::ROWB + c * 16) = val;
  }

  float oacc[HD / 8][4];
#pragma unroll
  for (int i = 0; i < HD / 8; ++i)
#pragma unroll
    for (int j = 0; j < 4; ++j) oacc[i][j] = 0.f;
  float m_run[2] = {-INFINITY, -INFINITY}, l_run[2] = {0.f, 0.f};  // rows lane/4 and lane/4 + 8 (l: this thread's share)

  const uint32_t q_addr = smem_u32(sQ) + (lane & 15) * C::ROWB + (lane >> 4) * 16;
  const int mi = lane >> 3;
  // K (non-transposed): matrices (keys 0-7, d 0-7), (keys 0-7, d 8-15), (keys 8-15, d 0-7), (keys 8-15, d 8-15)
  const uint32_t k_off = (uint32_t)((warp * 16 + (mi >> 1) * 8 + (lane & 7)) * C::ROWB + (mi & 1) * 16);
  // V (transposed on load): matrices (keys 0-7, d 0-7), (keys 8-15, d 0-7), (keys 0-7, d 8-15), (keys 8-15, d 8-15)
  const uint32_t v_off = (uint32_t)((warp * 16 + (mi & 1) * 8 + (lane & 7)) * C::ROWB + (mi >> 1) * 16);

  for (int tile = tile0; tile < tile1; ++tile) {
    cpa_wait<C::STAGES - 2>();  // this thread's copies of `tile` are done (one younger group may be in flight)
    __syncthreads();            // everybody's; and every warp is past tile - 1, whose stage the next copies overwrite
    stage_tile(tile + C::STAGES - 1);
    const uint32_t sK = smem_u32(ring + ((tile - tile0) % C::STAGES) * C::STAGE_BYTES);
    const uint32_t sV = sK + C::TILE_BYTES;
    const int key0 = tile * C::KT + warp * 16;
    if (key0 < Tk) {  // warp-uniform: a warp whose 16 keys are all past the end has nothing to add
      // ---- S[16 rows x 16 keys] = Q . K^T
      float sacc[2][4];
#pragma unroll
      for (int nb = 0; nb < 2; ++nb)
#pragma unroll
        for (int j = 0; j < 4; ++j) sacc[nb][j] = 0.f;
#pragma unroll
      for (int ks = 0; ks < HD / 16; ++ks) {
        uint32_t a[4], b0, b1, b2, b3;
        ldsm_x4(q_addr + ks * 32, a[0], a[1], a[2], a[3]);
        ldsm_x4(sK + k_off + ks * 32, b0, b1, b2, b3);
        mma_bf16(sacc[0], a, b0, b1);
        mma_bf16(sacc[1], a, b2, b3);
      }
      // ---- online softmax (base 2), rows lane/4 (+8), keys nb * 8 + 2 * (lane % 4) + {0, 1}
      float tmax[2] = {-INFINITY, -INFINITY};
#pragma unroll
      for (int nb = 0; nb < 2; ++nb)
#pragma unroll
        for (int j = 0; j < 4; ++j) {
          const int key = key0 + nb * 8 + 2 * (lane & 3) + (j & 1);
          const float s = key < Tk ? sacc[nb][j] * p.scale_log2 : -INFINITY;
          sacc[nb][j] = s;
          tmax[j >> 1] = fmaxf(tmax[j >> 1], s);
        }
      float alpha[2];
#pragma unroll
      for (int h = 0; h < 2; ++h) {
        tmax[h] = fmaxf(tmax[h], __shfl_xor_sync(0xffffffffu, tmax[h], 1));
        tmax[h] = fmaxf(tmax[h], __shfl_xor_sync(0xffffffffu, tmax[h], 2));
        const float mn = fmaxf(m_run[h], tmax[h]);  // finite: key0 < Tk, so at least one key of the 16 is real
        alpha[h] = exp2f(m_run[h] - mn);            // first tile: exp2(-inf) = 0
        m_run[h] = mn;
      }
      uint32_t pa[4];
      float psum[2] = {0.f, 0.f};
#pragma unroll
      for (int nb = 0; nb < 2; ++nb) {
        const float p0 = exp2f(sacc[nb][0] - m_run[0]), p1 = exp2f(sacc[nb][1] - m_run[0]);
        const float p2 = exp2f(sacc[nb][2] - m_run[1]), p3 = exp2f(sacc[nb][3] - m_run[1]);
        psum[0] += p0 + p1;
        psum[1] += p2 + p3;
        pa[nb * 2] = pack2_bf16(p0, p1);      // A fragment of P for the k16 step over these 16 keys
        pa[nb * 2 + 1] = pack2_bf16(p2, p3);
      }
      l_run[0] = l_run[0] * alpha[0] + psum[0];
      l_run[1] = l_run[1] * alpha[1] + psum[1];
#pragma unroll
      for (int i = 0; i < HD / 8; ++i) {
        oacc[i][0] *= alpha[0];
        oacc[i][1] *= alpha[0];
        oacc[i][2] *= alpha[1];
        oacc[i][3] *= alpha[1];
      }
      // ---- O[16 x hd] += P[16 x 16 keys] . V[16 keys x hd]
#pragma unroll
      for (int nb2 = 0; nb2 < HD / 16; ++nb2) {
        uint32_t b0, b1, b2, b3;
        ldsm_x4_t(sV + v_off + nb2 * 32, b0, b1, b2, b3);
        mma_bf16(oacc[2 * nb2], pa, b0, b1);
        mma_bf16(oacc[2 * nb2 + 1], pa, b2, b3);
      }
    }
  }
  cpa_wait<0>();
  __syncthreads();  // the ring is free: it becomes the merge buffer

  // ---- merge the four warps' (m, l, O), write this split's partial
  float* mo = reinterpret_cast<float*>(ring);                 // [4 warps][16 rows][HD]
  float* mml = mo + 4 * DM_ROWS * HD;                          // [4 warps][16 rows][2]
#pragma unroll
  for (int h = 0; h < 2; ++h) {
    l_run[h] += __shfl_xor_sync(0xffffffffu, l_run[h], 1);
    l_run[h] += __shfl_xor_sync(0xffffffffu, l_run[h], 2);
  }
  {
    const int r_lo = lane >> 2, c0 = 2 * (lane & 3);
#pragma unroll
    for (int i = 0; i < HD / 8; ++i) {
      float* d0 = mo + ((warp * DM_ROWS + r_lo) * HD) + i * 8 + c0;
      float* d1 = mo + ((warp * DM_ROWS + r_lo + 8) * HD) + i * 8 + c0;
      *reinterpret_cast<float2*>(d0) = make_float2(oacc[i][0], oacc[i][1]);
      *reinterpret_cast<float2*>(d1) = make_float2(oacc[i][2], oacc[i][3]);
    }
    if ((lane & 3) == 0) {
      mml[(warp * DM_ROWS + r_lo) * 2] = m_run[0];
      mml[(warp * DM_ROWS + r_lo) * 2 + 1] = l_run[0];
      mml[(warp * DM_ROWS + r_lo + 8) * 2] = m_run[1];
      mml[(warp * DM_ROWS + r_lo + 8) * 2 + 1] = l_run[1];
    }
  }
  __syncthreads();
  constexpr float LN2 = 0.6931471805599453f;
  for (int i = t; i < nr * HD; i += DM_THREADS) {
    const int r = i / HD, d = i % HD;
    float M = -INFINITY;
#pragma unroll
    for (int w = 0; w < 4; ++w) M = fmaxf(M, mml[(w * DM_ROWS + r) * 2]);
    float o = 0.f;
#pragma unroll
    for (int w = 0; w < 4; ++w) {
      const float mw = mml[(w * DM_ROWS + r) * 2];
      if (mw != -INFINITY) o += exp2f(mw - M) * mo[(w * DM_ROWS + r) * HD + d];
    }
    const int hq = g * n_rep + (r0 + r) / p.Tq, tq = (r0 + r) % p.Tq;
    const int64_t row = (((int64_t)b * p.Hq + hq) * p.Tq + tq) * p.splits + split;
    p.po[row * HD + d] = o;
  }
  if (t < nr) {
    const int r = t;
    float M = -INFINITY, Lsum = 0.f;
#pragma unroll
    for (int w = 0; w < 4; ++w) M = fmaxf(M, mml[(w * DM_ROWS + r) * 2]);
#pragma unroll
    for (int w = 0; w < 4; ++w) {
      const float mw = mml[(w * DM_ROWS + r) * 2];
      if (mw != -INFINITY) Lsum += exp2f(mw - M) * mml[(w * DM_ROWS + r) * 2 + 1];
    }
    const int hq = g * n_rep + (r0 + r) / p.Tq, tq = (r0 + r) % p.Tq;
    const int64_t row = (((int64_t)b * p.Hq + hq) * p.Tq + tq) * p.splits + split;
    p.pml[row * 2] = M * LN2;  // the combine kernel works in natural-log units; -inf (an empty split) stays -inf
    p.pml[row * 2 + 1] = Lsum;
  }
}

void mma_plan(int hd, int B, int Tq, int Hq, int Hkv, int Tk, int& row_groups, int& splits, int& tiles_per_split) {
  const int rtot = (Hq / Hkv) * Tq;
  row_groups = (rtot + DM_ROWS - 1) / DM_ROWS;
  const int tiles = (Tk + 63) / 64;
  const int smem = hd == 64 ? DecMma<64>::SMEM : (hd == 128 ? DecMma<128>::SMEM : DecMma<256>::SMEM);
  int per_sm = (227 * 1024) / (smem + 1024);
  if (per_sm < 1) per_sm = 1;
  if (per_sm > 4) per_sm = 4;
  const int64_t ctas = (int64_t)B * Hkv * row_groups;
  int64_t want = (int64_t)sm_count() * per_sm / ctas;  // key splits that still fit one wave of resident CTAs
  if (want < 1) want = 1;
  if (want > tiles) want = tiles;
  if (want > 64) want = 64;
  tiles_per_split = (int)((tiles + want - 1) / want);
  splits = (tiles + tiles_per_split - 1) / tiles_per_split;
}

template <int HD>
int launch_mma(const DecMmaArgs& a, cudaStream_t st) {
  SPB_CUDA(cudaFuncSetAttribute(attn_decode_mma_kernel<HD>, cudaFuncAttributeMaxDynamicSharedMemorySize, DecMma<HD>::SMEM));
  dim3 grid(a.splits, a.Hkv * a.row_groups, a.B);
  SPB_CUDA(launch_pdl(attn_decode_mma_kernel<HD>, grid, dim3(DM_THREADS), DecMma<HD>::SMEM, st, decode_pdl_enabled(), a));
  SPB_LAUNCH_CHECK();
  return 0;
}

}  // namespace

bool attn_decode_mma_supported(int dtype, int hd) {
  static const bool off = [] { const char* e = getenv("SPB_DECODE_MMA"); return e && e[0] == '0'; }();
  return !off && dtype == SPB_BF16 && (hd == 64 || hd == 128 || hd == 256);
}

size_t attn_decode_mma_scratch_bytes(int B, int Tq, int Hq, int Hkv, int hd, int Tk) {
  int rgs, splits, tps;
  mma_plan(hd, B, Tq, Hq, Hkv, Tk, rgs, splits, tps);
  return (size_t)B * Hq * Tq * splits * ((size_t)hd + 2) * sizeof(float) + 256;
}

int attn_decode_mma(int B, int Tq, int Hq, int Hkv, int hd, int Tk, float scale, const void* q, int64_t ldq, const void* k, int64_t ldk, int64_t kbs,
                    const void* v, int64_t ldv, int64_t vbs, void* o, int64_t ldo, float* lse, void* scratch, cudaStream_t st, const int* tk_dev,
                    int tk_add) {
  SPB_CHECK_ARG(ldq % 8 == 0 && (uintptr_t)q % 16 == 0, SPB_E_UNSUPPORTED, "attention decode: q needs 16-byte aligned rows");
  DecMmaArgs a;
  mma_plan(hd, B, Tq, Hq, Hkv, Tk, a.row_groups, a.splits, a.tiles_per_split);
  SPB_CHECK_ARG((int64_t)Hkv * a.row_groups <= 65535, SPB_E_LIMIT, "attention decode: too many query rows per kv head");
  a.q = (const __nv_bfloat16*)q; a.ldq = ldq;
  a.k = (const __nv_bfloat16*)k; a.v = (const __nv_bfloat16*)v;
  a.ldk = ldk; a.ldv = ldv; a.kbs = kbs; a.vbs = vbs;
  a.B = B; a.Tq = Tq; a.Hq = Hq; a.Hkv = Hkv; a.Tk = Tk;
  a.scale_log2 = scale * 1.4426950408889634f;
  a.tk_dev = tk_dev;
  a.tk_add = tk_add;
  a.po = (float*)scratch;
  a.pml = a.po + (size_t)B * Hq * Tq * a.splits * hd;
  int rc;
  if (hd == 64) rc = launch_mma<64>(a, st);
  else if (hd == 128) rc = launch_mma<128>(a, st);
  else rc = launch_mma<256>(a, st);
  if (rc) return rc;
  return attn_decode_combine(SPB_BF16, a.po, a.pml, o, ldo, lse, B, Tq, Hq, hd, a.splits, st);
}

}  // namespace spb
