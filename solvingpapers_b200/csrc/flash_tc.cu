// flash_tc.cu - K3: flash-style causal GQA attention forward on tcgen05 (bf16, hd in {64,128}); replaces the materialised
// scores of llama3:820-828 (and gpt:344-353, ViT's nn.MultiheadAttention).  Scores never leave the SM.
//
// One CTA = 256 query rows (two 128-row tiles) of one (batch, q head); the two tiles ping-pong on the tensor pipe:
//   warp 0    TMA producer: the two Q tiles once, then K / V tiles of 128 rows (K double buffered, V single)
//   warp 1    MMA issuer (one thread): per KV tile  S0 = Q0 K^T, S1 = Q1 K^T  (M128 N128 K=hd, each into its own TMEM S
//             buffer), then  O0 += P0 V, O1 += P1 V  (M128 N=hd K128, TMEM accumulators): while one tile is in its
//             softmax the tensor pipe works on the other
//   warps 2-5 softmax of tile 0, warps 6-9 softmax of tile 1: thread = one query row (TMEM lane).  Two passes over the S
//             row in TMEM (tcgen05.ld): pass 1 row max, pass 2 exp2 / sum / bf16 P into shared memory in the 128B-swizzled
//             K-major UMMA layout; 1/sqrt(hd)*log2(e) folded into one FFMA; lazy rescaling of O (only when the running
//             max grows by more than 2^8, decided per warp)
//   causal tile skipping: KV tiles above a tile's diagonal are neither loaded nor multiplied; diagonal and ragged last
//   tiles are masked by index (-1e9 additive / -1e4 replacing / -inf masks of the notebooks all give exp()==0 exactly).
//   GQA: kv head = q head / (Hq/Hkv) is just the TMA column coordinate (llama3:626-627 repeat_kv never materialised).
// TMEM: S0 [0,128) S1 [128,256) O0 [256,256+hd) O1 [384,384+hd).  Algorithmic FLOPs: 4*B*Hq*T^2*hd*m (m=(T+1)/2T causal).
#include <cstdlib>

#include "common.cuh"
#include "kernels.h"

namespace spb {

constexpr float LOG2E = 1.4426950408889634f;
constexpr float LN2 = 0.6931471805599453f;

__device__ __forceinline__ float fast_exp2(float x) {
  float y;
  asm volatile("ex2.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x));
  return y;
}

// PT = true: P (bf16) is written back into tensor memory over the S columns it came from and feeds the PV MMA as a TMEM
// A operand (no shared-memory round trip, which frees 64 KB for a second V stage).  PT = false keeps P in shared memory.
template <int HD, bool PT>
struct FaFwdCfg {
  static constexpr int NT = (HD <= 128) ? 2 : 1;    // query tiles per CTA (hd=256: O alone needs 256 TMEM columns -> one tile)
  static constexpr int KST = (HD <= 128) ? 2 : 1;   // K stages (hd=256: Q, K, V tiles are 64 KB each)
  static constexpr int VST = (HD <= 128 && PT) ? 2 : 1;
  static constexpr int CH = HD / 64;                // 64-column chunks per Q/K/V tile
  static constexpr int TILE_BYTES = 128 * HD * 2;   // 128 rows x HD bf16
  static constexpr int CHUNK_BYTES = 128 * 128;     // 128 rows x 128 B
  static constexpr int P_BYTES = 128 * 128 * 2;
  static constexpr int OFF_Q = 0;
  static constexpr int OFF_K = OFF_Q + NT * TILE_BYTES;
  static constexpr int OFF_V = OFF_K + KST * TILE_BYTES;
  static constexpr int OFF_P = OFF_V + VST * TILE_BYTES;
  static constexpr int OFF_BAR = OFF_P + (PT ? 0 : NT * P_BYTES);
  static constexpr int SMEM_BYTES = OFF_BAR + 256 + 1024;
  static constexpr int TMEM_COLS = 512;
  static constexpr int TM_O0 = (NT == 2) ? 256 : 128;     // O_t at TM_O0 + t*128 (NT=2) / one 256-column O at 128 (NT=1)
  static constexpr int THREADS = 64 + NT * 128;
};

struct FaFwdArgs {
  __nv_bfloat16* o;
  int64_t ldo;
  float* lse;
  int T, Hq, Hkv, causal;
  float scale_log2;  // scale * log2(e)
};

template <int HD, bool PT>
__global__ void __launch_bounds__(FaFwdCfg<HD, PT>::THREADS, 1)
    flash_fwd_tc_kernel(const __grid_constant__ CUtensorMap tmQ, const __grid_constant__ CUtensorMap tmK,
                        const __grid_constant__ CUtensorMap tmV, const FaFwdArgs p) {
  using Cfg = FaFwdCfg<HD, PT>;
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = smem_raw + ((1024u - (smem_u32(smem_raw) & 1023u)) & 1023u);
  uint8_t* sQ = smem + Cfg::OFF_Q;
  uint8_t* sK = smem + Cfg::OFF_K;
  uint8_t* sV = smem + Cfg::OFF_V;
  uint8_t* sP = smem + Cfg::OFF_P;
  uint64_t* bars = reinterpret_cast<uint64_t*>(smem + Cfg::OFF_BAR);
  uint64_t* q_full = bars;            // [1]
  uint64_t* k_full = bars + 1;        // [2] stages
  uint64_t* k_empty = bars + 3;       // [2]
  uint64_t* v_full = bars + 5;        // [2] stages
  uint64_t* v_empty = bars + 7;       // [2]
  uint64_t* s_full = bars + 9;        // [2] per q tile
  uint64_t* s_empty = bars + 11;      // [2]
  uint64_t* p_full = bars + 13;       // [2]
  uint64_t* pv_done = bars + 15;      // [2]
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(bars + 17);

  const int warp = __shfl_sync(0xffffffffu, (int)(threadIdx.x >> 5), 0);
  const int lane = threadIdx.x & 31;

  constexpr int NT = Cfg::NT, KST = Cfg::KST, VST = Cfg::VST;
  const int n_qp = (int)gridDim.x;
  const int qp = p.causal ? (n_qp - 1 - (int)blockIdx.x) : (int)blockIdx.x;  // heavy (late) causal tiles first
  const int h = blockIdx.y, b = blockIdx.z;
  const int kvh = h / (p.Hq / p.Hkv);
  const int q0 = qp * (NT * 128);
  const int n_kv_all = (p.T + 127) / 128;
  // per q tile: number of KV tiles it attends to (0 = the tile lies entirely past the end of the sequence)
  int n_kv_t[2] = {0, 0};
#pragma unroll
  for (int t = 0; t < NT; ++t) {
    const bool live = q0 + t * 128 < p.T;
    n_kv_t[t] = !live ? 0 : (p.causal ? min(n_kv_all, NT * qp + t + 1) : n_kv_all);
  }
  const int n_kv = max(n_kv_t[0], n_kv_t[1]);

  if (warp == 0 && lane == 0) {
    prefetch_tmap(&tmQ);
    prefetch_tmap(&tmK);
    prefetch_tmap(&tmV);
  }
  if (warp == 1) {
    if (lane == 0) {
      mbar_init(q_full, 1);
      for (int i = 0; i < 2; ++i) {
        mbar_init(&v_full[i], 1);
        mbar_init(&v_empty[i], 1);
        mbar_init(&k_full[i], 1);
        mbar_init(&k_empty[i], 1);
        mbar_init(&s_full[i], 1);
        mbar_init(&s_empty[i], 4);
        mbar_init(&p_full[i], 4);
        mbar_init(&pv_done[i], 1);
      }
      fence_barrier_init();
    }
    __syncwarp();
    tmem_alloc(tmem_slot, Cfg::TMEM_COLS);
    tmem_relinquish();
  }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;

  if (warp == 0) {
    // ------------------------------------------------------------------ TMA producer
    if (lane == 0) {
      const int n_live = (n_kv_t[0] > 0) + (n_kv_t[1] > 0);
      mbar_arrive_expect_tx(q_full, n_live * Cfg::TILE_BYTES);
#pragma unroll
      for (int t = 0; t < NT; ++t) {
        if (n_kv_t[t] > 0) {
#pragma unroll
          for (int c = 0; c < Cfg::CH; ++c)
            tma_load_2d(sQ + t * Cfg::TILE_BYTES + c * Cfg::CHUNK_BYTES, &tmQ, q_full, h * HD + c * 64, b * p.T + q0 + t * 128);
        }
      }
      for (int j = 0; j < n_kv; ++j) {
        const int bf = j % KST;
        const int row_kv = b * p.T + j * 128;
        mbar_wait(&k_empty[bf], ((j / KST) & 1) ^ 1);
        mbar_arrive_expect_tx(&k_full[bf], Cfg::TILE_BYTES);
#pragma unroll
        for (int c = 0; c < Cfg::CH; ++c)
          tma_load_2d(sK + bf * Cfg::TILE_BYTES + c * Cfg::CHUNK_BYTES, &tmK, &k_full[bf], kvh * HD + c * 64, row_kv);
        const int vf = j % VST;
        mbar_wait(&v_empty[vf], ((j / VST) & 1) ^ 1);
        mbar_arrive_expect_tx(&v_full[vf], Cfg::TILE_BYTES);
#pragma unroll
        for (int c = 0; c < Cfg::CH; ++c)
          tma_load_2d(sV + vf * Cfg::TILE_BYTES + c * Cfg::CHUNK_BYTES, &tmV, &v_full[vf], kvh * HD + c * 64, row_kv);
      }
    }
    __syncwarp();
  } else if (warp == 1) {
    // ------------------------------------------------------------------ MMA issuer
    if (lane == 0) {
      const uint32_t idesc_qk = umma_idesc_bf16(128, 128, 0, 0);  // A = Q (K-major), B = K (K-major)
      const uint32_t idesc_pv = umma_idesc_bf16(128, HD, 0, 1);   // A = P (K-major), B = V (MN-major: hd contiguous)
      mbar_wait(q_full, 0);
      auto qk_t = [&](int j, int t) {  // S_t = Q_t K(j)^T
        if (j >= n_kv_t[t]) return;
        // S_t is free: PT: PV_t(j-1), issued earlier on this in-order pipe, was the last reader of the P that aliases it;
        // smem P: the softmax group signalled that it has finished reading S_t of the previous KV tile
        if (!PT) mbar_wait(&s_empty[t], (j & 1) ^ 1);
        tc_fence_after();
        const uint32_t q_addr = smem_u32(sQ + t * Cfg::TILE_BYTES);
        const uint32_t k_addr = smem_u32(sK + (j % KST) * Cfg::TILE_BYTES);
#pragma unroll
        for (int kk = 0; kk < HD / 16; ++kk) {
          const uint32_t off = (kk >> 2) * Cfg::CHUNK_BYTES + (kk & 3) * 32;
          umma_bf16(tmem_base + t * 128, umma_desc_kmajor(q_addr + off), umma_desc_kmajor(k_addr + off), idesc_qk, kk > 0 ? 1u : 0u);
        }
        umma_commit(&s_full[t]);
      };
      auto pv_t = [&](int j, int t) {  // O_t += P_t V(j)
        if (j >= n_kv_t[t]) return;
        mbar_wait(&p_full[t], j & 1);
        tc_fence_after();
        const uint32_t v_addr = smem_u32(sV + (j % VST) * Cfg::TILE_BYTES);
        if (PT) {
#pragma unroll
          for (int kk = 0; kk < 8; ++kk)  // A = P_t from tensor memory: 16 bf16 of K per 8 columns
            umma_bf16_ts(tmem_base + Cfg::TM_O0 + t * 128, tmem_base + t * 128 + kk * 8, umma_desc_mnmajor(v_addr + kk * 2048, 128),
                         idesc_pv, (j > 0 || kk > 0) ? 1u : 0u);
        } else {
          const uint32_t p_addr = smem_u32(sP + t * Cfg::P_BYTES);
#pragma unroll
          for (int kk = 0; kk < 8; ++kk) {
            const uint32_t aoff = (kk >> 2) * Cfg::CHUNK_BYTES + (kk & 3) * 32;
            umma_bf16(tmem_base + Cfg::TM_O0 + t * 128, umma_desc_kmajor(p_addr + aoff), umma_desc_mnmajor(v_addr + kk * 2048, 128),
                      idesc_pv, (j > 0 || kk > 0) ? 1u : 0u);
          }
        }
        umma_commit(&pv_done[t]);
      };
      // issue order per KV tile j:  PV0(j) QK0(j+1) PV1(j) QK1(j+1)  - each tile's MMAs follow its own softmax, the other
      // tile's MMAs fill the tensor pipe meanwhile
      mbar_wait(&k_full[0], 0);
      qk_t(0, 0);
      if (NT == 2) qk_t(0, 1);
      umma_commit(&k_empty[0]);
      for (int j = 0; j < n_kv; ++j) {
        const bool more = j + 1 < n_kv;
        mbar_wait(&v_full[j % VST], (j / VST) & 1);
        pv_t(j, 0);
        if (more) {
          mbar_wait(&k_full[(j + 1) % KST], ((j + 1) / KST) & 1);
          qk_t(j + 1, 0);
        }
        if (NT == 2) pv_t(j, 1);
        if (more) {
          if (NT == 2) qk_t(j + 1, 1);
          umma_commit(&k_empty[(j + 1) % KST]);
        }
        umma_commit(&v_empty[j % VST]);
      }
    }
    __syncwarp();
  } else {
    // ------------------------------------------------------------------ softmax + epilogue: warps 2-5 tile 0, warps 6-9 tile 1
    const int t = (warp - 2) >> 2;
    const int qd = warp & 3;
    const int r = qd * 32 + lane;      // row inside the tile = TMEM lane
    const int q_idx = q0 + t * 128 + r;
    const uint32_t lane_base = tmem_base + ((uint32_t)(qd * 32) << 16);
    const uint32_t tS = lane_base + (uint32_t)(t * 128);
    const uint32_t tO = lane_base + (uint32_t)(Cfg::TM_O0 + t * 128);
    const int n_mine = n_kv_t[t];
    float m = -INFINITY, l = 0.f;
    for (int j = 0; j < n_mine; ++j) {
      mbar_wait(&s_full[t], j & 1);
      tc_fence_after();
      const int kv0 = j * 128;
      const bool need_mask = (p.causal && kv0 + 127 > q0 + t * 128) || (kv0 + 128 > p.T);
      // number of leading KV columns of this tile that row q_idx may attend to (everything else is masked)
      int n_valid = p.T - kv0;
      if (p.causal) n_valid = min(n_valid, q_idx - kv0 + 1);
      // ---- pass 1: row max of the raw scores (scale > 0, so scaling commutes with max); two 32-column TMEM loads in flight
      float mx0 = -INFINITY, mx1 = -INFINITY;
#pragma unroll 1
      for (int c2 = 0; c2 < 2; ++c2) {
        uint32_t ra[32], rb[32];
        SPB_TMEM_LD_32(tS + (uint32_t)(c2 * 64), ra);
        SPB_TMEM_LD_32(tS + (uint32_t)(c2 * 64 + 32), rb);
        tmem_ld_wait();
        if (need_mask) {
          const int lim = n_valid - c2 * 64;
#pragma unroll
          for (int i = 0; i < 32; ++i) {
            mx0 = fmaxf(mx0, i < lim ? __uint_as_float(ra[i]) : -INFINITY);
            mx1 = fmaxf(mx1, i + 32 < lim ? __uint_as_float(rb[i]) : -INFINITY);
          }
        } else {
#pragma unroll
          for (int i = 0; i < 32; ++i) {
            mx0 = fmaxf(mx0, __uint_as_float(ra[i]));
            mx1 = fmaxf(mx1, __uint_as_float(rb[i]));
          }
        }
      }
      const float tmax = fmaxf(mx0, mx1) * p.scale_log2;
      if (j == 0) {
        m = tmax;
      } else {
        // lazy rescale: keep the stale max while the true max is at most 2^8 above it (p <= 256, harmless in fp32/bf16)
        const bool grow = tmax > m + 8.0f;
        if (__any_sync(0xffffffffu, grow)) {
          float alpha = 1.0f;
          if (grow) {
            alpha = fast_exp2(m - tmax);
            m = tmax;
          }
          l *= alpha;
          mbar_wait(&pv_done[t], (j - 1) & 1);  // O_t must hold every PV issued so far
          tc_fence_after();
#pragma unroll 1
          for (int c = 0; c < HD / 32; ++c) {
            uint32_t o[32];
            SPB_TMEM_LD_32(tO + (uint32_t)(c * 32), o);
            tmem_ld_wait();
#pragma unroll
            for (int i = 0; i < 32; ++i) o[i] = __float_as_uint(__uint_as_float(o[i]) * alpha);
            SPB_TMEM_ST_32(tO + (uint32_t)(c * 32), o);
          }
          tmem_st_wait();
        }
      }
      // ---- pass 2: p = exp2(s*c - m) (packed fp32x2 FFMA), row sum, bf16 P into shared memory (UMMA K-major, 128B swizzle)
      if (!PT && j >= 1) mbar_wait(&pv_done[t], (j - 1) & 1);  // PV_t(j-1) has finished reading the smem P buffer
      uint8_t* prow = sP + t * Cfg::P_BYTES + r * 128;
      const float2 sc2 = make_float2(p.scale_log2, p.scale_log2);
      const float2 nm2 = make_float2(-m, -m);
      float2 sum2 = make_float2(0.f, 0.f);
#pragma unroll 1
      for (int c2 = 0; c2 < 2; ++c2) {
        uint32_t ra[32], rb[32];
        SPB_TMEM_LD_32(tS + (uint32_t)(c2 * 64), ra);
        SPB_TMEM_LD_32(tS + (uint32_t)(c2 * 64 + 32), rb);
        tmem_ld_wait();
        const int lim = need_mask ? n_valid - c2 * 64 : 64;
        uint32_t pk[32];  // 64 bf16
#pragma unroll
        for (int i = 0; i < 32; i += 2) {
          const float2 xa = __ffma2_rn(make_float2(__uint_as_float(ra[i]), __uint_as_float(ra[i + 1])), sc2, nm2);
          const float2 xb = __ffma2_rn(make_float2(__uint_as_float(rb[i]), __uint_as_float(rb[i + 1])), sc2, nm2);
          float2 pa = make_float2(fast_exp2(xa.x), fast_exp2(xa.y));
          float2 pb = make_float2(fast_exp2(xb.x), fast_exp2(xb.y));
          if (need_mask) {
            if (i >= lim) pa.x = 0.f;
            if (i + 1 >= lim) pa.y = 0.f;
            if (i + 32 >= lim) pb.x = 0.f;
            if (i + 33 >= lim) pb.y = 0.f;
          }
          sum2 = __fadd2_rn(sum2, __fadd2_rn(pa, pb));
          pk[i >> 1] = pack_bf16(pa.x, pa.y);
          pk[16 + (i >> 1)] = pack_bf16(pb.x, pb.y);
        }
        if (PT) {
          // P words 32*c2 .. 32*c2+31 overwrite S columns this thread has already consumed (S cols 0..63 are in registers
          // or done when c2 = 0; cols 32..63 were consumed in the c2 = 0 round)
          SPB_TMEM_ST_32(tS + (uint32_t)(c2 * 32), pk);
        } else {
#pragma unroll
          for (int u = 0; u < 8; ++u)
            *reinterpret_cast<uint4*>(prow + c2 * Cfg::CHUNK_BYTES + ((u ^ (r & 7)) << 4)) =
                make_uint4(pk[u * 4], pk[u * 4 + 1], pk[u * 4 + 2], pk[u * 4 + 3]);
        }
      }
      if (PT) tmem_st_wait();
      const float sum = sum2.x + sum2.y;
      l += sum;
      tc_fence_before();
      if (!PT) fence_proxy_async_smem();  // generic-proxy smem writes -> visible to the tensor core (async proxy)
      __syncwarp();
      if (lane == 0) {
        if (!PT) mbar_arrive(&s_empty[t]);  // S_t may be overwritten by QK^T of the next KV tile
        mbar_arrive(&p_full[t]);
      }
    }
    // epilogue: O / l -> bf16 -> global; lse = m*ln2 + ln(l)
    if (n_mine > 0) {
      mbar_wait(&pv_done[t], (n_mine - 1) & 1);
      tc_fence_after();
      const float inv = 1.0f / l;
      const bool valid = q_idx < p.T;
      __nv_bfloat16* orow = p.o + ((int64_t)b * p.T + q_idx) * p.ldo + (int64_t)h * HD;
#pragma unroll
      for (int c = 0; c < HD / 32; ++c) {
        uint32_t o[32];
        SPB_TMEM_LD_32(tO + (uint32_t)(c * 32), o);
        tmem_ld_wait();
        if (valid) {
#pragma unroll
          for (int i = 0; i < 32; i += 8) {
            uint4 v;
            v.x = pack_bf16(__uint_as_float(o[i]) * inv, __uint_as_float(o[i + 1]) * inv);
            v.y = pack_bf16(__uint_as_float(o[i + 2]) * inv, __uint_as_float(o[i + 3]) * inv);
            v.z = pack_bf16(__uint_as_float(o[i + 4]) * inv, __uint_as_float(o[i + 5]) * inv);
            v.w = pack_bf16(__uint_as_float(o[i + 6]) * inv, __uint_as_float(o[i + 7]) * inv);
            *reinterpret_cast<uint4*>(orow + c * 32 + i) = v;
          }
        }
      }
      if (valid) p.lse[((int64_t)b * p.Hq + h) * p.T + q_idx] = m * LN2 + logf(l);
    }
  }

  tc_fence_before();
  __syncthreads();
  if (warp == 1) tmem_dealloc(tmem_base, Cfg::TMEM_COLS);
}

bool flash_tc_supported(int dtype, int hd) { return dtype == SPB_BF16 && (hd == 64 || hd == 128 || hd == 256); }

static int fa_check_operand(const void* ptr, int64_t ld, const char* name) {
  SPB_CHECK_ARG(((uintptr_t)ptr % 16 == 0) && (ld % 8 == 0), SPB_E_UNSUPPORTED,
                "flash (tcgen05): %s must be 16-byte aligned with a leading dimension that is a multiple of 8 (ld=%lld)", name,
                (long long)ld);
  return 0;
}

template <int HD, bool PT>
static int flash_fwd_tc_launch(int B, int Tn, int Hq, int Hkv, int causal, float scale, const void* q, int64_t ldq, const void* k,
                               int64_t ldk, const void* v, int64_t ldv, void* o, int64_t ldo, float* lse, cudaStream_t st) {
  using Cfg = FaFwdCfg<HD, PT>;
  CUtensorMap tmQ, tmK, tmV;
  int rc;
  const uint64_t rows = (uint64_t)B * Tn;
  if ((rc = make_tmap_2d(&tmQ, q, 2, rows, (uint64_t)Hq * HD, (uint64_t)ldq, 128, 64, true, false))) return rc;
  if ((rc = make_tmap_2d(&tmK, k, 2, rows, (uint64_t)Hkv * HD, (uint64_t)ldk, 128, 64, true, false))) return rc;
  if ((rc = make_tmap_2d(&tmV, v, 2, rows, (uint64_t)Hkv * HD, (uint64_t)ldv, 128, 64, true, false))) return rc;
  FaFwdArgs a;
  a.o = (__nv_bfloat16*)o;
  a.ldo = ldo;
  a.lse = lse;
  a.T = Tn;
  a.Hq = Hq;
  a.Hkv = Hkv;
  a.causal = causal;
  a.scale_log2 = scale * LOG2E;
  SPB_CUDA(cudaFuncSetAttribute(flash_fwd_tc_kernel<HD, PT>, cudaFuncAttributeMaxDynamicSharedMemorySize, Cfg::SMEM_BYTES));
  SPB_CHECK_ARG(scale > 0.f, SPB_E_INVALID, "flash forward (tcgen05): scale must be positive");
  dim3 grid((Tn + Cfg::NT * 128 - 1) / (Cfg::NT * 128), Hq, B);
  flash_fwd_tc_kernel<HD, PT><<<grid, Cfg::THREADS, Cfg::SMEM_BYTES, st>>>(tmQ, tmK, tmV, a);
  SPB_LAUNCH_CHECK();
  return 0;
}

int flash_fwd_tc(int B, int Tn, int Hq, int Hkv, int hd, int causal, float scale, const void* q, int64_t ldq, const void* k,
                 int64_t ldk, const void* v, int64_t ldv, void* o, int64_t ldo, float* lse, cudaStream_t st) {
  int rc;
  if ((rc = fa_check_operand(q, ldq, "q")) || (rc = fa_check_operand(k, ldk, "k")) || (rc = fa_check_operand(v, ldv, "v")) ||
      (rc = fa_check_operand(o, ldo, "o")))
    return rc;
  static const bool p_smem = getenv("SPB_FA_PSMEM") != nullptr;  // A/B switch for timing experiments: keep P in shared memory
#define SPB_FA_GO(H)                                                                                                         \
  return p_smem ? flash_fwd_tc_launch<H, false>(B, Tn, Hq, Hkv, causal, scale, q, ldq, k, ldk, v, ldv, o, ldo, lse, st)          \
                : flash_fwd_tc_launch<H, true>(B, Tn, Hq, Hkv, causal, scale, q, ldq, k, ldk, v, ldv, o, ldo, lse, st)
  if (hd == 128) SPB_FA_GO(128);
  if (hd == 64) SPB_FA_GO(64);
  if (hd == 256) SPB_FA_GO(256);
#undef SPB_FA_GO
  SPB_CHECK_ARG(false, SPB_E_UNSUPPORTED, "flash forward (tcgen05): hd %d", hd);
}

}  // namespace spb
