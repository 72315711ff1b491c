// flash_tc.cu - K3: flash-style causal GQA attention forward on tcgen05 (bf16, hd in {64,128,256}); replaces the
// materialised scores of llama3:820-828 (and gpt:344-353, ViT's nn.MultiheadAttention).  Scores never leave the SM.
//
// One CTA = 256 query rows (two 128-row tiles; one tile at hd=256) of one (batch, q head):
//   warp 0    TMA producer: the Q tiles once, then K / V tiles of 128 rows (double buffered)
//   warp 1    MMA issuer (one thread): per KV tile  S_t = Q_t K^T (M128 N128 K=hd, own TMEM buffer per tile) and
//             O_t += P_t V (M128 N=hd K128; P_t is read from tensor memory, where it overwrote S_t)
//   warps 2-17 softmax (4*NSPLIT warps, NSPLIT = 4).  ALL of them work on ONE tile at a time and alternate between the two
//             tiles: warp w owns TMEM lane quadrant w%4 (32 query rows) and column part (w-2)/4 (128/NSPLIT of the 128 KV
//             columns), i.e. NSPLIT threads share a row (2 -> 4 threads per row: +2 % standalone, more warps per scheduler
//             to hide the MUFU / tcgen05.ld latencies).  So a tile's softmax takes half as long, and while the tensor pipe runs PV_t / S_t(next) the
//             whole softmax crew is busy with the other tile - the r01 kernel gave each tile its own four warps, both
//             tiles' softmaxes ran concurrently at half speed each and every tile waited for its own MMAs afterwards
//             (ablation: MMA time and softmax time were additive).
//             Per tile: 64 S values -> registers (two tcgen05.ld, one wait), row max (halves combined through shared
//             memory + one named barrier, which also orders every S load before any P store), exp2 with
//             1/sqrt(hd)*log2(e) folded into one packed FFMA, bf16 P back into tensor memory, lazy rescaling of O (only
//             when the running max grows by more than 2^8, decided per warp; each thread rescales its half of hd)
//   causal tile skipping: KV tiles above a tile's diagonal are neither loaded nor multiplied; diagonal and ragged last
//   tiles are classified per 32-column chunk and warp (fully visible: plain code; partly: per-element select; invisible:
//   zeros, no exp) - (-1e9 additive / -1e4 replacing / -inf masks of the notebooks all give exp()==0 exactly).
//   GQA: kv head = q head / (Hq/Hkv) is just the TMA column coordinate (llama3:626-627 repeat_kv never materialised).
// TMEM: S0 [0,128) S1 [128,256) O0 [256,256+hd) O1 [384,384+hd).  Algorithmic FLOPs: 4*B*Hq*T^2*hd*m (m=(T+1)/2T causal).
#include <cstdlib>

#include "common.cuh"
#include "kernels.h"

namespace spb {

// Timeline tracing for kernel development (tools/diag_trace_fwd.py): compiled in only with -DSPB_FA_TRACE, into a separate
// library; CTA (0,0,0) logs (clock, event id) per role.  The product build contains none of it.
#ifdef SPB_FA_TRACE
constexpr int FA_TRACE_ROLES = 4, FA_TRACE_EVENTS = 2048;
__device__ unsigned long long g_fa_trace[FA_TRACE_ROLES * FA_TRACE_EVENTS];
#define FA_TRACE_DECL(role) int tr_n = 0; const bool tr_on = (blockIdx.x == 0 && blockIdx.y == 0 && blockIdx.z == 0 && lane == 0); const int tr_role = (role)
#define FA_TRACE(id)                                                                                                  \
  do {                                                                                                                \
    if (tr_on && tr_n < FA_TRACE_EVENTS) g_fa_trace[tr_role * FA_TRACE_EVENTS + tr_n++] = ((unsigned long long)clock64() << 8) | (unsigned)(id); \
  } while (0)
#else
#define FA_TRACE_DECL(role)
#define FA_TRACE(id)
#endif

// exp2 on the FMA pipe for a pair of values (x <= ~2^7): Cody-Waite split with the 1.5*2^23 rounding trick, degree-3
// minimax polynomial of 2^f on [-0.5, 0.5] (max relative error 7.5e-5), exponent added with integer arithmetic.
#ifndef SPB_FA_NSPLIT
#define SPB_FA_NSPLIT 4  // softmax threads per query row: 2 (8 softmax warps) or 4 (16 softmax warps)
#endif
#ifndef SPB_FA_POLY_EVERY
#define SPB_FA_POLY_EVERY 2  // one pair of every 2nd group of four elements = 25 % of the exps; 0 disables
#endif
constexpr int FA_POLY_EVERY = SPB_FA_POLY_EVERY;
__device__ __forceinline__ float2 fa_exp2_poly2(float2 x) {
  x.x = fmaxf(x.x, -126.0f);
  x.y = fmaxf(x.y, -126.0f);
  const float2 magic = make_float2(12582912.0f, 12582912.0f);
  const float2 t = __fadd2_rn(x, magic);                                     // low mantissa bits = round(x)
  const float2 n = __fadd2_rn(t, make_float2(-12582912.0f, -12582912.0f));   // round(x) as a float
  const float2 f = __fadd2_rn(x, make_float2(-n.x, -n.y));                   // x - round(x) in [-0.5, 0.5]
  float2 r = __ffma2_rn(f, make_float2(0.0551716573536396f, 0.0551716573536396f), make_float2(0.2426111400127411f, 0.2426111400127411f));
  r = __ffma2_rn(r, f, make_float2(0.6932609677314758f, 0.6932609677314758f));
  r = __ffma2_rn(r, f, make_float2(0.9999280571937561f, 0.9999280571937561f));
  return make_float2(__uint_as_float(__float_as_uint(r.x) + (__float_as_uint(t.x) << 23)),
                     __uint_as_float(__float_as_uint(r.y) + (__float_as_uint(t.y) << 23)));
}

constexpr float LOG2E = 1.4426950408889634f;
constexpr float LN2 = 0.6931471805599453f;

__device__ __forceinline__ float fast_exp2(float x) {
  float y;
  asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x));
  return y;
}

template <int HD>
struct FaFwdCfg {
  static constexpr int NT = (HD <= 128) ? 2 : 1;    // query tiles per CTA (hd=256: O alone needs 256 TMEM columns -> one tile)
  static constexpr int KST = (HD <= 128) ? 2 : 1;   // K stages (hd=256: Q, K, V tiles are 64 KB each)
  static constexpr int VST = (HD <= 128) ? 2 : 1;
  static constexpr int CH = HD / 64;                // 64-column chunks per Q/K/V tile
  static constexpr int TILE_BYTES = 128 * HD * 2;   // 128 rows x HD bf16
  static constexpr int CHUNK_BYTES = 128 * 128;     // 128 rows x 128 B
  static constexpr int OFF_Q = 0;
  static constexpr int OFF_K = OFF_Q + NT * TILE_BYTES;
  static constexpr int OFF_V = OFF_K + KST * TILE_BYTES;
  static constexpr int OFF_X = OFF_V + VST * TILE_BYTES;   // row-max / row-sum exchange: [2 buffers][NSPLIT parts][128 rows] fp32
  static constexpr int NSPLIT = SPB_FA_NSPLIT;            // softmax threads per query row (each owns 128/NSPLIT KV columns)
  static constexpr int CW = 128 / NSPLIT;                 // S / P columns per softmax thread
  static constexpr int OFF_BAR = OFF_X + 2 * NSPLIT * 128 * 4;
  static constexpr int SMEM_BYTES = OFF_BAR + 256 + 1024;
  static constexpr int TMEM_COLS = 512;
  static constexpr int TM_O0 = (NT == 2) ? 256 : 128;     // O_t at TM_O0 + t*128 (NT=2) / one 256-column O at 128 (NT=1)
  static constexpr int SM_WARPS = 4 * NSPLIT;             // softmax warps
  static constexpr int THREADS = 64 + SM_WARPS * 32;
  static constexpr int OH = HD / NSPLIT;                  // O columns per softmax thread
};

struct FaFwdArgs {
  __nv_bfloat16* o;
  int64_t ldo;
  float* lse;
  int T, Hq, Hkv, causal;
  float scale_log2;  // scale * log2(e)
};

template <int HD>
__global__ void __launch_bounds__(FaFwdCfg<HD>::THREADS, 1)
    flash_fwd_tc_kernel(const __grid_constant__ CUtensorMap tmQ, const __grid_constant__ CUtensorMap tmK,
                        const __grid_constant__ CUtensorMap tmV, const FaFwdArgs p) {
  using Cfg = FaFwdCfg<HD>;
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = smem_raw + ((1024u - (smem_u32(smem_raw) & 1023u)) & 1023u);
  uint8_t* sQ = smem + Cfg::OFF_Q;
  uint8_t* sK = smem + Cfg::OFF_K;
  uint8_t* sV = smem + Cfg::OFF_V;
  float* sX = reinterpret_cast<float*>(smem + Cfg::OFF_X);
  uint64_t* bars = reinterpret_cast<uint64_t*>(smem + Cfg::OFF_BAR);
  uint64_t* q_full = bars;            // [1]
  uint64_t* k_full = bars + 1;        // [2] stages
  uint64_t* k_empty = bars + 3;       // [2]
  uint64_t* v_full = bars + 5;        // [2] stages
  uint64_t* v_empty = bars + 7;       // [2]
  uint64_t* s_full = bars + 9;        // [2] per q tile
  uint64_t* p_full = bars + 11;       // [2]
  uint64_t* pv_done = bars + 13;      // [2]
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(bars + 15);

  const int warp = __shfl_sync(0xffffffffu, (int)(threadIdx.x >> 5), 0);
  const int lane = threadIdx.x & 31;

  constexpr int NT = Cfg::NT, KST = Cfg::KST, VST = Cfg::VST;
  // grid (Hq, B, query tile pairs): the hardware hands out CTAs in linear blockIdx order, so with the tile pair in the
  // slowest dimension ALL heavy (late) causal pairs run first and the short ones fill the tail (longest-processing-time
  // order; with the pair in blockIdx.x a 31-unit CTA could start last: ~11 % of the kernel in a schedule simulation);
  // neighbouring CTAs are the q heads of one KV group at the same position, so K/V tiles are shared through L2
  const int n_qp = (int)gridDim.z;
  const int qp = p.causal ? (n_qp - 1 - (int)blockIdx.z) : (int)blockIdx.z;
  const int h = blockIdx.x, b = blockIdx.y;
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
        mbar_init(&p_full[i], Cfg::SM_WARPS);
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
    FA_TRACE_DECL(0);
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
        FA_TRACE(1);
        mbar_arrive_expect_tx(&k_full[bf], Cfg::TILE_BYTES);
#pragma unroll
        for (int c = 0; c < Cfg::CH; ++c)
          tma_load_2d(sK + bf * Cfg::TILE_BYTES + c * Cfg::CHUNK_BYTES, &tmK, &k_full[bf], kvh * HD + c * 64, row_kv);
        const int vf = j % VST;
        mbar_wait(&v_empty[vf], ((j / VST) & 1) ^ 1);
        FA_TRACE(2);
        mbar_arrive_expect_tx(&v_full[vf], Cfg::TILE_BYTES);
#pragma unroll
        for (int c = 0; c < Cfg::CH; ++c)
          tma_load_2d(sV + vf * Cfg::TILE_BYTES + c * Cfg::CHUNK_BYTES, &tmV, &v_full[vf], kvh * HD + c * 64, row_kv);
      }
    }
    __syncwarp();
  } else if (warp == 1) {
    // ------------------------------------------------------------------ MMA issuer (whole warp, uniform; one elected lane issues)
    FA_TRACE_DECL(1);
    {
      const uint32_t idesc_qk = umma_idesc_bf16(128, 128, 0, 0);  // A = Q (K-major), B = K (K-major)
      const uint32_t idesc_pv = umma_idesc_bf16(128, HD, 0, 1);   // A = P (tensor memory), B = V (MN-major: hd contiguous)
      mbar_wait(q_full, 0);
      auto qk_t = [&](int j, int t) {  // S_t = Q_t K(j)^T
        if (j >= n_kv_t[t]) return;
        // S_t is free: PV_t(j-1), issued earlier on this in-order pipe, was the last reader of the P that aliases it
        tc_fence_after();
        const uint64_t dq = umma_desc_kmajor(smem_u32(sQ + t * Cfg::TILE_BYTES));
        const uint64_t dk = umma_desc_kmajor(smem_u32(sK + (j % KST) * Cfg::TILE_BYTES));
#pragma unroll
        for (int kk = 0; kk < HD / 16; ++kk) {
          const uint32_t off = (kk >> 2) * Cfg::CHUNK_BYTES + (kk & 3) * 32;
          umma_bf16_w(tmem_base + t * 128, dq + (off >> 4), dk + (off >> 4), idesc_qk, kk > 0 ? 1u : 0u);
        }
        umma_commit_w(&s_full[t]);
        FA_TRACE(10 + t);  // QK_t issued
      };
      auto pv_t = [&](int j, int t) {  // O_t += P_t V(j)
        if (j >= n_kv_t[t]) return;
        mbar_wait(&p_full[t], j & 1);
        FA_TRACE(12 + t);  // p_full[t] seen
        tc_fence_after();
        const uint64_t dv = umma_desc_mnmajor(smem_u32(sV + (j % VST) * Cfg::TILE_BYTES), 128);
#pragma unroll
        for (int kk = 0; kk < 8; ++kk)  // A = P_t from tensor memory: 16 bf16 of K per 8 columns
          umma_bf16_ts_w(tmem_base + Cfg::TM_O0 + t * 128, tmem_base + t * 128 + kk * 8, dv + ((kk * 2048) >> 4), idesc_pv,
                         (j > 0 || kk > 0) ? 1u : 0u);
        umma_commit_w(&pv_done[t]);
        FA_TRACE(14 + t);  // PV_t issued
      };
      // issue order per KV tile j:  PV0(j) QK0(j+1) PV1(j) QK1(j+1)  - the softmax crew alternates between the tiles in
      // the same order, so a tile's MMAs run while the other tile is in its softmax
      mbar_wait(&k_full[0], 0);
      qk_t(0, 0);
      if (NT == 2) qk_t(0, 1);
      umma_commit_w(&k_empty[0]);
      for (int j = 0; j < n_kv; ++j) {
        const bool more = j + 1 < n_kv;
        mbar_wait(&v_full[j % VST], (j / VST) & 1);
        FA_TRACE(16);  // v_full seen
        pv_t(j, 0);
        if (more) {
          mbar_wait(&k_full[(j + 1) % KST], ((j + 1) / KST) & 1);
          FA_TRACE(17);  // k_full seen
          qk_t(j + 1, 0);
        }
        if (NT == 2) pv_t(j, 1);
        if (more) {
          if (NT == 2) qk_t(j + 1, 1);
          umma_commit_w(&k_empty[(j + 1) % KST]);
        }
        umma_commit_w(&v_empty[j % VST]);
      }
    }
    __syncwarp();
  } else {
    // ------------------------------------------------------------------ softmax + epilogue: warps 2-9, one tile at a time
    const int qd = warp & 3;            // TMEM lane quadrant this warp may touch (hardware rule: warp % 4)
    const int half = (warp - 2) >> 2;   // column part: which CW KV columns of S/P and which hd/NSPLIT columns of O this thread owns
    constexpr int NSPLIT = Cfg::NSPLIT, CW = Cfg::CW, SM_THREADS = Cfg::SM_WARPS * 32;
    const int r = qd * 32 + lane;       // row inside the tile = TMEM lane
    const uint32_t lane_base = tmem_base + ((uint32_t)(qd * 32) << 16);
    float m_t[2] = {-INFINITY, -INFINITY}, l_t[2] = {0.f, 0.f};
    int xbuf = 0;
    FA_TRACE_DECL(warp == 2 ? 2 : (warp == 6 ? 3 : 0));
#ifdef SPB_FA_TRACE
    const bool tr_keep = tr_on && (warp == 2 || warp == 6);
#define FA_TRACE_S(id) do { if (tr_keep) FA_TRACE(id); } while (0)
#else
#define FA_TRACE_S(id)
#endif
    // combine a per-row value over the row's NSPLIT threads (every softmax thread calls it): max or sum of all parts
    auto exchange = [&](float mine, bool is_max) -> float {
      float* buf = sX + xbuf * (NSPLIT * 128);
      buf[half * 128 + r] = mine;
      asm volatile("bar.sync 1, %0;" ::"n"(SM_THREADS) : "memory");
      xbuf ^= 1;
      float acc = buf[r];
#pragma unroll
      for (int pp = 1; pp < NSPLIT; ++pp) acc = is_max ? fmaxf(acc, buf[pp * 128 + r]) : acc + buf[pp * 128 + r];
      return acc;
    };
    // visible columns of this thread's 64 in unit (j,t): n_valid leading ones when the tile needs masking at all
    auto geometry = [&](int j, int t, bool& need_mask, int& n_valid) {
      const int kv0 = j * 128;
      need_mask = (p.causal && kv0 + 127 > q0 + t * 128) || (kv0 + 128 > p.T);
      n_valid = p.T - kv0;
      if (p.causal) n_valid = min(n_valid, q0 + t * 128 + r - kv0 + 1);
      n_valid -= half * CW;
    };
    auto row_max = [&](const uint32_t(&s)[CW], bool need_mask, int n_valid) -> float {
      // row max of the raw scores over the visible ones of CW columns (scale > 0, so scaling commutes with max)
      float mx[4] = {-INFINITY, -INFINITY, -INFINITY, -INFINITY};
      if (!need_mask) {
#pragma unroll
        for (int i = 0; i < CW; i += 8) {
#pragma unroll
          for (int c = 0; c < 4; ++c) mx[c] = fmaxf(mx[c], fmaxf(__uint_as_float(s[i + 2 * c]), __uint_as_float(s[i + 2 * c + 1])));
        }
      } else {
#pragma unroll
        for (int i = 0; i < CW; ++i) mx[i & 3] = fmaxf(mx[i & 3], i < n_valid ? __uint_as_float(s[i]) : -INFINITY);
      }
      return fmaxf(fmaxf(mx[0], mx[1]), fmaxf(mx[2], mx[3]));
    };
    for (int j = 0; j < n_kv; ++j) {
#pragma unroll
      for (int t = 0; t < NT; ++t) {
        if (j >= n_kv_t[t]) continue;  // CTA-uniform
        const uint32_t tS = lane_base + (uint32_t)(t * 128 + half * CW);
        const uint32_t tP = lane_base + (uint32_t)(t * 128 + half * (CW / 2));
        const uint32_t tO = lane_base + (uint32_t)(Cfg::TM_O0 + t * 128 + half * Cfg::OH);
        FA_TRACE_S(20 + t);  // start waiting for S_t
        mbar_wait(&s_full[t], j & 1);
        FA_TRACE_S(22 + t);  // S_t seen
        tc_fence_after();
        uint32_t s[CW];
        SPB_TMEM_LD_32(tS, s);
        if (CW > 32) SPB_TMEM_LD_32(tS + 32u, (s + (CW > 32 ? 32 : 0)));
        tmem_ld_wait();
        bool need_mask;
        int n_valid;
        geometry(j, t, need_mask, n_valid);
        const float mine = row_max(s, need_mask, n_valid);
        FA_TRACE_S(25);  // local max done
        // the barrier inside also orders every S load of this tile before any P store below (P aliases S)
        const float tmax = exchange(mine, true) * p.scale_log2;
        FA_TRACE_S(26);  // exchanged
        float m = m_t[t];
        if (j == 0) {
          m = tmax;
        } else {
          // lazy rescale: keep the stale max while the true max is at most 2^8 above it (p <= 256, harmless in fp32/bf16);
          // both threads of a row see the same tmax and take the same decision
          const bool grow = tmax > m + 8.0f;
          if (__any_sync(0xffffffffu, grow)) {
            float alpha = 1.0f;
            if (grow) {
              alpha = fast_exp2(m - tmax);
              m = tmax;
            }
            l_t[t] *= alpha;
            mbar_wait(&pv_done[t], (j - 1) & 1);  // O_t must hold every PV issued so far
            tc_fence_after();
#pragma unroll 1
            for (int c = 0; c < Cfg::OH / 16; ++c) {
              uint32_t o[16];
              SPB_TMEM_LD_16(tO + (uint32_t)(c * 16), o);
              tmem_ld_wait();
#pragma unroll
              for (int i = 0; i < 16; ++i) o[i] = __float_as_uint(__uint_as_float(o[i]) * alpha);
              SPB_TMEM_ST_16(tO + (uint32_t)(c * 16), o);
            }
            tmem_st_wait();
          }
        }
        m_t[t] = m;
        // ---- p = exp2(s*c - m) (packed fp32x2 FFMA), row sum, bf16 P -> tensor memory words [half*32, half*32+32) of S_t.
        // The exp phase is bound by the 16/clk/SM MUFU pipe, so every FA_POLY_EVERY-th group of four sends one pair through
        // a degree-3 polynomial on the FMA pipe instead (fa_exp2_poly2: 7.5e-5 relative error, far below P's bf16 rounding)
        const float2 sc2 = make_float2(p.scale_log2, p.scale_log2);
        const float2 nm2 = make_float2(-m, -m);
        float2 sum2 = make_float2(0.f, 0.f), sum2b = make_float2(0.f, 0.f);
#pragma unroll
        for (int q = 0; q < CW / 32; ++q) {  // 32 columns -> 16 packed words per tcgen05.st
          uint32_t pk[16];
          int mode = 0;  // 0 = all 32 visible, 1 = masked per element, 2 = none visible (warp-uniform)
          const int lq = n_valid - 32 * q;
          if (need_mask) mode = __all_sync(0xffffffffu, lq >= 32) ? 0 : (__all_sync(0xffffffffu, lq <= 0) ? 2 : 1);
          if (mode == 0) {
#pragma unroll
            for (int i = 0; i < 32; i += 4) {
              const float2 xa = __ffma2_rn(make_float2(__uint_as_float(s[32 * q + i]), __uint_as_float(s[32 * q + i + 1])), sc2, nm2);
              const float2 xb = __ffma2_rn(make_float2(__uint_as_float(s[32 * q + i + 2]), __uint_as_float(s[32 * q + i + 3])), sc2, nm2);
              const float2 pa = make_float2(fast_exp2(xa.x), fast_exp2(xa.y));
              const float2 pb = (FA_POLY_EVERY > 0 && ((i >> 2) % (FA_POLY_EVERY > 0 ? FA_POLY_EVERY : 1)) == 0)
                                    ? fa_exp2_poly2(xb)
                                    : make_float2(fast_exp2(xb.x), fast_exp2(xb.y));
              sum2 = __fadd2_rn(sum2, pa);
              sum2b = __fadd2_rn(sum2b, pb);
              pk[i >> 1] = pack_bf16(pa.x, pa.y);
              pk[(i >> 1) + 1] = pack_bf16(pb.x, pb.y);
            }
          } else if (mode == 1) {
#pragma unroll
            for (int i = 0; i < 32; i += 2) {
              const float2 xa = __ffma2_rn(make_float2(__uint_as_float(s[32 * q + i]), __uint_as_float(s[32 * q + i + 1])), sc2, nm2);
              float2 pa = make_float2(fast_exp2(xa.x), fast_exp2(xa.y));
              if (i >= lq) pa.x = 0.f;
              if (i + 1 >= lq) pa.y = 0.f;
              sum2 = __fadd2_rn(sum2, pa);
              pk[i >> 1] = pack_bf16(pa.x, pa.y);
            }
          } else {
#pragma unroll
            for (int i = 0; i < 16; ++i) pk[i] = 0u;
          }
          SPB_TMEM_ST_16(tP + (uint32_t)(q * 16), pk);
        }
        FA_TRACE_S(27);  // exp done
        tmem_st_wait();
        FA_TRACE_S(28);  // P stored
        sum2 = __fadd2_rn(sum2, sum2b);
        l_t[t] += sum2.x + sum2.y;
        tc_fence_before();
        __syncwarp();
        if (lane == 0) mbar_arrive(&p_full[t]);
        FA_TRACE_S(29);  // arrived
      }
    }
    // epilogue: O / l -> bf16 -> global (each thread its hd/2 columns); lse = m*ln2 + ln(l)
#pragma unroll
    for (int t = 0; t < NT; ++t) {
      const int n_mine = n_kv_t[t];
      if (n_mine == 0) continue;  // CTA-uniform
      const float l = exchange(l_t[t], false);
      const uint32_t tO = lane_base + (uint32_t)(Cfg::TM_O0 + t * 128 + half * Cfg::OH);
      const int q_idx = q0 + t * 128 + r;
      mbar_wait(&pv_done[t], (n_mine - 1) & 1);
      tc_fence_after();
      const float inv = 1.0f / l;
      const bool valid = q_idx < p.T;
      __nv_bfloat16* orow = p.o + ((int64_t)b * p.T + q_idx) * p.ldo + (int64_t)h * HD + half * Cfg::OH;
#pragma unroll
      for (int c = 0; c < Cfg::OH / 16; ++c) {
        uint32_t o[16];
        SPB_TMEM_LD_16(tO + (uint32_t)(c * 16), o);
        tmem_ld_wait();
        if (valid) {
#pragma unroll
          for (int i = 0; i < 16; i += 8) {
            uint4 v;
            v.x = pack_bf16(__uint_as_float(o[i]) * inv, __uint_as_float(o[i + 1]) * inv);
            v.y = pack_bf16(__uint_as_float(o[i + 2]) * inv, __uint_as_float(o[i + 3]) * inv);
            v.z = pack_bf16(__uint_as_float(o[i + 4]) * inv, __uint_as_float(o[i + 5]) * inv);
            v.w = pack_bf16(__uint_as_float(o[i + 6]) * inv, __uint_as_float(o[i + 7]) * inv);
            *reinterpret_cast<uint4*>(orow + c * 16 + i) = v;
          }
        }
      }
      if (valid && half == 0) p.lse[((int64_t)b * p.Hq + h) * p.T + q_idx] = m_t[t] * LN2 + logf(l);
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

template <int HD>
static int flash_fwd_tc_launch(int B, int Tn, int Hq, int Hkv, int causal, float scale, const void* q, int64_t ldq, const void* k,
                               int64_t ldk, const void* v, int64_t ldv, void* o, int64_t ldo, float* lse, cudaStream_t st) {
  using Cfg = FaFwdCfg<HD>;
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
  SPB_CUDA(cudaFuncSetAttribute(flash_fwd_tc_kernel<HD>, cudaFuncAttributeMaxDynamicSharedMemorySize, Cfg::SMEM_BYTES));
  SPB_CHECK_ARG(scale > 0.f, SPB_E_INVALID, "flash forward (tcgen05): scale must be positive");
  SPB_CHECK_ARG(B <= 65535 && (Tn + Cfg::NT * 128 - 1) / (Cfg::NT * 128) <= 65535, SPB_E_LIMIT, "flash forward (tcgen05): B or T/256 beyond 65535");
  dim3 grid(Hq, B, (Tn + Cfg::NT * 128 - 1) / (Cfg::NT * 128));
  flash_fwd_tc_kernel<HD><<<grid, Cfg::THREADS, Cfg::SMEM_BYTES, st>>>(tmQ, tmK, tmV, a);
  SPB_LAUNCH_CHECK();
  return 0;
}

int flash_fwd_tc(int B, int Tn, int Hq, int Hkv, int hd, int causal, float scale, const void* q, int64_t ldq, const void* k,
                 int64_t ldk, const void* v, int64_t ldv, void* o, int64_t ldo, float* lse, cudaStream_t st) {
  int rc;
  if ((rc = fa_check_operand(q, ldq, "q")) || (rc = fa_check_operand(k, ldk, "k")) || (rc = fa_check_operand(v, ldv, "v")) ||
      (rc = fa_check_operand(o, ldo, "o")))
    return rc;
  if (hd == 128) return flash_fwd_tc_launch<128>(B, Tn, Hq, Hkv, causal, scale, q, ldq, k, ldk, v, ldv, o, ldo, lse, st);
  if (hd == 64) return flash_fwd_tc_launch<64>(B, Tn, Hq, Hkv, causal, scale, q, ldq, k, ldk, v, ldv, o, ldo, lse, st);
  if (hd == 256) return flash_fwd_tc_launch<256>(B, Tn, Hq, Hkv, causal, scale, q, ldq, k, ldk, v, ldv, o, ldo, lse, st);
  SPB_CHECK_ARG(false, SPB_E_UNSUPPORTED, "flash forward (tcgen05): hd %d", hd);
}

}  // namespace spb

#ifdef SPB_FA_TRACE
extern "C" int spb_debug_fa_trace_read(unsigned long long* dst, int n) {
  const int total = spb::FA_TRACE_ROLES * spb::FA_TRACE_EVENTS;
  if (n > total) n = total;
  return (int)cudaMemcpyFromSymbol(dst, spb::g_fa_trace, sizeof(unsigned long long) * n);
}
extern "C" int spb_debug_fa_trace_clear() {
  void* p = nullptr;
  cudaError_t e = cudaGetSymbolAddress(&p, spb::g_fa_trace);
  if (e != cudaSuccess) return (int)e;
  return (int)cudaMemset(p, 0, sizeof(unsigned long long) * spb::FA_TRACE_ROLES * spb::FA_TRACE_EVENTS);
}
#endif
