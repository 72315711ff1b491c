// flash_bwd_tc.cu - K4: flash attention backward on tcgen05 (bf16, hd in {64,128}).  VJP of llama3:820-828; the notebook
// has no hand-written backward (jax.value_and_grad, llama3:995) and XLA's saves the [B,H,T,T] probabilities - here they
// are recomputed tile by tile from (q, k, v, lse).
//
// One CTA owns one KV tile (128 rows) of one (batch, kv head) and loops over the q heads of its group and the 64-row Q
// tiles that can see it (causal: from the diagonal on).  Everything is computed TRANSPOSED so that a thread owns a KV row
// in the element-wise phase and an hd column in the dQ phase:
//   (1) S^T  = K  Q^T          M=128(kv) N=64(q)  K=hd     A=K  (K-major)   B=Q_i  (K-major)
//   (2) dP^T = V  dO^T         M=128     N=64     K=hd     A=V  (K-major)   B=dO_i (K-major)
//       P^T = exp2(S^T*c - lse*log2e),  dS^T = P^T o (dP^T - delta) * scale      [threads; bf16 -> smem, UMMA layout]
//   (3) dV  += P^T  dO_i       M=128(kv) N=hd     K=64(q)  A=P^T (K-major)  B=dO_i (MN-major: same smem tile as (2))
//   (4) dK  += dS^T Q_i        M=128     N=hd     K=64     A=dS^T(K-major)  B=Q_i  (MN-major: same tile as (1))
//   (5) dQ^T = K^T  dS         M=128(hd) N=64(q)  K=128(kv) A=K (MN-major view of the resident K tile)
//                                                           B=dS^T buffer read MN-major
// dV and dK accumulate in TMEM over ALL iterations (the GQA group sum of repeat_kv's VJP happens there); dQ^T is drained
// every iteration: thread d owns column d of 64 query rows, the tile is transposed through shared memory and added to the
// fp32 dq accumulator with one TMA bulk reduce-add (cp.reduce.async.bulk.tensor).  TMEM: S^T 64 + dP^T 64 + dV hd + dK hd +
// dQ^T 64 <= 448 columns.
// Warps: 0 = TMA producer (K,V once; Q_i,dO_i double buffered), 1 = MMA issuer, 2-5 and 6-9 = two element-wise groups (each
// owns 32 of the 64 query columns of every KV row: P^T, dS^T into double-buffered smem; at the end one stores dV, the
// other dK), 10-13 = dQ-drain group - the three phases of an iteration overlap: the tensor pipe runs S(it+1) and G(it)
// back to back while the element-wise groups work on it+1 and the drain group reduces it.
// CTAs are ordered heaviest first (causal: KV tile 0 sees every Q tile), blockIdx.z = KV tile.
// Scratch: delta[B,Hq,T] (rowsum(dO*O), prep kernel) and dq_acc[N,Hq*hd] fp32 (zeroed per call, converted at the end).
#include <cstdlib>

#include <type_traits>

#include "common.cuh"
#include "kernels.h"

namespace spb {

// Timeline tracing for kernel development (tools/diag_trace_bwd.py): compiled in only with -DSPB_FB_TRACE, into a separate
// library; CTA (0,0,0) logs (clock, event id) per role.  The product build contains none of it.
#ifdef SPB_FB_TRACE
constexpr int FB_TRACE_ROLES = 4, FB_TRACE_EVENTS = 4096;
__device__ unsigned long long g_fb_trace[FB_TRACE_ROLES * FB_TRACE_EVENTS];
#define FB_TRACE_DECL(role, keep) int tr_n = 0; const bool tr_on = (blockIdx.x == 0 && blockIdx.y == 0 && blockIdx.z == 0 && lane == 0 && (keep)); const int tr_role = (role)
#define FB_TRACE(id)                                                                                                  \
  do {                                                                                                                \
    if (tr_on && tr_n < FB_TRACE_EVENTS) g_fb_trace[tr_role * FB_TRACE_EVENTS + tr_n++] = ((unsigned long long)clock64() << 8) | (unsigned)(id); \
  } while (0)
#else
#define FB_TRACE_DECL(role, keep)
#define FB_TRACE(id)
#endif

// Ablation switches for timing experiments (tools/diag_ablate.py): they skip MMAs or the dQ reduce, so the results become
// WRONG.  Compiled in only with -DSPB_FB_ABLATE (tools/build_variant.sh), into a separate library; the product build has
// neither the switches nor the getenv that reads them.
#ifdef SPB_FB_ABLATE
#define FB_ABL(bit) ((p.ablate & (bit)) != 0)
#else
#define FB_ABL(bit) false
#endif

constexpr int FB_AGROUPS = 2;                         // element-wise warpgroups; each owns 64 / FB_AGROUPS query columns
constexpr int FB_ACOLS = 64 / FB_AGROUPS;
constexpr int FB_ATHREADS = FB_AGROUPS * 128;
constexpr int FB_THREADS = 64 + FB_ATHREADS + 128;  // warp 0 TMA, warp 1 MMA, element-wise groups, dQ-drain group (last 4 warps)
constexpr float FB_LOG2E = 1.4426950408889634f;

__device__ __forceinline__ float fb_exp2(float x) {
  float y;
  asm volatile("ex2.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x));
  return y;
}

template <int HD>
struct FaBwdCfg {
  static constexpr int CH = HD / 64;
  // hd=256: dV + dK alone would fill TMEM, so two CTAs share a KV tile and each owns one 128-column half (z) of the dV, dK
  // and dQ outputs; both recompute the full S^T / dP^T (contraction over all 256 dims): 7 instead of 5 matmul units.
  static constexpr int HS = (HD == 256) ? 2 : 1;
  static constexpr int HDZ = HD / HS;              // output columns per CTA
  static constexpr int ZCH = HDZ / 64;             // 64-column chunks per output half
  static constexpr int K_BYTES = (CH < 2 ? 2 : CH) * 16384;  // hd=64: the second chunk is zeros (M=128 view of K^T)
  static constexpr int V_BYTES = CH * 16384;
  static constexpr int Q_BYTES = CH * 8192;        // 64 rows x HD
  static constexpr int PT_BYTES = 16384;           // 128 kv rows x 64 q columns
  static constexpr int OFF_K = 0;
  static constexpr int OFF_V = OFF_K + K_BYTES;
  // Q_i / dO_i ring: a TMA load takes ~1 us from issue to landed data, and a stage is only released when the MMA group of
  // its iteration has completed, so with 2 stages the ring - not the tensor pipe - set the pace (r01 ablation: 2750 clk per
  // iteration with all math switched off).  3 stages (4 at hd=64) fit because P^T / dS^T are single buffered: the next
  // iteration's element-wise phase reaches its shared-memory writes after the current MMA group has drained.
  static constexpr int NSTAGE = (HD == 128) ? 3 : (HD == 64 ? 4 : 1);  // hd=256: K, V (128 KB) leave room for one 64 KB stage
  static constexpr int OFF_Q = OFF_V + V_BYTES;
  static constexpr int OFF_DO = OFF_Q + NSTAGE * Q_BYTES;
  static constexpr int OFF_DST = OFF_DO + NSTAGE * Q_BYTES;   // dS^T in shared memory: only the dQ^T MMA reads it (as its B operand)
  // fp32 [DQS_ROWS q rows][HDZ] staging tile for the TMA reduce-add of dQ: all 64 rows at once, or - at hd=256, where K, V
  // and the Q/dO stage leave 16 KB - in two rounds of 32 rows (per-lane fp32 atomics, the previous hd=256 path, run at
  // ~1 element per clock and SM and made that backward 2x slower than the hd=128 one)
  static constexpr int OFF_DQS = OFF_DST + PT_BYTES;
  static constexpr int DQS_ROWS = (HD == 256) ? 32 : 64;
  static constexpr int DQS_BYTES = DQS_ROWS * HDZ * 4;
  static constexpr int OFF_LSE = OFF_DQS + DQS_BYTES;     // [2][64] lse*log2e, then [2][64] delta
  static constexpr int OFF_BAR = OFF_LSE + 4 * 64 * 4;
  static constexpr int SMEM_BYTES = OFF_BAR + 256 + 1024;
  static constexpr int TM_ST = 0, TM_DPT = 64, TM_DV = 128, TM_DK = 256, TM_DQT = 384;
  // bf16 P^T and dS^T (128 kv lanes x 64 q = 32 packed columns each): the A operands of the dV / dK MMAs come from tensor
  // memory, not shared memory - the kernel is bound by shared-memory bandwidth (~340 KB of operand, staging and TMA traffic
  // per iteration at ~128 B/clk; r01 ablations), and this takes 48 KB per iteration off it
  static constexpr int TM_P = 448, TM_DS = 480;
  static constexpr int TMEM_COLS = 512;
  static constexpr int DQ_WARPS = HDZ / 32;        // warps that own a valid hd row of dQ^T
};

struct FaBwdArgs {
  const float* lse;
  const float* delta;
  float* dq_acc;
  int64_t ld_acc;
  __nv_bfloat16* dk;
  int64_t lddk;
  __nv_bfloat16* dv;
  int64_t lddv;
  int T, Hq, Hkv, causal;
  float scale, scale_log2;
  const float4* rope;  // optional: apply the rotary VJP (M^T) to dK rows in the epilogue (dQ gets it in the convert kernel)
#ifdef SPB_FB_ABLATE
  int ablate;          // 1 skip dV MMAs, 2 skip dK MMAs, 4 skip dQ^T MMAs, 8 skip the dQ smem staging + TMA reduce,
                       // 16 skip the dP^T MMAs, 32 skip exp/dS math
#endif
};

template <int HD>
__global__ void __launch_bounds__(FB_THREADS, 1)
    flash_bwd_tc_kernel(const __grid_constant__ CUtensorMap tmK, const __grid_constant__ CUtensorMap tmV,
                        const __grid_constant__ CUtensorMap tmQ, const __grid_constant__ CUtensorMap tmDO,
                        const __grid_constant__ CUtensorMap tmDQ, const FaBwdArgs p) {
  using Cfg = FaBwdCfg<HD>;
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = smem_raw + ((1024u - (smem_u32(smem_raw) & 1023u)) & 1023u);
  uint8_t* sK = smem + Cfg::OFF_K;
  uint8_t* sV = smem + Cfg::OFF_V;
  uint8_t* sQ = smem + Cfg::OFF_Q;
  uint8_t* sDO = smem + Cfg::OFF_DO;
  uint8_t* sDST = smem + Cfg::OFF_DST;
  float* sDQ = reinterpret_cast<float*>(smem + Cfg::OFF_DQS);
  float* sLse = reinterpret_cast<float*>(smem + Cfg::OFF_LSE);  // [2][64]
  float* sDelta = sLse + 128;                                    // [2][64]
  uint64_t* bars = reinterpret_cast<uint64_t*>(smem + Cfg::OFF_BAR);
  constexpr int NS = Cfg::NSTAGE;
  uint64_t* kv_full = bars;               // [1]
  uint64_t* qdo_full = bars + 1;          // [NS]
  uint64_t* qdo_empty = bars + 1 + NS;    // [NS]
  uint64_t* sdp_full = bars + 1 + 2 * NS;
  uint64_t* sdp_empty = sdp_full + 1;
  uint64_t* pds_full = sdp_full + 2;
  uint64_t* pds_empty = sdp_full + 3;
  uint64_t* dq_full = sdp_full + 4;
  uint64_t* dq_empty = sdp_full + 5;
  uint64_t* dp_full = sdp_full + 6;       // dP^T complete (sdp_full: S^T complete - the exps start while dP^T is still running)
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(sdp_full + 7);

  const int warp = __shfl_sync(0xffffffffu, (int)(threadIdx.x >> 5), 0);
  const int lane = threadIdx.x & 31;

  const int nt = blockIdx.z, g = blockIdx.x / Cfg::HS, z = blockIdx.x % Cfg::HS, b = blockIdx.y;  // blockIdx.z slowest: heavy KV tiles first
  const int kv0 = nt * 128;
  const int n_rep = p.Hq / p.Hkv;
  const int n_q64 = (p.T + 63) / 64;
  const int i_start = p.causal ? (kv0 / 64) : 0;
  const int n_i = n_q64 - i_start;
  const int n_it = n_rep * n_i;

  if (warp == 0 && lane == 0) {
    prefetch_tmap(&tmK);
    prefetch_tmap(&tmV);
    prefetch_tmap(&tmQ);
    prefetch_tmap(&tmDO);
    prefetch_tmap(&tmDQ);
  }
  if (warp == 1) {
    if (lane == 0) {
      mbar_init(kv_full, 1);
      for (int i = 0; i < NS; ++i) {
        mbar_init(&qdo_full[i], 1);
        mbar_init(&qdo_empty[i], 1);
      }
      mbar_init(sdp_full, 1);
      mbar_init(dp_full, 1);
      mbar_init(sdp_empty, 4 * FB_AGROUPS);
      mbar_init(pds_full, 4 * FB_AGROUPS);
      mbar_init(pds_empty, 1);
      mbar_init(dq_full, 1);
      mbar_init(dq_empty, Cfg::DQ_WARPS);
      fence_barrier_init();
    }
    __syncwarp();
    tmem_alloc(tmem_slot, Cfg::TMEM_COLS);
    tmem_relinquish();
  }
  if (HD == 64 && warp >= 2) {
    // zero the second 64-column chunk of the K buffer: rows 64..127 of the M=128 view of K^T
    uint4* z = reinterpret_cast<uint4*>(sK + 16384);
    for (int i = threadIdx.x - 64; i < 16384 / 16; i += FB_THREADS - 64) z[i] = make_uint4(0, 0, 0, 0);
    fence_proxy_async_smem();
  }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;

  if (warp == 0) {
    // ------------------------------------------------------------------ TMA producer
    FB_TRACE_DECL(0, true);
    if (lane == 0) {
      const int row_kv = b * p.T + kv0;
      mbar_arrive_expect_tx(kv_full, Cfg::CH * 16384 * 2);
#pragma unroll
      for (int c = 0; c < Cfg::CH; ++c) {
        tma_load_2d(sK + c * 16384, &tmK, kv_full, g * HD + c * 64, row_kv);
        tma_load_2d(sV + c * 16384, &tmV, kv_full, g * HD + c * 64, row_kv);
      }
      for (int it = 0; it < n_it; ++it) {
        const int st = it % NS;
        const uint32_t u = (it / NS) & 1;
        const int h = g * n_rep + it / n_i;
        const int row_q = b * p.T + (i_start + it % n_i) * 64;
        mbar_wait(&qdo_empty[st], u ^ 1);
        FB_TRACE(1);
        mbar_arrive_expect_tx(&qdo_full[st], 2 * Cfg::Q_BYTES);
#pragma unroll
        for (int c = 0; c < Cfg::CH; ++c) {
          tma_load_2d(sQ + st * Cfg::Q_BYTES + c * 8192, &tmQ, &qdo_full[st], h * HD + c * 64, row_q);
          tma_load_2d(sDO + st * Cfg::Q_BYTES + c * 8192, &tmDO, &qdo_full[st], h * HD + c * 64, row_q);
        }
      }
    }
    __syncwarp();
  } else if (warp == 1) {
    // ------------------------------------------------------------------ MMA issuer (whole warp, uniform; one elected lane issues)
    FB_TRACE_DECL(1, true);
    {
      const uint32_t idesc_s = umma_idesc_bf16(128, 64, 0, 0);
      const uint32_t idesc_g = umma_idesc_bf16(128, Cfg::HDZ, 0, 1);
      const uint32_t idesc_q = umma_idesc_bf16(128, 64, 1, 1);
      const uint32_t k_addr = smem_u32(sK), v_addr = smem_u32(sV);
      // descriptors: built once; a k-step only adds a constant to the 14-bit start-address field (no carry out of it:
      // shared-memory addresses stay below 256 KB)
      const uint64_t dK_k = umma_desc_kmajor(k_addr), dV_k = umma_desc_kmajor(v_addr);
      const uint64_t dK_mn = umma_desc_mnmajor(k_addr + z * Cfg::ZCH * 16384, 128);
      const uint64_t dDST_mn = umma_desc_mnmajor(smem_u32(sDST), 128);
      mbar_wait(kv_full, 0);
      auto issue_s = [&](int it) {
        const int st = it % NS;
        mbar_wait(&qdo_full[st], (it / NS) & 1);
        FB_TRACE(10);
        mbar_wait(sdp_empty, (it & 1) ^ 1);
        FB_TRACE(11);
        tc_fence_after();
        const uint64_t dQ_k = umma_desc_kmajor(smem_u32(sQ + st * Cfg::Q_BYTES)), dDO_k = umma_desc_kmajor(smem_u32(sDO + st * Cfg::Q_BYTES));
#pragma unroll
        for (int kk = 0; kk < HD / 16; ++kk) {
          const uint32_t aoff = (kk >> 2) * 16384 + (kk & 3) * 32, boff = (kk >> 2) * 8192 + (kk & 3) * 32;
          umma_bf16_w(tmem_base + Cfg::TM_ST, dK_k + (aoff >> 4), dQ_k + (boff >> 4), idesc_s, kk > 0 ? 1u : 0u);
        }
        umma_commit_w(sdp_full);
        if (!FB_ABL(16))
#pragma unroll
        for (int kk = 0; kk < HD / 16; ++kk) {
          const uint32_t aoff = (kk >> 2) * 16384 + (kk & 3) * 32, boff = (kk >> 2) * 8192 + (kk & 3) * 32;
          umma_bf16_w(tmem_base + Cfg::TM_DPT, dV_k + (aoff >> 4), dDO_k + (boff >> 4), idesc_s, kk > 0 ? 1u : 0u);
        }
        umma_commit_w(dp_full);
        FB_TRACE(12);
      };
      auto issue_g = [&](int it) {
        const int st = it % NS;
        mbar_wait(pds_full, it & 1);
        FB_TRACE(13);
        mbar_wait(dq_empty, (it & 1) ^ 1);
        FB_TRACE(14);
        tc_fence_after();
        const uint64_t dDO_mn = umma_desc_mnmajor(smem_u32(sDO + st * Cfg::Q_BYTES) + z * Cfg::ZCH * 8192, 64);
        const uint64_t dQ_mn = umma_desc_mnmajor(smem_u32(sQ + st * Cfg::Q_BYTES) + z * Cfg::ZCH * 8192, 64);
        if (!FB_ABL(1))
#pragma unroll
        for (int kk = 0; kk < 4; ++kk)  // dV += P^T dO
          umma_bf16_ts_w(tmem_base + Cfg::TM_DV, tmem_base + Cfg::TM_P + kk * 8, dDO_mn + ((kk * 2048) >> 4), idesc_g, (it > 0 || kk > 0) ? 1u : 0u);
        if (!FB_ABL(2))
#pragma unroll
        for (int kk = 0; kk < 4; ++kk)  // dK += dS^T Q
          umma_bf16_ts_w(tmem_base + Cfg::TM_DK, tmem_base + Cfg::TM_DS + kk * 8, dQ_mn + ((kk * 2048) >> 4), idesc_g, (it > 0 || kk > 0) ? 1u : 0u);
        if (!FB_ABL(4))
#pragma unroll
        for (int kk = 0; kk < 8; ++kk)  // dQ^T = K^T dS
          umma_bf16_w(tmem_base + Cfg::TM_DQT, dK_mn + ((kk * 2048) >> 4), dDST_mn + ((kk * 2048) >> 4), idesc_q, kk > 0 ? 1u : 0u);
        umma_commit_w(&qdo_empty[st]);
        umma_commit_w(pds_empty);
        umma_commit_w(dq_full);
        FB_TRACE(15);
      };
      issue_s(0);
      for (int it = 0; it < n_it; ++it) {
        if (NS >= 2) {  // S^T / dP^T of the next iteration first: the tensor pipe stays busy while the element-wise groups work
          if (it + 1 < n_it) issue_s(it + 1);
          issue_g(it);
        } else {        // single Q/dO stage (hd=256): the stage of iteration it+1 only loads after G(it) has released it
          issue_g(it);
          if (it + 1 < n_it) issue_s(it + 1);
        }
      }
    }
    __syncwarp();
  } else if (warp < 2 + 4 * FB_AGROUPS) {
    // ------------------------------------------------------------------ element-wise groups (FB_AGROUPS x 4 warps, FB_ACOLS query columns each)
    const int wg = (warp - 2) >> 2;
    const int qd = warp & 3;
    const int r = qd * 32 + lane;     // kv row = TMEM lane
    const int ct = threadIdx.x - 64;  // 0 .. FB_ATHREADS-1
    const int kv_idx = kv0 + r;
    const int cbase = wg * FB_ACOLS;  // first of this group's query columns
    const uint32_t lane_base = tmem_base + ((uint32_t)(qd * 32) << 16);
    // lse*log2e / delta of the 64 query rows of an iteration are staged in shared memory (double buffered).  The global
    // loads for iteration it+1 are issued at the top of iteration it and land while it computes (their latency used to
    // sit on the critical path of every iteration).
    auto fetch_stat = [&](int it) -> float {
      if (ct >= 128 || it >= n_it) return 0.f;
      const int h = g * n_rep + it / n_i;
      const int qi = (i_start + it % n_i) * 64 + (ct & 63);
      if (qi >= p.T) return 0.f;
      return ((ct < 64) ? p.lse : p.delta)[((int64_t)b * p.Hq + h) * p.T + qi];  // consumed in put_stat, a whole iteration later
    };
    auto put_stat = [&](int it, float v) {
      if (ct < 64) sLse[(it & 1) * 64 + ct] = v * FB_LOG2E;
      else if (ct < 128) sDelta[(it & 1) * 64 + (ct - 64)] = v;
    };
    FB_TRACE_DECL(2, warp == 2);
    put_stat(0, fetch_stat(0));
    const float2 sc2 = make_float2(p.scale_log2, p.scale_log2);
    const float2 scale2 = make_float2(p.scale, p.scale);
    for (int it = 0; it < n_it; ++it) {
      const int st = it & 1;
      const int q0 = (i_start + it % n_i) * 64;
      const float stat_next = fetch_stat(it + 1);
      asm volatile("bar.sync 1, %0;" ::"n"(FB_ATHREADS) : "memory");  // stage st is complete; stage st^1 is no longer being read
      FB_TRACE(20);
      mbar_wait(sdp_full, it & 1);
      FB_TRACE(21);
      tc_fence_after();
      uint32_t rs[FB_ACOLS], rd[FB_ACOLS];
      if (FB_ACOLS == 32) {
        SPB_TMEM_LD_32(lane_base + (uint32_t)(Cfg::TM_ST + cbase), rs);
      } else {
        SPB_TMEM_LD_16(lane_base + (uint32_t)(Cfg::TM_ST + cbase), rs);
      }
      tmem_ld_wait();
      FB_TRACE(22);

      const bool need_mask = (p.causal && q0 < kv0 + 128) || (q0 + 64 > p.T) || (kv0 + 128 > p.T);
      const float2* lse2 = reinterpret_cast<const float2*>(sLse + st * 64 + cbase);
      const float2* del2 = reinterpret_cast<const float2*>(sDelta + st * 64 + cbase);
      uint32_t pkp[FB_ACOLS / 2], pkd[FB_ACOLS / 2];
      // Phase 1 (S^T only): P^T = exp2(S^T c - lse); the fp32 values go back into rs for phase 2, the bf16 ones into pkp.
      // Two explicit instantiations: left to one loop with `if (need_mask)` inside, the compiler if-converts the mask and
      // every tile pays its compares and selects (ncu r01h: 1.4 ISETP/PLOP3 per exp); only diagonal/ragged tiles need them
      auto probs = [&](auto masked) {
        constexpr bool MASKED = decltype(masked)::value;
#pragma unroll
        for (int i = 0; i < FB_ACOLS; i += 2) {
          const float2 l2 = lse2[i >> 1];
          const float2 x = __ffma2_rn(make_float2(__uint_as_float(rs[i]), __uint_as_float(rs[i + 1])), sc2, make_float2(-l2.x, -l2.y));
          float2 pv = make_float2(fb_exp2(x.x), fb_exp2(x.y));
          if (MASKED) {
            const int qi = q0 + cbase + i;
            const bool kv_dead = kv_idx >= p.T;
            if (kv_dead || qi >= p.T || (p.causal && kv_idx > qi)) pv.x = 0.f;
            if (kv_dead || qi + 1 >= p.T || (p.causal && kv_idx > qi + 1)) pv.y = 0.f;
          }
          rs[i] = __float_as_uint(pv.x);
          rs[i + 1] = __float_as_uint(pv.y);
          pkp[i >> 1] = pack_bf16(pv.x, pv.y);
        }
      };
      if (FB_ABL(32)) {
#pragma unroll
        for (int i = 0; i < FB_ACOLS / 2; ++i) pkp[i] = rs[i];
      } else if (need_mask) {
        probs(std::true_type{});
      } else {
        probs(std::false_type{});
      }
      // Phase 2: dP^T has been running on the tensor pipe meanwhile; dS^T = P^T o (dP^T - delta) * scale
      mbar_wait(dp_full, it & 1);
      tc_fence_after();
      if (FB_ACOLS == 32) {
        SPB_TMEM_LD_32(lane_base + (uint32_t)(Cfg::TM_DPT + cbase), rd);
      } else {
        SPB_TMEM_LD_16(lane_base + (uint32_t)(Cfg::TM_DPT + cbase), rd);
      }
      tmem_ld_wait();
      tc_fence_before();
      __syncwarp();
      if (lane == 0) mbar_arrive(sdp_empty);  // S^T / dP^T may be overwritten by iteration it+1
      if (FB_ABL(32)) {
#pragma unroll
        for (int i = 0; i < FB_ACOLS / 2; ++i) pkd[i] = rs[i] ^ rd[i];
      } else {
#pragma unroll
        for (int i = 0; i < FB_ACOLS; i += 2) {
          const float2 d2 = del2[i >> 1];
          const float2 dd = __fadd2_rn(make_float2(__uint_as_float(rd[i]), __uint_as_float(rd[i + 1])), make_float2(-d2.x, -d2.y));
          const float2 ds = __fmul2_rn(__fmul2_rn(make_float2(__uint_as_float(rs[i]), __uint_as_float(rs[i + 1])), dd), scale2);
          pkd[i >> 1] = pack_bf16(ds.x, ds.y);
        }
      }
      FB_TRACE(23);
      put_stat(it + 1, stat_next);  // visible to everyone after the next iteration's bar.sync
      mbar_wait(pds_empty, (it & 1) ^ 1);
      FB_TRACE(24);  // the MMA group of iteration it-1 has finished reading P^T / dS^T
      static_assert(FB_ACOLS == 32, "the tensor-memory stores below move 16 packed words per thread");
      // P^T and dS^T -> tensor memory (A operands of the dV / dK MMAs): lane = kv row, word w = query columns 2w, 2w+1
      SPB_TMEM_ST_16(lane_base + (uint32_t)(Cfg::TM_P + cbase / 2), pkp);
      SPB_TMEM_ST_16(lane_base + (uint32_t)(Cfg::TM_DS + cbase / 2), pkd);
      // dS^T also -> shared memory (UMMA K-major, 128B swizzle): B operand of the dQ^T MMA
      uint8_t* drow = sDST + r * 128;
#pragma unroll
      for (int c = 0; c < FB_ACOLS / 8; ++c) {
        const int slot = ((wg * (FB_ACOLS / 8) + c) ^ (r & 7)) << 4;
        *reinterpret_cast<uint4*>(drow + slot) = make_uint4(pkd[c * 4], pkd[c * 4 + 1], pkd[c * 4 + 2], pkd[c * 4 + 3]);
      }
      tmem_st_wait();
      fence_proxy_async_smem();
      tc_fence_before();
      __syncwarp();
      if (lane == 0) mbar_arrive(pds_full);
      FB_TRACE(25);
    }
    // ---- epilogue: group 0 stores dV, group 1 stores dK (rows -> bf16 -> global).  The last pds_empty commit covers
    // every MMA issued before it.
    mbar_wait(pds_empty, (n_it - 1) & 1);
    tc_fence_after();
    const bool valid = kv_idx < p.T && wg < 2;
    const int which = wg & 1;  // group 0: dV, group 1: dK (further groups only wait for completion)
    __nv_bfloat16* orow = (which ? p.dk + ((int64_t)b * p.T + kv_idx) * p.lddk : p.dv + ((int64_t)b * p.T + kv_idx) * p.lddv) + (int64_t)g * HD + z * Cfg::HDZ;
    const int tm = which ? Cfg::TM_DK : Cfg::TM_DV;
#pragma unroll
    for (int c = 0; c < Cfg::HDZ / 32; ++c) {
      uint32_t o[32];
      SPB_TMEM_LD_32(lane_base + (uint32_t)(tm + c * 32), o);
      tmem_ld_wait();
      if (valid) {
        if (which == 1 && p.rope != nullptr) {  // dK: transpose of the 2x2 rotary block, table row = position kv_idx
          const float4* tab = p.rope + (int64_t)kv_idx * (HD / 2) + z * (Cfg::HDZ / 2) + c * 16;
#pragma unroll
          for (int i = 0; i < 32; i += 2) {
            const float4 m = tab[i >> 1];
            const float xe = __uint_as_float(o[i]), xo = __uint_as_float(o[i + 1]);
            o[i] = __float_as_uint(m.x * xe + m.z * xo);
            o[i + 1] = __float_as_uint(m.y * xe + m.w * xo);
          }
        }
#pragma unroll
        for (int i = 0; i < 32; i += 8) {
          uint4 v;
          v.x = pack_bf16(__uint_as_float(o[i]), __uint_as_float(o[i + 1]));
          v.y = pack_bf16(__uint_as_float(o[i + 2]), __uint_as_float(o[i + 3]));
          v.z = pack_bf16(__uint_as_float(o[i + 4]), __uint_as_float(o[i + 5]));
          v.w = pack_bf16(__uint_as_float(o[i + 6]), __uint_as_float(o[i + 7]));
          *reinterpret_cast<uint4*>(orow + c * 32 + i) = v;
        }
      }
    }
  } else {
    // ------------------------------------------------------------------ drain group (warps 10..13): reduce dQ^T every iteration
    const int qd = warp & 3;
    const int r = qd * 32 + lane;  // hd column of dQ = TMEM lane of dQ^T
    const uint32_t lane_base = tmem_base + ((uint32_t)(qd * 32) << 16);
    FB_TRACE_DECL(3, qd == 0);
    if (qd < Cfg::DQ_WARPS) {  // warp-uniform (hd=64: only two warps hold real rows)
      for (int it = 0; it < n_it; ++it) {
        const int h = g * n_rep + it / n_i;
        const int q0 = (i_start + it % n_i) * 64;
        mbar_wait(dq_full, it & 1);
        FB_TRACE(30);
        tc_fence_after();
        float dq[64];
#pragma unroll
        for (int c = 0; c < 2; ++c) {
          uint32_t raw[32];
          SPB_TMEM_LD_32(lane_base + (uint32_t)(Cfg::TM_DQT + c * 32), raw);
          tmem_ld_wait();
#pragma unroll
          for (int i = 0; i < 32; ++i) dq[c * 32 + i] = __uint_as_float(raw[i]);
        }
        tc_fence_before();
        __syncwarp();
        if (lane == 0) mbar_arrive(dq_empty);  // dQ^T may be overwritten by iteration it+1 while the reduction is in flight
        FB_TRACE(31);
        // transpose through shared memory ([q][d] fp32, lanes write consecutive d: conflict-free) and let the TMA engine do
        // ONE bulk reduce-add of the 64 x hd tile into the fp32 dq accumulator (per-thread fp32 atomics were the v2
        // bottleneck: ~1 float atomic / clk / SM).  Rows past the sequence end hold exact zeros (their P^T is masked).
        if (FB_ABL(8)) continue;
        constexpr int NB = Cfg::DQ_WARPS * 32;
        const bool leader = (qd == 0 && lane == 0);
#pragma unroll
        for (int rd = 0; rd < 64 / Cfg::DQS_ROWS; ++rd) {
          if (leader && (it > 0 || rd > 0)) tma_store_wait_read<0>();  // previous reduce has finished reading the staging tile
          asm volatile("bar.sync 2, %0;" ::"n"(NB) : "memory");
#pragma unroll
          for (int i = 0; i < Cfg::DQS_ROWS; ++i) sDQ[i * Cfg::HDZ + r] = dq[rd * Cfg::DQS_ROWS + i];
          fence_proxy_async_smem();
          asm volatile("bar.sync 2, %0;" ::"n"(NB) : "memory");
          if (rd == 64 / Cfg::DQS_ROWS - 1) FB_TRACE(32);
          if (leader) {
            tma_reduce_add_2d(&tmDQ, sDQ, h * HD + z * Cfg::HDZ, b * p.T + q0 + rd * Cfg::DQS_ROWS);
            tma_store_commit();
          }
        }
        FB_TRACE(33);
      }
      if (qd == 0 && lane == 0) tma_store_wait_read<0>();
    }
  }

  tc_fence_before();
  __syncthreads();
  if (warp == 1) tmem_dealloc(tmem_base, Cfg::TMEM_COLS);
}

// delta[b,h,t] = sum_d dO[b,t,h,d] * O[b,t,h,d]; HD/8 lanes per (token, head) row, one 128-bit load per lane and operand
template <int HD>
__global__ void __launch_bounds__(256) flash_bwd_prep_kernel(const __nv_bfloat16* __restrict__ o, int64_t ldo,
                                                             const __nv_bfloat16* __restrict__ d_o, int64_t lddo,
                                                             float* __restrict__ delta, int B, int T, int Hq) {
  constexpr int TPR = HD / 8;  // lanes per row: 8, 16 or 32
  const int64_t idx = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  const int64_t w = idx / TPR;
  const int c = (int)(idx % TPR) * 8;
  const int64_t total = (int64_t)B * T * Hq;
  float s = 0.f;
  int h = 0;
  int64_t tok = 0;
  if (w < total) {
    h = (int)(w % Hq);
    tok = w / Hq;
    const uint4 a = *reinterpret_cast<const uint4*>(o + tok * ldo + (int64_t)h * HD + c);
    const uint4 g = *reinterpret_cast<const uint4*>(d_o + tok * lddo + (int64_t)h * HD + c);
    const float2 a0 = unpack_bf16(a.x), a1 = unpack_bf16(a.y), a2 = unpack_bf16(a.z), a3 = unpack_bf16(a.w);
    const float2 g0 = unpack_bf16(g.x), g1 = unpack_bf16(g.y), g2 = unpack_bf16(g.z), g3 = unpack_bf16(g.w);
    s = a0.x * g0.x + a0.y * g0.y + a1.x * g1.x + a1.y * g1.y + a2.x * g2.x + a2.y * g2.y + a3.x * g3.x + a3.y * g3.y;
  }
#pragma unroll
  for (int off = TPR / 2; off > 0; off >>= 1) s += __shfl_xor_sync(0xffffffffu, s, off);
  if (w < total && c == 0) {
    const int64_t bb = tok / T, t = tok % T;
    delta[(bb * Hq + h) * T + t] = s;
  }
}

// dq[n, c] = bf16(dq_acc[n, c]) with the rotary VJP fused; a thread owns (row n, one 8-column position inside a head) and
// walks over HPT heads, so the four table entries it needs are fetched once (they are the same for every head of a row).
template <int HPT>
__global__ void __launch_bounds__(256) flash_bwd_dq_convert_kernel(const float* __restrict__ acc, int64_t ld_acc,
                                                                   __nv_bfloat16* __restrict__ dq, int64_t lddq, int64_t N,
                                                                   int heads, const float4* __restrict__ rope, int T, int hd) {
  const int vph = hd / 8;  // 8-column vectors per head row
  const int64_t idx = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (idx >= N * vph) return;
  const int64_t n = idx / vph;
  const int v = (int)(idx % vph);
  float4 m0, m1, m2, m3;
  if (rope != nullptr) {
    const float4* tab = rope + (int64_t)(n % T) * (hd >> 1) + v * 4;
    m0 = tab[0]; m1 = tab[1]; m2 = tab[2]; m3 = tab[3];
  }
  const int h0 = blockIdx.y * HPT;
  const float* src = acc + n * ld_acc + (int64_t)h0 * hd + v * 8;
  __nv_bfloat16* dst = dq + n * lddq + (int64_t)h0 * hd + v * 8;
  float4 a[HPT], b[HPT];
#pragma unroll
  for (int h = 0; h < HPT; ++h) {
    if (h0 + h < heads) {
      a[h] = *reinterpret_cast<const float4*>(src + (int64_t)h * hd);
      b[h] = *reinterpret_cast<const float4*>(src + (int64_t)h * hd + 4);
    }
  }
#pragma unroll
  for (int h = 0; h < HPT; ++h) {
    if (h0 + h < heads) {
      float4 x = a[h], y = b[h];
      if (rope != nullptr) {  // transpose of the 2x2 rotary block
        float xe, xo;
        xe = x.x; xo = x.y; x.x = m0.x * xe + m0.z * xo; x.y = m0.y * xe + m0.w * xo;
        xe = x.z; xo = x.w; x.z = m1.x * xe + m1.z * xo; x.w = m1.y * xe + m1.w * xo;
        xe = y.x; xo = y.y; y.x = m2.x * xe + m2.z * xo; y.y = m2.y * xe + m2.w * xo;
        xe = y.z; xo = y.w; y.z = m3.x * xe + m3.z * xo; y.w = m3.y * xe + m3.w * xo;
      }
      uint4 o;
      o.x = pack_bf16(x.x, x.y);
      o.y = pack_bf16(x.z, x.w);
      o.z = pack_bf16(y.x, y.y);
      o.w = pack_bf16(y.z, y.w);
      *reinterpret_cast<uint4*>(dst + (int64_t)h * hd) = o;
    }
  }
}

// ------------------------------------------------------------------------------------------------ hd = 256: two ROLES per KV tile
// dV + dK at hd 256 are 512 TMEM columns: they cannot live in one CTA next to S^T, dP^T and dQ^T.  The first version above
// splits the OUTPUT columns over two CTAs (z = 0, 1), and both recompute the full S^T and dP^T (7 matmul units for 5) with a
// single Q/dO stage each (K and V resident: no shared memory for a second one), so every iteration is a serial chain
// TMA -> S^T/dP^T -> element-wise -> dV/dK/dQ^T (6100 clk, tensor pipe 29 % busy; profiles/README.md).
// Here the two CTAs of a KV tile split the WORK instead:
//   role V: dV (all 256 columns).  Needs only K resident (S^T) -> two Q/dO stages fit and the iteration pipelines like the
//           hd-128 kernel: S^T(it+1) runs while the element-wise groups turn S^T(it) into P^T.  No dP^T, no dS^T, no dQ.
//           2 matmul units per iteration.
//   role K: dK (all 256 columns) and dQ (both 128-row halves of dQ^T).  K and V resident, one stage; dS^T feeds the dK MMA
//           from shared memory (the tensor-memory columns the TS form would need hold the second dQ^T half).  4 units.
// 6 units instead of 7, and a third of them in a pipelined CTA.  SPB_BWD256_ROLES=0 keeps the column split.
template <int ROLE>
struct Fb256Cfg {
  static constexpr int HD = 256, CH = 4;
  static constexpr int NS = (ROLE == 1) ? 2 : 1;
  static constexpr int K_BYTES = 65536;
  static constexpr int V_BYTES = (ROLE == 2) ? 65536 : 0;
  static constexpr int Q_BYTES = 32768;                    // 64 rows x 256
  static constexpr int OFF_K = 0;
  static constexpr int OFF_V = OFF_K + K_BYTES;
  static constexpr int OFF_Q = OFF_V + V_BYTES;
  static constexpr int OFF_DO = OFF_Q + NS * Q_BYTES;
  static constexpr int OFF_DST = OFF_DO + NS * Q_BYTES;    // role K: dS^T [128 kv][64 q] bf16, UMMA K-major, 128B swizzle
  static constexpr int DST_BYTES = (ROLE == 2) ? 16384 : 0;
  static constexpr int OFF_DQS = OFF_DST + DST_BYTES;      // role K: fp32 [32 q rows][128 hd] staging of a quarter dQ tile
  static constexpr int DQS_BYTES = (ROLE == 2) ? 16384 : 0;
  static constexpr int OFF_LSE = OFF_DQS + DQS_BYTES;      // [2][64] lse*log2e, [2][64] delta
  static constexpr int OFF_BAR = OFF_LSE + 4 * 64 * 4;
  static constexpr int SMEM_BYTES = OFF_BAR + 256 + 1024;
  static constexpr int TM_ST = 0, TM_DPT = 64, TM_ACC = 128 /* dV or dK: 256 columns */, TM_DQT = 384 /* role K: 2 x 64 */, TM_P = 448 /* role V */;
};

template <int ROLE>
__device__ __forceinline__ void fb256_body(const CUtensorMap& tmK, const CUtensorMap& tmV, const CUtensorMap& tmQ, const CUtensorMap& tmDO,
                                           const CUtensorMap& tmDQ, const FaBwdArgs& p, uint8_t* smem, const int g, const int rank, const int S) {
  using Cfg = Fb256Cfg<ROLE>;
  constexpr int HD = 256;
  constexpr int NS = Cfg::NS;
  uint8_t* sK = smem + Cfg::OFF_K;
  uint8_t* sV = smem + Cfg::OFF_V;
  uint8_t* sQ = smem + Cfg::OFF_Q;
  uint8_t* sDO = smem + Cfg::OFF_DO;
  uint8_t* sDST = smem + Cfg::OFF_DST;
  float* sDQ = reinterpret_cast<float*>(smem + Cfg::OFF_DQS);
  float* sLse = reinterpret_cast<float*>(smem + Cfg::OFF_LSE);
  float* sDelta = sLse + 128;
  uint64_t* bars = reinterpret_cast<uint64_t*>(smem + Cfg::OFF_BAR);
  // The Q_i and dO_i halves of a stage are tracked apart, because they die at different times.  Role K (one stage, a serial
  // chain): dO_i is free as soon as dP^T is done (reloaded under the element-wise phase), Q_i as soon as the dK MMAs are done
  // (reloaded under the dQ^T MMAs), and the element-wise groups start on S^T while dP^T is still running.  Role V (two
  // stages): Q_i is free right after S^T - long before the dV MMAs release dO_i - so its reload starts a phase earlier; the
  // two streams have their own producer threads (warp 0: K, Q; first drain warp: dO).
  uint64_t* kv_full = bars;
  uint64_t* q_full = bars + 1;            // [NS]
  uint64_t* q_empty = bars + 1 + NS;      // [NS]
  uint64_t* do_full = bars + 1 + 2 * NS;  // [NS]
  uint64_t* do_empty = bars + 1 + 3 * NS; // [NS]
  uint64_t* s_full = bars + 1 + 4 * NS;   // S^T complete
  uint64_t* dp_full = s_full + 1;         // role K: dP^T complete
  uint64_t* sdp_empty = s_full + 2;
  uint64_t* pds_full = s_full + 3;
  uint64_t* pds_empty = s_full + 4;
  uint64_t* dq_full = s_full + 5;
  uint64_t* dq_empty = s_full + 6;
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(s_full + 7);

  const int warp = __shfl_sync(0xffffffffu, (int)(threadIdx.x >> 5), 0);
  const int lane = threadIdx.x & 31;
  const int nt = blockIdx.z, b = blockIdx.y;
  const int kv0 = nt * 128;
  const int n_rep = p.Hq / p.Hkv;
  const int n_q64 = (p.T + 63) / 64;
  const int i_start = p.causal ? (kv0 / 64) : 0;
  const int n_i = n_q64 - i_start;
  // S > 1: the S CTAs of a cluster share this role of this KV tile; rank r takes the iterations r, r+S, ... (local index j,
  // global iteration rank + j*S) and the partial dV / dK of ranks 1.. are added into rank 0's through distributed shared memory
  const int n_it = (n_rep * n_i - rank + S - 1) / S;
  auto giter = [&](int j) { return rank + j * S; };

  if (warp == 0 && lane == 0) {
    prefetch_tmap(&tmK);
    prefetch_tmap(&tmQ);
    prefetch_tmap(&tmDO);
    if (ROLE == 2) {
      prefetch_tmap(&tmV);
      prefetch_tmap(&tmDQ);
    }
  }
  if (warp == 1) {
    if (lane == 0) {
      mbar_init(kv_full, 1);
      for (int i = 0; i < NS; ++i) {
        mbar_init(&q_full[i], 1);
        mbar_init(&q_empty[i], 1);
        mbar_init(&do_full[i], 1);
        mbar_init(&do_empty[i], 1);
      }
      mbar_init(s_full, 1);
      mbar_init(dp_full, 1);
      mbar_init(sdp_empty, 4 * FB_AGROUPS);
      mbar_init(pds_full, 4 * FB_AGROUPS);
      mbar_init(pds_empty, 1);
      mbar_init(dq_full, 1);
      mbar_init(dq_empty, 4);
      fence_barrier_init();
    }
    __syncwarp();
    tmem_alloc(tmem_slot, 512);
    tmem_relinquish();
  }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;

  if (warp == 0) {
    // ------------------------------------------------------------------ TMA producer
    FB_TRACE_DECL(0, true);
    if (lane == 0) {
      const int row_kv = b * p.T + kv0;
      mbar_arrive_expect_tx(kv_full, (ROLE == 2 ? 2 : 1) * 4 * 16384);
#pragma unroll
      for (int c = 0; c < 4; ++c) {
        tma_load_2d(sK + c * 16384, &tmK, kv_full, g * HD + c * 64, row_kv);
        if (ROLE == 2) tma_load_2d(sV + c * 16384, &tmV, kv_full, g * HD + c * 64, row_kv);
      }
      for (int it = 0; it < n_it; ++it) {
        const int st = it % NS;
        const uint32_t u = (it / NS) & 1;
        const int h = g * n_rep + giter(it) / n_i;
        const int row_q = b * p.T + (i_start + giter(it) % n_i) * 64;
        const bool more = it + 1 < n_it;
        const int h_n = g * n_rep + giter(it + 1) / n_i;
        const int row_n = b * p.T + (i_start + giter(it + 1) % n_i) * 64;
        if (ROLE == 1) {  // the Q stream (the dO stream has its own thread below)
          mbar_wait(&q_empty[st], u ^ 1);
          FB_TRACE(1);
          mbar_arrive_expect_tx(&q_full[st], Cfg::Q_BYTES);
#pragma unroll
          for (int c = 0; c < 4; ++c) tma_load_2d(sQ + st * Cfg::Q_BYTES + c * 8192, &tmQ, &q_full[st], h * HD + c * 64, row_q);
        } else {
          // Role K keeps ONE stage, but its two 32 KB halves swap jobs every iteration: Q_it sits in half (it & 1), dO_it in
          // the other.  dO_it dies early (after dP^T) - Q_{it+1} is loaded over it while the element-wise phase runs - and Q_it
          // dies after the dK MMAs - dO_{it+1} is loaded over it and lands under S^T_{it+1}.  Neither load is on the chain.
          if (it == 0) {
            mbar_arrive_expect_tx(&q_full[0], Cfg::Q_BYTES);
            mbar_arrive_expect_tx(&do_full[0], Cfg::Q_BYTES);
#pragma unroll
            for (int c = 0; c < 4; ++c) {
              tma_load_2d(sQ + c * 8192, &tmQ, &q_full[0], h * HD + c * 64, row_q);
              tma_load_2d(sQ + Cfg::Q_BYTES + c * 8192, &tmDO, &do_full[0], h * HD + c * 64, row_q);
            }
          }
          if (more) {
            uint8_t* half_q = sQ + (it & 1) * Cfg::Q_BYTES;         // holds Q_it
            uint8_t* half_do = sQ + ((it & 1) ^ 1) * Cfg::Q_BYTES;  // holds dO_it
            mbar_wait(&do_empty[0], it & 1);  // dP^T of iteration it is done
            FB_TRACE(1);
            mbar_arrive_expect_tx(&q_full[0], Cfg::Q_BYTES);
#pragma unroll
            for (int c = 0; c < 4; ++c) tma_load_2d(half_do + c * 8192, &tmQ, &q_full[0], h_n * HD + c * 64, row_n);
            mbar_wait(&q_empty[0], it & 1);   // the dK MMAs of iteration it are done
            FB_TRACE(2);
            mbar_arrive_expect_tx(&do_full[0], Cfg::Q_BYTES);
#pragma unroll
            for (int c = 0; c < 4; ++c) tma_load_2d(half_q + c * 8192, &tmDO, &do_full[0], h_n * HD + c * 64, row_n);
          }
        }
      }
    }
    __syncwarp();
  } else if (warp == 1) {
    // ------------------------------------------------------------------ MMA issuer
    FB_TRACE_DECL(1, true);
    const uint32_t idesc_s = umma_idesc_bf16(128, 64, 0, 0);
    const uint32_t idesc_g = umma_idesc_bf16(128, 256, 0, 1);   // dV / dK: N = all 256 head columns, B MN-major (the Q / dO stage)
    const uint32_t idesc_q = umma_idesc_bf16(128, 64, 1, 1);
    const uint32_t k_addr = smem_u32(sK), v_addr = smem_u32(sV);
    const uint64_t dK_k = umma_desc_kmajor(k_addr), dV_k = umma_desc_kmajor(v_addr);
    const uint64_t dDST_k = umma_desc_kmajor(smem_u32(sDST));
    const uint64_t dDST_mn = umma_desc_mnmajor(smem_u32(sDST), 128);
    mbar_wait(kv_full, 0);
    auto issue_s = [&](int it) {
      const int st = it % NS;
      mbar_wait(&q_full[st], (it / NS) & 1);
      FB_TRACE(10);
      mbar_wait(sdp_empty, (it & 1) ^ 1);
      FB_TRACE(11);
      tc_fence_after();
      // role K: the two halves of its single stage swap jobs every iteration (see the producer)
      const uint8_t* bq = ROLE == 2 ? sQ + (it & 1) * Cfg::Q_BYTES : sQ + st * Cfg::Q_BYTES;
      const uint8_t* bdo = ROLE == 2 ? sQ + ((it & 1) ^ 1) * Cfg::Q_BYTES : sDO + st * Cfg::Q_BYTES;
      const uint64_t dQ_k = umma_desc_kmajor(smem_u32(bq)), dDO_k = umma_desc_kmajor(smem_u32(bdo));
#pragma unroll
      for (int kk = 0; kk < HD / 16; ++kk) {
        const uint32_t aoff = (kk >> 2) * 16384 + (kk & 3) * 32, boff = (kk >> 2) * 8192 + (kk & 3) * 32;
        umma_bf16_w(tmem_base + Cfg::TM_ST, dK_k + (aoff >> 4), dQ_k + (boff >> 4), idesc_s, kk > 0 ? 1u : 0u);
      }
      umma_commit_w(s_full);
      if (ROLE == 1) umma_commit_w(&q_empty[st]);  // role V reads Q_i only here
      FB_TRACE(12);
      if (ROLE == 2) {
        mbar_wait(&do_full[0], it & 1);
        FB_TRACE(16);
        tc_fence_after();
#pragma unroll
        for (int kk = 0; kk < HD / 16; ++kk) {
          const uint32_t aoff = (kk >> 2) * 16384 + (kk & 3) * 32, boff = (kk >> 2) * 8192 + (kk & 3) * 32;
          umma_bf16_w(tmem_base + Cfg::TM_DPT, dV_k + (aoff >> 4), dDO_k + (boff >> 4), idesc_s, kk > 0 ? 1u : 0u);
        }
        umma_commit_w(dp_full);
        umma_commit_w(&do_empty[0]);  // nothing else in this role reads dO_i
        FB_TRACE(17);
      }
    };
    auto issue_g = [&](int it) {
      const int st = it % NS;
      mbar_wait(pds_full, it & 1);
      FB_TRACE(13);
      if (ROLE == 2) mbar_wait(dq_empty, (it & 1) ^ 1);
      FB_TRACE(14);
      tc_fence_after();
      if (ROLE == 1) {
        mbar_wait(&do_full[st], (it / NS) & 1);
        tc_fence_after();
        const uint64_t dDO_mn = umma_desc_mnmajor(smem_u32(sDO + st * Cfg::Q_BYTES), 64);
#pragma unroll
        for (int kk = 0; kk < 4; ++kk)  // dV += P^T dO, P^T from tensor memory
          umma_bf16_ts_w(tmem_base + Cfg::TM_ACC, tmem_base + Cfg::TM_P + kk * 8, dDO_mn + ((kk * 2048) >> 4), idesc_g, (it > 0 || kk > 0) ? 1u : 0u);
        umma_commit_w(&do_empty[st]);
        umma_commit_w(pds_empty);
      } else {
        const uint64_t dQ_mn = umma_desc_mnmajor(smem_u32(sQ + (it & 1) * Cfg::Q_BYTES), 64);
#pragma unroll
        for (int kk = 0; kk < 4; ++kk)  // dK += dS^T Q, dS^T from shared memory (K-major, 16 q columns = 32 B per k-step)
          umma_bf16_w(tmem_base + Cfg::TM_ACC, dDST_k + ((kk * 32) >> 4), dQ_mn + ((kk * 2048) >> 4), idesc_g, (it > 0 || kk > 0) ? 1u : 0u);
        umma_commit_w(&q_empty[0]);  // Q_i is free: its reload runs under the dQ^T MMAs
#pragma unroll
        for (int hh = 0; hh < 2; ++hh) {  // dQ^T = K^T dS, head dims hh*128 .. hh*128+127 as the M rows
          const uint64_t dK_mn = umma_desc_mnmajor(k_addr + hh * 2 * 16384, 128);
#pragma unroll
          for (int kk = 0; kk < 8; ++kk)
            umma_bf16_w(tmem_base + Cfg::TM_DQT + hh * 64, dK_mn + ((kk * 2048) >> 4), dDST_mn + ((kk * 2048) >> 4), idesc_q, kk > 0 ? 1u : 0u);
        }
        umma_commit_w(pds_empty);
        umma_commit_w(dq_full);
      }
      FB_TRACE(15);
    };
    if (n_it > 0) issue_s(0);
    for (int it = 0; it < n_it; ++it) {
      if (NS >= 2) {
        if (it + 1 < n_it) issue_s(it + 1);
        issue_g(it);
      } else {
        issue_g(it);
        if (it + 1 < n_it) issue_s(it + 1);
      }
    }
    __syncwarp();
  } else if (warp < 2 + 4 * FB_AGROUPS) {
    // ------------------------------------------------------------------ element-wise groups
    const int wg = (warp - 2) >> 2;
    const int qd = warp & 3;
    const int r = qd * 32 + lane;
    const int ct = threadIdx.x - 64;
    const int kv_idx = kv0 + r;
    const int cbase = wg * FB_ACOLS;
    const uint32_t lane_base = tmem_base + ((uint32_t)(qd * 32) << 16);
    auto fetch_stat = [&](int it) -> float {
      if (ct >= (ROLE == 2 ? 128 : 64) || it >= n_it) return 0.f;
      const int h = g * n_rep + giter(it) / n_i;
      const int qi = (i_start + giter(it) % n_i) * 64 + (ct & 63);
      if (qi >= p.T) return 0.f;
      return ((ct < 64) ? p.lse : p.delta)[((int64_t)b * p.Hq + h) * p.T + qi];
    };
    auto put_stat = [&](int it, float v) {
      if (ct < 64) sLse[(it & 1) * 64 + ct] = v * FB_LOG2E;
      else if (ct < 128) sDelta[(it & 1) * 64 + (ct - 64)] = v;
    };
    FB_TRACE_DECL(2, warp == 2);
    put_stat(0, fetch_stat(0));
    const float2 sc2 = make_float2(p.scale_log2, p.scale_log2);
    const float2 scale2 = make_float2(p.scale, p.scale);
    for (int it = 0; it < n_it; ++it) {
      const int st = it & 1;
      const int q0 = (i_start + giter(it) % n_i) * 64;
      const float stat_next = fetch_stat(it + 1);
      asm volatile("bar.sync 1, %0;" ::"n"(FB_ATHREADS) : "memory");
      FB_TRACE(20);
      mbar_wait(s_full, it & 1);
      FB_TRACE(21);
      tc_fence_after();
      uint32_t rs[FB_ACOLS], rd[FB_ACOLS];
      SPB_TMEM_LD_32(lane_base + (uint32_t)(Cfg::TM_ST + cbase), rs);
      tmem_ld_wait();
      if (ROLE == 1) {
        tc_fence_before();
        __syncwarp();
        if (lane == 0) mbar_arrive(sdp_empty);
      }

      const bool need_mask = (p.causal && q0 < kv0 + 128) || (q0 + 64 > p.T) || (kv0 + 128 > p.T);
      const float2* lse2 = reinterpret_cast<const float2*>(sLse + st * 64 + cbase);
      const float2* del2 = reinterpret_cast<const float2*>(sDelta + st * 64 + cbase);
      uint32_t pk[FB_ACOLS / 2];  // role V: P^T; role K: dS^T
      float2 pvv[FB_ACOLS / 2];   // role K: P^T in fp32 until dP^T has arrived
      auto probs = [&](auto masked) {
        constexpr bool MASKED = decltype(masked)::value;
#pragma unroll
        for (int i = 0; i < FB_ACOLS; i += 2) {
          const float2 l2 = lse2[i >> 1];
          const float2 x = __ffma2_rn(make_float2(__uint_as_float(rs[i]), __uint_as_float(rs[i + 1])), sc2, make_float2(-l2.x, -l2.y));
          float2 pv = make_float2(fb_exp2(x.x), fb_exp2(x.y));
          if (MASKED) {
            const int qi = q0 + cbase + i;
            const bool kv_dead = kv_idx >= p.T;
            if (kv_dead || qi >= p.T || (p.causal && kv_idx > qi)) pv.x = 0.f;
            if (kv_dead || qi + 1 >= p.T || (p.causal && kv_idx > qi + 1)) pv.y = 0.f;
          }
          if (ROLE == 1) pk[i >> 1] = pack_bf16(pv.x, pv.y);
          else pvv[i >> 1] = pv;
        }
      };
      if (need_mask) probs(std::true_type{});
      else probs(std::false_type{});
      FB_TRACE(26);
      if (ROLE == 2) {  // the exps above ran while the tensor pipe was still on dP^T
        mbar_wait(dp_full, it & 1);
        FB_TRACE(27);
        tc_fence_after();
        SPB_TMEM_LD_32(lane_base + (uint32_t)(Cfg::TM_DPT + cbase), rd);
        tmem_ld_wait();
        tc_fence_before();
        __syncwarp();
        if (lane == 0) mbar_arrive(sdp_empty);
#pragma unroll
        for (int i = 0; i < FB_ACOLS; i += 2) {
          const float2 d2 = del2[i >> 1];
          const float2 dd = __fadd2_rn(make_float2(__uint_as_float(rd[i]), __uint_as_float(rd[i + 1])), make_float2(-d2.x, -d2.y));
          const float2 ds = __fmul2_rn(__fmul2_rn(pvv[i >> 1], dd), scale2);
          pk[i >> 1] = pack_bf16(ds.x, ds.y);
        }
      }
      FB_TRACE(23);
      put_stat(it + 1, stat_next);
      mbar_wait(pds_empty, (it & 1) ^ 1);  // the MMA group of iteration it-1 has finished reading P^T / dS^T
      FB_TRACE(24);
      if (ROLE == 1) {
        SPB_TMEM_ST_16(lane_base + (uint32_t)(Cfg::TM_P + cbase / 2), pk);
        tmem_st_wait();
      } else {
        uint8_t* drow = sDST + r * 128;
#pragma unroll
        for (int c = 0; c < FB_ACOLS / 8; ++c) {
          const int slot = ((wg * (FB_ACOLS / 8) + c) ^ (r & 7)) << 4;
          *reinterpret_cast<uint4*>(drow + slot) = make_uint4(pk[c * 4], pk[c * 4 + 1], pk[c * 4 + 2], pk[c * 4 + 3]);
        }
        fence_proxy_async_smem();
      }
      tc_fence_before();
      __syncwarp();
      if (lane == 0) mbar_arrive(pds_full);
      FB_TRACE(25);
    }
    if (n_it > 0) {
      mbar_wait(pds_empty, (n_it - 1) & 1);  // the last commit covers every MMA issued before it
      tc_fence_after();
    }
  } else if (ROLE == 1) {
    // ------------------------------------------------------------------ role V: the first of the (otherwise idle) drain warps streams dO_i
    if (warp == 2 + 4 * FB_AGROUPS && lane == 0) {
      for (int it = 0; it < n_it; ++it) {
        const int st = it % NS;
        const int h = g * n_rep + giter(it) / n_i;
        const int row_q = b * p.T + (i_start + giter(it) % n_i) * 64;
        mbar_wait(&do_empty[st], ((it / NS) & 1) ^ 1);
        mbar_arrive_expect_tx(&do_full[st], Cfg::Q_BYTES);
#pragma unroll
        for (int c = 0; c < 4; ++c) tma_load_2d(sDO + st * Cfg::Q_BYTES + c * 8192, &tmDO, &do_full[st], h * HD + c * 64, row_q);
      }
    }
    __syncwarp();
  } else {
    // ------------------------------------------------------------------ drain group (role K): reduce both halves of dQ^T every iteration
    const int qd = warp & 3;
    const int r = qd * 32 + lane;  // head dim inside the half = TMEM lane of dQ^T
    const uint32_t lane_base = tmem_base + ((uint32_t)(qd * 32) << 16);
    const bool leader = (qd == 0 && lane == 0);
    FB_TRACE_DECL(3, qd == 0);
    for (int it = 0; it < n_it; ++it) {
      const int h = g * n_rep + giter(it) / n_i;
      const int q0 = (i_start + giter(it) % n_i) * 64;
      mbar_wait(dq_full, it & 1);
      FB_TRACE(30);
      tc_fence_after();
#pragma unroll 1
      for (int hh = 0; hh < 2; ++hh) {
        float dq[64];
#pragma unroll
        for (int c = 0; c < 2; ++c) {
          uint32_t raw[32];
          SPB_TMEM_LD_32(lane_base + (uint32_t)(Cfg::TM_DQT + hh * 64 + c * 32), raw);
          tmem_ld_wait();
#pragma unroll
          for (int i = 0; i < 32; ++i) dq[c * 32 + i] = __uint_as_float(raw[i]);
        }
        if (hh == 1) {  // both halves are in registers / on their way: the next iteration may overwrite dQ^T
          tc_fence_before();
          __syncwarp();
          if (lane == 0) mbar_arrive(dq_empty);
          FB_TRACE(31);
        }
#pragma unroll
        for (int rd = 0; rd < 2; ++rd) {  // 32 query rows at a time through the 16 KB staging tile
          if (leader && (it > 0 || hh > 0 || rd > 0)) tma_store_wait_read<0>();
          asm volatile("bar.sync 2, 128;" ::: "memory");
#pragma unroll
          for (int i = 0; i < 32; ++i) sDQ[i * 128 + r] = dq[rd * 32 + i];
          fence_proxy_async_smem();
          asm volatile("bar.sync 2, 128;" ::: "memory");
          if (leader) {
            tma_reduce_add_2d(&tmDQ, sDQ, h * HD + hh * 128, b * p.T + q0 + rd * 32);
            tma_store_commit();
          }
        }
      }
      FB_TRACE(33);
    }
    if (leader) tma_store_wait_read<0>();
  }

  // ---- epilogue: the role's accumulator (dV or dK); element-wise group wg owns head columns wg*128 .. wg*128+127 of kv row r.
  // The exchange buffer is rank 0's K | V (role V: K | first stage) region, 128 KB, free by now: float4 j of row r at
  // (j * 128 + r) * 16 B, so a warp's accesses are consecutive on both sides.
  const bool ew = warp >= 2 && warp < 2 + 4 * FB_AGROUPS;
  const int wg = (warp - 2) >> 2, qd = warp & 3;
  const int r = qd * 32 + lane;
  const int kv_idx = kv0 + r;
  const uint32_t lane_base = tmem_base + ((uint32_t)(qd * 32) << 16);
  // S > 1: ranks 1 .. S-1 ship their partial sums, one rank per round, into rank 0's shared memory; rank 0 adds each round into
  // its tensor-memory accumulator.  The barrier at the top of a round also says that rank 0 is done reading the previous one.
  for (int src = 1; src < S; ++src) {
    cluster_sync_all();  // round 1: every MMA of every CTA has completed (the element-wise warps arrive after their last pds_empty wait)
    if (rank == src && ew) {
      tc_fence_after();
      uint32_t remote;
      asm volatile("mapa.shared::cluster.u32 %0, %1, 0;" : "=r"(remote) : "r"(smem_u32(smem)));
#pragma unroll
      for (int c = 0; c < 4; ++c) {
        uint32_t o[32];
        if (n_it > 0) {
          SPB_TMEM_LD_32(lane_base + (uint32_t)(Cfg::TM_ACC + wg * 128 + c * 32), o);
          tmem_ld_wait();
        } else {
#pragma unroll
          for (int i = 0; i < 32; ++i) o[i] = 0u;
        }
#pragma unroll
        for (int i = 0; i < 32; i += 4) {
          const uint32_t j = (uint32_t)(wg * 32 + c * 8 + (i >> 2));
          asm volatile("st.shared::cluster.v4.b32 [%0], {%1, %2, %3, %4};" ::"r"(remote + (j * 128u + (uint32_t)r) * 16u), "r"(o[i]), "r"(o[i + 1]),
                       "r"(o[i + 2]), "r"(o[i + 3])
                       : "memory");
        }
      }
    }
    cluster_sync_all();  // release / acquire: the stores are visible to rank 0
    if (rank == 0 && ew) {
      tc_fence_after();
#pragma unroll
      for (int c = 0; c < 4; ++c) {
        uint32_t o[32];
        SPB_TMEM_LD_32(lane_base + (uint32_t)(Cfg::TM_ACC + wg * 128 + c * 32), o);
        tmem_ld_wait();
#pragma unroll
        for (int i = 0; i < 32; i += 4) {
          const int j = wg * 32 + c * 8 + (i >> 2);
          const float4 add = *reinterpret_cast<const float4*>(smem + ((size_t)j * 128 + r) * 16);
          o[i] = __float_as_uint(__uint_as_float(o[i]) + add.x);
          o[i + 1] = __float_as_uint(__uint_as_float(o[i + 1]) + add.y);
          o[i + 2] = __float_as_uint(__uint_as_float(o[i + 2]) + add.z);
          o[i + 3] = __float_as_uint(__uint_as_float(o[i + 3]) + add.w);
        }
        SPB_TMEM_ST_16(lane_base + (uint32_t)(Cfg::TM_ACC + wg * 128 + c * 32), o);
        SPB_TMEM_ST_16(lane_base + (uint32_t)(Cfg::TM_ACC + wg * 128 + c * 32 + 16), (o + 16));
      }
      tmem_st_wait();
      tc_fence_before();
    }
  }
  if (ew && rank == 0) {
    tc_fence_after();
    const bool valid = kv_idx < p.T;
    __nv_bfloat16* orow = (ROLE == 2 ? p.dk + ((int64_t)b * p.T + kv_idx) * p.lddk : p.dv + ((int64_t)b * p.T + kv_idx) * p.lddv) + (int64_t)g * HD + wg * 128;
#pragma unroll
    for (int c = 0; c < 4; ++c) {
      uint32_t o[32];
      SPB_TMEM_LD_32(lane_base + (uint32_t)(Cfg::TM_ACC + wg * 128 + c * 32), o);
      tmem_ld_wait();
      if (valid) {
        if (ROLE == 2 && p.rope != nullptr) {  // dK: transpose of the 2x2 rotary block, table row = position kv_idx
          const float4* tab = p.rope + (int64_t)kv_idx * (HD / 2) + wg * 64 + c * 16;
#pragma unroll
          for (int i = 0; i < 32; i += 2) {
            const float4 m = tab[i >> 1];
            const float xe = __uint_as_float(o[i]), xo = __uint_as_float(o[i + 1]);
            o[i] = __float_as_uint(m.x * xe + m.z * xo);
            o[i + 1] = __float_as_uint(m.y * xe + m.w * xo);
          }
        }
#pragma unroll
        for (int i = 0; i < 32; i += 8) {
          uint4 v;
          v.x = pack_bf16(__uint_as_float(o[i]), __uint_as_float(o[i + 1]));
          v.y = pack_bf16(__uint_as_float(o[i + 2]), __uint_as_float(o[i + 3]));
          v.z = pack_bf16(__uint_as_float(o[i + 4]), __uint_as_float(o[i + 5]));
          v.w = pack_bf16(__uint_as_float(o[i + 6]), __uint_as_float(o[i + 7]));
          *reinterpret_cast<uint4*>(orow + c * 32 + i) = v;
        }
      }
    }
  }

  tc_fence_before();
  __syncthreads();
  if (warp == 1) tmem_dealloc(tmem_base, 512);
}

__global__ void __launch_bounds__(FB_THREADS, 1)
    flash_bwd256_roles_kernel(const __grid_constant__ CUtensorMap tmK, const __grid_constant__ CUtensorMap tmV,
                              const __grid_constant__ CUtensorMap tmQ, const __grid_constant__ CUtensorMap tmDO,
                              const __grid_constant__ CUtensorMap tmDQ, const FaBwdArgs p, const int S) {
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = smem_raw + ((1024u - (smem_u32(smem_raw) & 1023u)) & 1023u);
  // blockIdx.x = (kv head * 2 + role bit) * S + rank; role K first: it is the longer of the two
  const int rank = (int)(blockIdx.x % (unsigned)S);
  const int rg = (int)(blockIdx.x / (unsigned)S);
#ifdef SPB_FB_TRACE_ROLE_V  // development: the traced CTA (0,0,0) becomes a role-V one
  if ((rg & 1) == 1) fb256_body<2>(tmK, tmV, tmQ, tmDO, tmDQ, p, smem, rg >> 1, rank, S);
  else fb256_body<1>(tmK, tmV, tmQ, tmDO, tmDQ, p, smem, rg >> 1, rank, S);
#else
  if ((rg & 1) == 0) fb256_body<2>(tmK, tmV, tmQ, tmDO, tmDQ, p, smem, rg >> 1, rank, S);
  else fb256_body<1>(tmK, tmV, tmQ, tmDO, tmDQ, p, smem, rg >> 1, rank, S);
#endif
}

bool flash_bwd_tc_supported(int dtype, int hd) { return dtype == SPB_BF16 && (hd == 64 || hd == 128 || hd == 256); }

template <int HD>
static int flash_bwd_tc_launch(int B, int Tn, int Hq, int Hkv, int causal, float scale, const void* q, int64_t ldq, const void* k,
                               int64_t ldk, const void* v, int64_t ldv, const void* o, int64_t ldo, const void* d_o, int64_t lddo,
                               const float* lse, float* delta, void* dq, int64_t lddq, void* dk, int64_t lddk, void* dv, int64_t lddv,
                               float* dq_acc, cudaStream_t st, const float* rope, int phase) {
  using Cfg = FaBwdCfg<HD>;
  const int64_t N = (int64_t)B * Tn;
  const int64_t ld_acc = (int64_t)Hq * HD;
  // phase 1 (FB_PHASE_PRE): the scratch the main kernel needs - dq accumulator zeroed, delta = rowsum(dO * O).  It only
  // depends on o and d_o, so the block backward runs it on a helper stream beside the dwo GEMM.
  if (phase & FB_PHASE_PRE) {
    SPB_CUDA(cudaMemsetAsync(dq_acc, 0, (size_t)N * ld_acc * sizeof(float), st));
    const int64_t lanes = N * Hq * (HD / 8);
    flash_bwd_prep_kernel<HD><<<(unsigned)((lanes + 255) / 256), 256, 0, st>>>((const __nv_bfloat16*)o, ldo, (const __nv_bfloat16*)d_o, lddo,
                                                                          delta, B, Tn, Hq);
    SPB_LAUNCH_CHECK();
  }
  if (!(phase & FB_PHASE_MAIN)) return 0;
  CUtensorMap tmK, tmV, tmQ, tmDO;
  int rc;
  if ((rc = make_tmap_2d(&tmK, k, 2, (uint64_t)N, (uint64_t)Hkv * HD, (uint64_t)ldk, 128, 64, true, false))) return rc;
  if ((rc = make_tmap_2d(&tmV, v, 2, (uint64_t)N, (uint64_t)Hkv * HD, (uint64_t)ldv, 128, 64, true, false))) return rc;
  if ((rc = make_tmap_2d(&tmQ, q, 2, (uint64_t)N, (uint64_t)Hq * HD, (uint64_t)ldq, 64, 64, true, false))) return rc;
  if ((rc = make_tmap_2d(&tmDO, d_o, 2, (uint64_t)N, (uint64_t)Hq * HD, (uint64_t)lddo, 64, 64, true, false))) return rc;
  CUtensorMap tmDQ;
  if ((rc = make_tmap_2d(&tmDQ, dq_acc, 4, (uint64_t)N, (uint64_t)Hq * HD, (uint64_t)ld_acc, Cfg::DQS_ROWS, Cfg::HDZ, false, true))) return rc;
  FaBwdArgs a;
  a.lse = lse;
  a.delta = delta;
  a.dq_acc = dq_acc;
  a.ld_acc = ld_acc;
  a.dk = (__nv_bfloat16*)dk;
  a.lddk = lddk;
  a.dv = (__nv_bfloat16*)dv;
  a.lddv = lddv;
  a.T = Tn;
  a.Hq = Hq;
  a.Hkv = Hkv;
  a.causal = causal;
  a.scale = scale;
  a.scale_log2 = scale * FB_LOG2E;
  a.rope = (const float4*)rope;
#ifdef SPB_FB_ABLATE
  {
    const char* ab = getenv("SPB_FB_ABLATE");
    a.ablate = ab ? atoi(ab) : 0;
  }
#endif
  dim3 grid(Hkv * Cfg::HS, B, (Tn + 127) / 128);
  static const bool roles_off = [] { const char* e = getenv("SPB_BWD256_ROLES"); return e && e[0] == '0'; }();
  if (HD == 256 && !roles_off) {  // two roles per KV tile (dV | dK + dQ) instead of two column halves; same dQ staging box
    constexpr int SM = Fb256Cfg<2>::SMEM_BYTES > Fb256Cfg<1>::SMEM_BYTES ? Fb256Cfg<2>::SMEM_BYTES : Fb256Cfg<1>::SMEM_BYTES;
    SPB_CUDA(cudaFuncSetAttribute(flash_bwd256_roles_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, SM));
    // few KV tiles (one CTA per role and tile would leave the machine to the heaviest role-K CTA): each role of a tile is a
    // cluster of two or four CTAs that take alternate iterations and merge their partial dV / dK through distributed shared memory
    static const int force_split = [] { const char* e = getenv("SPB_BWD256_SPLIT"); return e ? atoi(e) : 0; }();
    const int64_t role_ctas = (int64_t)grid.x * grid.y * grid.z;
    const int S = force_split ? (force_split == 4 ? 4 : (force_split == 2 ? 2 : 1))
                              : (role_ctas < 4 * (int64_t)sm_count() ? 2 : 1);  // clusters of four measured slower (C4: 1.59 vs 1.33 ms)
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = dim3(grid.x * S, grid.y, grid.z);
    cfg.blockDim = dim3(FB_THREADS);
    cfg.dynamicSmemBytes = SM;
    cfg.stream = st;
    cudaLaunchAttribute attr[1];
    attr[0].id = cudaLaunchAttributeClusterDimension;
    attr[0].val.clusterDim.x = S;
    attr[0].val.clusterDim.y = 1;
    attr[0].val.clusterDim.z = 1;
    cfg.attrs = attr;
    cfg.numAttrs = 1;
    SPB_CUDA(cudaLaunchKernelEx(&cfg, flash_bwd256_roles_kernel, tmK, tmV, tmQ, tmDO, tmDQ, a, S));
  } else {
    SPB_CUDA(cudaFuncSetAttribute(flash_bwd_tc_kernel<HD>, cudaFuncAttributeMaxDynamicSharedMemorySize, Cfg::SMEM_BYTES));
    flash_bwd_tc_kernel<HD><<<grid, FB_THREADS, Cfg::SMEM_BYTES, st>>>(tmK, tmV, tmQ, tmDO, tmDQ, a);
  }
  SPB_LAUNCH_CHECK();
  {
    constexpr int HPT = 4;
    const int64_t total = N * (HD / 8);
    dim3 cgrid((unsigned)((total + 255) / 256), (unsigned)((Hq + HPT - 1) / HPT));
    flash_bwd_dq_convert_kernel<HPT><<<cgrid, 256, 0, st>>>(dq_acc, ld_acc, (__nv_bfloat16*)dq, lddq, N, Hq, (const float4*)rope, Tn, HD);
    SPB_LAUNCH_CHECK();
  }
  return 0;
}

int flash_bwd_tc(int B, int Tn, int Hq, int Hkv, int hd, int causal, float scale, const void* q, int64_t ldq, const void* k,
                 int64_t ldk, const void* v, int64_t ldv, const void* o, int64_t ldo, const void* d_o, int64_t lddo,
                 const float* lse, float* delta, void* dq, int64_t lddq, void* dk, int64_t lddk, void* dv, int64_t lddv,
                 float* dq_acc, cudaStream_t st, const float* rope, int phase) {
  const void* ptrs[8] = {q, k, v, o, d_o, dq, dk, dv};
  const int64_t lds[8] = {ldq, ldk, ldv, ldo, lddo, lddq, lddk, lddv};
  for (int i = 0; i < 8; ++i)
    SPB_CHECK_ARG(((uintptr_t)ptrs[i] % 16 == 0) && (lds[i] % 8 == 0), SPB_E_UNSUPPORTED,
                  "flash backward (tcgen05): operand %d must be 16-byte aligned with a leading dimension multiple of 8", i);
  SPB_CHECK_ARG((uintptr_t)dq_acc % 16 == 0, SPB_E_INVALID, "flash backward (tcgen05): dq_acc must be 16-byte aligned");
  if (hd == 128)
    return flash_bwd_tc_launch<128>(B, Tn, Hq, Hkv, causal, scale, q, ldq, k, ldk, v, ldv, o, ldo, d_o, lddo, lse, delta, dq, lddq, dk, lddk,
                                    dv, lddv, dq_acc, st, rope, phase);
  if (hd == 256)
    return flash_bwd_tc_launch<256>(B, Tn, Hq, Hkv, causal, scale, q, ldq, k, ldk, v, ldv, o, ldo, d_o, lddo, lse, delta, dq, lddq, dk, lddk,
                                    dv, lddv, dq_acc, st, rope, phase);
  if (hd == 64)
    return flash_bwd_tc_launch<64>(B, Tn, Hq, Hkv, causal, scale, q, ldq, k, ldk, v, ldv, o, ldo, d_o, lddo, lse, delta, dq, lddq, dk, lddk, dv,
                                   lddv, dq_acc, st, rope, phase);
  SPB_CHECK_ARG(false, SPB_E_UNSUPPORTED, "flash backward (tcgen05): hd %d", hd);
}

}  // namespace spb

#ifdef SPB_FB_TRACE
extern "C" int spb_debug_fb_trace_read(unsigned long long* dst, int n) {
  const int total = spb::FB_TRACE_ROLES * spb::FB_TRACE_EVENTS;
  if (n > total) n = total;
  return (int)cudaMemcpyFromSymbol(dst, spb::g_fb_trace, sizeof(unsigned long long) * n);
}
extern "C" int spb_debug_fb_trace_clear() {
  void* p = nullptr;
  cudaError_t e = cudaGetSymbolAddress(&p, spb::g_fb_trace);
  if (e != cudaSuccess) return (int)e;
  return (int)cudaMemset(p, 0, sizeof(unsigned long long) * spb::FB_TRACE_ROLES * spb::FB_TRACE_EVENTS);
}
#endif
