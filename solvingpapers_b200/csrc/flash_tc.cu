// flash_tc.cu - K3/K4: flash-style causal GQA attention on tcgen05 (bf16, hd in {64,128}); replaces the materialised
// scores of llama3:820-828 (and gpt:344-353, ViT's nn.MultiheadAttention).  Scores never leave the SM.
//
// Forward (one CTA = 128 query rows of one (batch, q head)):
//   warp 0   TMA producer: Q tile once, then K / V tiles of 128 rows, double buffered (mbarrier full/empty rings)
//   warp 1   MMA issuer (one thread): S = Q K^T (M128 N128 K=hd) into one of two TMEM S buffers, O += P V (M128 N=hd
//            K128) into the TMEM O accumulator; QK^T of tile j+1 is issued BEFORE PV of tile j so the tensor pipe works
//            while the softmax warps are busy with tile j
//   warps 2-5 softmax: thread = one query row (TMEM lane): tcgen05.ld the S row, online softmax in registers (exp2 with
//            the 1/sqrt(hd)*log2(e) factor folded in), lazy rescaling of O (only when the running max grows by > 2^8,
//            decided per warp), P written as bf16 into shared memory in the 128B-swizzled K-major UMMA layout
//   causal tile skipping: KV tiles above the diagonal are never loaded; the diagonal tile and the ragged last tile are
//   masked by index (-1e9 additive / -1e4 replacing / -inf masks of the notebooks all give exp()==0 exactly).
//   GQA: kv head = q head / (Hq/Hkv) is just the TMA column coordinate (llama3:626-627 repeat_kv never materialised).
// Algorithmic FLOPs: 4*B*Hq*T^2*hd*m forward (m = (T+1)/2T causal).
#include "common.cuh"
#include "kernels.h"

namespace spb {

constexpr int FA_THREADS = 192;
constexpr float LOG2E = 1.4426950408889634f;
constexpr float LN2 = 0.6931471805599453f;

__device__ __forceinline__ float fast_exp2(float x) {
  float y;
  asm volatile("ex2.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x));
  return y;
}

template <int HD>
struct FaFwdCfg {
  static constexpr int CH = HD / 64;                // 64-column chunks per Q/K/V tile
  static constexpr int TILE_BYTES = 128 * HD * 2;   // 128 rows x HD bf16
  static constexpr int CHUNK_BYTES = 128 * 128;     // 128 rows x 128 B
  static constexpr int P_BYTES = 128 * 128 * 2;
  static constexpr int OFF_Q = 0;
  static constexpr int OFF_K = OFF_Q + TILE_BYTES;
  static constexpr int OFF_V = OFF_K + 2 * TILE_BYTES;
  static constexpr int OFF_P = OFF_V + 2 * TILE_BYTES;
  static constexpr int OFF_BAR = OFF_P + 2 * P_BYTES;
  static constexpr int SMEM_BYTES = OFF_BAR + 256 + 1024;
  static constexpr int TMEM_COLS = 512;  // S0 [0,128) S1 [128,256) O [256, 256+HD)
};

struct FaFwdArgs {
  __nv_bfloat16* o;
  int64_t ldo;
  float* lse;
  int T, Hq, Hkv, causal;
  float scale_log2;  // scale * log2(e)
};

template <int HD>
__global__ void __launch_bounds__(FA_THREADS, 1)
    flash_fwd_tc_kernel(const __grid_constant__ CUtensorMap tmQ, const __grid_constant__ CUtensorMap tmK,
                        const __grid_constant__ CUtensorMap tmV, const FaFwdArgs p) {
  using Cfg = FaFwdCfg<HD>;
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = smem_raw + ((1024u - (smem_u32(smem_raw) & 1023u)) & 1023u);
  uint8_t* sQ = smem + Cfg::OFF_Q;
  uint8_t* sK = smem + Cfg::OFF_K;
  uint8_t* sV = smem + Cfg::OFF_V;
  uint8_t* sP = smem + Cfg::OFF_P;
  uint64_t* bars = reinterpret_cast<uint64_t*>(smem + Cfg::OFF_BAR);
  uint64_t* q_full = bars;            // [1]
  uint64_t* k_full = bars + 1;        // [2]
  uint64_t* k_empty = bars + 3;       // [2]
  uint64_t* v_full = bars + 5;        // [2]
  uint64_t* v_empty = bars + 7;       // [2]
  uint64_t* s_full = bars + 9;        // [2]
  uint64_t* s_empty = bars + 11;      // [2]
  uint64_t* p_full = bars + 13;       // [2]
  uint64_t* pv_done = bars + 15;      // [2]
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(bars + 17);

  const int warp = __shfl_sync(0xffffffffu, (int)(threadIdx.x >> 5), 0);
  const int lane = threadIdx.x & 31;

  const int n_qt = (int)gridDim.x;
  const int qt = p.causal ? (n_qt - 1 - (int)blockIdx.x) : (int)blockIdx.x;  // heavy (late) causal tiles first
  const int h = blockIdx.y, b = blockIdx.z;
  const int kvh = h / (p.Hq / p.Hkv);
  const int q0 = qt * 128;
  const int n_kv_all = (p.T + 127) / 128;
  const int n_kv = p.causal ? min(n_kv_all, qt + 1) : n_kv_all;

  if (warp == 0 && lane == 0) {
    prefetch_tmap(&tmQ);
    prefetch_tmap(&tmK);
    prefetch_tmap(&tmV);
  }
  if (warp == 1) {
    if (lane == 0) {
      mbar_init(q_full, 1);
      for (int i = 0; i < 2; ++i) {
        mbar_init(&k_full[i], 1);
        mbar_init(&k_empty[i], 1);
        mbar_init(&v_full[i], 1);
        mbar_init(&v_empty[i], 1);
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
  const uint32_t tmem_O = tmem_base + 256;

  if (warp == 0) {
    // ------------------------------------------------------------------ TMA producer
    if (lane == 0) {
      const int row_q = b * p.T + q0;
      mbar_arrive_expect_tx(q_full, Cfg::TILE_BYTES);
#pragma unroll
      for (int c = 0; c < Cfg::CH; ++c) tma_load_2d(sQ + c * Cfg::CHUNK_BYTES, &tmQ, q_full, h * HD + c * 64, row_q);
      for (int j = 0; j < n_kv; ++j) {
        const int bf = j & 1;
        const uint32_t u = (j >> 1) & 1;
        const int row_kv = b * p.T + j * 128;
        mbar_wait(&k_empty[bf], u ^ 1);
        mbar_arrive_expect_tx(&k_full[bf], Cfg::TILE_BYTES);
#pragma unroll
        for (int c = 0; c < Cfg::CH; ++c)
          tma_load_2d(sK + bf * Cfg::TILE_BYTES + c * Cfg::CHUNK_BYTES, &tmK, &k_full[bf], kvh * HD + c * 64, row_kv);
        mbar_wait(&v_empty[bf], u ^ 1);
        mbar_arrive_expect_tx(&v_full[bf], Cfg::TILE_BYTES);
#pragma unroll
        for (int c = 0; c < Cfg::CH; ++c)
          tma_load_2d(sV + bf * Cfg::TILE_BYTES + c * Cfg::CHUNK_BYTES, &tmV, &v_full[bf], kvh * HD + c * 64, row_kv);
      }
    }
    __syncwarp();
  } else if (warp == 1) {
    // ------------------------------------------------------------------ MMA issuer
    if (lane == 0) {
      const uint32_t idesc_qk = umma_idesc_bf16(128, 128, 0, 0);  // A = Q (K-major), B = K (K-major)
      const uint32_t idesc_pv = umma_idesc_bf16(128, HD, 0, 1);   // A = P (K-major), B = V (MN-major: hd contiguous)
      const uint32_t q_addr = smem_u32(sQ);
      mbar_wait(q_full, 0);
      auto issue_qk = [&](int j) {
        const int bf = j & 1;
        const uint32_t u = (j >> 1) & 1;
        mbar_wait(&k_full[bf], u);
        mbar_wait(&s_empty[bf], u ^ 1);
        tc_fence_after();
        const uint32_t k_addr = smem_u32(sK + bf * Cfg::TILE_BYTES);
#pragma unroll
        for (int kk = 0; kk < HD / 16; ++kk) {
          const uint32_t off = (kk >> 2) * Cfg::CHUNK_BYTES + (kk & 3) * 32;
          umma_bf16(tmem_base + bf * 128, umma_desc_kmajor(q_addr + off), umma_desc_kmajor(k_addr + off), idesc_qk, kk > 0 ? 1u : 0u);
        }
        umma_commit(&k_empty[bf]);
        umma_commit(&s_full[bf]);
      };
      auto issue_pv = [&](int j) {
        const int bf = j & 1;
        const uint32_t u = (j >> 1) & 1;
        mbar_wait(&v_full[bf], u);
        mbar_wait(&p_full[bf], u);
        tc_fence_after();
        const uint32_t p_addr = smem_u32(sP + bf * Cfg::P_BYTES);
        const uint32_t v_addr = smem_u32(sV + bf * Cfg::TILE_BYTES);
#pragma unroll
        for (int kk = 0; kk < 8; ++kk) {
          const uint32_t aoff = (kk >> 2) * Cfg::CHUNK_BYTES + (kk & 3) * 32;
          umma_bf16(tmem_O, umma_desc_kmajor(p_addr + aoff), umma_desc_mnmajor(v_addr + kk * 2048, 128), idesc_pv,
                    (j > 0 || kk > 0) ? 1u : 0u);
        }
        umma_commit(&v_empty[bf]);
        umma_commit(&pv_done[bf]);
      };
      issue_qk(0);
      for (int j = 0; j < n_kv; ++j) {
        if (j + 1 < n_kv) issue_qk(j + 1);
        issue_pv(j);
      }
    }
    __syncwarp();
  } else {
    // ------------------------------------------------------------------ softmax + epilogue (warps 2..5 -> lane quarters 2,3,0,1)
    const int qd = warp & 3;
    const int r = qd * 32 + lane;      // row inside the tile = TMEM lane
    const int q_idx = q0 + r;
    const uint32_t lane_base = tmem_base + ((uint32_t)(qd * 32) << 16);
    float m = -INFINITY, l = 0.f;
    float s[128];
    for (int j = 0; j < n_kv; ++j) {
      const int bf = j & 1;
      const uint32_t u = (j >> 1) & 1;
      mbar_wait(&s_full[bf], u);
      tc_fence_after();
#pragma unroll
      for (int c = 0; c < 4; ++c) {
        uint32_t raw[32];
        SPB_TMEM_LD_32(lane_base + (uint32_t)(bf * 128 + c * 32), raw);
        tmem_ld_wait();
#pragma unroll
        for (int i = 0; i < 32; ++i) s[c * 32 + i] = __uint_as_float(raw[i]);
      }
      tc_fence_before();
      __syncwarp();
      if (lane == 0) mbar_arrive(&s_empty[bf]);  // S buffer may be overwritten by QK^T of tile j+2

      const int kv0 = j * 128;
      const bool need_mask = (p.causal && j == qt) || (kv0 + 128 > p.T);
      float tmax = -INFINITY;
      if (need_mask) {
#pragma unroll
        for (int i = 0; i < 128; ++i) {
          const int kv = kv0 + i;
          const bool dead = (p.causal && kv > q_idx) || (kv >= p.T);
          s[i] = dead ? -INFINITY : s[i] * p.scale_log2;
          tmax = fmaxf(tmax, s[i]);
        }
      } else {
#pragma unroll
        for (int i = 0; i < 128; ++i) {
          s[i] *= p.scale_log2;
          tmax = fmaxf(tmax, s[i]);
        }
      }
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
          mbar_wait(&pv_done[(j - 1) & 1], ((j - 1) >> 1) & 1);  // O must hold every PV issued so far
          tc_fence_after();
#pragma unroll
          for (int c = 0; c < HD / 32; ++c) {
            uint32_t o[32];
            SPB_TMEM_LD_32(lane_base + (uint32_t)(256 + c * 32), o);
            tmem_ld_wait();
#pragma unroll
            for (int i = 0; i < 32; ++i) o[i] = __float_as_uint(__uint_as_float(o[i]) * alpha);
            SPB_TMEM_ST_32(lane_base + (uint32_t)(256 + c * 32), o);
          }
          tmem_st_wait();
        }
      }
      float sum = 0.f;
#pragma unroll
      for (int i = 0; i < 128; ++i) {
        s[i] = fast_exp2(s[i] - m);
        sum += s[i];
      }
      l += sum;
      if (j >= 2) mbar_wait(&pv_done[bf], ((j - 2) >> 1) & 1);  // PV(j-2) has finished reading this P buffer
      uint8_t* prow = sP + bf * Cfg::P_BYTES + r * 128;
#pragma unroll
      for (int t = 0; t < 2; ++t) {
#pragma unroll
        for (int c = 0; c < 8; ++c) {
          const float* f = &s[t * 64 + c * 8];
          uint4 v;
          v.x = pack_bf16(f[0], f[1]);
          v.y = pack_bf16(f[2], f[3]);
          v.z = pack_bf16(f[4], f[5]);
          v.w = pack_bf16(f[6], f[7]);
          *reinterpret_cast<uint4*>(prow + t * Cfg::CHUNK_BYTES + ((c ^ (r & 7)) << 4)) = v;
        }
      }
      fence_proxy_async_smem();  // generic-proxy smem writes -> visible to the tensor core (async proxy)
      tc_fence_before();
      __syncwarp();
      if (lane == 0) mbar_arrive(&p_full[bf]);
    }
    // epilogue: O / l -> bf16 -> global; lse = m*ln2 + ln(l)
    mbar_wait(&pv_done[(n_kv - 1) & 1], ((n_kv - 1) >> 1) & 1);
    tc_fence_after();
    const float inv = 1.0f / l;
    const bool valid = q_idx < p.T;
    __nv_bfloat16* orow = p.o + ((int64_t)b * p.T + q_idx) * p.ldo + (int64_t)h * HD;
#pragma unroll
    for (int c = 0; c < HD / 32; ++c) {
      uint32_t o[32];
      SPB_TMEM_LD_32(lane_base + (uint32_t)(256 + c * 32), o);
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

  tc_fence_before();
  __syncthreads();
  if (warp == 1) tmem_dealloc(tmem_base, Cfg::TMEM_COLS);
}

bool flash_tc_supported(int dtype, int hd) { return dtype == SPB_BF16 && (hd == 64 || hd == 128); }

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
  dim3 grid((Tn + 127) / 128, Hq, B);
  flash_fwd_tc_kernel<HD><<<grid, FA_THREADS, Cfg::SMEM_BYTES, st>>>(tmQ, tmK, tmV, a);
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
  SPB_CHECK_ARG(false, SPB_E_UNSUPPORTED, "flash forward (tcgen05): hd %d", hd);
}

}  // namespace spb
