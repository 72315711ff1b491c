// attn_ffma.cu - fp32 attention forward / backward on the CUDA cores, register-tiled (fp32 parity mode: the notebooks' own
// default dtype, llama3:349-358 / gpt:293-302; BASELINE configs C1 and C2-literal).  Replaces the warp-per-row kernels of
// attn_simt.cu for hd in {64, 128, 256}: those re-read K and V from L2 for every query row (1-3 TFLOP/s); here a CTA of 256
// threads (16 x 16) owns a TS x TS score tile (TS = 64; 32 at hd 256), every thread an MR x MR micro-tile (MR = TS/16), and the
// three kinds of products - S = Q K^T, O += P V and their backward counterparts - run from shared-memory tiles with 128-bit
// loads at 8 FMAs per shared-memory load.  The tolerance of this mode is 1e-4 (max|a-b|/max|b|): plain fp32 FMAs, exp2f-free
// __expf, deterministic (no atomics: dQ and dK/dV come from two kernels, as in attn_simt.cu).
//   forward   : CTA = TS query rows of one (batch, q head); loop over the KV tiles it can see: online softmax, O += P V
//   backward 1: CTA = TS query rows; per KV tile S, dP -> dS; dQ += dS K
//   backward 2: CTA = TS key rows of one (batch, kv head); loops over the q heads of the group (repeat_kv's VJP) and the Q
//               tiles that see it; per tile S, dP -> P^T, dS^T; dV += P^T dO, dK += dS^T Q
#include "common.cuh"
#include "kernels.h"

namespace spb {

constexpr int AF_THREADS = 256;

template <int HD>
struct AfCfg {
  static constexpr int TS = (HD <= 128) ? 64 : 32;  // score tile; hd 256: 32, so that five [TS, HD] fp32 tiles fit shared memory
  static constexpr int MR = TS / 16;                // micro-tile edge
  static constexpr int LDA = HD + 4;                // row stride of "A-type" tiles (rows read with 128-bit loads, broadcast per row)
  static constexpr int LDS_ = TS + 4;               // row stride of the score / probability tiles
  static constexpr int NC = HD / 16;                // output columns per thread in the [TS, HD] products
};

// ---- tile loaders (all 256 threads) ------------------------------------------------------------------------------------
// natural: dst[r][c] = src[(row0 + r) * ld + c], rows past `rows_valid` are zero
template <int HD, int LD>
__device__ __forceinline__ void af_load_nat(float* dst, const float* src, int64_t ld, int rows_valid, int TS) {
  for (int i = threadIdx.x; i < TS * (HD / 4); i += AF_THREADS) {
    const int r = i / (HD / 4), c = (i % (HD / 4)) * 4;
    float4 v = make_float4(0.f, 0.f, 0.f, 0.f);
    if (r < rows_valid) v = __ldg(reinterpret_cast<const float4*>(src + (int64_t)r * ld + c));
    *reinterpret_cast<float4*>(dst + r * LD + c) = v;
  }
}
// transposed: dst[c][r] = src[(row0 + r) * ld + c]; consecutive threads take consecutive rows: conflict-free stores
template <int HD, int TS>
__device__ __forceinline__ void af_load_tr(float* dst, const float* src, int64_t ld, int rows_valid) {
  for (int i = threadIdx.x; i < TS * (HD / 4); i += AF_THREADS) {
    const int r = i % TS, c = (i / TS) * 4;
    float4 v = make_float4(0.f, 0.f, 0.f, 0.f);
    if (r < rows_valid) v = __ldg(reinterpret_cast<const float4*>(src + (int64_t)r * ld + c));
    dst[(c + 0) * TS + r] = v.x;
    dst[(c + 1) * TS + r] = v.y;
    dst[(c + 2) * TS + r] = v.z;
    dst[(c + 3) * TS + r] = v.w;
  }
}

// ---- micro-GEMMs ---------------------------------------------------------------------------------------------------------
// c[i][j] = sum_d A[(ty*MR + i)][d] * Bt[d][tx*MR + j]      (A natural with row stride LDA, Bt transposed [KD][TS])
template <int MR, int KD, int LDA, int TS>
__device__ __forceinline__ void af_mm_nt(const float* __restrict__ A, const float* __restrict__ Bt, int ty, int tx, float (&c)[MR][MR]) {
#pragma unroll
  for (int i = 0; i < MR; ++i)
#pragma unroll
    for (int j = 0; j < MR; ++j) c[i][j] = 0.f;
#pragma unroll 4
  for (int d0 = 0; d0 < KD; d0 += 4) {
    float4 a[MR];
#pragma unroll
    for (int i = 0; i < MR; ++i) a[i] = *reinterpret_cast<const float4*>(A + (ty * MR + i) * LDA + d0);
#pragma unroll
    for (int dd = 0; dd < 4; ++dd) {
      float b[MR];
      if constexpr (MR == 4) {
        const float4 bv = *reinterpret_cast<const float4*>(Bt + (d0 + dd) * TS + tx * 4);
        b[0] = bv.x; b[1] = bv.y; b[2] = bv.z; b[3] = bv.w;
      } else {
        const float2 bv = *reinterpret_cast<const float2*>(Bt + (d0 + dd) * TS + tx * 2);
        b[0] = bv.x; b[1] = bv.y;
      }
#pragma unroll
      for (int i = 0; i < MR; ++i) {
        const float av = dd == 0 ? a[i].x : (dd == 1 ? a[i].y : (dd == 2 ? a[i].z : a[i].w));
#pragma unroll
        for (int j = 0; j < MR; ++j) c[i][j] = fmaf(av, b[j], c[i][j]);
      }
    }
  }
}
// acc[i][4*g + e] += sum_k P[(ty*MR + i)][k] * B[k][tx*4 + 64*g + e]     (P [TS][LDP], B natural [TS][LDB], HD = 64*NG columns)
template <int MR, int TS, int LDP, int LDB, int NC>
__device__ __forceinline__ void af_mm_nn(const float* __restrict__ P, const float* __restrict__ B, int ty, int tx, float (&acc)[MR][NC]) {
#pragma unroll 2
  for (int k0 = 0; k0 < TS; k0 += 4) {
    float4 p[MR];
#pragma unroll
    for (int i = 0; i < MR; ++i) p[i] = *reinterpret_cast<const float4*>(P + (ty * MR + i) * LDP + k0);
#pragma unroll
    for (int kk = 0; kk < 4; ++kk) {
#pragma unroll
      for (int g = 0; g < NC / 4; ++g) {
        const float4 b = *reinterpret_cast<const float4*>(B + (k0 + kk) * LDB + tx * 4 + 64 * g);
#pragma unroll
        for (int i = 0; i < MR; ++i) {
          const float pv = kk == 0 ? p[i].x : (kk == 1 ? p[i].y : (kk == 2 ? p[i].z : p[i].w));
          acc[i][4 * g + 0] = fmaf(pv, b.x, acc[i][4 * g + 0]);
          acc[i][4 * g + 1] = fmaf(pv, b.y, acc[i][4 * g + 1]);
          acc[i][4 * g + 2] = fmaf(pv, b.z, acc[i][4 * g + 2]);
          acc[i][4 * g + 3] = fmaf(pv, b.w, acc[i][4 * g + 3]);
        }
      }
    }
  }
}
// reduce over the 16 threads (consecutive lanes) that share a row of the score tile
__device__ __forceinline__ float af_row_max(float v) {
#pragma unroll
  for (int o = 8; o > 0; o >>= 1) v = fmaxf(v, __shfl_xor_sync(0xffffffffu, v, o));
  return v;
}
__device__ __forceinline__ float af_row_sum(float v) {
#pragma unroll
  for (int o = 8; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
  return v;
}

struct AfArgs {
  const float *q, *k, *v, *o, *d_o;
  float *out, *lse, *delta, *dq, *dk, *dv;
  int64_t ldq, ldk, ldv, ldo, lddo, lddq, lddk, lddv;
  int T, Hq, Hkv, causal;
  float scale;
};

// ------------------------------------------------------------------------------------------------ forward
template <int HD>
__global__ void __launch_bounds__(AF_THREADS) attn_ffma_fwd_kernel(const AfArgs p) {
  using C = AfCfg<HD>;
  constexpr int TS = C::TS, MR = C::MR, LDA = C::LDA, LDP = C::LDS_, NC = C::NC;
  extern __shared__ __align__(16) float sm[];
  float* Qs = sm;                    // [TS][LDA]
  float* Kt = Qs + TS * LDA;         // [HD][TS]
  float* Vs = Kt + HD * TS;          // [TS][HD]
  float* Ps = Vs + TS * HD;          // [TS][LDP]
  const int tx = threadIdx.x & 15, ty = threadIdx.x >> 4;
  const int qt = (int)gridDim.x - 1 - (int)blockIdx.x;  // late (heavy under a causal mask) query tiles first
  const int h = blockIdx.y, b = blockIdx.z, g = h / (p.Hq / p.Hkv);
  const int q0 = qt * TS;
  const float* qp = p.q + ((int64_t)b * p.T + q0) * p.ldq + (int64_t)h * HD;
  af_load_nat<HD, LDA>(Qs, qp, p.ldq, min(TS, p.T - q0), TS);
  float m_run[MR], l_run[MR], acc[MR][NC];
#pragma unroll
  for (int i = 0; i < MR; ++i) {
    m_run[i] = -INFINITY;
    l_run[i] = 0.f;
#pragma unroll
    for (int c = 0; c < NC; ++c) acc[i][c] = 0.f;
  }
  const int kv_end = p.causal ? min(p.T, q0 + TS) : p.T;
  for (int k0 = 0; k0 < kv_end; k0 += TS) {
    __syncthreads();  // previous tile's readers are done with Kt, Vs, Ps (and Qs is written)
    const int kn = min(TS, p.T - k0);
    af_load_tr<HD, TS>(Kt, p.k + ((int64_t)b * p.T + k0) * p.ldk + (int64_t)g * HD, p.ldk, kn);
    af_load_nat<HD, HD>(Vs, p.v + ((int64_t)b * p.T + k0) * p.ldv + (int64_t)g * HD, p.ldv, kn, TS);
    __syncthreads();
    float s[MR][MR];
    af_mm_nt<MR, HD, LDA, TS>(Qs, Kt, ty, tx, s);
#pragma unroll
    for (int i = 0; i < MR; ++i) {
      const int qi = q0 + ty * MR + i;
      float mx = -INFINITY;
#pragma unroll
      for (int j = 0; j < MR; ++j) {
        const int kj = k0 + tx * MR + j;
        s[i][j] = (kj < p.T && (!p.causal || kj <= qi)) ? s[i][j] * p.scale : -INFINITY;
        mx = fmaxf(mx, s[i][j]);
      }
      mx = af_row_max(mx);
      const float mn = fmaxf(m_run[i], mx);
      const float alpha = (m_run[i] == -INFINITY) ? 0.f : __expf(m_run[i] - mn);
      float rs = 0.f;
      float pr[MR];
#pragma unroll
      for (int j = 0; j < MR; ++j) {
        pr[j] = (s[i][j] == -INFINITY) ? 0.f : __expf(s[i][j] - mn);
        rs += pr[j];
      }
      rs = af_row_sum(rs);
      m_run[i] = mn;
      l_run[i] = l_run[i] * alpha + rs;
#pragma unroll
      for (int c = 0; c < NC; ++c) acc[i][c] *= alpha;
#pragma unroll
      for (int j = 0; j < MR; ++j) Ps[(ty * MR + i) * LDP + tx * MR + j] = pr[j];
    }
    __syncthreads();
    af_mm_nn<MR, TS, LDP, HD, NC>(Ps, Vs, ty, tx, acc);
  }
#pragma unroll
  for (int i = 0; i < MR; ++i) {
    const int qi = q0 + ty * MR + i;
    if (qi >= p.T) continue;
    const float inv = 1.0f / l_run[i];
    float* op = p.out + ((int64_t)b * p.T + qi) * p.ldo + (int64_t)h * HD;
#pragma unroll
    for (int gq = 0; gq < NC / 4; ++gq)
      *reinterpret_cast<float4*>(op + tx * 4 + 64 * gq) =
          make_float4(acc[i][4 * gq] * inv, acc[i][4 * gq + 1] * inv, acc[i][4 * gq + 2] * inv, acc[i][4 * gq + 3] * inv);
    if (tx == 0) p.lse[((int64_t)b * p.Hq + h) * p.T + qi] = m_run[i] + logf(l_run[i]);
  }
}

// ------------------------------------------------------------------------------------------------ backward 1: dQ
template <int HD>
__global__ void __launch_bounds__(AF_THREADS) attn_ffma_bwd_dq_kernel(const AfArgs p) {
  using C = AfCfg<HD>;
  constexpr int TS = C::TS, MR = C::MR, LDA = C::LDA, LDP = C::LDS_, NC = C::NC;
  extern __shared__ __align__(16) float sm[];
  float* Qs = sm;                     // [TS][LDA]
  float* dOs = Qs + TS * LDA;         // [TS][LDA]
  float* Kt = dOs + TS * LDA;         // [HD][TS]
  float* Vt = Kt + HD * TS;           // [HD][TS]
  float* Kn = Vt + HD * TS;           // [TS][HD]
  float* dSs = Kn + TS * HD;          // [TS][LDP]
  const int tx = threadIdx.x & 15, ty = threadIdx.x >> 4;
  const int qt = (int)gridDim.x - 1 - (int)blockIdx.x;
  const int h = blockIdx.y, b = blockIdx.z, g = h / (p.Hq / p.Hkv);
  const int q0 = qt * TS;
  const int qn = min(TS, p.T - q0);
  af_load_nat<HD, LDA>(Qs, p.q + ((int64_t)b * p.T + q0) * p.ldq + (int64_t)h * HD, p.ldq, qn, TS);
  af_load_nat<HD, LDA>(dOs, p.d_o + ((int64_t)b * p.T + q0) * p.lddo + (int64_t)h * HD, p.lddo, qn, TS);
  float lse_r[MR], del_r[MR], acc[MR][NC];
#pragma unroll
  for (int i = 0; i < MR; ++i) {
    const int qi = q0 + ty * MR + i;
    lse_r[i] = qi < p.T ? p.lse[((int64_t)b * p.Hq + h) * p.T + qi] : 0.f;
    del_r[i] = qi < p.T ? p.delta[((int64_t)b * p.Hq + h) * p.T + qi] : 0.f;
#pragma unroll
    for (int c = 0; c < NC; ++c) acc[i][c] = 0.f;
  }
  const int kv_end = p.causal ? min(p.T, q0 + TS) : p.T;
  for (int k0 = 0; k0 < kv_end; k0 += TS) {
    __syncthreads();
    const int kn = min(TS, p.T - k0);
    const float* kp = p.k + ((int64_t)b * p.T + k0) * p.ldk + (int64_t)g * HD;
    af_load_tr<HD, TS>(Kt, kp, p.ldk, kn);
    af_load_tr<HD, TS>(Vt, p.v + ((int64_t)b * p.T + k0) * p.ldv + (int64_t)g * HD, p.ldv, kn);
    af_load_nat<HD, HD>(Kn, kp, p.ldk, kn, TS);
    __syncthreads();
    float s[MR][MR], dp[MR][MR];
    af_mm_nt<MR, HD, LDA, TS>(Qs, Kt, ty, tx, s);
    af_mm_nt<MR, HD, LDA, TS>(dOs, Vt, ty, tx, dp);
#pragma unroll
    for (int i = 0; i < MR; ++i) {
      const int qi = q0 + ty * MR + i;
#pragma unroll
      for (int j = 0; j < MR; ++j) {
        const int kj = k0 + tx * MR + j;
        const bool vis = qi < p.T && kj < p.T && (!p.causal || kj <= qi);
        const float pv = vis ? __expf(s[i][j] * p.scale - lse_r[i]) : 0.f;
        dSs[(ty * MR + i) * LDP + tx * MR + j] = pv * (dp[i][j] - del_r[i]) * p.scale;
      }
    }
    __syncthreads();
    af_mm_nn<MR, TS, LDP, HD, NC>(dSs, Kn, ty, tx, acc);
  }
#pragma unroll
  for (int i = 0; i < MR; ++i) {
    const int qi = q0 + ty * MR + i;
    if (qi >= p.T) continue;
    float* op = p.dq + ((int64_t)b * p.T + qi) * p.lddq + (int64_t)h * HD;
#pragma unroll
    for (int gq = 0; gq < NC / 4; ++gq)
      *reinterpret_cast<float4*>(op + tx * 4 + 64 * gq) = make_float4(acc[i][4 * gq], acc[i][4 * gq + 1], acc[i][4 * gq + 2], acc[i][4 * gq + 3]);
  }
}

// ------------------------------------------------------------------------------------------------ backward 2: dK, dV
template <int HD>
__global__ void __launch_bounds__(AF_THREADS) attn_ffma_bwd_dkv_kernel(const AfArgs p) {
  using C = AfCfg<HD>;
  constexpr int TS = C::TS, MR = C::MR, LDA = C::LDA, LDP = C::LDS_, NC = C::NC;
  extern __shared__ __align__(16) float sm[];
  float* Kt = sm;                     // [HD][TS] resident
  float* Vt = Kt + HD * TS;           // [HD][TS] resident
  float* Qn = Vt + HD * TS;           // [TS][LDA]: A operand of S = Q K^T and B operand of dK += dS^T Q
  float* dOn = Qn + TS * LDA;         // [TS][LDA]: A operand of dP = dO V^T and B operand of dV += P^T dO
  float* Pt = dOn + TS * LDA;         // [TS keys][LDP]: P^T
  float* dSt = Pt + TS * LDP;         // [TS keys][LDP]: dS^T
  const int tx = threadIdx.x & 15, ty = threadIdx.x >> 4;
  const int kt = blockIdx.x, g = blockIdx.y, b = blockIdx.z;  // kv tile 0 is the heaviest under a causal mask and comes first
  const int n_rep = p.Hq / p.Hkv;
  const int k0 = kt * TS;
  const int kn = min(TS, p.T - k0);
  af_load_tr<HD, TS>(Kt, p.k + ((int64_t)b * p.T + k0) * p.ldk + (int64_t)g * HD, p.ldk, kn);
  af_load_tr<HD, TS>(Vt, p.v + ((int64_t)b * p.T + k0) * p.ldv + (int64_t)g * HD, p.ldv, kn);
  float dk[MR][NC], dv[MR][NC];
#pragma unroll
  for (int i = 0; i < MR; ++i)
#pragma unroll
    for (int c = 0; c < NC; ++c) dk[i][c] = dv[i][c] = 0.f;
  const int q_first = p.causal ? k0 : 0;  // (k0 is a multiple of TS)
  for (int hh = 0; hh < n_rep; ++hh) {
    const int h = g * n_rep + hh;
    for (int q0 = q_first; q0 < p.T; q0 += TS) {
      __syncthreads();
      const int qn = min(TS, p.T - q0);
      af_load_nat<HD, LDA>(Qn, p.q + ((int64_t)b * p.T + q0) * p.ldq + (int64_t)h * HD, p.ldq, qn, TS);
      af_load_nat<HD, LDA>(dOn, p.d_o + ((int64_t)b * p.T + q0) * p.lddo + (int64_t)h * HD, p.lddo, qn, TS);
      __syncthreads();
      float s[MR][MR], dp[MR][MR];  // rows = query (ty), columns = key (tx)
      af_mm_nt<MR, HD, LDA, TS>(Qn, Kt, ty, tx, s);
      af_mm_nt<MR, HD, LDA, TS>(dOn, Vt, ty, tx, dp);
#pragma unroll
      for (int j = 0; j < MR; ++j) {
        const int kj = k0 + tx * MR + j;
        float pr[MR], dr[MR];
#pragma unroll
        for (int i = 0; i < MR; ++i) {
          const int qi = q0 + ty * MR + i;
          const bool vis = qi < p.T && kj < p.T && (!p.causal || kj <= qi);
          const int64_t si = ((int64_t)b * p.Hq + h) * p.T + (qi < p.T ? qi : 0);
          const float pv = vis ? __expf(s[i][j] * p.scale - p.lse[si]) : 0.f;
          pr[i] = pv;
          dr[i] = pv * (dp[i][j] - p.delta[si]) * p.scale;
        }
        // transposed store: row = key, MR consecutive query columns
        float* pp = Pt + (tx * MR + j) * LDP + ty * MR;
        float* dd = dSt + (tx * MR + j) * LDP + ty * MR;
        if constexpr (MR == 4) {
          *reinterpret_cast<float4*>(pp) = make_float4(pr[0], pr[1], pr[2], pr[3]);
          *reinterpret_cast<float4*>(dd) = make_float4(dr[0], dr[1], dr[2], dr[3]);
        } else {
          *reinterpret_cast<float2*>(pp) = make_float2(pr[0], pr[1]);
          *reinterpret_cast<float2*>(dd) = make_float2(dr[0], dr[1]);
        }
      }
      __syncthreads();
      af_mm_nn<MR, TS, LDP, LDA, NC>(Pt, dOn, ty, tx, dv);   // dV[key][d] += sum_q P^T[key][q] dO[q][d]
      af_mm_nn<MR, TS, LDP, LDA, NC>(dSt, Qn, ty, tx, dk);   // dK[key][d] += sum_q dS^T[key][q] Q[q][d]
    }
  }
#pragma unroll
  for (int i = 0; i < MR; ++i) {
    const int ki = k0 + ty * MR + i;
    if (ki >= p.T) continue;
    float* kp = p.dk + ((int64_t)b * p.T + ki) * p.lddk + (int64_t)g * HD;
    float* vp = p.dv + ((int64_t)b * p.T + ki) * p.lddv + (int64_t)g * HD;
#pragma unroll
    for (int gq = 0; gq < NC / 4; ++gq) {
      *reinterpret_cast<float4*>(kp + tx * 4 + 64 * gq) = make_float4(dk[i][4 * gq], dk[i][4 * gq + 1], dk[i][4 * gq + 2], dk[i][4 * gq + 3]);
      *reinterpret_cast<float4*>(vp + tx * 4 + 64 * gq) = make_float4(dv[i][4 * gq], dv[i][4 * gq + 1], dv[i][4 * gq + 2], dv[i][4 * gq + 3]);
    }
  }
}

// delta[b, h, t] = sum_c dO[n, h*hd + c] * O[n, h*hd + c]: one warp per (token, head)
__global__ void __launch_bounds__(256) attn_ffma_delta_kernel(const float* __restrict__ o, int64_t ldo, const float* __restrict__ d_o, int64_t lddo,
                                                              float* __restrict__ delta, int B, int T, int Hq, int hd) {
  const int64_t w = ((int64_t)blockIdx.x * blockDim.x + threadIdx.x) >> 5;
  const int lane = threadIdx.x & 31;
  if (w >= (int64_t)B * T * Hq) return;
  const int h = (int)(w % Hq);
  const int64_t tok = w / Hq;
  const float* a = o + tok * ldo + (int64_t)h * hd;
  const float* g = d_o + tok * lddo + (int64_t)h * hd;
  float s = 0.f;
  for (int c = lane * 4; c < hd; c += 128) {
    const float4 x = *reinterpret_cast<const float4*>(a + c), y = *reinterpret_cast<const float4*>(g + c);
    s = fmaf(x.x, y.x, fmaf(x.y, y.y, fmaf(x.z, y.z, fmaf(x.w, y.w, s))));
  }
  s = warp_sum(s);
  if (lane == 0) delta[((tok / T) * Hq + h) * T + tok % T] = s;
}

bool attn_ffma_supported(int dtype, int hd, const void* const* ptrs, const int64_t* lds, int n) {
  if (dtype != SPB_F32 || !(hd == 64 || hd == 128 || hd == 256)) return false;
  for (int i = 0; i < n; ++i)
    if (((uintptr_t)ptrs[i] % 16) != 0 || (lds[i] % 4) != 0) return false;  // 128-bit loads / stores
  return true;
}

template <int HD>
static int attn_ffma_fwd_t(const AfArgs& a, int B, cudaStream_t st) {
  using C = AfCfg<HD>;
  const size_t smem = (size_t)(C::TS * C::LDA + HD * C::TS + C::TS * HD + C::TS * C::LDS_) * sizeof(float);
  SPB_CUDA(cudaFuncSetAttribute(attn_ffma_fwd_kernel<HD>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
  dim3 grid((a.T + C::TS - 1) / C::TS, a.Hq, B);
  attn_ffma_fwd_kernel<HD><<<grid, AF_THREADS, smem, st>>>(a);
  SPB_LAUNCH_CHECK();
  return 0;
}

template <int HD>
static int attn_ffma_bwd_t(const AfArgs& a, int B, cudaStream_t st) {
  using C = AfCfg<HD>;
  const int64_t warps = (int64_t)B * a.T * a.Hq;
  attn_ffma_delta_kernel<<<(unsigned)((warps * 32 + 255) / 256), 256, 0, st>>>(a.o, a.ldo, a.d_o, a.lddo, a.delta, B, a.T, a.Hq, HD);
  SPB_LAUNCH_CHECK();
  const size_t smem_q = (size_t)(2 * C::TS * C::LDA + 2 * HD * C::TS + C::TS * HD + C::TS * C::LDS_) * sizeof(float);
  SPB_CUDA(cudaFuncSetAttribute(attn_ffma_bwd_dq_kernel<HD>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem_q));
  dim3 gq((a.T + C::TS - 1) / C::TS, a.Hq, B);
  attn_ffma_bwd_dq_kernel<HD><<<gq, AF_THREADS, smem_q, st>>>(a);
  SPB_LAUNCH_CHECK();
  const size_t smem_k = (size_t)(2 * HD * C::TS + 2 * C::TS * C::LDA + 2 * C::TS * C::LDS_) * sizeof(float);
  SPB_CUDA(cudaFuncSetAttribute(attn_ffma_bwd_dkv_kernel<HD>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem_k));
  dim3 gk((a.T + C::TS - 1) / C::TS, a.Hkv, B);
  attn_ffma_bwd_dkv_kernel<HD><<<gk, AF_THREADS, smem_k, st>>>(a);
  SPB_LAUNCH_CHECK();
  return 0;
}

int attn_ffma_fwd(int B, int Tn, int Hq, int Hkv, int hd, int causal, float scale, const void* q, int64_t ldq, const void* k, int64_t ldk,
                  const void* v, int64_t ldv, void* o, int64_t ldo, float* lse, cudaStream_t st) {
  SPB_CHECK_ARG(B <= 65535 && Hq <= 65535, SPB_E_LIMIT, "attention (fp32): B or Hq beyond 65535");
  AfArgs a = {};
  a.q = (const float*)q; a.k = (const float*)k; a.v = (const float*)v; a.out = (float*)o; a.lse = lse;
  a.ldq = ldq; a.ldk = ldk; a.ldv = ldv; a.ldo = ldo;
  a.T = Tn; a.Hq = Hq; a.Hkv = Hkv; a.causal = causal; a.scale = scale;
  if (hd == 64) return attn_ffma_fwd_t<64>(a, B, st);
  if (hd == 128) return attn_ffma_fwd_t<128>(a, B, st);
  return attn_ffma_fwd_t<256>(a, B, st);
}

int attn_ffma_bwd(int B, int Tn, int Hq, int Hkv, int hd, int causal, float scale, const void* q, int64_t ldq, const void* k, int64_t ldk,
                  const void* v, int64_t ldv, const void* o, int64_t ldo, const void* d_o, int64_t lddo, const float* lse, float* delta, void* dq,
                  int64_t lddq, void* dk, int64_t lddk, void* dv, int64_t lddv, cudaStream_t st) {
  SPB_CHECK_ARG(B <= 65535 && Hq <= 65535, SPB_E_LIMIT, "attention (fp32): B or Hq beyond 65535");
  AfArgs a = {};
  a.q = (const float*)q; a.k = (const float*)k; a.v = (const float*)v; a.o = (const float*)o; a.d_o = (const float*)d_o;
  a.lse = const_cast<float*>(lse); a.delta = delta; a.dq = (float*)dq; a.dk = (float*)dk; a.dv = (float*)dv;
  a.ldq = ldq; a.ldk = ldk; a.ldv = ldv; a.ldo = ldo; a.lddo = lddo; a.lddq = lddq; a.lddk = lddk; a.lddv = lddv;
  a.T = Tn; a.Hq = Hq; a.Hkv = Hkv; a.causal = causal; a.scale = scale;
  if (hd == 64) return attn_ffma_bwd_t<64>(a, B, st);
  if (hd == 128) return attn_ffma_bwd_t<128>(a, B, st);
  return attn_ffma_bwd_t<256>(a, B, st);
}

}  // namespace spb
