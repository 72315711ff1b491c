// norm_rope.cu - K1: the memory-bound kernels of the block.
//   * norm forward  (RMSNorm llama3:536-538 / gemma:139-150, LayerNorm flax + torch): 1-8 warps per row, the row is
//     read ONCE with 128-bit loads and kept packed in registers, reduced with warp shuffles in fp32, written once.
//   * norm backward (dx: same structure; dweight/dbias: coalesced column-sum kernel + deterministic 2nd stage).
//   * rotary (llama3:592-601 adjacent-pair rotation; gemma:182-200 "rotary" = general 2x2 per (t,pair) table),
//     in place on the q|k column range of the packed qkv buffer, 128-bit loads, and its transpose for the VJP.
// Algorithmic bytes: norm fwd 2*N*C*e (+8N stats); norm bwd dx 3*N*C*e; rope 2*N*cols*e.
#include <mutex>

#include "common.cuh"

namespace spb {

template <typename T>
struct Vec;  // 128-bit vector of T
template <>
struct Vec<float> {
  static constexpr int N = 4;
  float4 v;
  __device__ __forceinline__ void load(const float* p) { v = *reinterpret_cast<const float4*>(p); }
  __device__ __forceinline__ void store(float* p) const { *reinterpret_cast<float4*>(p) = v; }
  __device__ __forceinline__ void unpack(float* f) const { f[0] = v.x; f[1] = v.y; f[2] = v.z; f[3] = v.w; }
  __device__ __forceinline__ void pack(const float* f) { v = make_float4(f[0], f[1], f[2], f[3]); }
};
template <>
struct Vec<__nv_bfloat16> {
  static constexpr int N = 8;
  uint4 v;
  __device__ __forceinline__ void load(const __nv_bfloat16* p) { v = *reinterpret_cast<const uint4*>(p); }
  __device__ __forceinline__ void store(__nv_bfloat16* p) const { *reinterpret_cast<uint4*>(p) = v; }
  __device__ __forceinline__ void unpack(float* f) const {
    float2 a = unpack_bf16(v.x), b = unpack_bf16(v.y), c = unpack_bf16(v.z), d = unpack_bf16(v.w);
    f[0] = a.x; f[1] = a.y; f[2] = b.x; f[3] = b.y; f[4] = c.x; f[5] = c.y; f[6] = d.x; f[7] = d.y;
  }
  __device__ __forceinline__ void pack(const float* f) {
    v.x = pack_bf16(f[0], f[1]); v.y = pack_bf16(f[2], f[3]); v.z = pack_bf16(f[4], f[5]); v.w = pack_bf16(f[6], f[7]);
  }
};

constexpr int NORM_WARPS = 4;

// ------------------------------------------------------------------------------------------------ norm forward
// 256-thread CTA = 8 warps; WPR warps share one row (8/WPR rows per CTA), every thread keeps VPT 128-bit vectors of the row
// PACKED in registers (<= 16 registers), so occupancy stays high and the whole machine has loads in flight: the row is
// read from HBM once and written once.  Row reductions: warp shuffles, then (WPR > 1) one shared-memory exchange.
constexpr int NR_THREADS = 256;

template <int WPR>
__device__ __forceinline__ float row_reduce(float v, float* red, int warp) {
  v = warp_sum(v);
  if (WPR > 1) {
    if ((threadIdx.x & 31) == 0) red[warp] = v;
    __syncthreads();
    const int base = (warp / WPR) * WPR;
    v = 0.f;
#pragma unroll
    for (int j = 0; j < WPR; ++j) v += red[base + j];
  }
  return v;
}

template <typename T, int VPT, int WPR>
__global__ void __launch_bounds__(NR_THREADS) norm_fwd_kernel(const T* __restrict__ x, const T* __restrict__ w,
                                                               const T* __restrict__ b, T* __restrict__ y,
                                                               float* __restrict__ stats, int64_t N, int C, float eps,
                                                               int kind, int upcast) {
  constexpr int VN = Vec<T>::N;
  constexpr int RPC = 8 / WPR;
  __shared__ float red[2][8];
  pdl_launch_dependents();  // a kernel launched with launch_pdl behind this one may begin its own prologue
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int w_in = warp % WPR;
  const int64_t row = (int64_t)blockIdx.x * RPC + warp / WPR;
  const bool active = row < N;
  const T* xr = x + row * C;
  Vec<T> xv[VPT];
#pragma unroll
  for (int k = 0; k < VPT; ++k) {
    const int c = ((k * WPR + w_in) * 32 + lane) * VN;
    if (active && c < C) xv[k].load(xr + c);
  }
  float mean = 0.f;
  if (kind == SPB_NORM_LAYER) {
    float s = 0.f;
#pragma unroll
    for (int k = 0; k < VPT; ++k) {
      const int c = ((k * WPR + w_in) * 32 + lane) * VN;
      if (active && c < C) {
        float f[VN];
        xv[k].unpack(f);
#pragma unroll
        for (int e = 0; e < VN; ++e) s += f[e];
      }
    }
    mean = row_reduce<WPR>(s, red[0], warp) / (float)C;
  }
  float ss = 0.f;
#pragma unroll
  for (int k = 0; k < VPT; ++k) {
    const int c = ((k * WPR + w_in) * 32 + lane) * VN;
    if (active && c < C) {
      float f[VN];
      xv[k].unpack(f);
#pragma unroll
      for (int e = 0; e < VN; ++e) {
        const float d = f[e] - mean;
        ss += d * d;
      }
    }
  }
  const float var = row_reduce<WPR>(ss, red[1], warp) / (float)C;
  const float rstd = 1.0f / sqrtf(var + eps);
  if (!active) return;
  if (stats != nullptr && lane == 0 && w_in == 0) {
    stats[2 * row] = mean;
    stats[2 * row + 1] = rstd;
  }
  T* yr = y + row * C;
#pragma unroll
  for (int k = 0; k < VPT; ++k) {
    const int c = ((k * WPR + w_in) * 32 + lane) * VN;
    if (c < C) {
      Vec<T> wv;
      wv.load(w + c);
      float f[VN], wf[VN], o[VN];
      xv[k].unpack(f);
      wv.unpack(wf);
      if (kind == SPB_NORM_LAYER) {
        float bf[VN];
        if (b != nullptr) {
          Vec<T> bv;
          bv.load(b + c);
          bv.unpack(bf);
        } else {
#pragma unroll
          for (int e = 0; e < VN; ++e) bf[e] = 0.f;
        }
#pragma unroll
        for (int e = 0; e < VN; ++e) o[e] = (f[e] - mean) * rstd * wf[e] + bf[e];
      } else if (upcast) {  // gemma:148-150: round the normalised value to T first, then multiply by the weight
#pragma unroll
        for (int e = 0; e < VN; ++e) o[e] = to_f32<T>(from_f32<T>(f[e] * rstd)) * wf[e];
      } else {  // llama3:538: (x * weight) * 1/sqrt(var + eps)
#pragma unroll
        for (int e = 0; e < VN; ++e) o[e] = (f[e] * wf[e]) * rstd;
      }
      Vec<T> ov;
      ov.pack(o);
      ov.store(yr + c);
    }
  }
}

// Generic (any C, scalar) fallback shape: one warp per row, two passes over the row (second pass hits L1/L2).
template <typename T>
__global__ void __launch_bounds__(NORM_WARPS * 32) norm_fwd_generic_kernel(const T* __restrict__ x, const T* __restrict__ w,
                                                                            const T* __restrict__ b, T* __restrict__ y,
                                                                            float* __restrict__ stats, int64_t N, int C,
                                                                            float eps, int kind, int upcast) {
  const int lane = threadIdx.x & 31;
  const int64_t row = (int64_t)blockIdx.x * NORM_WARPS + (threadIdx.x >> 5);
  if (row >= N) return;
  const T* xr = x + row * C;
  float mean = 0.f;
  if (kind == SPB_NORM_LAYER) {
    float s = 0.f;
    for (int c = lane; c < C; c += 32) s += to_f32<T>(xr[c]);
    mean = warp_sum(s) / (float)C;
  }
  float ss = 0.f;
  for (int c = lane; c < C; c += 32) {
    const float d = to_f32<T>(xr[c]) - mean;
    ss += d * d;
  }
  const float rstd = 1.0f / sqrtf(warp_sum(ss) / (float)C + eps);
  if (stats != nullptr && lane == 0) {
    stats[2 * row] = mean;
    stats[2 * row + 1] = rstd;
  }
  T* yr = y + row * C;
  for (int c = lane; c < C; c += 32) {
    const float xv = to_f32<T>(xr[c]), wv = to_f32<T>(w[c]);
    float o;
    if (kind == SPB_NORM_LAYER) o = (xv - mean) * rstd * wv + (b ? to_f32<T>(b[c]) : 0.f);
    else if (upcast) o = to_f32<T>(from_f32<T>(xv * rstd)) * wv;
    else o = (xv * wv) * rstd;
    yr[c] = from_f32<T>(o);
  }
}

// ------------------------------------------------------------------------------------------------ norm backward (dx)
// RMS  : y = x*w*r,            g = dy*w,  dx = r*g - x * r^3 * mean(g*x)
// LAYER: y = (x-m)*r*w + b,    g = dy*w,  xh = (x-m)*r,  dx = r*(g - mean(g) - xh*mean(g*xh))
template <typename T, int VPT, int WPR>
__global__ void __launch_bounds__(NR_THREADS) norm_bwd_dx_kernel(const T* __restrict__ x, const T* __restrict__ w,
                                                                  const float* __restrict__ stats, const T* __restrict__ dy,
                                                                  T* __restrict__ dx, int64_t N, int C, int kind,
                                                                  const T* __restrict__ add) {
  constexpr int VN = Vec<T>::N;
  constexpr int RPC = 8 / WPR;
  __shared__ float red[2][8];
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int w_in = warp % WPR;
  const int64_t row = (int64_t)blockIdx.x * RPC + warp / WPR;
  const bool active = row < N;
  float mean = 0.f, rstd = 0.f;
  if (active) {
    mean = stats[2 * row];
    rstd = stats[2 * row + 1];
  }
  // x and dy of the row stay packed in registers (read from HBM once); w is re-read from L1/L2
  Vec<T> xv[VPT], dv[VPT];
#pragma unroll
  for (int k = 0; k < VPT; ++k) {
    const int c = ((k * WPR + w_in) * 32 + lane) * VN;
    if (active && c < C) {
      xv[k].load(x + row * C + c);
      dv[k].load(dy + row * C + c);
    }
  }
  float s_g = 0.f, s_gx = 0.f;
#pragma unroll
  for (int k = 0; k < VPT; ++k) {
    const int c = ((k * WPR + w_in) * 32 + lane) * VN;
    if (active && c < C) {
      Vec<T> wv;
      wv.load(w + c);
      float xf[VN], g[VN], wf[VN];
      xv[k].unpack(xf);
      dv[k].unpack(g);
      wv.unpack(wf);
#pragma unroll
      for (int e = 0; e < VN; ++e) {
        const float gv = g[e] * wf[e];
        const float xn = (kind == SPB_NORM_LAYER) ? (xf[e] - mean) * rstd : xf[e];
        s_g += gv;
        s_gx += gv * xn;
      }
    }
  }
  s_g = row_reduce<WPR>(s_g, red[0], warp) / (float)C;
  s_gx = row_reduce<WPR>(s_gx, red[1], warp) / (float)C;
  if (!active) return;
  const float r3 = rstd * rstd * rstd;
#pragma unroll
  for (int k = 0; k < VPT; ++k) {
    const int c = ((k * WPR + w_in) * 32 + lane) * VN;
    if (c < C) {
      Vec<T> wv;
      wv.load(w + c);
      float xf[VN], g[VN], wf[VN], o[VN];
      xv[k].unpack(xf);
      dv[k].unpack(g);
      wv.unpack(wf);
#pragma unroll
      for (int e = 0; e < VN; ++e) {
        const float gv = g[e] * wf[e];
        if (kind == SPB_NORM_LAYER) o[e] = rstd * (gv - s_g - (xf[e] - mean) * rstd * s_gx);
        else o[e] = rstd * gv - xf[e] * r3 * s_gx;
      }
      if (add != nullptr) {  // gradient of the residual branch (h = x + block(x)): dx += dy_block
        Vec<T> av;
        av.load(add + row * C + c);
        float af[VN];
        av.unpack(af);
#pragma unroll
        for (int e = 0; e < VN; ++e) o[e] += af[e];
      }
      Vec<T> ov;
      ov.pack(o);
      ov.store(dx + row * C + c);
    }
  }
}

// Fused backward: dx AND the per-CTA partial column sums of dweight (and dbias for LayerNorm) in ONE pass over x and dy
// (3*N*C*e bytes instead of 5).  A CTA walks `groups_per_cta` consecutive row groups; a thread always owns the same
// columns, and its share of dw/db lives in SHARED memory (thread-private float4 slots, lane-consecutive: conflict-free),
// not in registers - the first version kept 32-64 accumulators per thread, fell to 2 CTAs/SM and ran at 1.4 TB/s; this one
// keeps the register footprint of the plain dx kernel.  At the end the CTA's row lanes are combined and one row of
// partial[] is written (2nd stage: colsum_final_kernel).
template <typename T, int VPT, int WPR, bool LAYER>
__global__ void __launch_bounds__(NR_THREADS, 4) norm_bwd_fused_kernel(const T* __restrict__ x, const T* __restrict__ w,
                                                                        const float* __restrict__ stats, const T* __restrict__ dy,
                                                                        T* __restrict__ dx, int64_t N, int C, int64_t row_groups,
                                                                        int groups_per_cta, float* __restrict__ partial,
                                                                        const T* __restrict__ add) {
  constexpr int VN = Vec<T>::N;
  constexpr int RPC = 8 / WPR;
  constexpr int Q = VN / 4;                        // float4 slots per vector
  constexpr int PLANE = 8 * VPT * Q * 32;          // float4 slots of one plane (dw; LayerNorm adds a db plane)
  extern __shared__ float4 fused_acc[];
  __shared__ float red[4][8];  // [iteration parity][s_g, s_gx]: a slow warp may still read the previous row's sums
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int w_in = warp % WPR, rlane = warp / WPR;
  for (int i = threadIdx.x; i < (LAYER ? 2 : 1) * PLANE; i += NR_THREADS) fused_acc[i] = make_float4(0.f, 0.f, 0.f, 0.f);
  // (thread-private slots: no barrier needed before the thread's own read-modify-writes)
  float4* mine = fused_acc + (warp * VPT * Q) * 32 + lane;  // slot (k, h) at mine[(k*Q + h) * 32]
  __syncthreads();
  const int64_t g0 = (int64_t)blockIdx.x * groups_per_cta;
  const int64_t g1 = g0 + groups_per_cta < row_groups ? g0 + groups_per_cta : row_groups;
  int par = 0;
  for (int64_t rg = g0; rg < g1; ++rg) {
    const int64_t row = rg * RPC + rlane;
    const bool active = row < N;
    float mean = 0.f, rstd = 0.f;
    if (active) {
      if (LAYER) mean = stats[2 * row];
      rstd = stats[2 * row + 1];
    }
    Vec<T> xv[VPT], dv[VPT];
#pragma unroll
    for (int k = 0; k < VPT; ++k) {
      const int c = ((k * WPR + w_in) * 32 + lane) * VN;
      if (active && c < C) {
        xv[k].load(x + row * C + c);
        dv[k].load(dy + row * C + c);
      }
    }
    float s_g = 0.f, s_gx = 0.f;
#pragma unroll
    for (int k = 0; k < VPT; ++k) {
      const int c = ((k * WPR + w_in) * 32 + lane) * VN;
      if (active && c < C) {
        Vec<T> wv;
        wv.load(w + c);
        float xf[VN], g[VN], wf[VN];
        xv[k].unpack(xf);
        dv[k].unpack(g);
        wv.unpack(wf);
#pragma unroll
        for (int h = 0; h < Q; ++h) {
          float4 a = mine[(k * Q + h) * 32];
          float* af = reinterpret_cast<float*>(&a);
#pragma unroll
          for (int e = 0; e < 4; ++e) {
            const float xh = LAYER ? (xf[4 * h + e] - mean) * rstd : xf[4 * h + e] * rstd;
            const float gv = g[4 * h + e] * wf[4 * h + e];
            af[e] += g[4 * h + e] * xh;
            s_g += gv;
            s_gx += gv * xh;
          }
          mine[(k * Q + h) * 32] = a;
          if (LAYER) {
            float4 bsum = mine[PLANE + (k * Q + h) * 32];
            bsum.x += g[4 * h];
            bsum.y += g[4 * h + 1];
            bsum.z += g[4 * h + 2];
            bsum.w += g[4 * h + 3];
            mine[PLANE + (k * Q + h) * 32] = bsum;
          }
        }
      }
    }
    if (LAYER) s_g = row_reduce<WPR>(s_g, red[par], warp) / (float)C;
    s_gx = row_reduce<WPR>(s_gx, red[par + 1], warp) / (float)C;
    par ^= 2;
#pragma unroll
    for (int k = 0; k < VPT; ++k) {
      const int c = ((k * WPR + w_in) * 32 + lane) * VN;
      if (active && c < C) {
        Vec<T> wv;
        wv.load(w + c);
        float xf[VN], g[VN], wf[VN], o[VN];
        xv[k].unpack(xf);
        dv[k].unpack(g);
        wv.unpack(wf);
#pragma unroll
        for (int e = 0; e < VN; ++e) {
          const float gv = g[e] * wf[e];
          const float xh = LAYER ? (xf[e] - mean) * rstd : xf[e] * rstd;
          o[e] = LAYER ? rstd * (gv - s_g - xh * s_gx) : rstd * (gv - xh * s_gx);
        }
        if (add != nullptr) {  // gradient of the residual branch (h = x + block(x)): dx += dy_block
          Vec<T> av;
          av.load(add + row * C + c);
          float af[VN];
          av.unpack(af);
#pragma unroll
          for (int e = 0; e < VN; ++e) o[e] += af[e];
        }
        Vec<T> ov;
        ov.pack(o);
        ov.store(dx + row * C + c);
      }
    }
  }
  // combine the CTA's row lanes: the warps of row lane 0 add the slots of the same columns in the other lanes and write this
  // CTA's row of the partial sums - partial[blockIdx][c] (dw) and partial[gridDim + blockIdx][c] (db)
  __syncthreads();
  if (rlane == 0) {
#pragma unroll
    for (int pl = 0; pl < (LAYER ? 2 : 1); ++pl) {
      float* out = partial + ((int64_t)pl * gridDim.x + blockIdx.x) * (int64_t)C;
#pragma unroll
      for (int k = 0; k < VPT; ++k) {
        const int c = ((k * WPR + w_in) * 32 + lane) * VN;
        if (c < C) {
#pragma unroll
          for (int h = 0; h < Q; ++h) {
            float4 a = mine[pl * PLANE + (k * Q + h) * 32];
#pragma unroll
            for (int l = 1; l < RPC; ++l) {
              const float4 t = mine[pl * PLANE + (l * WPR * VPT * Q) * 32 + (k * Q + h) * 32];
              a.x += t.x; a.y += t.y; a.z += t.z; a.w += t.w;
            }
            *reinterpret_cast<float4*>(out + c + 4 * h) = a;
          }
        }
      }
    }
  }
}

template <typename T>
__global__ void __launch_bounds__(NORM_WARPS * 32) norm_bwd_dx_generic_kernel(const T* __restrict__ x, const T* __restrict__ w,
                                                                               const float* __restrict__ stats,
                                                                               const T* __restrict__ dy, T* __restrict__ dx,
                                                                               int64_t N, int C, int kind, const T* __restrict__ add) {
  const int lane = threadIdx.x & 31;
  const int64_t row = (int64_t)blockIdx.x * NORM_WARPS + (threadIdx.x >> 5);
  if (row >= N) return;
  const float mean = stats[2 * row], rstd = stats[2 * row + 1];
  const T* xr = x + row * C;
  const T* dr = dy + row * C;
  float s_g = 0.f, s_gx = 0.f;
  for (int c = lane; c < C; c += 32) {
    const float gv = to_f32<T>(dr[c]) * to_f32<T>(w[c]);
    const float xv = (kind == SPB_NORM_LAYER) ? (to_f32<T>(xr[c]) - mean) * rstd : to_f32<T>(xr[c]);
    s_g += gv;
    s_gx += gv * xv;
  }
  s_g = warp_sum(s_g) / (float)C;
  s_gx = warp_sum(s_gx) / (float)C;
  const float r3 = rstd * rstd * rstd;
  for (int c = lane; c < C; c += 32) {
    const float gv = to_f32<T>(dr[c]) * to_f32<T>(w[c]);
    float o;
    if (kind == SPB_NORM_LAYER) {
      const float xh = (to_f32<T>(xr[c]) - mean) * rstd;
      o = rstd * (gv - s_g - xh * s_gx);
    } else {
      o = rstd * gv - to_f32<T>(xr[c]) * r3 * s_gx;
    }
    if (add != nullptr) o += to_f32<T>(add[row * C + c]);
    dx[row * C + c] = from_f32<T>(o);
  }
}

// ------------------------------------------------------------------------------------------------ column sums
// partial[chunk][c]          = sum over the chunk's rows of  a[row][c]                       (MODE 0: bias grads)
// partial[chunk][c]          = sum of dy[row][c] * xn[row][c]   with xn = x*rstd  (RMS) or (x-mean)*rstd (LAYER)
// partial[chunks + chunk][c] = sum of dy[row][c]                                            (MODE 1: norm dw, db)
// Threads run along columns (coalesced); each thread keeps its column's running sum in a register.
template <typename T, int MODE>
__global__ void __launch_bounds__(256) colsum_partial_kernel(const T* __restrict__ a, int64_t lda, const T* __restrict__ x,
                                                             const float* __restrict__ stats, int kind, int64_t N, int C,
                                                             int rows_per_chunk, int chunks, float* __restrict__ partial) {
  const int c = blockIdx.x * blockDim.x + threadIdx.x;
  const int chunk = blockIdx.y;
  if (c >= C) return;
  const int64_t r0 = (int64_t)chunk * rows_per_chunk;
  int64_t r1 = r0 + rows_per_chunk;
  if (r1 > N) r1 = N;
  float s0 = 0.f, s1 = 0.f;
  for (int64_t r = r0; r < r1; ++r) {
    const float av = to_f32<T>(a[r * lda + c]);
    if (MODE == 1) {
      const float mean = (kind == SPB_NORM_LAYER) ? stats[2 * r] : 0.f;
      const float rstd = stats[2 * r + 1];
      s0 += av * ((to_f32<T>(x[r * (int64_t)C + c]) - mean) * rstd);
      s1 += av;
    } else {
      s0 += av;
    }
  }
  partial[(int64_t)chunk * C + c] = s0;
  if (MODE == 1) partial[((int64_t)chunks + chunk) * C + c] = s1;
}

// Vectorised variant: each thread owns Vec<T>::N consecutive columns (one 128-bit load per row and operand).  The 256
// threads of a block are arranged as TPR column-threads x 256/TPR row-lanes, so narrow matrices (C = 256: 32 column
// threads) still keep every thread busy; the row-lanes are combined through shared memory at the end.
template <typename T, int MODE>
__global__ void __launch_bounds__(256) colsum_partial_vec_kernel(const T* __restrict__ a, int64_t lda, const T* __restrict__ x,
                                                                 const float* __restrict__ stats, int kind, int64_t N, int C,
                                                                 int rows_per_chunk, int chunks, int tpr,
                                                                 float* __restrict__ partial) {
  constexpr int VN = Vec<T>::N;
  __shared__ float red[(MODE == 1 ? 2 : 1) * 256 * VN];
  const int tx = threadIdx.x % tpr, ty = threadIdx.x / tpr, rlanes = 256 / tpr;
  const int c = (blockIdx.x * tpr + tx) * VN;
  const int chunk = blockIdx.y;
  const bool live = c < C;
  const int64_t r0 = (int64_t)chunk * rows_per_chunk;
  int64_t r1 = r0 + rows_per_chunk;
  if (r1 > N) r1 = N;
  float s0[VN], s1[VN];
#pragma unroll
  for (int e = 0; e < VN; ++e) s0[e] = s1[e] = 0.f;
  if (live) {
#pragma unroll 2
    for (int64_t r = r0 + ty; r < r1; r += rlanes) {
      Vec<T> av;
      av.load(a + r * lda + c);
      float af[VN];
      av.unpack(af);
      if (MODE == 1) {
        Vec<T> xv;
        xv.load(x + r * (int64_t)C + c);
        float xf[VN];
        xv.unpack(xf);
        const float mean = (kind == SPB_NORM_LAYER) ? stats[2 * r] : 0.f;
        const float rstd = stats[2 * r + 1];
#pragma unroll
        for (int e = 0; e < VN; ++e) {
          s0[e] += af[e] * ((xf[e] - mean) * rstd);
          s1[e] += af[e];
        }
      } else {
#pragma unroll
        for (int e = 0; e < VN; ++e) s0[e] += af[e];
      }
    }
  }
  if (rlanes > 1) {  // block-uniform
#pragma unroll
    for (int e = 0; e < VN; ++e) {
      red[(e * 256) + threadIdx.x] = s0[e];
      if (MODE == 1) red[(VN + e) * 256 + threadIdx.x] = s1[e];
    }
    __syncthreads();
    if (ty == 0) {
      for (int l = 1; l < rlanes; ++l) {
#pragma unroll
        for (int e = 0; e < VN; ++e) {
          s0[e] += red[(e * 256) + l * tpr + tx];
          if (MODE == 1) s1[e] += red[(VN + e) * 256 + l * tpr + tx];
        }
      }
    }
  }
  if (!live || ty != 0) return;
  float* p0 = partial + (int64_t)chunk * C + c;
#pragma unroll
  for (int e = 0; e < VN; e += 4) *reinterpret_cast<float4*>(p0 + e) = make_float4(s0[e], s0[e + 1], s0[e + 2], s0[e + 3]);
  if (MODE == 1) {
    float* p1 = partial + ((int64_t)chunks + chunk) * C + c;
#pragma unroll
    for (int e = 0; e < VN; e += 4) *reinterpret_cast<float4*>(p1 + e) = make_float4(s1[e], s1[e + 1], s1[e + 2], s1[e + 3]);
  }
}

// column threads per block for a matrix of nvec 128-bit column vectors: power of two in [8, 256]
static int colsum_tpr(int nvec) {
  int t = 8;
  while (t < 256 && t < nvec) t <<= 1;
  return t;
}

// 2nd stage: out[c] = sum over the partial rows.  32 columns x 8 row lanes per block (the r01i launch list showed the
// one-thread-per-column version - 16 blocks walking 592 rows serially - taking as long as the fused first stage: 43 us).
__global__ void __launch_bounds__(256) colsum_final_kernel(const float* __restrict__ partial, int chunks, int C,
                                                           float* __restrict__ out0, float* __restrict__ out1) {
  __shared__ float red[2][8][33];
  const int tx = threadIdx.x & 31, ty = threadIdx.x >> 5;
  const int c = blockIdx.x * 32 + tx;
  float s0 = 0.f, s1 = 0.f;
  if (c < C) {
    if (out0 != nullptr) {
#pragma unroll 4
      for (int k = ty; k < chunks; k += 8) s0 += partial[(int64_t)k * C + c];
    }
    if (out1 != nullptr) {
#pragma unroll 4
      for (int k = ty; k < chunks; k += 8) s1 += partial[((int64_t)chunks + k) * C + c];
    }
  }
  red[0][ty][tx] = s0;
  red[1][ty][tx] = s1;
  __syncthreads();
  if (ty == 0 && c < C) {
#pragma unroll
    for (int l = 1; l < 8; ++l) {
      s0 += red[0][l][tx];
      s1 += red[1][l][tx];
    }
    if (out0 != nullptr) out0[c] = s0;
    if (out1 != nullptr) out1[c] = s1;
  }
}

// ------------------------------------------------------------------------------------------------ rotary
// x[N, heads*hd] (row stride ld) in place; one thread per 128-bit vector (bf16: 4 pairs, f32: 2 pairs).
// table[t][j] = (a,b,c,d):  o_even = a*x_even + b*x_odd,  o_odd = c*x_even + d*x_odd;  transpose swaps b and c.
template <typename T>
__global__ void __launch_bounds__(256) rope_kernel(T* __restrict__ x, int64_t ld, int64_t N, int T_len, int heads, int hd,
                                                   const float4* __restrict__ table, int transpose) {
  constexpr int VN = Vec<T>::N;
  const int vec_per_row = heads * hd / VN;
  const int64_t idx = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (idx >= N * vec_per_row) return;
  const int64_t n = idx / vec_per_row;
  const int col = (int)(idx % vec_per_row) * VN;
  const int t = (int)(n % T_len);
  const int pair0 = (col % hd) >> 1;
  T* p = x + n * ld + col;
  Vec<T> v;
  v.load(p);
  float f[VN];
  v.unpack(f);
#pragma unroll
  for (int e = 0; e < VN; e += 2) {
    const float4 m = table[(int64_t)t * (hd >> 1) + pair0 + (e >> 1)];
    const float bb = transpose ? m.z : m.y, cc = transpose ? m.y : m.z;
    const float xe = f[e], xo = f[e + 1];
    f[e] = m.x * xe + bb * xo;
    f[e + 1] = cc * xe + m.w * xo;
  }
  v.pack(f);
  v.store(p);
}

// Head-looping variant: a thread owns (row n, one 128-bit vector position inside a head) and walks over a group of heads.
// Every head of a row uses the same table entries, so they are fetched ONCE into registers; the plain kernel above reads
// 64 B of table per 16 B of data (bf16), which made it run at ~2.9 TB/s.  HPT heads per thread keep enough loads in flight.
template <typename T, int HPT>
__global__ void __launch_bounds__(256) rope_heads_kernel(T* __restrict__ x, int64_t ld, int64_t N, int T_len, int heads, int hd,
                                                         const float4* __restrict__ table, int transpose) {
  constexpr int VN = Vec<T>::N;
  const int vph = hd / VN;  // vectors per head row
  const int64_t idx = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (idx >= N * vph) return;
  const int64_t n = idx / vph;
  const int v = (int)(idx % vph);
  const int t = (int)(n % T_len);
  float4 m[VN / 2];
#pragma unroll
  for (int e = 0; e < VN / 2; ++e) {
    m[e] = table[(int64_t)t * (hd >> 1) + v * (VN / 2) + e];
    if (transpose) {
      const float tmp = m[e].y;
      m[e].y = m[e].z;
      m[e].z = tmp;
    }
  }
  const int h0 = blockIdx.y * HPT;
  T* p = x + n * ld + (int64_t)h0 * hd + v * VN;
  Vec<T> vec[HPT];
#pragma unroll
  for (int h = 0; h < HPT; ++h)
    if (h0 + h < heads) vec[h].load(p + (int64_t)h * hd);
#pragma unroll
  for (int h = 0; h < HPT; ++h) {
    if (h0 + h < heads) {
      float f[VN];
      vec[h].unpack(f);
#pragma unroll
      for (int e = 0; e < VN; e += 2) {
        const float4 mm = m[e >> 1];
        const float xe = f[e], xo = f[e + 1];
        f[e] = mm.x * xe + mm.y * xo;
        f[e + 1] = mm.z * xe + mm.w * xo;
      }
      vec[h].pack(f);
      vec[h].store(p + (int64_t)h * hd);
    }
  }
}

template <typename T>
__global__ void __launch_bounds__(256) rope_generic_kernel(T* __restrict__ x, int64_t ld, int64_t N, int T_len, int heads,
                                                           int hd, const float4* __restrict__ table, int transpose) {
  const int pairs_per_row = heads * hd / 2;
  const int64_t idx = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (idx >= N * pairs_per_row) return;
  const int64_t n = idx / pairs_per_row;
  const int pr = (int)(idx % pairs_per_row);
  const int t = (int)(n % T_len);
  const int j = pr % (hd >> 1);
  T* p = x + n * ld + 2 * pr;
  const float4 m = table[(int64_t)t * (hd >> 1) + j];
  const float bb = transpose ? m.z : m.y, cc = transpose ? m.y : m.z;
  const float xe = to_f32<T>(p[0]), xo = to_f32<T>(p[1]);
  p[0] = from_f32<T>(m.x * xe + bb * xo);
  p[1] = from_f32<T>(cc * xe + m.w * xo);
}

// ------------------------------------------------------------------------------------------------ host launchers
constexpr int COLSUM_MAX_CHUNKS = 592;  // 4 * 148; upper bound of every column-sum plan (sizes the partial scratch)
void colsum_plan(int64_t N, int* chunks, int* rows_per_chunk, int col_blocks = 1);

// (VPT, WPR) for a row of nv 128-bit vectors: at most 4 vectors per thread, up to 8 warps per row
#define SPB_NORM_DISPATCH(nv, CALL) \
  do {                              \
    if ((nv) <= 32) CALL(1, 1);     \
    else if ((nv) <= 64) CALL(2, 1);  \
    else if ((nv) <= 128) CALL(4, 1); \
    else if ((nv) <= 256) CALL(4, 2); \
    else if ((nv) <= 512) CALL(4, 4); \
    else CALL(4, 8);                \
  } while (0)

template <typename T>
static int norm_fwd_launch(const void* x, const void* w, const void* b, void* y, float* stats, int64_t N, int C, float eps,
                           int kind, int upcast, cudaStream_t st) {
  constexpr int VN = Vec<T>::N;
  const bool vec_ok = (C % VN == 0) && (C / VN <= 1024) && ((uintptr_t)x % 16 == 0) && ((uintptr_t)y % 16 == 0) &&
                      ((uintptr_t)w % 16 == 0) && (b == nullptr || (uintptr_t)b % 16 == 0);
  if (vec_ok) {
    const int nv = C / VN;
#define SPB_NF(V, W)                                                                                                      \
  norm_fwd_kernel<T, V, W><<<(unsigned)((N + (8 / W) - 1) / (8 / W)), NR_THREADS, 0, st>>>((const T*)x, (const T*)w, (const T*)b, \
                                                                                           (T*)y, stats, N, C, eps, kind, upcast)
    SPB_NORM_DISPATCH(nv, SPB_NF);
#undef SPB_NF
  } else {
    const unsigned grid = (unsigned)((N + NORM_WARPS - 1) / NORM_WARPS);
    norm_fwd_generic_kernel<T><<<grid, NORM_WARPS * 32, 0, st>>>((const T*)x, (const T*)w, (const T*)b, (T*)y, stats, N, C,
                                                                 eps, kind, upcast);
  }
  SPB_LAUNCH_CHECK();
  return 0;
}

template <typename T>
static int norm_bwd_launch(const void* x, const void* w, const float* stats, const void* dy, void* dx, float* dw, float* db,
                           void* partial, int64_t N, int C, int kind, cudaStream_t st, const void* add) {
  constexpr int VN = Vec<T>::N;
  const bool vec_ok = (C % VN == 0) && (C / VN <= 1024) && ((uintptr_t)x % 16 == 0) && ((uintptr_t)dy % 16 == 0) &&
                      ((uintptr_t)dx % 16 == 0) && ((uintptr_t)w % 16 == 0) && ((uintptr_t)add % 16 == 0);
  if (dx != nullptr && vec_ok && (dw != nullptr || db != nullptr) && partial != nullptr && (uintptr_t)partial % 16 == 0 &&
      (kind == SPB_NORM_LAYER || db == nullptr)) {
    // one fused pass: dx + per-CTA partial sums of dw (db), then the deterministic 2nd stage
    const int nv = C / VN;
    int64_t row_groups = 0;
    int grid = 0, gpc = 1;
#define SPB_NBF_ONE(V, W, L)                                                                                                    \
  do {                                                                                                                          \
    const size_t smem = (size_t)(L ? 2 : 1) * 8 * V * (VN / 4) * 32 * sizeof(float4);                                           \
    if (smem > 48 * 1024)                                                                                                       \
      SPB_CUDA(cudaFuncSetAttribute(norm_bwd_fused_kernel<T, V, W, L>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem)); \
    row_groups = (N + (8 / W) - 1) / (8 / W);                                                                                   \
    /* one full wave of CTAs when there are enough rows; never more partial rows than the scratch holds */                  \
    int occ = 1;                                                                                                                \
    SPB_CUDA(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&occ, norm_bwd_fused_kernel<T, V, W, L>, NR_THREADS, smem));         \
    int64_t want = (int64_t)(occ < 1 ? 1 : occ) * sm_count();                                                                   \
    if (want > COLSUM_MAX_CHUNKS) want = COLSUM_MAX_CHUNKS;                                                                     \
    gpc = (int)((row_groups + want - 1) / want);                                                                                \
    if (gpc < 1) gpc = 1;                                                                                                       \
    grid = (int)((row_groups + gpc - 1) / gpc);                                                                                 \
    norm_bwd_fused_kernel<T, V, W, L><<<grid, NR_THREADS, smem, st>>>((const T*)x, (const T*)w, stats, (const T*)dy, (T*)dx, N, C, \
                                                                      row_groups, gpc, (float*)partial, (const T*)add);         \
  } while (0)
#define SPB_NBF(V, W)                                          \
  do {                                                         \
    if (kind == SPB_NORM_LAYER) SPB_NBF_ONE(V, W, true);       \
    else SPB_NBF_ONE(V, W, false);                             \
  } while (0)
    SPB_NORM_DISPATCH(nv, SPB_NBF);
#undef SPB_NBF
#undef SPB_NBF_ONE
    SPB_LAUNCH_CHECK();
    colsum_final_kernel<<<(C + 31) / 32, 256, 0, st>>>((const float*)partial, grid, C, dw, kind == SPB_NORM_LAYER ? db : nullptr);
    SPB_LAUNCH_CHECK();
    return 0;
  }
  if (dx != nullptr) {
    if (vec_ok) {
      const int nv = C / VN;
#define SPB_NB(V, W)                                                                                                         \
  norm_bwd_dx_kernel<T, V, W><<<(unsigned)((N + (8 / W) - 1) / (8 / W)), NR_THREADS, 0, st>>>((const T*)x, (const T*)w, stats, \
                                                                                              (const T*)dy, (T*)dx, N, C, kind, (const T*)add)
      SPB_NORM_DISPATCH(nv, SPB_NB);
#undef SPB_NB
    } else {
      const unsigned grid = (unsigned)((N + NORM_WARPS - 1) / NORM_WARPS);
      norm_bwd_dx_generic_kernel<T><<<grid, NORM_WARPS * 32, 0, st>>>((const T*)x, (const T*)w, stats, (const T*)dy, (T*)dx,
                                                                      N, C, kind, (const T*)add);
    }
    SPB_LAUNCH_CHECK();
  }
  if (dw != nullptr || db != nullptr) {
    SPB_CHECK_ARG(partial != nullptr, SPB_E_INVALID, "norm_bwd: partial scratch is NULL");
    int chunks, rpc;
    colsum_plan(N, &chunks, &rpc);
    if (C % VN == 0 && (uintptr_t)dy % 16 == 0 && (uintptr_t)x % 16 == 0 && (uintptr_t)partial % 16 == 0) {
      const int tpr = colsum_tpr(C / VN);
      colsum_plan(N, &chunks, &rpc, (C / VN + tpr - 1) / tpr);
      dim3 grid((C / VN + tpr - 1) / tpr, chunks);
      colsum_partial_vec_kernel<T, 1><<<grid, 256, 0, st>>>((const T*)dy, C, (const T*)x, stats, kind, N, C, rpc, chunks, tpr,
                                                            (float*)partial);
    } else {
      dim3 grid((C + 255) / 256, chunks);
      colsum_partial_kernel<T, 1><<<grid, 256, 0, st>>>((const T*)dy, C, (const T*)x, stats, kind, N, C, rpc, chunks,
                                                        (float*)partial);
    }
    SPB_LAUNCH_CHECK();
    colsum_final_kernel<<<(C + 31) / 32, 256, 0, st>>>((const float*)partial, chunks, C, dw, db);
    SPB_LAUNCH_CHECK();
  }
  return 0;
}

void colsum_plan(int64_t N, int* chunks, int* rows_per_chunk, int col_blocks) {
  // enough chunks to fill the machine for narrow matrices, few enough that the 2nd stage (one thread per column looping
  // over the chunks) stays negligible for wide ones: about 4 waves of blocks in total
  int64_t ch = (N + 31) / 32;
  int64_t cap = 592 / (col_blocks < 1 ? 1 : col_blocks);  // 4 * 148 blocks
  if (cap < 16) cap = 16;
  if (cap > 592) cap = 592;
  if (ch > cap) ch = cap;
  if (ch < 1) ch = 1;
  const int64_t rpc = (N + ch - 1) / ch;
  *rows_per_chunk = (int)(rpc < 1 ? 1 : rpc);
  *chunks = (int)((N + *rows_per_chunk - 1) / *rows_per_chunk);
  if (*chunks < 1) *chunks = 1;
}

template <typename T>
static int colsum_launch(const void* x, int64_t ld, int64_t N, int cols, float* out, void* partial, cudaStream_t st) {
  constexpr int VN = Vec<T>::N;
  int chunks, rpc;
  colsum_plan(N, &chunks, &rpc);
  if (cols % VN == 0 && ld % VN == 0 && (uintptr_t)x % 16 == 0 && (uintptr_t)partial % 16 == 0) {
    const int tpr = colsum_tpr(cols / VN);
    colsum_plan(N, &chunks, &rpc, (cols / VN + tpr - 1) / tpr);
    dim3 grid((cols / VN + tpr - 1) / tpr, chunks);
    colsum_partial_vec_kernel<T, 0><<<grid, 256, 0, st>>>((const T*)x, ld, nullptr, nullptr, 0, N, cols, rpc, chunks, tpr,
                                                          (float*)partial);
  } else {
    dim3 grid((cols + 255) / 256, chunks);
    colsum_partial_kernel<T, 0><<<grid, 256, 0, st>>>((const T*)x, ld, nullptr, nullptr, 0, N, cols, rpc, chunks, (float*)partial);
  }
  SPB_LAUNCH_CHECK();
  colsum_final_kernel<<<(cols + 31) / 32, 256, 0, st>>>((const float*)partial, chunks, cols, out, nullptr);
  SPB_LAUNCH_CHECK();
  return 0;
}

template <typename T>
static int rope_launch(void* x, int64_t ld, int64_t N, int T_len, int heads, int hd, const float* table, int transpose,
                       cudaStream_t st) {
  constexpr int VN = Vec<T>::N;
  const bool vec_ok = (hd % VN == 0) && (ld % VN == 0) && ((uintptr_t)x % 16 == 0);
  if (vec_ok && heads >= 4) {
    constexpr int HPT = 8;
    const int64_t total = N * (hd / VN);
    dim3 grid((unsigned)((total + 255) / 256), (unsigned)((heads + HPT - 1) / HPT));
    rope_heads_kernel<T, HPT><<<grid, 256, 0, st>>>((T*)x, ld, N, T_len, heads, hd, (const float4*)table, transpose);
  } else if (vec_ok) {
    const int64_t total = N * (heads * hd / VN);
    rope_kernel<T><<<(unsigned)((total + 255) / 256), 256, 0, st>>>((T*)x, ld, N, T_len, heads, hd, (const float4*)table,
                                                                     transpose);
  } else {
    const int64_t total = N * (heads * hd / 2);
    rope_generic_kernel<T><<<(unsigned)((total + 255) / 256), 256, 0, st>>>((T*)x, ld, N, T_len, heads, hd,
                                                                             (const float4*)table, transpose);
  }
  SPB_LAUNCH_CHECK();
  return 0;
}

// ------------------------------------------------------------------------------------------------ allocation-free column sum
// Scratch for spb_colsum_static lives in the library image itself, so a caller that cannot obtain device memory can still
// run the two-stage column sum.  4 MiB per GPU that loads the library, resident until the process exits.
constexpr size_t COLSUM_STATIC_FLOATS = (size_t)1 << 20;
constexpr int COLSUM_STATIC_MAX_COLS = 8192;  // => at least 128 row chunks always fit
constexpr int COLSUM_STATIC_MAX_DEVICES = 64;
static __device__ __align__(256) float g_colsum_static_scratch[COLSUM_STATIC_FLOATS];

// In host code `g_colsum_static_scratch` is only a handle for the symbol, never an address a kernel may use: the device
// address is looked up once per device (with that device current) and cached.
static int colsum_static_scratch_addr(float** out) {
  static std::mutex mu;
  static float* addr[COLSUM_STATIC_MAX_DEVICES] = {};
  int dev = 0;
  SPB_CUDA(cudaGetDevice(&dev));
  SPB_CHECK_ARG(dev >= 0 && dev < COLSUM_STATIC_MAX_DEVICES, SPB_E_LIMIT, "spb_colsum_static: device ordinal %d beyond %d", dev,
                COLSUM_STATIC_MAX_DEVICES);
  std::lock_guard<std::mutex> lk(mu);
  if (addr[dev] == nullptr) {
    float* p = nullptr;
    SPB_CUDA(cudaGetSymbolAddress((void**)&p, g_colsum_static_scratch));
    SPB_CHECK_ARG(p != nullptr, SPB_E_STATE, "spb_colsum_static: cudaGetSymbolAddress returned a null address");
    addr[dev] = p;
  }
  *out = addr[dev];
  return 0;
}

template <typename T>
static int colsum_static_launch(const void* x, int64_t ld, int64_t N, int cols, float* out, float* scratch, cudaStream_t st) {
  int chunks, rpc;
  colsum_plan(N, &chunks, &rpc);
  const int max_chunks = (int)(COLSUM_STATIC_FLOATS / (size_t)cols);  // >= 128 because cols <= 8192
  if (chunks > max_chunks) {
    rpc = (int)((N + max_chunks - 1) / max_chunks);
    chunks = (int)((N + rpc - 1) / rpc);
  }
  dim3 grid((cols + 255) / 256, chunks);
  colsum_partial_kernel<T, 0><<<grid, 256, 0, st>>>((const T*)x, ld, nullptr, nullptr, 0, N, cols, rpc, chunks, scratch);
  SPB_LAUNCH_CHECK();
  colsum_final_kernel<<<(cols + 31) / 32, 256, 0, st>>>(scratch, chunks, cols, out, nullptr);
  SPB_LAUNCH_CHECK();
  return 0;
}

// spb_norm_bwd with an optional term added to dx: `add` [N,C] (same dtype) is the gradient arriving through the residual
// branch h = x + block(x) (llama3:885), fused into the dx store instead of a separate 3-pass add kernel.
int norm_bwd_add(int dtype, int norm_kind, int upcast, const void* x, const void* w, const float* stats, const void* dy, void* dx,
                 float* dw, float* db, void* partial, int64_t N, int32_t C, const void* add, cudaStream_t st) {
  (void)upcast;  // the VJP ignores the intermediate rounding of the gemma variant
  SPB_CHECK_ARG(N >= 0 && C > 0, SPB_E_INVALID, "spb_norm_bwd: bad shape");
  SPB_CHECK_ARG(norm_kind == SPB_NORM_RMS || norm_kind == SPB_NORM_LAYER, SPB_E_INVALID, "spb_norm_bwd: bad norm_kind %d", norm_kind);
  SPB_CHECK_ARG(N > 0, SPB_E_INVALID, "spb_norm_bwd: empty input (the weight gradient of an empty batch is undefined here)");
  SPB_CHECK_ARG(x && w && stats && dy, SPB_E_INVALID, "spb_norm_bwd: null pointer");
  if (dtype == SPB_F32) return norm_bwd_launch<float>(x, w, stats, dy, dx, dw, db, partial, N, C, norm_kind, st, add);
  if (dtype == SPB_BF16) return norm_bwd_launch<__nv_bfloat16>(x, w, stats, dy, dx, dw, db, partial, N, C, norm_kind, st, add);
  SPB_CHECK_ARG(false, SPB_E_UNSUPPORTED, "spb_norm_bwd: dtype %d", dtype);
}

}  // namespace spb



// ------------------------------------------------------------------------------------------------ C ABI
extern "C" {

int spb_norm_fwd(int dtype, int norm_kind, int upcast, float eps, const void* x, const void* w, const void* b, void* y,
                 float* stats, int64_t N, int32_t C, spb_stream_t stream) {
  SPB_CHECK_ARG(N >= 0 && C > 0, SPB_E_INVALID, "spb_norm_fwd: bad shape N=%lld C=%d", (long long)N, C);
  SPB_CHECK_ARG(norm_kind == SPB_NORM_RMS || norm_kind == SPB_NORM_LAYER, SPB_E_INVALID, "spb_norm_fwd: bad norm_kind %d", norm_kind);
  if (N == 0) return 0;  // empty input: nothing to do (and its pointers may legitimately be NULL)
  SPB_CHECK_ARG(x && w && y, SPB_E_INVALID, "spb_norm_fwd: null pointer");
  cudaStream_t st = (cudaStream_t)stream;
  if (dtype == SPB_F32) return spb::norm_fwd_launch<float>(x, w, b, y, stats, N, C, eps, norm_kind, upcast, st);
  if (dtype == SPB_BF16) return spb::norm_fwd_launch<__nv_bfloat16>(x, w, b, y, stats, N, C, eps, norm_kind, upcast, st);
  SPB_CHECK_ARG(false, SPB_E_UNSUPPORTED, "spb_norm_fwd: dtype %d", dtype);
}

size_t spb_norm_bwd_partial_bytes(int64_t N, int32_t C) {
  // 2 planes (dweight, dbias) x the most partial rows any plan uses: the column-sum chunks or the fused kernel's CTAs
  int chunks, rpc;
  spb::colsum_plan(N < 1 ? 1 : N, &chunks, &rpc);
  const int64_t groups = N < 1 ? 1 : N;  // fused kernel: <= one CTA per row group, capped
  int64_t rows = groups < spb::COLSUM_MAX_CHUNKS ? groups : spb::COLSUM_MAX_CHUNKS;
  if (rows < chunks) rows = chunks;
  return (size_t)2 * (size_t)rows * (size_t)C * sizeof(float);
}

int spb_norm_bwd(int dtype, int norm_kind, int upcast, const void* x, const void* w, const float* stats, const void* dy,
                 void* dx, float* dw, float* db, void* partial, int64_t N, int32_t C, spb_stream_t stream) {
  return spb::norm_bwd_add(dtype, norm_kind, upcast, x, w, stats, dy, dx, dw, db, partial, N, C, nullptr, (cudaStream_t)stream);
}

int spb_colsum(int dtype, const void* x, int64_t ld, int64_t N, int32_t cols, float* out, void* partial, spb_stream_t stream) {
  SPB_CHECK_ARG(x && out && partial, SPB_E_INVALID, "spb_colsum: null pointer");
  SPB_CHECK_ARG(N > 0 && cols > 0 && ld >= cols, SPB_E_INVALID, "spb_colsum: bad shape");
  cudaStream_t st = (cudaStream_t)stream;
  if (dtype == SPB_F32) return spb::colsum_launch<float>(x, ld, N, cols, out, partial, st);
  if (dtype == SPB_BF16) return spb::colsum_launch<__nv_bfloat16>(x, ld, N, cols, out, partial, st);
  SPB_CHECK_ARG(false, SPB_E_UNSUPPORTED, "spb_colsum: dtype %d", dtype);
}

int spb_rope_apply(int dtype, void* x, int64_t ld, int64_t N, int32_t T, int32_t heads, int32_t hd, const float* table,
                   int transpose, spb_stream_t stream) {
  SPB_CHECK_ARG(x && table, SPB_E_INVALID, "spb_rope_apply: null pointer");
  SPB_CHECK_ARG(N >= 0 && T > 0 && heads > 0 && hd > 0 && hd % 2 == 0 && ld >= (int64_t)heads * hd, SPB_E_INVALID,
                "spb_rope_apply: bad shape (hd must be even)");
  SPB_CHECK_ARG((uintptr_t)table % 16 == 0, SPB_E_INVALID, "spb_rope_apply: table must be 16-byte aligned");
  if (N == 0) return 0;
  cudaStream_t st = (cudaStream_t)stream;
  if (dtype == SPB_F32) return spb::rope_launch<float>(x, ld, N, T, heads, hd, table, transpose, st);
  if (dtype == SPB_BF16) return spb::rope_launch<__nv_bfloat16>(x, ld, N, T, heads, hd, table, transpose, st);
  SPB_CHECK_ARG(false, SPB_E_UNSUPPORTED, "spb_rope_apply: dtype %d", dtype);
}

// Column sum for callers that have no way to get scratch memory: out[c] = sum_n x[n, c], fp32.  See include/spb200.h.
int spb_colsum_static(int dtype, const void* x, int64_t ld, int64_t N, int32_t cols, float* out, spb_stream_t stream) {
  SPB_CHECK_ARG(x && out, SPB_E_INVALID, "spb_colsum_static: null pointer");
  SPB_CHECK_ARG(N > 0 && cols > 0 && ld >= cols, SPB_E_INVALID, "spb_colsum_static: bad shape");
  SPB_CHECK_ARG(cols <= spb::COLSUM_STATIC_MAX_COLS, SPB_E_LIMIT,
                "spb_colsum_static: %d columns is beyond the fixed scratch (at most %d columns; use spb_colsum with your own "
                "scratch for wider matrices)", cols, spb::COLSUM_STATIC_MAX_COLS);
  float* scratch = nullptr;
  int rc = spb::colsum_static_scratch_addr(&scratch);  // first call on a device resolves the symbol's device address
  if (rc) return rc;
  cudaStream_t st = (cudaStream_t)stream;
  if (dtype == SPB_F32) return spb::colsum_static_launch<float>(x, ld, N, cols, out, scratch, st);
  if (dtype == SPB_BF16) return spb::colsum_static_launch<__nv_bfloat16>(x, ld, N, cols, out, scratch, st);
  SPB_CHECK_ARG(false, SPB_E_UNSUPPORTED, "spb_colsum_static: dtype %d", dtype);
}

}  // extern "C"
