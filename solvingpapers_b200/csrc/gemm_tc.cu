// gemm_tc.cu - K2: the projections (llama3:812-814,829) and their VJPs as ONE hand-written tcgen05 GEMM.
//   D[M,N] = opA(A) . opB(B) (+bias) (+D),  bf16 operands, fp32 accumulation in TMEM, bf16 or fp32 output.
// Design (sm_100a):
//   * persistent CTAs (grid = min(work items, #SM)), 128 x BN output tile (BN = 256 or 128), BK = 64;
//   * warp 0 = TMA producer (cp.async.bulk.tensor, 128-byte swizzle, 4-6 stage mbarrier ring),
//     warp 1 = MMA issuer (one thread, tcgen05.mma kind::f16, M=128 N=BN K=16) and TMEM owner,
//     warps 2-5 = epilogue (tcgen05.ld 32x32b -> bias / accumulate / convert -> global), double-buffered TMEM
//     accumulators (2 x BN columns) so the epilogue of tile i overlaps the main loop of tile i+1;
//   * both operand majors without transpose copies: the [in,out] JAX weights (B is N-contiguous), X^T / dY^T for dW
//     (both MN-major) are expressed by the UMMA major bits + a TMA box per 64-column chunk (SURVEY 8b weight-layout note);
//   * split-K (fp32 atomic accumulate into a zeroed output) when a weight-gradient GEMM has fewer tiles than SMs
//     (d=256 models: dW is 2x3 tiles with K = 65536 tokens).
// Roofline: tensor pipe. Algorithmic FLOPs 2*M*N*K.
#include <cstdlib>

#include "common.cuh"
#include "kernels.h"

namespace spb {

constexpr int GT_BM = 128, GT_BK = 64, GT_THREADS = 192;

struct GemmTcArgs {
  void* D;
  int64_t ldd;
  const __nv_bfloat16* bias;
  const float4* rope;  // optional fused rotary epilogue (llama3:592-601): table [T, hd/2] of (a,b,c,d), applied to columns < rope_cols
  int rope_cols, rope_T, rope_hd;
  int out_f32, accumulate, atomic, vec_ok;
  int tma_store;       // plain (non-accumulating) output tiles leave through shared memory + TMA bulk stores (see gemm_tc_tma_store_chunk)
  int bias_vec;        // bias pointer is 16-byte aligned
  int bias_f32;        // bias elements are fp32 (else bf16)
  int group_m;         // tile raster: row-blocks per group (see gemm_tc_tile_coords)
  const void* addsrc;  // accumulate: what is added to the product - D itself (NULL) or a separate [M,N] array of D's type
  int64_t ld_add;      //            (the residual stream: y = x + attn . wo)
  int M, N;
  int a_mn, b_mn;
  // optional attention-backward preparation fused into the d(attn_o) = dy . wo^T GEMM (GemmTcFuse): while a thread holds the
  // fp32 accumulators of its row it also forms delta[b,h,t] = sum_c dO[n,c] * O[n,c] per head and zeroes the fp32 dq accumulator
  const __nv_bfloat16* fuse_o;
  int64_t fuse_ldo;
  float* fuse_delta;
  float* fuse_zero;
  int64_t fuse_ldz;
  int fuse_hd, fuse_T, fuse_H;
  // loss-head epilogues (GemmTcHead): mode 1 = no D store, per (row, 256-column tile) online (max, sum exp) partials +
  // the target's logit; mode 2 = D = (exp(acc - lse[row]) - [col == target[row]]) * gscale (dlogits, bf16)
  int head_mode;
  float2* head_part;
  float* head_tgt;
  const int64_t* head_targets;
  const float* head_lse;
  float head_gscale;
  // work items: the first `full_tiles` tiles (whole waves) run over the whole K range; the remaining tail tiles are each
  // split along K into `tail_parts` parts that add into a pre-zeroed fp32 output with atomics, so the last, partial wave
  // costs 1/tail_parts of a wave (also covers few-tile weight gradients: full_tiles = 0).
  int num_m, num_n, kb_total, full_tiles, tail_parts, kb_per_part, total_items;
};

struct GemmTcItem {
  int m0, n0, kb0, kb1, atomic;
};

// Tile raster: groups of GT_GROUP_M row-blocks, row-block fastest inside a group, so that the ~148 concurrently resident
// tiles share a few A row-panels and B column-panels that stay in the 126 MB L2 (ncu r01b: 5.7x DRAM re-reads without it).
constexpr int GT_GROUP_M = 16;
static int gemm_group_m(int dflt) {  // SPB_GEMM_GROUP_M: timing experiments
  static const int env = [] {
    const char* e = getenv("SPB_GEMM_GROUP_M");
    return e ? atoi(e) : 0;
  }();
  return env > 0 ? env : dflt;
}
__device__ __forceinline__ void gemm_tc_tile_coords(const GemmTcArgs& p, int tile, int& m_blk, int& n_blk) {
  const int group_sz = p.group_m * p.num_n;
  const int gid = tile / group_sz;
  const int first_m = gid * p.group_m;
  const int gm = min(p.num_m - first_m, p.group_m);
  const int in_group = tile - gid * group_sz;
  m_blk = first_m + in_group % gm;
  n_blk = in_group / gm;
}
template <int BN>
__device__ __forceinline__ GemmTcItem gemm_tc_item(const GemmTcArgs& p, int item) {
  GemmTcItem it;
  int tile = item, part = 0;
  it.atomic = 0;
  it.kb0 = 0;
  it.kb1 = p.kb_total;
  if (item >= p.full_tiles) {
    const int j = item - p.full_tiles;
    tile = p.full_tiles + j / p.tail_parts;
    part = j % p.tail_parts;
    it.kb0 = part * p.kb_per_part;
    it.kb1 = min(it.kb0 + p.kb_per_part, p.kb_total);
    it.atomic = p.tail_parts > 1;
  }
  int m_blk, n_blk;
  gemm_tc_tile_coords(p, tile, m_blk, n_blk);
  it.m0 = m_blk * GT_BM;
  it.n0 = n_blk * BN;
  return it;
}

template <int BN>
struct GemmTcCfg {
  static constexpr int STAGES = (BN == 256) ? 4 : (BN == 192 ? 4 : 6);
  static constexpr int A_BYTES = GT_BM * GT_BK * 2;
  static constexpr int B_BYTES = BN * GT_BK * 2;
  static constexpr int STAGE_BYTES = A_BYTES + B_BYTES;
  static constexpr int BAR_BYTES = 256;
  // rotary epilogue: each epilogue warp stages the table slice of its 32 rows x 16 pairs, pair-major with an XOR swizzle:
  // entry (row r, pair i) at 16-byte slot i*32 + (r ^ i) - both the cooperative writes and the per-row reads are conflict-free
  static constexpr int ROPE_WARP_BYTES = 32 * 16 * 16;
  static constexpr int ROPE_BYTES = 4 * ROPE_WARP_BYTES;
  static constexpr int SMEM_BYTES = STAGES * STAGE_BYTES + BAR_BYTES + ROPE_BYTES + 1024;  // +1024: manual 1 KB alignment
  static constexpr int TMEM_COLS = (BN == 192) ? 512 : 2 * BN;  // power of two; accumulator stage a sits at column a*BN
};

// Output through shared memory + TMA (cp.async.bulk.tensor store).  A thread owns one output ROW of the chunk, so direct stores
// are 32 scattered 16-byte pieces per warp instruction - partial sectors on 32 different lines - and the epilogue, not the
// main loop, sets the pace of short-K GEMMs (K = 768: 13 k clk per 256x256 tile against a 6 k clk main loop).  Here the warp
// stages its 32 rows in its shared-memory buffer (rows of 128 B, 128B-swizzled: conflict-free) and one TMA store writes
// whole lines; the TMA unit clips rows / columns past M / N.  bf16: two chunks (64 columns) per store; fp32: one (32 columns).
__device__ __forceinline__ void gemm_tc_tma_store_chunk(const GemmTcArgs& p, const CUtensorMap* tmD, int row0, int col0, const float* v,
                                                        uint8_t* wbuf2, int c, int lane, bool last) {
  if (p.out_f32) {
    uint8_t* wbuf = wbuf2 + (c & 1) * 4096;
    if (lane == 0) tma_store_wait_read<1>();  // the store issued two chunks ago has finished reading this buffer
    __syncwarp();
    uint8_t* rowp = wbuf + lane * 128;
#pragma unroll
    for (int u = 0; u < 8; ++u)
      *reinterpret_cast<float4*>(rowp + ((u ^ (lane & 7)) << 4)) = make_float4(v[4 * u], v[4 * u + 1], v[4 * u + 2], v[4 * u + 3]);
    fence_proxy_async_smem();
    __syncwarp();
    if (lane == 0) {
      tma_store_2d(tmD, wbuf, col0, row0);
      tma_store_commit();
    }
  } else {
    const int half = c & 1;  // which 64-byte half of the 128-byte staged row this chunk fills
    uint8_t* wbuf = wbuf2 + ((c >> 1) & 1) * 4096;
    if (half == 0) {
      if (lane == 0) tma_store_wait_read<1>();
      __syncwarp();
    }
    uint8_t* rowp = wbuf + lane * 128;
#pragma unroll
    for (int u = 0; u < 4; ++u) {
      uint4 o;
      o.x = pack_bf16(v[8 * u], v[8 * u + 1]);
      o.y = pack_bf16(v[8 * u + 2], v[8 * u + 3]);
      o.z = pack_bf16(v[8 * u + 4], v[8 * u + 5]);
      o.w = pack_bf16(v[8 * u + 6], v[8 * u + 7]);
      *reinterpret_cast<uint4*>(rowp + (((half * 4 + u) ^ (lane & 7)) << 4)) = o;
    }
    if (half == 1 || last) {  // `last` with half == 0: the tile ends inside this 64-column box; the stale right half lies past N and is clipped
      fence_proxy_async_smem();
      __syncwarp();
      if (lane == 0) {
        tma_store_2d(tmD, wbuf, col0 - half * 32, row0);
        tma_store_commit();
      }
    }
  }
}

__device__ __forceinline__ void gemm_tc_store_chunk(const GemmTcArgs& p, int row, int col0, const uint32_t* r,
                                                    const uint8_t* rope_row, int lane_in_warp, int atomic, const CUtensorMap* tmD = nullptr,
                                                    uint8_t* wbuf2 = nullptr, int c = 0, int row0 = 0, bool last = false,
                                                    const uint4* bias_pre = nullptr) {
  // 32 consecutive fp32 accumulator columns of one output row
  const int ncol = min(32, p.N - col0);
  float v[32];
#pragma unroll
  for (int i = 0; i < 32; ++i) v[i] = __uint_as_float(r[i]);
  if (p.bias != nullptr && p.bias_f32) {  // fp32 mode (gemm_split.cu): the bias stays fp32
    const float* bf = reinterpret_cast<const float*>(p.bias) + col0;
#pragma unroll
    for (int i = 0; i < 32; ++i)
      if (i < ncol) v[i] += __ldg(bf + i);
  } else if (p.bias != nullptr) {
    if (ncol == 32 && p.bias_vec) {  // 16-byte aligned bias: four 128-bit loads instead of 32 scalar ones
      const uint4* bp = reinterpret_cast<const uint4*>(p.bias + col0);
#pragma unroll
      for (int i = 0; i < 4; ++i) {
        const uint4 b = bias_pre != nullptr ? bias_pre[i] : __ldg(bp + i);  // preloaded while the accumulators were on their way
        const float2 b0 = unpack_bf16(b.x), b1 = unpack_bf16(b.y), b2 = unpack_bf16(b.z), b3 = unpack_bf16(b.w);
        v[8 * i] += b0.x; v[8 * i + 1] += b0.y; v[8 * i + 2] += b1.x; v[8 * i + 3] += b1.y;
        v[8 * i + 4] += b2.x; v[8 * i + 5] += b2.y; v[8 * i + 6] += b3.x; v[8 * i + 7] += b3.y;
      }
    } else {
#pragma unroll
      for (int i = 0; i < 32; ++i)
        if (i < ncol) v[i] += __bfloat162float(p.bias[col0 + i]);
    }
  }
  if (rope_row != nullptr) {
    // adjacent pairs (2j, 2j+1) of one head live in this thread's registers; rotate the fp32 accumulators in place.
    // rope_row = the warp's staged table slice (32 rows x 16 pairs, swizzled), loaded coalesced by the warp.
#pragma unroll
    for (int i = 0; i < 32; i += 2) {
      if (col0 + i < p.rope_cols && i + 1 < ncol) {
        const float4 m = *reinterpret_cast<const float4*>(rope_row + ((((i >> 1) << 5) + (lane_in_warp ^ (i >> 1))) << 4));
        const float xe = v[i], xo = v[i + 1];
        v[i] = m.x * xe + m.y * xo;
        v[i + 1] = m.z * xe + m.w * xo;
      }
    }
  }
  if (p.tma_store) {
    gemm_tc_tma_store_chunk(p, tmD, row0, col0, v, wbuf2, c, lane_in_warp, last);
    return;
  }
  if (p.out_f32) {
    float* d = reinterpret_cast<float*>(p.D) + (int64_t)row * p.ldd + col0;
    const float* a = p.addsrc ? reinterpret_cast<const float*>(p.addsrc) + (int64_t)row * p.ld_add + col0 : d;
    if (atomic) {
      // handled by gemm_tc_atomic_chunk (warp-cooperative, coalesced); never reached
    } else if (p.vec_ok && ncol == 32) {
#pragma unroll
      for (int i = 0; i < 32; i += 4) {
        float4 o = make_float4(v[i], v[i + 1], v[i + 2], v[i + 3]);
        if (p.accumulate) {
          const float4 old = *reinterpret_cast<const float4*>(a + i);
          o.x += old.x; o.y += old.y; o.z += old.z; o.w += old.w;
        }
        *reinterpret_cast<float4*>(d + i) = o;
      }
    } else {
#pragma unroll
      for (int i = 0; i < 32; ++i)
        if (i < ncol) d[i] = p.accumulate ? a[i] + v[i] : v[i];
    }
  } else {
    __nv_bfloat16* d = reinterpret_cast<__nv_bfloat16*>(p.D) + (int64_t)row * p.ldd + col0;
    const __nv_bfloat16* a = p.addsrc ? reinterpret_cast<const __nv_bfloat16*>(p.addsrc) + (int64_t)row * p.ld_add + col0 : d;
    if (p.vec_ok && ncol == 32) {
#pragma unroll
      for (int i = 0; i < 32; i += 8) {
        if (p.accumulate) {
          const uint4 old = *reinterpret_cast<const uint4*>(a + i);
          const float2 a0 = unpack_bf16(old.x), b = unpack_bf16(old.y), c = unpack_bf16(old.z), e = unpack_bf16(old.w);
          v[i] += a0.x; v[i + 1] += a0.y; v[i + 2] += b.x; v[i + 3] += b.y;
          v[i + 4] += c.x; v[i + 5] += c.y; v[i + 6] += e.x; v[i + 7] += e.y;
        }
        uint4 o;
        o.x = pack_bf16(v[i], v[i + 1]);
        o.y = pack_bf16(v[i + 2], v[i + 3]);
        o.z = pack_bf16(v[i + 4], v[i + 5]);
        o.w = pack_bf16(v[i + 6], v[i + 7]);
        *reinterpret_cast<uint4*>(d + i) = o;
      }
    } else {
#pragma unroll
      for (int i = 0; i < 32; ++i)
        if (i < ncol) d[i] = __float2bfloat16_rn(p.accumulate ? __bfloat162float(a[i]) + v[i] : v[i]);
    }
  }
}

// K-split tail tiles: the warp's 32x32 fp32 chunk goes to shared memory (rows of 128 B, 128B-swizzled: conflict-free) and
// ONE TMA bulk reduce-add (cp.reduce.async.bulk.tensor) adds it to the zeroed output tile; the TMA unit clips rows and
// columns past M / N.  Per-lane fp32 atomics run at ~1 element per clock and SM: with them a half-K tail part took as long
// as a full tile and the split bought nothing (dwo at C3: 4096x4096x8192, 3.46 waves, 1293 TFLOP/s = the 4-wave time).
__device__ __forceinline__ void gemm_tc_reduce_chunk(const CUtensorMap* tmD, int row0, int col0, const uint32_t* r, uint8_t* wbuf2, int c,
                                                     int lane) {
  uint8_t* wbuf = wbuf2 + (c & 1) * 4096;   // two 4 KB staging buffers per warp: a reduce is in flight while the next chunk is staged
  if (lane == 0) tma_store_wait_read<1>();  // the reduce issued two chunks ago has finished reading this buffer
  __syncwarp();
  uint8_t* rowp = wbuf + lane * 128;
#pragma unroll
  for (int u = 0; u < 8; ++u)
    *reinterpret_cast<uint4*>(rowp + ((u ^ (lane & 7)) << 4)) = make_uint4(r[4 * u], r[4 * u + 1], r[4 * u + 2], r[4 * u + 3]);
  fence_proxy_async_smem();
  __syncwarp();
  if (lane == 0) {
    tma_reduce_add_2d(tmD, wbuf, col0, row0);
    tma_store_commit();
  }
}
// fallback when the output is not TMA-addressable (unaligned base or row pitch): coalesced fp32 atomics through a pitch-33
// shared-memory transpose
__device__ __forceinline__ void gemm_tc_atomic_chunk(const GemmTcArgs& p, int row0, int col0, const uint32_t* r, float* wbuf, int lane) {
  __syncwarp();
#pragma unroll
  for (int i = 0; i < 32; ++i) wbuf[lane * 33 + i] = __uint_as_float(r[i]);
  __syncwarp();
  const int col = col0 + lane;
  if (col < p.N) {
    float* d = reinterpret_cast<float*>(p.D) + (int64_t)row0 * p.ldd + col;
#pragma unroll 4
    for (int rr = 0; rr < 32; ++rr)
      if (row0 + rr < p.M) atomicAdd(d + (int64_t)rr * p.ldd, wbuf[rr * 33 + lane]);
  }
}

// Drain one 128 x NCOLS fp32 accumulator (this warp: TMEM lanes q*32 .. q*32+31 = tile rows) through the epilogue.
// m0 = first output row of the CTA's 128-row slab, n0 = first output column; warp_buf = this warp's 8 KB staging buffer.
template <int NCOLS>
__device__ __forceinline__ void gemm_tc_drain(const GemmTcArgs& p, const CUtensorMap* tmD, int atomic, int m0, int n0, uint32_t taddr,
                                              uint8_t* warp_buf, int q, int lane) {
  const int row = m0 + q * 32 + lane;
  if (m0 + q * 32 >= p.M) return;  // warp-uniform: this warp's 32 rows lie past the last row (small-M problems)
  if (p.tma_store || atomic) {
    // the staging halves alternate with the chunk index; a tile that ended on an odd count (N edge, 192-wide tiles) would hand
    // the next tile a half whose store may still be reading it - start every tile with both halves free
    if (lane == 0) tma_store_wait_read<0>();
    __syncwarp();
  }
  float dsum = 0.f;  // running sum_c dO * O of the current head (fuse_o != NULL)
  float hm = -INFINITY, hs = 0.f;  // loss-head mode 1: online (max, sum exp) of this row over the tile's columns
#pragma unroll 1
      for (int c = 0; c < NCOLS / 32; ++c) {
        const int col0 = n0 + c * 32;
        if (col0 >= p.N) break;  // warp-uniform
        uint32_t r[32];
        SPB_TMEM_LD_32(taddr + (uint32_t)(c * 32), r);
        uint4 ov[4];
        const bool fuse = p.fuse_o != nullptr && row < p.M;  // N % 32 == 0 and 16-byte aligned rows are checked by the launcher
        if (fuse) {  // the O chunk of this row: in flight while the accumulators arrive from tensor memory
          const uint4* op = reinterpret_cast<const uint4*>(p.fuse_o + (int64_t)row * p.fuse_ldo + col0);
#pragma unroll
          for (int i = 0; i < 4; ++i) ov[i] = __ldg(op + i);
          float4* zp = reinterpret_cast<float4*>(p.fuse_zero + (int64_t)row * p.fuse_ldz + col0);
#pragma unroll
          for (int i = 0; i < 8; ++i) zp[i] = make_float4(0.f, 0.f, 0.f, 0.f);
        }
        const uint8_t* rope_row = nullptr;
        if (p.rope != nullptr && col0 < p.rope_cols) {  // warp-uniform
          // stage table[t(row)][pair0 .. pair0+15] for the warp's 32 rows: 2 rows (2 x 256 B) per load instruction
          uint8_t* wbuf = warp_buf;
          const int pair0 = (col0 % p.rope_hd) >> 1;
          const int half = lane >> 4, l16 = lane & 15;
          // phase 1: all 16 coalesced global loads in flight (the table is L2 resident); phase 2: swizzled smem stores
          float4 tm[16];
          const bool pair_ok = pair0 + l16 < (p.rope_hd >> 1);
#pragma unroll
          for (int k = 0; k < 16; ++k) {
            const int grow = m0 + q * 32 + 2 * k + half;
            tm[k] = make_float4(1.f, 0.f, 0.f, 1.f);
            if (grow < p.M && pair_ok) tm[k] = __ldg(p.rope + (int64_t)(grow % p.rope_T) * (p.rope_hd >> 1) + pair0 + l16);
          }
          __syncwarp();  // previous chunk's readers are done with the buffer
#pragma unroll
          for (int k = 0; k < 16; ++k)
            *reinterpret_cast<float4*>(wbuf + (((l16 << 5) + ((2 * k + half) ^ l16)) << 4)) = tm[k];
          __syncwarp();
          rope_row = wbuf;
        }
        uint4 bpre[4];
        const bool bias_early = p.bias != nullptr && p.bias_vec && !p.bias_f32 && col0 + 32 <= p.N;  // warp-uniform
        if (bias_early) {
          const uint4* bp = reinterpret_cast<const uint4*>(p.bias + col0);
#pragma unroll
          for (int i = 0; i < 4; ++i) bpre[i] = __ldg(bp + i);
        }
        tmem_ld_wait();
        if (fuse) {
          // the same rounding the attention backward sees: dO as the bf16 value that is stored
#pragma unroll
          for (int i = 0; i < 4; ++i) {
            const uint32_t w[4] = {ov[i].x, ov[i].y, ov[i].z, ov[i].w};
#pragma unroll
            for (int j = 0; j < 4; ++j) {
              const float2 o2 = unpack_bf16(w[j]);
              const float2 d2 = unpack_bf16(pack_bf16(__uint_as_float(r[i * 8 + 2 * j]), __uint_as_float(r[i * 8 + 2 * j + 1])));
              dsum = fmaf(o2.x, d2.x, fmaf(o2.y, d2.y, dsum));
            }
          }
          if ((col0 + 32) % p.fuse_hd == 0) {  // last chunk of a head
            const int h = col0 / p.fuse_hd;
            p.fuse_delta[((int64_t)(row / p.fuse_T) * p.fuse_H + h) * p.fuse_T + row % p.fuse_T] = dsum;
            dsum = 0.f;
          }
        }
        if (p.head_mode != 0) {
          if (row < p.M) {
            const int ncolv = min(32, p.N - col0);
            const int tgt = (int)p.head_targets[row];
            if (p.head_mode == 1) {
              float cm = -INFINITY;
#pragma unroll
              for (int i = 0; i < 32; ++i)
                if (i < ncolv) cm = fmaxf(cm, __uint_as_float(r[i]));
              if (cm > hm) {
                hs = (hm == -INFINITY) ? 0.f : hs * __expf(hm - cm);
                hm = cm;
              }
              if (hm != -INFINITY) {
#pragma unroll
                for (int i = 0; i < 32; ++i)
                  if (i < ncolv) hs += __expf(__uint_as_float(r[i]) - hm);
              }
              if (tgt >= col0 && tgt < col0 + ncolv) {
#pragma unroll
                for (int i = 0; i < 32; ++i)
                  if (col0 + i == tgt) p.head_tgt[row] = __uint_as_float(r[i]);
              }
            } else {
              const float l = p.head_lse[row];
#pragma unroll
              for (int i = 0; i < 32; ++i)
                r[i] = __float_as_uint((__expf(__uint_as_float(r[i]) - l) - ((col0 + i == tgt) ? 1.0f : 0.0f)) * p.head_gscale);
            }
          }
          if (p.head_mode == 1) continue;  // warp-uniform: nothing is stored
        }
        if (atomic && p.vec_ok) gemm_tc_reduce_chunk(tmD, m0 + q * 32, col0, r, warp_buf, c, lane);
        else if (atomic) gemm_tc_atomic_chunk(p, m0 + q * 32, col0, r, reinterpret_cast<float*>(warp_buf), lane);
        else if (p.tma_store)  // whole warp: the rows past M are clipped by the TMA unit
          gemm_tc_store_chunk(p, row, col0, r, rope_row, lane, 0, tmD, warp_buf, c, m0 + q * 32, (c == NCOLS / 32 - 1) || (col0 + 32 >= p.N),
                              bias_early ? bpre : nullptr);
        else if (row < p.M)
          gemm_tc_store_chunk(p, row, col0, r, rope_row, lane, 0, nullptr, nullptr, 0, 0, false, bias_early ? bpre : nullptr);
      }
  if (p.head_mode == 1 && row < p.M) p.head_part[(int64_t)row * p.num_n + n0 / NCOLS] = make_float2(hm, hs);
}

template <int BN>
__global__ void __launch_bounds__(GT_THREADS, 1)
    gemm_tc_kernel(const __grid_constant__ CUtensorMap tmA, const __grid_constant__ CUtensorMap tmB,
                   const __grid_constant__ CUtensorMap tmD, const GemmTcArgs p) {
  using Cfg = GemmTcCfg<BN>;
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = smem_raw + ((1024u - (smem_u32(smem_raw) & 1023u)) & 1023u);
  uint8_t* sA = smem;
  uint8_t* sB = smem + Cfg::STAGES * Cfg::A_BYTES;
  uint64_t* full = reinterpret_cast<uint64_t*>(smem + Cfg::STAGES * Cfg::STAGE_BYTES + 4 * Cfg::ROPE_WARP_BYTES);
  uint64_t* empty = full + Cfg::STAGES;
  uint64_t* tfull = empty + Cfg::STAGES;
  uint64_t* tempty = tfull + 2;
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(tempty + 2);
  uint8_t* sRope = smem + Cfg::STAGES * Cfg::STAGE_BYTES;  // 4 x 8 KB warp staging buffers, 1 KB aligned (TMA 128B swizzle)

  const int warp = __shfl_sync(0xffffffffu, (int)(threadIdx.x >> 5), 0);
  const int lane = threadIdx.x & 31;

  if (warp == 0 && lane == 0) {
    prefetch_tmap(&tmA);
    prefetch_tmap(&tmB);
  }
  if (warp == 1) {
    if (lane == 0) {
      for (int s = 0; s < Cfg::STAGES; ++s) {
        mbar_init(&full[s], 1);
        mbar_init(&empty[s], 1);
      }
      for (int a = 0; a < 2; ++a) {
        mbar_init(&tfull[a], 1);
        mbar_init(&tempty[a], 4);
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

  const int total_items = p.total_items;

  if (warp == 0) {
    // ------------------------------------------------------------------ TMA producer
    if (lane == 0) {
      int stage = 0;
      uint32_t phase = 0;
      for (int item = blockIdx.x; item < total_items; item += gridDim.x) {
        const GemmTcItem wi = gemm_tc_item<BN>(p, item);
        const int m0 = wi.m0, n0 = wi.n0;
        for (int kb = wi.kb0; kb < wi.kb1; ++kb) {
          mbar_wait(&empty[stage], phase ^ 1);
          mbar_arrive_expect_tx(&full[stage], Cfg::STAGE_BYTES);
          uint8_t* a = sA + stage * Cfg::A_BYTES;
          uint8_t* b = sB + stage * Cfg::B_BYTES;
          if (!p.a_mn) {
            tma_load_2d(a, &tmA, &full[stage], kb * GT_BK, m0);
          } else {
#pragma unroll
            for (int c = 0; c < GT_BM / 64; ++c) tma_load_2d(a + c * (GT_BK * 128), &tmA, &full[stage], m0 + c * 64, kb * GT_BK);
          }
          if (!p.b_mn) {
            tma_load_2d(b, &tmB, &full[stage], kb * GT_BK, n0);
          } else {
#pragma unroll
            for (int c = 0; c < BN / 64; ++c) tma_load_2d(b + c * (GT_BK * 128), &tmB, &full[stage], n0 + c * 64, kb * GT_BK);
          }
          if (++stage == Cfg::STAGES) { stage = 0; phase ^= 1; }
        }
      }
    }
    __syncwarp();
  } else if (warp == 1) {
    // ------------------------------------------------------------------ MMA issuer (whole warp, uniform; one elected lane issues)
    {
      const uint32_t idesc = umma_idesc_bf16(GT_BM, BN, p.a_mn, p.b_mn);
      int stage = 0, acc = 0;
      uint32_t phase = 0, acc_phase = 0;
      for (int item = blockIdx.x; item < total_items; item += gridDim.x) {
        const GemmTcItem wi = gemm_tc_item<BN>(p, item);
        const int kb0 = wi.kb0, kb1 = wi.kb1;
        mbar_wait(&tempty[acc], acc_phase ^ 1);
        tc_fence_after();
        const uint32_t d_tmem = tmem_base + (uint32_t)(acc * BN);
        for (int kb = kb0; kb < kb1; ++kb) {
          mbar_wait(&full[stage], phase);
          tc_fence_after();
          const uint32_t a_addr = smem_u32(sA + stage * Cfg::A_BYTES);
          const uint32_t b_addr = smem_u32(sB + stage * Cfg::B_BYTES);
          // descriptor of the stage's first k-step; a k-step adds a constant to the start-address field
          const uint64_t adesc0 = p.a_mn ? umma_desc_mnmajor(a_addr, GT_BK) : umma_desc_kmajor(a_addr);
          const uint64_t bdesc0 = p.b_mn ? umma_desc_mnmajor(b_addr, GT_BK) : umma_desc_kmajor(b_addr);
          const uint32_t astep = p.a_mn ? (2048u >> 4) : (32u >> 4), bstep = p.b_mn ? (2048u >> 4) : (32u >> 4);
#pragma unroll
          for (int k = 0; k < GT_BK / 16; ++k)
            umma_bf16_w(d_tmem, adesc0 + (uint64_t)(k * astep), bdesc0 + (uint64_t)(k * bstep), idesc, (kb > kb0 || k > 0) ? 1u : 0u);
          umma_commit_w(&empty[stage]);  // frees the smem slot when these MMAs have read it
          if (++stage == Cfg::STAGES) { stage = 0; phase ^= 1; }
        }
        umma_commit_w(&tfull[acc]);  // accumulator complete -> epilogue
        acc ^= 1;
        if (acc == 0) acc_phase ^= 1;
      }
    }
    __syncwarp();
  } else {
    // ------------------------------------------------------------------ epilogue (warps 2..5 -> TMEM lane quarters 2,3,0,1)
    const int q = warp & 3;
    int acc = 0;
    uint32_t acc_phase = 0;
    for (int item = blockIdx.x; item < total_items; item += gridDim.x) {
      const GemmTcItem wi = gemm_tc_item<BN>(p, item);
      const int m0 = wi.m0, n0 = wi.n0;
      mbar_wait(&tfull[acc], acc_phase);
      tc_fence_after();
      gemm_tc_drain<BN>(p, &tmD, wi.atomic, m0, n0, tmem_base + ((uint32_t)(q * 32) << 16) + (uint32_t)(acc * BN), sRope + (warp - 2) * Cfg::ROPE_WARP_BYTES, q, lane);
      tc_fence_before();
      __syncwarp();
      if (lane == 0) mbar_arrive(&tempty[acc]);
      acc ^= 1;
      if (acc == 0) acc_phase ^= 1;
    }
    if (lane == 0) tma_store_wait_read<0>();  // bulk reduce-adds of a K-split tail tile still read this warp's staging buffer
  }

  tc_fence_before();
  __syncthreads();
  if (warp == 1) tmem_dealloc(tmem_base, Cfg::TMEM_COLS);
}

template <int BN, int BMT>
__global__ void gemm_tc_zero_tail_kernel(const GemmTcArgs p);

// ------------------------------------------------------------------------------------------------ CTA-pair variant
// 256 x 256 output tile per pair of CTAs (cluster (2,1,1), tcgen05 cta_group::2): each CTA stages its own 128 rows of A and
// only HALF of B (128 of the 256 N rows) per k-block - 32 KB instead of 48 KB per 128 x 256 x 64 MACs - which cuts the
// shared-memory fill traffic and the L2->SM bytes per FLOP by a third and leaves room for 6 pipeline stages.  The leader
// CTA (rank 0) issues every MMA for the pair; commits are multicast to both CTAs' barriers; each CTA drains its own
// 128 x 256 TMEM accumulator.  BN = 256.  Whole waves run full-K tiles; the tail wave of an fp32 weight-gradient GEMM is split
// along K and leaves through TMA bulk reduce-adds into pre-zeroed tiles (gemm_tc2_item / gemm_tc_reduce_chunk).
struct GemmTc2Cfg {
  static constexpr int BN = 256;
  static constexpr int STAGES = 6;
  static constexpr int A_BYTES = GT_BM * GT_BK * 2;        // this CTA's 128 rows
  static constexpr int B_BYTES = (BN / 2) * GT_BK * 2;     // this CTA's half of the N rows
  static constexpr int STAGE_BYTES = A_BYTES + B_BYTES;    // per CTA
  static constexpr int BAR_BYTES = 256;
  static constexpr int ROPE_WARP_BYTES = 32 * 16 * 16;
  static constexpr int SMEM_BYTES = STAGES * STAGE_BYTES + BAR_BYTES + 4 * ROPE_WARP_BYTES + 1024;
  static constexpr int TMEM_COLS = 512;
};

// work item -> (tile, k-block range, atomic): whole waves run full-K tiles, the tail wave's tiles are split along K
__device__ __forceinline__ void gemm_tc2_item(const GemmTcArgs& p, int item, int& tile, int& kb0, int& kb1, int& atomic) {
  tile = item;
  kb0 = 0;
  kb1 = p.kb_total;
  atomic = 0;
  if (item >= p.full_tiles) {
    const int j = item - p.full_tiles;
    tile = p.full_tiles + j / p.tail_parts;
    kb0 = (j % p.tail_parts) * p.kb_per_part;
    kb1 = min(kb0 + p.kb_per_part, p.kb_total);
    atomic = p.tail_parts > 1;
  }
}

__global__ void __launch_bounds__(GT_THREADS, 1)
    gemm_tc2_kernel(const __grid_constant__ CUtensorMap tmA, const __grid_constant__ CUtensorMap tmB,
                    const __grid_constant__ CUtensorMap tmD, const GemmTcArgs p) {
  using Cfg = GemmTc2Cfg;
  constexpr int BN = Cfg::BN;
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = smem_raw + ((1024u - (smem_u32(smem_raw) & 1023u)) & 1023u);
  uint8_t* sA = smem;
  uint8_t* sB = smem + Cfg::STAGES * Cfg::A_BYTES;
  uint64_t* full = reinterpret_cast<uint64_t*>(smem + Cfg::STAGES * Cfg::STAGE_BYTES + 4 * Cfg::ROPE_WARP_BYTES);
  uint64_t* empty = full + Cfg::STAGES;
  uint64_t* tfull = empty + Cfg::STAGES;
  uint64_t* tempty = tfull + 2;
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(tempty + 2);
  uint8_t* sRope = smem + Cfg::STAGES * Cfg::STAGE_BYTES;  // 4 x 8 KB warp staging buffers, 1 KB aligned (TMA 128B swizzle)

  const int warp = __shfl_sync(0xffffffffu, (int)(threadIdx.x >> 5), 0);
  const int lane = threadIdx.x & 31;
  const uint32_t rank = cluster_ctarank();
  const int pair = blockIdx.x >> 1, npairs = gridDim.x >> 1;

  if (warp == 0 && lane == 0) {
    prefetch_tmap(&tmA);
    prefetch_tmap(&tmB);
  }
  if (warp == 1) {
    if (lane == 0) {
      for (int s = 0; s < Cfg::STAGES; ++s) {
        mbar_init(&full[s], 2);   // leader: its own expect_tx arrive + the peer's remote arrive
        mbar_init(&empty[s], 1);  // multicast commit from the leader's MMA thread
      }
      for (int a = 0; a < 2; ++a) {
        mbar_init(&tfull[a], 1);
        mbar_init(&tempty[a], 8);  // leader: 4 epilogue warps of each CTA
      }
      fence_barrier_init();
    }
    __syncwarp();
    tmem_alloc_2sm(tmem_slot, Cfg::TMEM_COLS);
    tmem_relinquish_2sm();
  }
  tc_fence_before();
  cluster_sync_all();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;

  const int total_items = p.total_items;  // 256 x 256 tiles

  if (warp == 0) {
    // ------------------------------------------------------------------ TMA producer (both CTAs, own halves)
    if (lane == 0) {
      int stage = 0;
      uint32_t phase = 0;
      for (int item = pair; item < total_items; item += npairs) {
        int tile, kb0, kb1, atomic, m_blk, n_blk;
        gemm_tc2_item(p, item, tile, kb0, kb1, atomic);
        gemm_tc_tile_coords(p, tile, m_blk, n_blk);
        const int m0 = m_blk * 256 + (int)rank * 128, n0 = n_blk * BN + (int)rank * (BN / 2);
        for (int kb = kb0; kb < kb1; ++kb) {
          mbar_wait(&empty[stage], phase ^ 1);
          if (rank == 0) mbar_arrive_expect_tx(&full[stage], 2 * Cfg::STAGE_BYTES);
          else mbar_arrive_remote(&full[stage], 0);
          uint8_t* a = sA + stage * Cfg::A_BYTES;
          uint8_t* b = sB + stage * Cfg::B_BYTES;
          if (!p.a_mn) {
            tma_load_2d_2sm(a, &tmA, &full[stage], kb * GT_BK, m0);
          } else {
#pragma unroll
            for (int c = 0; c < 2; ++c) tma_load_2d_2sm(a + c * (GT_BK * 128), &tmA, &full[stage], m0 + c * 64, kb * GT_BK);
          }
          if (!p.b_mn) {
            tma_load_2d_2sm(b, &tmB, &full[stage], kb * GT_BK, n0);
          } else {
#pragma unroll
            for (int c = 0; c < 2; ++c) tma_load_2d_2sm(b + c * (GT_BK * 128), &tmB, &full[stage], n0 + c * 64, kb * GT_BK);
          }
          if (++stage == Cfg::STAGES) { stage = 0; phase ^= 1; }
        }
      }
    }
    __syncwarp();
  } else if (warp == 1) {
    // ------------------------------------------------------------------ MMA issuer (leader CTA only; whole warp, one elected lane issues)
    if (rank == 0) {
      const uint32_t idesc = umma_idesc_bf16(256, BN, p.a_mn, p.b_mn);
      int stage = 0, acc = 0;
      uint32_t phase = 0, acc_phase = 0;
      for (int item = pair; item < total_items; item += npairs) {
        int tile, kb0, kb1, atomic;
        gemm_tc2_item(p, item, tile, kb0, kb1, atomic);
        mbar_wait(&tempty[acc], acc_phase ^ 1);
        tc_fence_after();
        const uint32_t d_tmem = tmem_base + (uint32_t)(acc * BN);
        for (int kb = kb0; kb < kb1; ++kb) {
          mbar_wait(&full[stage], phase);
          tc_fence_after();
          const uint32_t a_addr = smem_u32(sA + stage * Cfg::A_BYTES);
          const uint32_t b_addr = smem_u32(sB + stage * Cfg::B_BYTES);
          const uint64_t adesc0 = p.a_mn ? umma_desc_mnmajor(a_addr, GT_BK) : umma_desc_kmajor(a_addr);
          const uint64_t bdesc0 = p.b_mn ? umma_desc_mnmajor(b_addr, GT_BK) : umma_desc_kmajor(b_addr);
          const uint32_t astep = p.a_mn ? (2048u >> 4) : (32u >> 4), bstep = p.b_mn ? (2048u >> 4) : (32u >> 4);
#pragma unroll
          for (int k = 0; k < GT_BK / 16; ++k)
            umma_bf16_2sm_w(d_tmem, adesc0 + (uint64_t)(k * astep), bdesc0 + (uint64_t)(k * bstep), idesc, (kb > kb0 || k > 0) ? 1u : 0u);
          umma_commit_2sm_mc_w(&empty[stage], 3);  // frees the slot in BOTH CTAs
          if (++stage == Cfg::STAGES) { stage = 0; phase ^= 1; }
        }
        umma_commit_2sm_mc_w(&tfull[acc], 3);  // accumulators complete -> both epilogues
        acc ^= 1;
        if (acc == 0) acc_phase ^= 1;
      }
    }
    __syncwarp();
  } else {
    // ------------------------------------------------------------------ epilogue: each CTA drains its own 128 x 256 accumulator
    const int q = warp & 3;
    int acc = 0;
    uint32_t acc_phase = 0;
    for (int item = pair; item < total_items; item += npairs) {
      int tile, kb0, kb1, atomic, m_blk, n_blk;
      gemm_tc2_item(p, item, tile, kb0, kb1, atomic);
      gemm_tc_tile_coords(p, tile, m_blk, n_blk);
      const int m0 = m_blk * 256 + (int)rank * 128, n0 = n_blk * BN;
      mbar_wait(&tfull[acc], acc_phase);
      tc_fence_after();
      gemm_tc_drain<BN>(p, &tmD, atomic, m0, n0, tmem_base + ((uint32_t)(q * 32) << 16) + (uint32_t)(acc * BN),
                        sRope + (warp - 2) * Cfg::ROPE_WARP_BYTES, q, lane);
      tc_fence_before();
      __syncwarp();
      if (lane == 0) {
        if (rank == 0) mbar_arrive(&tempty[acc]);
        else mbar_arrive_remote(&tempty[acc], 0);
      }
      acc ^= 1;
      if (acc == 0) acc_phase ^= 1;
    }
    if (lane == 0) tma_store_wait_read<0>();  // bulk reduce-adds of a K-split tail tile still read this warp's staging buffer
  }

  tc_fence_before();
  cluster_sync_all();
  if (warp == 1) tmem_dealloc_2sm(tmem_base, Cfg::TMEM_COLS);
}

// Launch hint (thread-local, set by the block backward around the GEMMs that run while an all-reduce holds some SMs).
// The persistent kernel hands tiles out statically (tile = pair, pair + npairs, ...): a cluster that cannot become resident
// until the collective's CTAs leave still owns its full share, so the GEMM ends "collective time + own time" later
// (8 GPUs, r02g: dWqkv 0.29 -> 0.52 ms).  With the hint the grid has one cluster per `tiles_per_cluster` tiles and the
// hardware's block scheduler does the balancing: late SMs simply take fewer clusters.
static thread_local int g_bias_f32 = 0;  // set by gemm_f32_split around its call: the bias pointer holds fp32 values
void gemm_tc_hint_bias_f32(int on) { g_bias_f32 = on; }
static thread_local int g_shared_sm_tiles = 0;
void gemm_tc_hint_shared_sms(int tiles_per_cluster) { g_shared_sm_tiles = tiles_per_cluster; }

static int gemm_tc2_launch(const void* A, int64_t lda, const void* B, int64_t ldb, int64_t K, GemmTcArgs args, bool may_split,
                           cudaStream_t st) {
  using Cfg = GemmTc2Cfg;
  const int G = sm_count();
  args.num_m = (args.M + 255) / 256;          // 256-row tiles
  args.group_m = gemm_group_m(GT_GROUP_M);  // 4 / 8 / 12 / 16 measured within +-2 % of each other on the C3 GEMMs (r01i)
  args.num_n = (args.N + Cfg::BN - 1) / Cfg::BN;
  const int tiles = args.num_m * args.num_n;
  int pairs = G / 2;
  if (pairs > tiles) pairs = tiles;
  if (g_shared_sm_tiles > 0 && tiles > pairs) {  // SMs are shared with a collective: many short-lived clusters instead of a persistent grid
    pairs = (tiles + g_shared_sm_tiles - 1) / g_shared_sm_tiles;
    may_split = false;
  }
  args.full_tiles = tiles;
  args.tail_parts = 1;
  args.kb_per_part = args.kb_total;
  if (may_split) {  // fp32 weight gradients: split the tail wave's tiles along K (coalesced fp32 atomics into zeroed tiles)
    const int full = (tiles / pairs) * pairs, R = tiles - full;
    if (R > 0 && R * 2 <= pairs && args.kb_total >= 16) {
      int parts = pairs / R;
      const int max_parts = args.kb_total / 8;
      if (parts > max_parts) parts = max_parts;
      if (parts > 1) {
        args.kb_per_part = (args.kb_total + parts - 1) / parts;
        args.tail_parts = (args.kb_total + args.kb_per_part - 1) / args.kb_per_part;
        args.full_tiles = full;
      }
    }
  }
  args.total_items = args.full_tiles + (tiles - args.full_tiles) * args.tail_parts;
  CUtensorMap tmA, tmB;
  int rc;
  if (!args.a_mn) rc = make_tmap_2d(&tmA, A, 2, (uint64_t)args.M, (uint64_t)K, (uint64_t)lda, GT_BM, GT_BK, true, false);
  else rc = make_tmap_2d(&tmA, A, 2, (uint64_t)K, (uint64_t)args.M, (uint64_t)lda, GT_BK, 64, true, false);
  if (rc) return rc;
  if (!args.b_mn) rc = make_tmap_2d(&tmB, B, 2, (uint64_t)args.N, (uint64_t)K, (uint64_t)ldb, Cfg::BN / 2, GT_BK, true, false);
  else rc = make_tmap_2d(&tmB, B, 2, (uint64_t)K, (uint64_t)args.N, (uint64_t)ldb, GT_BK, 64, true, false);
  if (rc) return rc;
  CUtensorMap tmD = tmA;  // dereferenced by K-split tail parts (fp32 reduce-add) and by the TMA-store epilogue
  if (args.out_f32 && args.vec_ok && (args.tail_parts > 1 || args.tma_store)) {
    if ((rc = make_tmap_2d(&tmD, args.D, 4, (uint64_t)args.M, (uint64_t)args.N, (uint64_t)args.ldd, 32, 32, true, true))) return rc;
  } else if (args.tma_store) {
    if ((rc = make_tmap_2d(&tmD, args.D, 2, (uint64_t)args.M, (uint64_t)args.N, (uint64_t)args.ldd, 32, 64, true, false))) return rc;
  }
  if (args.tail_parts > 1) {
    gemm_tc_zero_tail_kernel<Cfg::BN, 256><<<dim3(tiles - args.full_tiles, 256 / 32), 256, 0, st>>>(args);
    SPB_LAUNCH_CHECK();
  }
  SPB_CUDA(cudaFuncSetAttribute(gemm_tc2_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, Cfg::SMEM_BYTES));
  cudaLaunchConfig_t cfg{};
  cfg.gridDim = dim3(2 * pairs);
  cfg.blockDim = dim3(GT_THREADS);
  cfg.dynamicSmemBytes = Cfg::SMEM_BYTES;
  cfg.stream = st;
  cudaLaunchAttribute attr[1];
  attr[0].id = cudaLaunchAttributeClusterDimension;
  attr[0].val.clusterDim.x = 2;
  attr[0].val.clusterDim.y = 1;
  attr[0].val.clusterDim.z = 1;
  cfg.attrs = attr;
  cfg.numAttrs = 1;
  SPB_CUDA(cudaLaunchKernelEx(&cfg, gemm_tc2_kernel, tmA, tmB, tmD, args));
  SPB_LAUNCH_CHECK();
  return 0;
}

// zero the tail tiles that the K-split parts will reduce into: blockIdx.x = tail tile, blockIdx.y = 32-row slab of it
template <int BN, int BMT = GT_BM>
__global__ void __launch_bounds__(256) gemm_tc_zero_tail_kernel(const GemmTcArgs p) {
  int m_blk, n_blk;
  gemm_tc_tile_coords(p, p.full_tiles + (int)blockIdx.x, m_blk, n_blk);
  float* d = reinterpret_cast<float*>(p.D);
  const int m0 = m_blk * BMT + (int)blockIdx.y * 32, n0 = n_blk * BN;
  if (p.vec_ok) {  // 16-byte aligned rows: float4 stores
    for (int e = threadIdx.x; e < 32 * (BN / 4); e += 256) {
      const int r = m0 + e / (BN / 4), c = n0 + (e % (BN / 4)) * 4;
      if (r >= p.M || c >= p.N) continue;
      float* q = d + (int64_t)r * p.ldd + c;
      if (c + 4 <= p.N) *reinterpret_cast<float4*>(q) = make_float4(0.f, 0.f, 0.f, 0.f);
      else for (int i = 0; c + i < p.N; ++i) q[i] = 0.f;
    }
    return;
  }
  for (int e = threadIdx.x; e < 32 * BN; e += 256) {
    const int r = m0 + e / BN, c = n0 + e % BN;
    if (r < p.M && c < p.N) d[(int64_t)r * p.ldd + c] = 0.f;
  }
}

template <int BN>
static int gemm_tc_launch(const void* A, int64_t lda, const void* B, int64_t ldb, int64_t K, GemmTcArgs args, bool may_split,
                          cudaStream_t st) {
  using Cfg = GemmTcCfg<BN>;
  args.group_m = gemm_group_m(GT_GROUP_M);
  const int G = sm_count();
  args.num_n = (args.N + BN - 1) / BN;
  const int tiles = args.num_m * args.num_n;
  args.full_tiles = tiles;
  args.tail_parts = 1;
  args.kb_per_part = args.kb_total;
  if (may_split) {
    const int full = (tiles / G) * G, R = tiles - full;
    if (R > 0 && (R * 2 <= G || full == 0) && args.kb_total >= 16) {
      int parts = G / R;
      const int max_parts = args.kb_total / 8;  // at least 8 k-blocks (512 k) per part
      if (full == 0) {
        // fewer tiles than SMs (weight gradients of narrow models: K = tokens is long, the output small): every tile is split, and
        // the part count is the one whose items fill whole waves best (54 tiles: 2 parts = 108 items = 73 % of a wave; 11 parts = 4.01 waves)
        int best = parts < 1 ? 1 : parts;
        double best_eff = 0.0;
        const int hi = 6 * best < max_parts ? 6 * best : max_parts;
        for (int c = best; c <= hi; ++c) {  // more parts = more reduce-add traffic: a later candidate must fill the waves 10 % better
          const int items = R * c;
          const double eff = (double)items / ((double)((items + G - 1) / G) * G);
          if (eff > best_eff + (c == best ? 0.0 : 0.10)) {
            best_eff = eff;
            best = c;
          }
        }
        parts = best;
      }
      if (parts > max_parts) parts = max_parts;
      if (parts > 1) {
        args.kb_per_part = (args.kb_total + parts - 1) / parts;
        args.tail_parts = (args.kb_total + args.kb_per_part - 1) / args.kb_per_part;
        args.full_tiles = full;
      }
    }
  }
  args.total_items = args.full_tiles + (tiles - args.full_tiles) * args.tail_parts;
  CUtensorMap tmA, tmB;
  int rc;
  if (!args.a_mn) rc = make_tmap_2d(&tmA, A, 2, (uint64_t)args.M, (uint64_t)K, (uint64_t)lda, GT_BM, GT_BK, true, false);
  else rc = make_tmap_2d(&tmA, A, 2, (uint64_t)K, (uint64_t)args.M, (uint64_t)lda, GT_BK, 64, true, false);
  if (rc) return rc;
  if (!args.b_mn) rc = make_tmap_2d(&tmB, B, 2, (uint64_t)args.N, (uint64_t)K, (uint64_t)ldb, (uint32_t)BN, GT_BK, true, false);
  else rc = make_tmap_2d(&tmB, B, 2, (uint64_t)K, (uint64_t)args.N, (uint64_t)ldb, GT_BK, 64, true, false);
  if (rc) return rc;
  CUtensorMap tmD = tmA;  // dereferenced by K-split tail parts (fp32 reduce-add) and by the TMA-store epilogue
  if (args.out_f32 && args.vec_ok && (args.tail_parts > 1 || args.tma_store)) {
    if ((rc = make_tmap_2d(&tmD, args.D, 4, (uint64_t)args.M, (uint64_t)args.N, (uint64_t)args.ldd, 32, 32, true, true))) return rc;
  } else if (args.tma_store) {
    if ((rc = make_tmap_2d(&tmD, args.D, 2, (uint64_t)args.M, (uint64_t)args.N, (uint64_t)args.ldd, 32, 64, true, false))) return rc;
  }
  if (args.tail_parts > 1) {
    gemm_tc_zero_tail_kernel<BN><<<dim3(tiles - args.full_tiles, GT_BM / 32), 256, 0, st>>>(args);
    SPB_LAUNCH_CHECK();
  }
  SPB_CUDA(cudaFuncSetAttribute(gemm_tc_kernel<BN>, cudaFuncAttributeMaxDynamicSharedMemorySize, Cfg::SMEM_BYTES));
  const int grid = args.total_items < G ? args.total_items : G;
  gemm_tc_kernel<BN><<<grid, GT_THREADS, Cfg::SMEM_BYTES, st>>>(tmA, tmB, tmD, args);
  SPB_LAUNCH_CHECK();
  return 0;
}

static double gemm_tc_wave_eff(int64_t M, int64_t N, int bn, int G) {
  const int64_t tiles = ((M + GT_BM - 1) / GT_BM) * ((N + bn - 1) / bn);
  const int64_t waves = (tiles + G - 1) / G;
  const double useful = (double)M * (double)N;  // output elements actually needed
  return useful / ((double)waves * G * GT_BM * bn);
}

int gemm_tc(int out_dtype, int64_t M, int64_t N, int64_t K, const void* A, int64_t lda, int transA, const void* B, int64_t ldb,
            int transB, void* D, int64_t ldd, const void* bias, int accumulate, cudaStream_t st, const float* rope, int rope_cols,
            int rope_T, int rope_hd, const void* residual, int64_t ldr, const GemmTcFuse* fuse, const GemmTcHead* head) {
  SPB_CHECK_ARG(M < (1LL << 31) && N < (1LL << 31) && K < (1LL << 31), SPB_E_LIMIT, "gemm_tc: dimension beyond 2^31");
  SPB_CHECK_ARG(M > 0 && N > 0 && K > 0, SPB_E_INVALID, "gemm_tc: empty problem");
  SPB_CHECK_ARG(((uintptr_t)A % 16 == 0) && ((uintptr_t)B % 16 == 0) && (lda % 8 == 0) && (ldb % 8 == 0), SPB_E_UNSUPPORTED,
                "gemm_tc: TMA needs 16-byte aligned operands and leading dimensions that are multiples of 8 (lda=%lld ldb=%lld)",
                (long long)lda, (long long)ldb);
  SPB_CHECK_ARG(out_dtype == SPB_F32 || out_dtype == SPB_BF16, SPB_E_UNSUPPORTED, "gemm_tc: out_dtype %d", out_dtype);
  GemmTcArgs args{};
  args.D = D;
  args.ldd = ldd;
  args.bias = (const __nv_bfloat16*)bias;
  args.bias_vec = bias != nullptr && (uintptr_t)bias % 16 == 0;
  args.bias_f32 = g_bias_f32;
  args.rope = (const float4*)rope;
  args.rope_cols = rope_cols;
  args.rope_T = rope_T > 0 ? rope_T : 1;
  args.rope_hd = rope_hd > 0 ? rope_hd : 2;
  SPB_CHECK_ARG(rope == nullptr || (!accumulate && rope_hd % 32 == 0 && (uintptr_t)rope % 16 == 0), SPB_E_INVALID,
                "gemm_tc: the fused rotary epilogue needs a head dim that is a multiple of 32, a 16-byte aligned table and no accumulation");
  args.out_f32 = (out_dtype == SPB_F32);
  args.accumulate = accumulate || residual != nullptr;
  args.addsrc = residual;
  args.ld_add = ldr;
  args.M = (int)M;
  args.N = (int)N;
  if (fuse != nullptr) {
    SPB_CHECK_ARG(out_dtype == SPB_BF16 && bias == nullptr && rope == nullptr && !accumulate && residual == nullptr, SPB_E_INVALID,
                  "gemm_tc: the attention-backward preparation epilogue takes a plain bf16 GEMM");
    SPB_CHECK_ARG(fuse->hd % 32 == 0 && N % fuse->hd == 0 && (256 % fuse->hd == 0 || fuse->hd % 256 == 0) && M % fuse->T == 0 && N == (int64_t)fuse->H * fuse->hd,
                  SPB_E_INVALID, "gemm_tc: fused delta needs hd %% 32 == 0, N = H*hd and M = B*T");
    SPB_CHECK_ARG((uintptr_t)fuse->o % 16 == 0 && fuse->ldo % 8 == 0 && (uintptr_t)fuse->zero % 16 == 0 && fuse->ldz % 4 == 0, SPB_E_INVALID,
                  "gemm_tc: fused delta operands must be 16-byte aligned");
    args.fuse_o = (const __nv_bfloat16*)fuse->o;
    args.fuse_ldo = fuse->ldo;
    args.fuse_delta = fuse->delta;
    args.fuse_zero = fuse->zero;
    args.fuse_ldz = fuse->ldz;
    args.fuse_hd = fuse->hd;
    args.fuse_T = fuse->T;
    args.fuse_H = fuse->H;
  }
  if (head != nullptr) {
    SPB_CHECK_ARG(bias == nullptr && rope == nullptr && !accumulate && residual == nullptr && fuse == nullptr && N > 128, SPB_E_INVALID,
                  "gemm_tc: the loss-head epilogues take a plain GEMM with more than 128 output columns");
    SPB_CHECK_ARG(head->mode == 1 || (head->mode == 2 && out_dtype == SPB_BF16), SPB_E_INVALID, "gemm_tc: loss-head mode %d", head->mode);
    args.head_mode = head->mode;
    args.head_part = (float2*)head->part;
    args.head_tgt = head->tgt_logit;
    args.head_targets = head->targets;
    args.head_lse = head->lse;
    args.head_gscale = head->gscale;
  }
  // transA=0: A stored [M,K] -> K-major.  transA=1: stored [K,M] -> MN-major.
  // transB=0: B stored [K,N] -> MN-major (N contiguous).  transB=1: stored [N,K] -> K-major.
  args.a_mn = transA ? 1 : 0;
  args.b_mn = transB ? 0 : 1;
  args.num_m = (int)((M + GT_BM - 1) / GT_BM);
  args.kb_total = (int)((K + GT_BK - 1) / GT_BK);
  const int esz = args.out_f32 ? 4 : 2;
  args.vec_ok = ((uintptr_t)D % 16 == 0) && ((ldd * esz) % 16 == 0) &&
                (residual == nullptr || (((uintptr_t)residual % 16 == 0) && ((ldr * esz) % 16 == 0)));
  // TMA-store epilogue: plain stores only (no read-modify-write of D, no loss-head partial mode), output addressable by TMA
  // (16-byte aligned rows), and the warp staging buffer not taken by the fused rotary table.  SPB_GEMM_TMA_STORE=0: direct stores.
  static const bool tma_store_off = [] { const char* e = getenv("SPB_GEMM_TMA_STORE"); return e && e[0] == '0'; }();
  args.tma_store = (!tma_store_off && args.vec_ok && !args.accumulate && residual == nullptr && rope == nullptr &&
                    !(head != nullptr && head->mode == 1)) ? 1 : 0;
  // K-split of the tail wave needs a plain fp32 output (weight gradients)
  const bool may_split = args.out_f32 && bias == nullptr && rope == nullptr && !args.accumulate && head == nullptr;
  // tile width: 256 unless a 192-wide tile fills the waves of this problem clearly better (e.g. 8192 x 6144: 94 % -> 99 %)
  int BN = (N > 128) ? 256 : 128;
  if (BN == 256 && !may_split && head == nullptr) {  // (the loss-head partials are laid out per 256-column tile)
    const int G = sm_count();
    // measured (r01d): the 192-wide tile runs at 1219 TF/s vs 1342 TF/s for 256 on 8192x6144x4096 despite 99 % vs 94 % wave
    // efficiency (lower arithmetic intensity per smem byte), so it is only taken for a much larger gain
    if (gemm_tc_wave_eff(M, N, 192, G) > gemm_tc_wave_eff(M, N, 256, G) + 0.15) BN = 192;
  }
  // few-tile problems (e.g. the [T, 768] outputs of the wide-head attention: 48 tiles of 128 x 256 on 148 SMs): the 128-wide tile
  // doubles the number of CTAs; taken when that fills the machine clearly better and the pair kernel is out of reach
  if (BN == 256 && head == nullptr && ((M + 255) / 256) * ((N + 255) / 256) < sm_count() / 2) {
    const int G = sm_count();
    if (gemm_tc_wave_eff(M, N, 128, G) > gemm_tc_wave_eff(M, N, 256, G) + 0.15) BN = 128;
  }
  // CTA-pair kernel: large problems with at least one 256 x 256 tile per pair (SPB_GEMM_2CTA=0 forces the single-CTA kernel)
  static const int use_pair = [] { const char* e = getenv("SPB_GEMM_2CTA"); return e ? atoi(e) : 1; }();
  if (use_pair && BN == 256 && M >= 256 && N >= 256 && ((M + 255) / 256) * ((N + 255) / 256) >= sm_count() / 2)
    return gemm_tc2_launch(A, lda, B, ldb, K, args, may_split, st);
  if (BN == 256) return gemm_tc_launch<256>(A, lda, B, ldb, K, args, may_split, st);
  if (BN == 192) return gemm_tc_launch<192>(A, lda, B, ldb, K, args, may_split, st);
  return gemm_tc_launch<128>(A, lda, B, ldb, K, args, may_split, st);
}

}  // namespace spb
