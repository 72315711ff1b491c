// api.cu - the C ABI of libspb200.so: argument checking, dispatch, and the block-level forward/backward that chains
// the kernels on one stream.  No allocation, no synchronisation, no CPU fallback.
#include <atomic>
#include <cstdarg>
#include <cstring>

#include "common.cuh"
#include "kernels.h"

namespace spb {

static thread_local char g_err[512] = "";
static std::atomic<long long> g_launches{0};

void set_error(const char* fmt, ...) {
  va_list ap;
  va_start(ap, fmt);
  vsnprintf(g_err, sizeof(g_err), fmt, ap);
  va_end(ap);
}
void count_launch(int n) { g_launches.fetch_add(n, std::memory_order_relaxed); }

int sm_count() {
  static thread_local int cached_dev = -1, cached = 0;
  int dev = 0;
  if (cudaGetDevice(&dev) != cudaSuccess) return 148;
  if (dev != cached_dev) {
    int n = 0;
    if (cudaDeviceGetAttribute(&n, cudaDevAttrMultiProcessorCount, dev) != cudaSuccess || n <= 0) n = 148;
    cached = n;
    cached_dev = dev;
  }
  return cached;
}

typedef CUresult (*EncodeTiledFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*, const cuuint64_t*,
                                  const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave, CUtensorMapSwizzle,
                                  CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

static EncodeTiledFn encode_fn() {
  // resolved through the runtime so the library has no link-time dependency on libcuda.so.1 (absent on CPU-only hosts)
  static EncodeTiledFn fn = [] {
    void* p = nullptr;
    cudaDriverEntryPointQueryResult q;
    if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &p, cudaEnableDefault, &q) != cudaSuccess ||
        q != cudaDriverEntryPointSuccess)
      p = nullptr;
    return (EncodeTiledFn)p;
  }();
  return fn;
}

int make_tmap_2d(CUtensorMap* map, const void* gptr, int elem_bytes, uint64_t rows, uint64_t cols, uint64_t ld, uint32_t box_rows,
                 uint32_t box_cols, bool swizzle128, bool is_float) {
  EncodeTiledFn fn = encode_fn();
  SPB_CHECK_ARG(fn != nullptr, SPB_E_STATE, "cuTensorMapEncodeTiled driver entry point not available");
  SPB_CHECK_ARG(((uintptr_t)gptr % 16 == 0) && ((ld * elem_bytes) % 16 == 0), SPB_E_UNSUPPORTED,
                "TMA operand must be 16-byte aligned with a 16-byte multiple row pitch (ptr=%p ld=%llu)", gptr,
                (unsigned long long)ld);
  const cuuint64_t dims[2] = {cols, rows};
  const cuuint64_t strides[1] = {ld * (uint64_t)elem_bytes};
  const cuuint32_t box[2] = {box_cols, box_rows};
  const cuuint32_t estr[2] = {1, 1};
  CUtensorMapDataType dt = is_float ? CU_TENSOR_MAP_DATA_TYPE_FLOAT32 : CU_TENSOR_MAP_DATA_TYPE_BFLOAT16;
  CUresult r = fn(map, dt, 2, const_cast<void*>(gptr), dims, strides, box, estr, CU_TENSOR_MAP_INTERLEAVE_NONE,
                  swizzle128 ? CU_TENSOR_MAP_SWIZZLE_128B : CU_TENSOR_MAP_SWIZZLE_NONE, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
                  CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  SPB_CHECK_ARG(r == CUDA_SUCCESS, 700 + (int)r, "cuTensorMapEncodeTiled failed (CUresult %d; rows=%llu cols=%llu ld=%llu box=%ux%u)",
                (int)r, (unsigned long long)rows, (unsigned long long)cols, (unsigned long long)ld, box_rows, box_cols);
  return 0;
}

// ---- optional stage timing (bench.py's per-kernel roofline): CUDA events recorded on the caller's stream at stage
// boundaries of spb_block_fwd / spb_block_bwd.  Off by default; events are created once in spb_prof_enable.
constexpr int PROF_MARKS = 40;
struct Prof {
  bool on = false;
  int max_steps = 0, step = -1;
  cudaEvent_t* ev = nullptr;      // [max_steps][PROF_MARKS]
  const char** names = nullptr;   // [max_steps][PROF_MARKS]
  int* count = nullptr;           // [max_steps]
};
static Prof g_prof;

static void prof_new_step() {
  if (!g_prof.on) return;
  g_prof.step += 1;
  if (g_prof.step < g_prof.max_steps) g_prof.count[g_prof.step] = 0;
}
static void prof_mark(const char* name, cudaStream_t st) {
  if (!g_prof.on || g_prof.step < 0 || g_prof.step >= g_prof.max_steps) return;
  int& n = g_prof.count[g_prof.step];
  if (n >= PROF_MARKS) return;
  cudaEventRecord(g_prof.ev[g_prof.step * PROF_MARKS + n], st);
  g_prof.names[g_prof.step * PROF_MARKS + n] = name;
  ++n;
}

static inline size_t align_up(size_t v, size_t a) { return (v + a - 1) / a * a; }
static inline size_t esize(int dtype) { return dtype == SPB_BF16 ? 2 : 4; }

static int check_desc(const spb_block_desc* d) {
  SPB_CHECK_ARG(d != nullptr, SPB_E_INVALID, "block desc is NULL");
  SPB_CHECK_ARG(d->B > 0 && d->T > 0 && d->C > 0 && d->Hq > 0 && d->Hkv > 0 && d->hd > 0, SPB_E_INVALID,
                "block desc: non-positive dimension (B=%d T=%d C=%d Hq=%d Hkv=%d hd=%d)", d->B, d->T, d->C, d->Hq, d->Hkv, d->hd);
  SPB_CHECK_ARG(d->Hq % d->Hkv == 0, SPB_E_INVALID, "block desc: Hq %d not a multiple of Hkv %d", d->Hq, d->Hkv);
  SPB_CHECK_ARG(d->dtype == SPB_F32 || d->dtype == SPB_BF16, SPB_E_UNSUPPORTED, "block desc: dtype %d", d->dtype);
  SPB_CHECK_ARG(d->norm_kind >= SPB_NORM_NONE && d->norm_kind <= SPB_NORM_LAYER, SPB_E_INVALID, "block desc: norm_kind %d", d->norm_kind);
  SPB_CHECK_ARG(d->rope_kind == SPB_ROPE_NONE || d->rope_kind == SPB_ROPE_TABLE, SPB_E_INVALID, "block desc: rope_kind %d", d->rope_kind);
  SPB_CHECK_ARG(d->rope_kind == SPB_ROPE_NONE || d->hd % 2 == 0, SPB_E_INVALID, "block desc: rotary needs an even head dim");
  SPB_CHECK_ARG(d->weight_layout == SPB_W_IN_OUT || d->weight_layout == SPB_W_OUT_IN, SPB_E_INVALID, "block desc: weight_layout %d",
                d->weight_layout);
  SPB_CHECK_ARG(d->residual == 0 || d->residual == 1, SPB_E_INVALID, "block desc: residual %d", d->residual);
  return 0;
}

static int pick_gemm_impl(int impl, int dtype) {
  if (impl == SPB_IMPL_AUTO) return dtype == SPB_BF16 ? SPB_IMPL_TCGEN05 : SPB_IMPL_SIMT;
  return impl;
}
static int pick_attn_impl(int impl, int dtype, int hd) {
  if (impl == SPB_IMPL_AUTO) return flash_tc_supported(dtype, hd) ? SPB_IMPL_TCGEN05 : SPB_IMPL_SIMT;
  return impl;
}
static int pick_attn_bwd_impl(int impl, int dtype, int hd) {
  if (impl == SPB_IMPL_AUTO) return flash_bwd_tc_supported(dtype, hd) ? SPB_IMPL_TCGEN05 : SPB_IMPL_SIMT;
  return impl;
}

// bf16 blocks whose head dim exceeds what the fused attention kernels hold on chip take the tcgen05 GEMM pipeline of attn_wide.cu
static bool block_uses_wide_attn(const spb_block_desc* d) {
  return (d->attn_impl == SPB_IMPL_AUTO || d->attn_impl == SPB_IMPL_TCGEN05) && attn_wide_supported(d->dtype, d->hd) &&
         !flash_tc_supported(d->dtype, d->hd);
}

// scratch of the fp32 split GEMM: set for the duration of a block call by the host thread that runs it (SplitScratch guard)
static thread_local void* g_split_ws = nullptr;
static thread_local size_t g_split_bytes = 0;
struct SplitScratch {
  SplitScratch(void* ws, size_t bytes) { g_split_ws = bytes ? ws : nullptr; g_split_bytes = bytes; }
  ~SplitScratch() { g_split_ws = nullptr; g_split_bytes = 0; }
};

// GEMM dispatch shared by spb_gemm and the block: with `rope` the rotary runs as the in-place kernel right after the GEMM
// (default) or in the tcgen05 epilogue on the fp32 accumulators (SPB_ROPE_FUSE=1).
static int gemm_any(int in_dtype, int out_dtype, int impl, int64_t M, int64_t N, int64_t K, const void* A, int64_t lda, int transA,
                    const void* Bm, int64_t ldb, int transB, void* D, int64_t ldd, const void* bias, int accumulate, cudaStream_t st,
                    const float* rope, int rope_cols, int rope_T, int rope_hd, const void* residual = nullptr, int64_t ldr = 0) {
  SPB_CHECK_ARG(residual == nullptr || rope == nullptr, SPB_E_INVALID, "gemm: residual and rotary epilogues are exclusive");
  // fp32 mode inside a block that reserved split scratch: the tcgen05 GEMM on two-term bf16 splits of both operands
  // (gemm_split.cu; SPB_F32_TC=0 keeps the FFMA GEMM)
  if (in_dtype == SPB_F32 && out_dtype == SPB_F32 && impl != SPB_IMPL_SIMT && g_split_ws != nullptr) {
    static const bool f32_tc = [] { const char* e = getenv("SPB_F32_TC"); return e == nullptr || atoi(e) != 0; }();
    if (f32_tc && gemm_f32_split_ok(M, N, K, A, lda, transA, Bm, ldb, transB) && g_split_bytes >= gemm_f32_split_scratch_bytes(M, N, K) &&
        (uintptr_t)D % 16 == 0 && ldd % 4 == 0) {
      int rc = gemm_f32_split(M, N, K, A, lda, transA, Bm, ldb, transB, D, ldd, bias, accumulate, residual, ldr, g_split_ws, g_split_bytes, st);
      if (rc || rope == nullptr) return rc;
      return spb_rope_apply(out_dtype, D, ldd, M, rope_T, rope_cols / rope_hd, rope_hd, rope, 0, (spb_stream_t)st);
    }
  }
  const int use = pick_gemm_impl(impl, in_dtype);
  if (use == SPB_IMPL_TCGEN05) {
    SPB_CHECK_ARG(in_dtype == SPB_BF16, SPB_E_UNSUPPORTED, "spb_gemm: the tcgen05 GEMM takes bf16 operands (fp32 mode uses the FFMA GEMM)");
    // Measured at C3 after the GEMM reached ~1.6 PFLOP/s in-step: GEMM + the head-looping in-place rotary kernel (0.303 ms)
    // beats the fused epilogue (0.328 ms) - the GEMM main loop saturates shared-memory bandwidth, and the epilogue's table
    // staging competes for it.  SPB_ROPE_FUSE=1 selects the fused epilogue (kept for hosts where the extra pass costs more).
    static const bool fuse_rope = [] {
      const char* e = getenv("SPB_ROPE_FUSE");
      return e != nullptr && atoi(e) != 0;
    }();
    if (rope == nullptr || (rope_hd % 32 == 0 && fuse_rope))
      return gemm_tc(out_dtype, M, N, K, A, lda, transA, Bm, ldb, transB, D, ldd, bias, accumulate, st, rope, rope_cols, rope_T, rope_hd,
                     residual, ldr);
    // head dims that do not tile the 32-column epilogue chunk (e.g. 16): GEMM, then the in-place rotary kernel
    int rc = gemm_tc(out_dtype, M, N, K, A, lda, transA, Bm, ldb, transB, D, ldd, bias, accumulate, st);
    if (rc) return rc;
    return spb_rope_apply(out_dtype, D, ldd, M, rope_T, rope_cols / rope_hd, rope_hd, rope, 0, (spb_stream_t)st);
  }
  SPB_CHECK_ARG(use == SPB_IMPL_SIMT, SPB_E_INVALID, "spb_gemm: impl %d", impl);
  int rc = gemm_simt(in_dtype, out_dtype, M, N, K, A, lda, transA, Bm, ldb, transB, D, ldd, bias, accumulate, st, residual, ldr);
  if (rc || rope == nullptr) return rc;
  return spb_rope_apply(out_dtype, D, ldd, M, rope_T, rope_cols / rope_hd, rope_hd, rope, 0, (spb_stream_t)st);
}

}  // namespace spb

using namespace spb;

extern "C" {

const char* spb_version(void) { return "spb200 0.1 (sm_100a)"; }
const char* spb_last_error(void) { return g_err; }
int64_t spb_launch_count(void) { return (int64_t)g_launches.load(); }

int spb_device_info(int32_t* sm, int32_t* cc_major, int32_t* cc_minor, size_t* smem_optin) {
  int dev = 0;
  SPB_CUDA(cudaGetDevice(&dev));
  int v = 0;
  if (sm) { SPB_CUDA(cudaDeviceGetAttribute(&v, cudaDevAttrMultiProcessorCount, dev)); *sm = v; }
  if (cc_major) { SPB_CUDA(cudaDeviceGetAttribute(&v, cudaDevAttrComputeCapabilityMajor, dev)); *cc_major = v; }
  if (cc_minor) { SPB_CUDA(cudaDeviceGetAttribute(&v, cudaDevAttrComputeCapabilityMinor, dev)); *cc_minor = v; }
  if (smem_optin) { SPB_CUDA(cudaDeviceGetAttribute(&v, cudaDevAttrMaxSharedMemoryPerBlockOptin, dev)); *smem_optin = (size_t)v; }
  return 0;
}

int spb_prof_enable(int max_steps) {
  if (g_prof.ev != nullptr) {
    for (int i = 0; i < g_prof.max_steps * PROF_MARKS; ++i) cudaEventDestroy(g_prof.ev[i]);
    delete[] g_prof.ev;
    delete[] g_prof.names;
    delete[] g_prof.count;
    g_prof = Prof();
  }
  if (max_steps <= 0) return 0;
  SPB_CHECK_ARG(max_steps <= 4096, SPB_E_LIMIT, "spb_prof_enable: at most 4096 steps");
  g_prof.ev = new cudaEvent_t[(size_t)max_steps * PROF_MARKS];
  g_prof.names = new const char*[(size_t)max_steps * PROF_MARKS];
  g_prof.count = new int[max_steps];
  for (int i = 0; i < max_steps * PROF_MARKS; ++i) SPB_CUDA(cudaEventCreate(&g_prof.ev[i]));
  for (int i = 0; i < max_steps; ++i) g_prof.count[i] = 0;
  g_prof.max_steps = max_steps;
  g_prof.step = -1;
  g_prof.on = true;
  return 0;
}

int spb_prof_steps(void) { return g_prof.on ? (g_prof.step + 1 < g_prof.max_steps ? g_prof.step + 1 : g_prof.max_steps) : 0; }

int spb_prof_read(int step, int max_n, const char** names, float* ms) {
  SPB_CHECK_ARG(g_prof.on && step >= 0 && step < g_prof.max_steps && step <= g_prof.step, SPB_E_STATE, "spb_prof_read: no such step");
  const int n = g_prof.count[step];
  int out = 0;
  for (int i = 0; i + 1 < n && out < max_n; ++i) {
    float t = 0.f;
    SPB_CUDA(cudaEventElapsedTime(&t, g_prof.ev[step * PROF_MARKS + i], g_prof.ev[step * PROF_MARKS + i + 1]));
    names[out] = g_prof.names[step * PROF_MARKS + i];
    ms[out] = t;
    ++out;
  }
  return out;  // number of stages written (>= 0)
}

int spb_gemm(int in_dtype, int out_dtype, int impl, int64_t M, int64_t N, int64_t K, const void* A, int64_t lda, int transA,
             const void* Bm, int64_t ldb, int transB, void* D, int64_t ldd, const void* bias, int accumulate, spb_stream_t stream) {
  SPB_CHECK_ARG(A && Bm && D, SPB_E_INVALID, "spb_gemm: null pointer");
  SPB_CHECK_ARG(M >= 0 && N >= 0 && K > 0, SPB_E_INVALID, "spb_gemm: bad shape M=%lld N=%lld K=%lld", (long long)M, (long long)N, (long long)K);
  SPB_CHECK_ARG(lda >= (transA ? M : K) && ldb >= (transB ? K : N) && ldd >= N, SPB_E_INVALID,
                "spb_gemm: leading dimension smaller than the row length (lda=%lld ldb=%lld ldd=%lld)", (long long)lda, (long long)ldb,
                (long long)ldd);
  if (M == 0 || N == 0) return 0;
  return gemm_any(in_dtype, out_dtype, impl, M, N, K, A, lda, transA, Bm, ldb, transB, D, ldd, bias, accumulate, (cudaStream_t)stream,
                  nullptr, 0, 0, 0);
}

int spb_flash_fwd(int dtype, int impl, int32_t B, int32_t T, int32_t Hq, int32_t Hkv, int32_t hd, int causal, float scale,
                  const void* q, int64_t ldq, const void* k, int64_t ldk, const void* v, int64_t ldv, void* o, int64_t ldo, float* lse,
                  spb_stream_t stream) {
  SPB_CHECK_ARG(q && k && v && o && lse, SPB_E_INVALID, "spb_flash_fwd: null pointer");
  SPB_CHECK_ARG(B > 0 && T > 0 && Hq > 0 && Hkv > 0 && hd > 0 && Hq % Hkv == 0, SPB_E_INVALID, "spb_flash_fwd: bad shape");
  SPB_CHECK_ARG(ldq >= (int64_t)Hq * hd && ldk >= (int64_t)Hkv * hd && ldv >= (int64_t)Hkv * hd && ldo >= (int64_t)Hq * hd, SPB_E_INVALID,
                "spb_flash_fwd: leading dimension smaller than heads*hd");
  cudaStream_t st = (cudaStream_t)stream;
  const int use = pick_attn_impl(impl, dtype, hd);
  if (use == SPB_IMPL_TCGEN05) {
    SPB_CHECK_ARG(flash_tc_supported(dtype, hd), SPB_E_UNSUPPORTED, "spb_flash_fwd: tcgen05 path needs bf16 and hd in {64,128,256} (dtype=%d hd=%d)", dtype, hd);
    return flash_fwd_tc(B, T, Hq, Hkv, hd, causal, scale, q, ldq, k, ldk, v, ldv, o, ldo, lse, st);
  }
  SPB_CHECK_ARG(use == SPB_IMPL_SIMT, SPB_E_INVALID, "spb_flash_fwd: impl %d", impl);
  return attn_simt_fwd(dtype, B, T, Hq, Hkv, hd, causal, scale, q, ldq, k, ldk, v, ldv, o, ldo, lse, st);
}

// `rope` != NULL: also apply the rotary VJP to dq and dk (fused into the tcgen05 kernels; separate in-place kernels otherwise)
static int flash_bwd_any(int dtype, int impl, int32_t B, int32_t T, int32_t Hq, int32_t Hkv, int32_t hd, int causal, float scale,
                         const void* q, int64_t ldq, const void* k, int64_t ldk, const void* v, int64_t ldv, const void* o, int64_t ldo,
                         const void* d_o, int64_t lddo, const float* lse, float* delta, void* dq, int64_t lddq, void* dk, int64_t lddk,
                         void* dv, int64_t lddv, float* dq_acc, spb_stream_t stream, const float* rope, int phase = 3);

int spb_flash_bwd(int dtype, int impl, int32_t B, int32_t T, int32_t Hq, int32_t Hkv, int32_t hd, int causal, float scale,
                  const void* q, int64_t ldq, const void* k, int64_t ldk, const void* v, int64_t ldv, const void* o, int64_t ldo,
                  const void* d_o, int64_t lddo, const float* lse, float* delta, void* dq, int64_t lddq, void* dk, int64_t lddk,
                  void* dv, int64_t lddv, float* dq_acc, spb_stream_t stream) {
  return flash_bwd_any(dtype, impl, B, T, Hq, Hkv, hd, causal, scale, q, ldq, k, ldk, v, ldv, o, ldo, d_o, lddo, lse, delta, dq, lddq, dk,
                       lddk, dv, lddv, dq_acc, stream, nullptr);
}

static int flash_bwd_any(int dtype, int impl, int32_t B, int32_t T, int32_t Hq, int32_t Hkv, int32_t hd, int causal, float scale,
                         const void* q, int64_t ldq, const void* k, int64_t ldk, const void* v, int64_t ldv, const void* o, int64_t ldo,
                         const void* d_o, int64_t lddo, const float* lse, float* delta, void* dq, int64_t lddq, void* dk, int64_t lddk,
                         void* dv, int64_t lddv, float* dq_acc, spb_stream_t stream, const float* rope, int phase) {
  SPB_CHECK_ARG(q && k && v && o && d_o && lse && delta && dq && dk && dv, SPB_E_INVALID, "spb_flash_bwd: null pointer");
  SPB_CHECK_ARG(B > 0 && T > 0 && Hq > 0 && Hkv > 0 && hd > 0 && Hq % Hkv == 0, SPB_E_INVALID, "spb_flash_bwd: bad shape");
  cudaStream_t st = (cudaStream_t)stream;
  const int use = pick_attn_bwd_impl(impl, dtype, hd);
  if (use == SPB_IMPL_TCGEN05) {
    SPB_CHECK_ARG(flash_bwd_tc_supported(dtype, hd), SPB_E_UNSUPPORTED, "spb_flash_bwd: tcgen05 path needs bf16 and hd in {64,128,256} (dtype=%d hd=%d)", dtype, hd);
    SPB_CHECK_ARG(dq_acc != nullptr, SPB_E_INVALID, "spb_flash_bwd: the tcgen05 path needs the fp32 dq_acc scratch");
    return flash_bwd_tc(B, T, Hq, Hkv, hd, causal, scale, q, ldq, k, ldk, v, ldv, o, ldo, d_o, lddo, lse, delta, dq, lddq, dk, lddk, dv,
                        lddv, dq_acc, st, rope, phase);
  }
  SPB_CHECK_ARG(use == SPB_IMPL_SIMT, SPB_E_INVALID, "spb_flash_bwd: impl %d", impl);
  if (!(phase & FB_PHASE_MAIN)) return 0;  // the FFMA path has no separate preparation
  int rc = attn_simt_bwd(dtype, B, T, Hq, Hkv, hd, causal, scale, q, ldq, k, ldk, v, ldv, o, ldo, d_o, lddo, lse, delta, dq, lddq, dk,
                         lddk, dv, lddv, st);
  if (rc || rope == nullptr) return rc;
  rc = spb_rope_apply(dtype, dq, lddq, (int64_t)B * T, T, Hq, hd, rope, 1, stream);
  if (rc) return rc;
  return spb_rope_apply(dtype, dk, lddk, (int64_t)B * T, T, Hkv, hd, rope, 1, stream);
}

size_t spb_attn_wide_scratch_bytes(int32_t B, int32_t T, int32_t Hq, int32_t Hkv, int32_t hd) {
  if (B <= 0 || T <= 0 || Hq <= 0 || Hkv <= 0 || hd <= 0) return 0;
  return attn_wide_scratch_bytes(B, T, Hq, Hkv, hd);
}

int spb_attn_wide_fwd(int32_t B, int32_t T, int32_t Hq, int32_t Hkv, int32_t hd, int causal, float scale, const void* q, int64_t ldq, const void* k,
                      int64_t ldk, const void* v, int64_t ldv, void* o, int64_t ldo, float* lse, void* scratch, spb_stream_t stream) {
  SPB_CHECK_ARG(q && k && v && o && lse && scratch, SPB_E_INVALID, "spb_attn_wide_fwd: null pointer");
  SPB_CHECK_ARG(B > 0 && T > 0 && Hq > 0 && Hkv > 0 && hd > 0 && Hq % Hkv == 0, SPB_E_INVALID, "spb_attn_wide_fwd: bad shape");
  SPB_CHECK_ARG(attn_wide_supported(SPB_BF16, hd), SPB_E_UNSUPPORTED, "spb_attn_wide_fwd: bf16 with hd > 256 and hd %% 8 == 0 (hd=%d); smaller heads use spb_flash_fwd", hd);
  return attn_wide_fwd(B, T, Hq, Hkv, hd, causal, scale, q, ldq, k, ldk, v, ldv, o, ldo, lse, scratch, (cudaStream_t)stream);
}

int spb_attn_wide_bwd(int32_t B, int32_t T, int32_t Hq, int32_t Hkv, int32_t hd, int causal, float scale, const void* q, int64_t ldq, const void* k,
                      int64_t ldk, const void* v, int64_t ldv, const void* o, int64_t ldo, const void* d_o, int64_t lddo, const float* lse,
                      float* delta, void* dq, int64_t lddq, void* dk, int64_t lddk, void* dv, int64_t lddv, void* scratch, spb_stream_t stream) {
  SPB_CHECK_ARG(q && k && v && o && d_o && lse && delta && dq && dk && dv && scratch, SPB_E_INVALID, "spb_attn_wide_bwd: null pointer");
  SPB_CHECK_ARG(B > 0 && T > 0 && Hq > 0 && Hkv > 0 && hd > 0 && Hq % Hkv == 0, SPB_E_INVALID, "spb_attn_wide_bwd: bad shape");
  SPB_CHECK_ARG(attn_wide_supported(SPB_BF16, hd), SPB_E_UNSUPPORTED, "spb_attn_wide_bwd: bf16 with hd > 256 and hd %% 8 == 0 (hd=%d)", hd);
  return attn_wide_bwd(B, T, Hq, Hkv, hd, causal, scale, q, ldq, k, ldk, v, ldv, o, ldo, d_o, lddo, lse, delta, dq, lddq, dk, lddk, dv, lddv, scratch,
                       (cudaStream_t)stream);
}

// ------------------------------------------------------------------------------------------------ block
int spb_block_ws_layout(const spb_block_desc* d, spb_ws_layout* out) {
  int rc = check_desc(d);
  if (rc) return rc;
  SPB_CHECK_ARG(out != nullptr, SPB_E_INVALID, "spb_block_ws_layout: out is NULL");
  const size_t N = (size_t)d->B * d->T, e = esize(d->dtype);
  const size_t Nq = (size_t)d->Hq * d->hd, Nkv = (size_t)d->Hkv * d->hd, Ntot = Nq + 2 * Nkv;
  size_t off = 0;
  auto take = [&](size_t bytes) {
    size_t o = off;
    off = align_up(off + bytes, 256);
    return o;
  };
  memset(out, 0, sizeof(*out));
  out->ld_qkv = (int64_t)Ntot;
  out->xhat = take(N * d->C * e);
  out->stats = take(N * 2 * sizeof(float));
  out->qkv = take(N * Ntot * e);
  out->attn_o = take(N * Nq * e);
  out->lse = take((size_t)d->B * d->Hq * d->T * sizeof(float));
  out->delta = take((size_t)d->B * d->Hq * d->T * sizeof(float));
  out->dattn_o = take(N * Nq * e);
  out->dqkv = take(N * Ntot * e);
  out->dq_acc = take(N * Nq * sizeof(float));
  out->dxhat = take(N * d->C * e);
  size_t widest = Ntot > (size_t)d->C ? Ntot : (size_t)d->C;
  out->partial_bytes = spb_norm_bwd_partial_bytes((int64_t)N, (int32_t)widest);
  out->partial = take(out->partial_bytes);
  if (block_uses_wide_attn(d)) {
    out->attn_scratch_bytes = attn_wide_scratch_bytes(d->B, d->T, d->Hq, d->Hkv, d->hd);
    out->attn_scratch = take(out->attn_scratch_bytes);
  }
  if (d->dtype == SPB_F32 && d->gemm_impl != SPB_IMPL_SIMT) {  // fp32 mode: split scratch for the largest GEMM of the block
    const int64_t Nn = (int64_t)N, Cc = d->C, Nt = (int64_t)Ntot, Nqq = (int64_t)Nq;
    const int64_t shapes[6][3] = {{Nn, Nt, Cc}, {Nn, Cc, Nqq}, {Nn, Nqq, Cc}, {Nqq, Cc, Nn}, {Cc, Nt, Nn}, {Nn, Cc, Nt}};
    size_t worst = 0;
    for (auto& sh : shapes) {
      const size_t b = gemm_f32_split_scratch_bytes(sh[0], sh[1], sh[2]);
      if (b > worst) worst = b;
    }
    out->gemm_scratch_bytes = worst;
    out->gemm_scratch = take(worst);
  }
  out->total = off;
  return 0;
}

// Pointer and alignment checks shared by spb_block_fwd and spb_block_bwd: both directions take the same descriptor,
// weights and workspace, and a mismatch between the two calls must be an error return, not a device fault.
static int check_block_params(const char* who, const spb_block_desc* d, const spb_block_params* p, const void* x, const void* ws) {
  SPB_CHECK_ARG(p && x && ws, SPB_E_INVALID, "%s: null pointer", who);
  SPB_CHECK_ARG(p->wq && p->wk && p->wv && p->wo, SPB_E_INVALID, "%s: null weight", who);
  SPB_CHECK_ARG(d->norm_kind == SPB_NORM_NONE || p->norm_w, SPB_E_INVALID, "%s: norm weight is NULL", who);
  SPB_CHECK_ARG(d->rope_kind == SPB_ROPE_NONE || p->rope, SPB_E_INVALID, "%s: rope table is NULL", who);
  SPB_CHECK_ARG(d->rope_kind == SPB_ROPE_NONE || (uintptr_t)p->rope % 16 == 0, SPB_E_INVALID, "%s: rope table must be 16-byte aligned", who);
  SPB_CHECK_ARG(!d->qkv_bias || (p->bq && p->bk && p->bv), SPB_E_INVALID, "%s: qkv_bias set but a bias pointer is NULL", who);
  SPB_CHECK_ARG(!d->out_bias || p->bo, SPB_E_INVALID, "%s: out_bias set but bo is NULL", who);
  SPB_CHECK_ARG((uintptr_t)ws % 256 == 0, SPB_E_INVALID, "%s: workspace must be 256-byte aligned", who);
  return 0;
}

int spb_block_fwd(const spb_block_desc* d, const spb_block_params* p, const void* x, void* y, void* ws, spb_stream_t stream) {
  int rc = check_desc(d);
  if (rc) return rc;
  SPB_CHECK_ARG(y != nullptr, SPB_E_INVALID, "spb_block_fwd: y is NULL");
  rc = check_block_params("spb_block_fwd", d, p, x, ws);
  if (rc) return rc;
  spb_ws_layout L;
  rc = spb_block_ws_layout(d, &L);
  if (rc) return rc;
  char* w8 = (char*)ws;
  SplitScratch split_guard(w8 + L.gemm_scratch, L.gemm_scratch_bytes);
  const int64_t N = (int64_t)d->B * d->T;
  const int64_t Nq = (int64_t)d->Hq * d->hd, Nkv = (int64_t)d->Hkv * d->hd, Ntot = Nq + 2 * Nkv;
  const size_t e = esize(d->dtype);
  const int dt = d->dtype;
  const int gimpl = d->gemm_impl;
  const bool in_out = d->weight_layout == SPB_W_IN_OUT;

  cudaStream_t pst = (cudaStream_t)stream;
  prof_new_step();
  prof_mark("norm_fwd", pst);
  // 1. norm
  const void* xin = x;
  if (d->norm_kind != SPB_NORM_NONE) {
    rc = spb_norm_fwd(dt, d->norm_kind, d->norm_upcast, d->norm_eps, x, p->norm_w, p->norm_b, w8 + L.xhat, (float*)(w8 + L.stats), N,
                      d->C, stream);
    if (rc) return rc;
    xin = w8 + L.xhat;
  }
  // 2. QKV projection (one GEMM when the three weights are one packed array)
  prof_mark("gemm_qkv", pst);
  char* qkv = w8 + L.qkv;
  const char *wq = (const char*)p->wq, *wk = (const char*)p->wk, *wv = (const char*)p->wv;
  bool packed = p->ldw_q == p->ldw_k && p->ldw_k == p->ldw_v;
  if (in_out) packed = packed && wk == wq + Nq * e && wv == wk + Nkv * e;
  else packed = packed && wk == wq + Nq * p->ldw_q * e && wv == wk + Nkv * p->ldw_q * e;
  if (d->qkv_bias) packed = packed && (const char*)p->bk == (const char*)p->bq + Nq * e && (const char*)p->bv == (const char*)p->bk + Nkv * e;
  const int tB = in_out ? 0 : 1;
  // 3. rotary on the q|k column range: fused into the projection (tcgen05 epilogue / kernel right after the FFMA GEMM)
  const float* rope = d->rope_kind == SPB_ROPE_TABLE ? p->rope : nullptr;
  if (packed) {
    rc = gemm_any(dt, dt, gimpl, N, Ntot, d->C, xin, d->C, 0, wq, p->ldw_q, tB, qkv, Ntot, d->qkv_bias ? p->bq : nullptr, 0, pst, rope,
                  (int)(Nq + Nkv), d->T, d->hd);
    if (rc) return rc;
  } else {
    rc = gemm_any(dt, dt, gimpl, N, Nq, d->C, xin, d->C, 0, wq, p->ldw_q, tB, qkv, Ntot, d->qkv_bias ? p->bq : nullptr, 0, pst, rope,
                  (int)Nq, d->T, d->hd);
    if (rc) return rc;
    rc = gemm_any(dt, dt, gimpl, N, Nkv, d->C, xin, d->C, 0, wk, p->ldw_k, tB, qkv + Nq * e, Ntot, d->qkv_bias ? p->bk : nullptr, 0, pst,
                  rope, (int)Nkv, d->T, d->hd);
    if (rc) return rc;
    rc = spb_gemm(dt, dt, gimpl, N, Nkv, d->C, xin, d->C, 0, wv, p->ldw_v, tB, qkv + (Nq + Nkv) * e, Ntot, d->qkv_bias ? p->bv : nullptr, 0,
                  stream);
    if (rc) return rc;
  }
  // 4. attention
  prof_mark("flash_fwd", pst);
  if (block_uses_wide_attn(d))
    rc = attn_wide_fwd(d->B, d->T, d->Hq, d->Hkv, d->hd, d->causal, d->scale, qkv, Ntot, qkv + Nq * e, Ntot, qkv + (Nq + Nkv) * e, Ntot,
                       w8 + L.attn_o, Nq, (float*)(w8 + L.lse), w8 + L.attn_scratch, pst);
  else
    rc = spb_flash_fwd(dt, d->attn_impl, d->B, d->T, d->Hq, d->Hkv, d->hd, d->causal, d->scale, qkv, Ntot, qkv + Nq * e, Ntot,
                       qkv + (Nq + Nkv) * e, Ntot, w8 + L.attn_o, Nq, (float*)(w8 + L.lse), stream);
  if (rc) return rc;
  // 5. output projection
  prof_mark("gemm_out", pst);
  // residual = 1: y = x + attn . wo (llama3:885 `h = x + attention(...)`), added in the GEMM epilogue
  rc = gemm_any(dt, dt, gimpl, N, d->C, Nq, w8 + L.attn_o, Nq, 0, p->wo, p->ldw_o, tB, y, d->C, d->out_bias ? p->bo : nullptr, 0, pst, nullptr,
                0, 0, 0, d->residual ? x : nullptr, d->C);
  prof_mark("fwd_end", pst);
  return rc;
}

static int block_bwd_impl(const spb_block_desc* d, const spb_block_params* p, const void* x, const void* dy, void* ws,
                          const spb_block_grads* g, void* comm, spb_stream_t stream);

int spb_block_bwd(const spb_block_desc* d, const spb_block_params* p, const void* x, const void* dy, void* ws,
                  const spb_block_grads* g, void* comm, spb_stream_t stream) {
  const int rc = block_bwd_impl(d, p, x, dy, ws, g, comm, stream);
  // an error return after a side-stream all-reduce was issued must not leave that work un-joined
  if (rc != 0 && comm != nullptr) comm_side_cleanup(comm, (cudaStream_t)stream);
  return rc;
}

static int block_bwd_impl(const spb_block_desc* d, const spb_block_params* p, const void* x, const void* dy, void* ws,
                          const spb_block_grads* g, void* comm, spb_stream_t stream) {
  int rc = check_desc(d);
  if (rc) return rc;
  SPB_CHECK_ARG(dy && g, SPB_E_INVALID, "spb_block_bwd: null pointer");
  rc = check_block_params("spb_block_bwd", d, p, x, ws);
  if (rc) return rc;
  spb_ws_layout L;
  rc = spb_block_ws_layout(d, &L);
  if (rc) return rc;
  char* w8 = (char*)ws;
  SplitScratch split_guard(w8 + L.gemm_scratch, L.gemm_scratch_bytes);
  const int64_t N = (int64_t)d->B * d->T;
  const int64_t Nq = (int64_t)d->Hq * d->hd, Nkv = (int64_t)d->Hkv * d->hd, Ntot = Nq + 2 * Nkv;
  const size_t e = esize(d->dtype);
  const int dt = d->dtype;
  const int gimpl = d->gemm_impl;
  const bool in_out = d->weight_layout == SPB_W_IN_OUT;
  const void* xin = (d->norm_kind != SPB_NORM_NONE) ? (const void*)(w8 + L.xhat) : x;
  char* qkv = w8 + L.qkv;
  char* dqkv = w8 + L.dqkv;
  void* partial = w8 + L.partial;
  cudaStream_t st = (cudaStream_t)stream;

  // 1. d(attn_o) = dy . wo^T.  When both the GEMM and the attention backward are the tcgen05 kernels, the backward's
  //    preparation (delta = rowsum(dO * O) per head, zeroed fp32 dq accumulator) rides in this GEMM's epilogue: its threads
  //    hold each dO row anyway, the epilogue has > 10x slack under the main loop, and two passes over HBM (a 134 MB memset and a
  //    kernel over o and dO: ~50 us of the C3 step) disappear.  SPB_BWD_FUSE_PREP=0 keeps them separate (A/B timing).
  prof_mark("gemm_dattn", st);
  static const bool fuse_prep_env = [] {
    const char* e = getenv("SPB_BWD_FUSE_PREP");
    return e == nullptr || atoi(e) != 0;
  }();
  const bool fuse_prep = fuse_prep_env && dt == SPB_BF16 && pick_gemm_impl(gimpl, dt) == SPB_IMPL_TCGEN05 &&
                         pick_attn_bwd_impl(d->attn_impl, dt, d->hd) == SPB_IMPL_TCGEN05 && flash_bwd_tc_supported(dt, d->hd);
  if (fuse_prep) {
    GemmTcFuse fz;
    fz.o = w8 + L.attn_o;
    fz.ldo = Nq;
    fz.delta = (float*)(w8 + L.delta);
    fz.zero = (float*)(w8 + L.dq_acc);
    fz.ldz = Nq;
    fz.hd = d->hd;
    fz.T = d->T;
    fz.H = d->Hq;
    rc = gemm_tc(dt, N, Nq, d->C, dy, d->C, 0, p->wo, p->ldw_o, in_out ? 1 : 0, w8 + L.dattn_o, Nq, nullptr, 0, st, nullptr, 0, 0, 0, nullptr, 0, &fz);
  } else {
    rc = spb_gemm(dt, dt, gimpl, N, Nq, d->C, dy, d->C, 0, p->wo, p->ldw_o, in_out ? 1 : 0, w8 + L.dattn_o, Nq, nullptr, 0, stream);
  }
  if (rc) return rc;
  static const bool dwo_early = [] {
    const char* e = getenv("SPB_BWD_DWO_EARLY");
    return e == nullptr || atoi(e) != 0;
  }();
  // 2. dwo, dbo (+ their all-reduce on the side stream).  Two positions (SPB_BWD_DWO_EARLY, timing experiments / tuning):
  //    early (default): right after the dattn GEMM - its all-reduce hides under the (non-persistent) attention backward
  //                     (8 GPUs: 2.59-2.60 ms per step vs 2.62-2.64 ms late; 1 GPU: no difference);
  //    late           : between the dWqkv and dx GEMMs - d(attn_o) goes straight into the attention backward.
  auto do_dwo = [&]() -> int {
    prof_mark("gemm_dwo", st);
    if (g->dwo) {
      if (in_out) rc = spb_gemm(dt, SPB_F32, gimpl, Nq, d->C, N, w8 + L.attn_o, Nq, 1, dy, d->C, 0, g->dwo, g->ldw_o, nullptr, 0, stream);
      else rc = spb_gemm(dt, SPB_F32, gimpl, d->C, Nq, N, dy, d->C, 1, w8 + L.attn_o, Nq, 0, g->dwo, g->ldw_o, nullptr, 0, stream);
      if (rc) return rc;
    }
    prof_mark("bias_grad", st);
    if (g->dbo) {
      rc = spb_colsum(dt, dy, d->C, N, d->C, g->dbo, partial, stream);
      if (rc) return rc;
    }
    // data parallelism: dwo's all-reduce runs on the side stream under the attention backward (SURVEY 8e "Overlap")
    if (comm != nullptr && (g->dwo || g->dbo)) {
      float* ptrs[2];
      size_t counts[2];
      int n = 0;
      if (g->dwo) {
        const int64_t rows = in_out ? Nq : d->C, cols = in_out ? d->C : Nq;
        SPB_CHECK_ARG(g->ldw_o == cols, SPB_E_UNSUPPORTED, "spb_block_bwd: all-reduce needs a contiguous dwo");
        ptrs[n] = g->dwo;
        counts[n++] = (size_t)(rows * cols);
      }
      if (g->dbo) {
        ptrs[n] = g->dbo;
        counts[n++] = (size_t)d->C;
      }
      rc = comm_allreduce_side(comm, ptrs, counts, n, st);
      if (rc) return rc;
    }
    return 0;
  };
  if (dwo_early) {
    rc = do_dwo();
    if (rc) return rc;
  }
  // 3. attention backward -> dq|dk|dv (post-rotary gradients)
  prof_mark("flash_bwd", st);
  // 4. rotary VJP on dq, dk: fused into the attention backward (dK epilogue + dQ fp32->bf16 conversion)
  if (block_uses_wide_attn(d)) {
    rc = attn_wide_bwd(d->B, d->T, d->Hq, d->Hkv, d->hd, d->causal, d->scale, qkv, Ntot, qkv + Nq * e, Ntot, qkv + (Nq + Nkv) * e, Ntot,
                       w8 + L.attn_o, Nq, w8 + L.dattn_o, Nq, (const float*)(w8 + L.lse), (float*)(w8 + L.delta), dqkv, Ntot, dqkv + Nq * e, Ntot,
                       dqkv + (Nq + Nkv) * e, Ntot, w8 + L.attn_scratch, st);
    if (rc) return rc;
    if (d->rope_kind == SPB_ROPE_TABLE) {  // rotary VJP on dq and dk, in place
      rc = spb_rope_apply(dt, dqkv, Ntot, N, d->T, d->Hq, d->hd, p->rope, 1, stream);
      if (rc) return rc;
      rc = spb_rope_apply(dt, dqkv + Nq * e, Ntot, N, d->T, d->Hkv, d->hd, p->rope, 1, stream);
      if (rc) return rc;
    }
  } else
  rc = flash_bwd_any(dt, d->attn_impl, d->B, d->T, d->Hq, d->Hkv, d->hd, d->causal, d->scale, qkv, Ntot, qkv + Nq * e, Ntot,
                     qkv + (Nq + Nkv) * e, Ntot, w8 + L.attn_o, Nq, w8 + L.dattn_o, Nq, (const float*)(w8 + L.lse),
                     (float*)(w8 + L.delta), dqkv, Ntot, dqkv + Nq * e, Ntot, dqkv + (Nq + Nkv) * e, Ntot, (float*)(w8 + L.dq_acc), stream,
                     d->rope_kind == SPB_ROPE_TABLE ? p->rope : nullptr, fuse_prep ? FB_PHASE_MAIN : (FB_PHASE_PRE | FB_PHASE_MAIN));
  if (rc) return rc;
  // 5. qkv bias gradients
  prof_mark("bias_grad", st);
  if (g->dbq && g->dbk && g->dbv && g->dbk == g->dbq + Nq && g->dbv == g->dbk + Nkv) {  // packed bias gradient: one pass over dqkv
    rc = spb_colsum(dt, dqkv, Ntot, N, (int32_t)Ntot, g->dbq, partial, stream);
    if (rc) return rc;
  } else {
    if (g->dbq) { rc = spb_colsum(dt, dqkv, Ntot, N, (int32_t)Nq, g->dbq, partial, stream); if (rc) return rc; }
    if (g->dbk) { rc = spb_colsum(dt, dqkv + Nq * e, Ntot, N, (int32_t)Nkv, g->dbk, partial, stream); if (rc) return rc; }
    if (g->dbv) { rc = spb_colsum(dt, dqkv + (Nq + Nkv) * e, Ntot, N, (int32_t)Nkv, g->dbv, partial, stream); if (rc) return rc; }
  }
  const char *wq = (const char*)p->wq, *wk = (const char*)p->wk, *wv = (const char*)p->wv;
  bool packed = p->ldw_q == p->ldw_k && p->ldw_k == p->ldw_v;
  if (in_out) packed = packed && wk == wq + Nq * e && wv == wk + Nkv * e;
  else packed = packed && wk == wq + Nq * p->ldw_q * e && wv == wk + Nkv * p->ldw_q * e;
  // 7. dWq, dWk, dWv
  prof_mark("gemm_dwqkv", st);
  bool gpacked = g->dwq && g->dwk && g->dwv && g->ldw_q == g->ldw_k && g->ldw_k == g->ldw_v;
  if (gpacked) {
    if (in_out) gpacked = g->dwk == g->dwq + Nq && g->dwv == g->dwk + Nkv;
    else gpacked = g->dwk == g->dwq + Nq * g->ldw_q && g->dwv == g->dwk + Nkv * g->ldw_q;
  }
  auto dw_gemm = [&](float* dw, int64_t ldw, const char* dcols, int64_t width) -> int {
    if (dw == nullptr) return 0;
    if (in_out) return spb_gemm(dt, SPB_F32, gimpl, d->C, width, N, xin, d->C, 1, dcols, Ntot, 0, dw, ldw, nullptr, 0, stream);
    return spb_gemm(dt, SPB_F32, gimpl, width, d->C, N, dcols, Ntot, 1, xin, d->C, 0, dw, ldw, nullptr, 0, stream);
  };
  // (measured and not adopted: cutting the packed dWqkv GEMM into two row halves so that the first half's all-reduce starts
  // earlier - at 8 GPUs the exposed all-reduce fell from 0.23 to 0.05 ms but the dWqkv and dx GEMMs, now both sharing the GPU
  // with NCCL, got 0.06 + 0.2 ms slower: 2.588 vs 2.593 ms per step)
  if (gpacked) {
    rc = dw_gemm(g->dwq, g->ldw_q, dqkv, Ntot);
    if (rc) return rc;
  } else {
    rc = dw_gemm(g->dwq, g->ldw_q, dqkv, Nq);
    if (rc) return rc;
    rc = dw_gemm(g->dwk, g->ldw_k, dqkv + Nq * e, Nkv);
    if (rc) return rc;
    rc = dw_gemm(g->dwv, g->ldw_v, dqkv + (Nq + Nkv) * e, Nkv);
    if (rc) return rc;
  }
  // data parallelism: the (large) dWqkv all-reduce goes to the side stream and overlaps the dx GEMM and the norm VJP
  if (comm != nullptr) {
    float* ptrs[8];
    size_t counts[8];
    int n = 0;
    if (gpacked) {
      const int64_t rows = in_out ? d->C : Ntot, cols = in_out ? Ntot : d->C;
      SPB_CHECK_ARG(g->ldw_q == cols, SPB_E_UNSUPPORTED, "spb_block_bwd: all-reduce needs a contiguous packed dWqkv");
      ptrs[n] = g->dwq;
      counts[n++] = (size_t)(rows * cols);
    } else {
      float* dws[3] = {g->dwq, g->dwk, g->dwv};
      const int64_t lds[3] = {g->ldw_q, g->ldw_k, g->ldw_v};
      const int64_t widths[3] = {Nq, Nkv, Nkv};
      for (int i = 0; i < 3; ++i) {
        if (!dws[i]) continue;
        const int64_t rows = in_out ? d->C : widths[i], cols = in_out ? widths[i] : d->C;
        SPB_CHECK_ARG(lds[i] == cols, SPB_E_UNSUPPORTED, "spb_block_bwd: all-reduce needs contiguous dWq/dWk/dWv");
        ptrs[n] = dws[i];
        counts[n++] = (size_t)(rows * cols);
      }
    }
    if (g->dbq) { ptrs[n] = g->dbq; counts[n++] = (size_t)Nq; }
    if (g->dbk) { ptrs[n] = g->dbk; counts[n++] = (size_t)Nkv; }
    if (g->dbv) { ptrs[n] = g->dbv; counts[n++] = (size_t)Nkv; }
    if (n > 0) {
      rc = comm_allreduce_side(comm, ptrs, counts, n, st);
      if (rc) return rc;
    }
  }
  if (!dwo_early) {
    rc = do_dwo();
    if (rc) return rc;
  }
  // 6. d(xhat) = dqkv . Wqkv^T.  With a communicator this GEMM shares the SMs with dWqkv's all-reduce on the side stream:
  //    short-lived clusters instead of the persistent grid (8 GPUs, r02h: 0.365 -> 0.31 ms; the weight-gradient GEMMs keep
  //    the persistent kernel and its K-split tail - the high-priority side stream gets dWo's all-reduce out of their way).
  struct SharedSmHint {
    bool on;
    explicit SharedSmHint(bool o) : on(o) {
      static const int tiles = [] { const char* e = getenv("SPB_GEMM_SHARED_TILES"); return e ? atoi(e) : 2; }();
      if (on) gemm_tc_hint_shared_sms(tiles);
    }
    ~SharedSmHint() { if (on) gemm_tc_hint_shared_sms(0); }
  };
  prof_mark("gemm_dx", st);
  void* dxhat = (d->norm_kind != SPB_NORM_NONE) ? (void*)(w8 + L.dxhat) : g->dx;
  const int tBx = in_out ? 1 : 0;
  // residual = 1: dx = d(block)/dx + dy; without a norm the dx GEMM's epilogue adds dy, with one the norm VJP's dx store does
  const void* dx_res = (d->residual && d->norm_kind == SPB_NORM_NONE) ? dy : nullptr;
  if (dxhat != nullptr) {
    SharedSmHint shared_sm_hint(comm != nullptr);
    if (packed) {
      rc = gemm_any(dt, dt, gimpl, N, d->C, Ntot, dqkv, Ntot, 0, wq, p->ldw_q, tBx, dxhat, d->C, nullptr, 0, st, nullptr, 0, 0, 0, dx_res, d->C);
      if (rc) return rc;
    } else {
      rc = gemm_any(dt, dt, gimpl, N, d->C, Nq, dqkv, Ntot, 0, wq, p->ldw_q, tBx, dxhat, d->C, nullptr, 0, st, nullptr, 0, 0, 0, dx_res, d->C);
      if (rc) return rc;
      rc = spb_gemm(dt, dt, gimpl, N, d->C, Nkv, dqkv + Nq * e, Ntot, 0, wk, p->ldw_k, tBx, dxhat, d->C, nullptr, 1, stream);
      if (rc) return rc;
      rc = spb_gemm(dt, dt, gimpl, N, d->C, Nkv, dqkv + (Nq + Nkv) * e, Ntot, 0, wv, p->ldw_v, tBx, dxhat, d->C, nullptr, 1, stream);
      if (rc) return rc;
    }
  }
  // 8. norm VJP
  prof_mark("norm_bwd", st);
  if (d->norm_kind != SPB_NORM_NONE) {
    rc = norm_bwd_add(dt, d->norm_kind, d->norm_upcast, x, p->norm_w, (const float*)(w8 + L.stats), w8 + L.dxhat, g->dx, g->dnorm_w,
                      g->dnorm_b, partial, N, d->C, d->residual ? dy : nullptr, st);
    if (rc) return rc;
  }
  // 9. the small norm gradients on the main stream, then join the side stream
  prof_mark("allreduce", st);
  if (comm != nullptr) {
    float* ptrs[2];
    size_t counts[2];
    int n = 0;
    if (g->dnorm_w && d->norm_kind != SPB_NORM_NONE) { ptrs[n] = g->dnorm_w; counts[n++] = (size_t)d->C; }
    if (g->dnorm_b && d->norm_kind == SPB_NORM_LAYER) { ptrs[n] = g->dnorm_b; counts[n++] = (size_t)d->C; }
    rc = comm_allreduce_join(comm, ptrs, counts, n, st);
    if (rc) return rc;
  }
  prof_mark("bwd_end", st);
  return 0;
}


// ------------------------------------------------------------------------------------------------ feed-forward sub-block
static int check_ffn_desc(const spb_ffn_desc* d) {
  SPB_CHECK_ARG(d != nullptr, SPB_E_INVALID, "ffn desc is NULL");
  SPB_CHECK_ARG(d->N > 0 && d->C > 0 && d->H > 0, SPB_E_INVALID, "ffn desc: non-positive dimension (N=%d C=%d H=%d)", d->N, d->C, d->H);
  SPB_CHECK_ARG(d->dtype == SPB_F32 || d->dtype == SPB_BF16, SPB_E_UNSUPPORTED, "ffn desc: dtype %d", d->dtype);
  SPB_CHECK_ARG(d->norm_kind >= SPB_NORM_NONE && d->norm_kind <= SPB_NORM_LAYER, SPB_E_INVALID, "ffn desc: norm_kind %d", d->norm_kind);
  SPB_CHECK_ARG(d->act == SPB_ACT_SWIGLU || d->act == SPB_ACT_GEGLU || d->act == SPB_ACT_GELU_TANH, SPB_E_UNSUPPORTED,
                "ffn desc: activation %d (SPB_ACT_SWIGLU, SPB_ACT_GEGLU, SPB_ACT_GELU_TANH)", d->act);
  SPB_CHECK_ARG(d->weight_layout == SPB_W_IN_OUT || d->weight_layout == SPB_W_OUT_IN, SPB_E_INVALID, "ffn desc: weight_layout %d",
                d->weight_layout);
  SPB_CHECK_ARG(d->residual == 0 || d->residual == 1, SPB_E_INVALID, "ffn desc: residual %d", d->residual);
  return 0;
}

int spb_ffn_ws_layout(const spb_ffn_desc* d, spb_ffn_ws* out) {
  int rc = check_ffn_desc(d);
  if (rc) return rc;
  SPB_CHECK_ARG(out != nullptr, SPB_E_INVALID, "spb_ffn_ws_layout: out is NULL");
  const size_t N = (size_t)d->N, e = esize(d->dtype), C = (size_t)d->C, H = (size_t)d->H;
  size_t off = 0;
  auto take = [&](size_t bytes) {
    size_t o = off;
    off = align_up(off + bytes, 256);
    return o;
  };
  memset(out, 0, sizeof(*out));
  out->xhat = take(N * C * e);
  out->stats = take(N * 2 * sizeof(float));
  out->g13 = take(N * 2 * H * e);
  out->act = take(N * H * e);
  out->dact = take(N * H * e);
  out->dg13 = take(N * 2 * H * e);
  out->dxhat = take(N * C * e);
  out->partial_bytes = spb_norm_bwd_partial_bytes((int64_t)N, d->C > d->H ? d->C : d->H);  // also the bias column sums ([H] wide)
  out->partial = take(out->partial_bytes);
  out->total = off;
  return 0;
}

// is [w1 | w3] one packed array?  in_out: [C, 2H] (w3 = w1 + H);  out_in: [2H, C] (w3 = w1 + H rows)
static bool ffn_packed(const spb_ffn_desc* d, const void* w1, const void* w3, int64_t ld) {
  const size_t e = esize(d->dtype);
  if (d->weight_layout == SPB_W_IN_OUT) return (const char*)w3 == (const char*)w1 + (size_t)d->H * e;
  return (const char*)w3 == (const char*)w1 + (size_t)d->H * (size_t)ld * e;
}

int spb_ffn_fwd(const spb_ffn_desc* d, const spb_ffn_params* p, const void* x, void* y, void* ws, spb_stream_t stream) {
  int rc = check_ffn_desc(d);
  if (rc) return rc;
  const bool gated = d->act != SPB_ACT_GELU_TANH;
  SPB_CHECK_ARG(p && x && y && ws && p->w1 && (p->w3 || !gated) && p->w2, SPB_E_INVALID, "spb_ffn_fwd: null pointer");
  spb_ffn_ws L;
  rc = spb_ffn_ws_layout(d, &L);
  if (rc) return rc;
  char* w8 = (char*)ws;
  cudaStream_t st = (cudaStream_t)stream;
  const int dt = d->dtype, gimpl = d->gemm_impl;
  const int64_t N = d->N, C = d->C, H = d->H;
  const size_t e = esize(dt);
  const bool in_out = d->weight_layout == SPB_W_IN_OUT;
  const int tB = in_out ? 0 : 1;
  // 1. ffn_norm (llama3:886 rms_norm(h, params['ffn_norm']))
  const void* xin = x;
  if (d->norm_kind != SPB_NORM_NONE) {
    SPB_CHECK_ARG(p->norm_w != nullptr, SPB_E_INVALID, "spb_ffn_fwd: norm_w is NULL");
    rc = spb_norm_fwd(dt, d->norm_kind, d->norm_upcast, d->norm_eps, x, p->norm_w, p->norm_b, w8 + L.xhat, (float*)(w8 + L.stats), N, d->C,
                      stream);
    if (rc) return rc;
    xin = w8 + L.xhat;
  }
  // 2. pre-activations: gated kinds g13 = [x.w1 | x.w3] (llama3:855: the two inner dots); plain MLP g = x.w1 + b1 (gpt:386)
  char* g13 = w8 + L.g13;
  const int64_t G = gated ? 2 * H : H;
  if (!gated) {
    rc = gemm_any(dt, dt, gimpl, N, H, C, xin, C, 0, p->w1, p->ldw_13, tB, g13, G, p->b1, 0, st, nullptr, 0, 0, 0);
    if (rc) return rc;
  } else if (ffn_packed(d, p->w1, p->w3, p->ldw_13)) {
    rc = gemm_any(dt, dt, gimpl, N, 2 * H, C, xin, C, 0, p->w1, p->ldw_13, tB, g13, G, nullptr, 0, st, nullptr, 0, 0, 0);
    if (rc) return rc;
  } else {
    rc = gemm_any(dt, dt, gimpl, N, H, C, xin, C, 0, p->w1, p->ldw_13, tB, g13, G, nullptr, 0, st, nullptr, 0, 0, 0);
    if (rc) return rc;
    rc = gemm_any(dt, dt, gimpl, N, H, C, xin, C, 0, p->w3, p->ldw_13, tB, g13 + H * e, G, nullptr, 0, st, nullptr, 0, 0, 0);
    if (rc) return rc;
  }
  // 3. a = act(gate) * value   /   a = act(g)
  rc = ffn_act_fwd(dt, d->act, g13, G, w8 + L.act, H, N, (int)H, st);
  if (rc) return rc;
  // 4. y = a . w2 (+ h: llama3:890 out = h + ffn_output, in the epilogue)
  return gemm_any(dt, dt, gimpl, N, C, H, w8 + L.act, H, 0, p->w2, p->ldw_2, tB, y, C, p->b2, 0, st, nullptr, 0, 0, 0,
                  d->residual ? x : nullptr, C);
}

static int ffn_bwd_impl(const spb_ffn_desc* d, const spb_ffn_params* p, const void* x, const void* dy, void* ws, const spb_ffn_grads* g,
                        void* comm, spb_stream_t stream);

int spb_ffn_bwd(const spb_ffn_desc* d, const spb_ffn_params* p, const void* x, const void* dy, void* ws, const spb_ffn_grads* g,
                void* comm, spb_stream_t stream) {
  const int rc = ffn_bwd_impl(d, p, x, dy, ws, g, comm, stream);
  if (rc != 0 && comm != nullptr) comm_side_cleanup(comm, (cudaStream_t)stream);
  return rc;
}

static int ffn_bwd_impl(const spb_ffn_desc* d, const spb_ffn_params* p, const void* x, const void* dy, void* ws, const spb_ffn_grads* g,
                        void* comm, spb_stream_t stream) {
  int rc = check_ffn_desc(d);
  if (rc) return rc;
  const bool gated = d->act != SPB_ACT_GELU_TANH;
  SPB_CHECK_ARG(p && x && dy && ws && g && p->w1 && (p->w3 || !gated) && p->w2, SPB_E_INVALID, "spb_ffn_bwd: null pointer");
  SPB_CHECK_ARG(d->norm_kind == SPB_NORM_NONE || p->norm_w, SPB_E_INVALID, "spb_ffn_bwd: norm_w is NULL");
  SPB_CHECK_ARG((uintptr_t)ws % 256 == 0, SPB_E_INVALID, "spb_ffn_bwd: workspace must be 256-byte aligned");
  spb_ffn_ws L;
  rc = spb_ffn_ws_layout(d, &L);
  if (rc) return rc;
  char* w8 = (char*)ws;
  cudaStream_t st = (cudaStream_t)stream;
  const int dt = d->dtype, gimpl = d->gemm_impl;
  const int64_t N = d->N, C = d->C, H = d->H;
  const size_t e = esize(dt);
  const bool in_out = d->weight_layout == SPB_W_IN_OUT;
  const void* xin = (d->norm_kind != SPB_NORM_NONE) ? (const void*)(w8 + L.xhat) : x;
  char* g13 = w8 + L.g13;
  char* dg13 = w8 + L.dg13;
  const bool packed = gated && ffn_packed(d, p->w1, p->w3, p->ldw_13);
  const int64_t G = gated ? 2 * H : H;
  void* partial = w8 + L.partial;
  // 0. bias gradient of the down projection: db2 = column sums of dy
  if (g->db2) {
    rc = spb_colsum(dt, dy, C, N, (int32_t)C, g->db2, partial, stream);
    if (rc) return rc;
  }
  // 1. da = dy . w2^T
  rc = gemm_any(dt, dt, gimpl, N, H, C, dy, C, 0, p->w2, p->ldw_2, in_out ? 1 : 0, w8 + L.dact, H, nullptr, 0, st, nullptr, 0, 0, 0);
  if (rc) return rc;
  // 2. dg13 from da and the saved pre-activations
  rc = ffn_act_bwd(dt, d->act, g13, G, w8 + L.dact, H, dg13, G, N, (int)H, st);
  if (rc) return rc;
  if (g->db1) {  // bias gradient of the up projection: column sums of dg
    rc = spb_colsum(dt, dg13, G, N, (int32_t)H, g->db1, partial, stream);
    if (rc) return rc;
  }
  // 3. weight gradients (fp32): dw13 = xin^T . dg13 first (the larger all-reduce gets the most work behind it), then dw2 = a^T . dy
  bool gpacked = gated && g->dw1 && g->dw3 && (in_out ? g->dw3 == g->dw1 + H : g->dw3 == g->dw1 + H * g->ldw_13);
  auto dw13 = [&](float* dw, const char* dcols, int64_t width) -> int {
    if (dw == nullptr) return 0;
    if (in_out) return spb_gemm(dt, SPB_F32, gimpl, C, width, N, xin, C, 1, dcols, G, 0, dw, g->ldw_13, nullptr, 0, stream);
    return spb_gemm(dt, SPB_F32, gimpl, width, C, N, dcols, G, 1, xin, C, 0, dw, g->ldw_13, nullptr, 0, stream);
  };
  if (gpacked) {
    rc = dw13(g->dw1, dg13, 2 * H);
    if (rc) return rc;
  } else {
    rc = dw13(g->dw1, dg13, H);
    if (rc) return rc;
    if (gated) {
      rc = dw13(g->dw3, dg13 + H * e, H);
      if (rc) return rc;
    }
  }
  if (comm != nullptr && (g->dw1 || (gated && g->dw3))) {
    float* ptrs[2];
    size_t counts[2];
    int n = 0;
    const int64_t rows = in_out ? C : H, cols = in_out ? H : C;
    if (gpacked) {
      SPB_CHECK_ARG(g->ldw_13 == (in_out ? 2 * H : C), SPB_E_UNSUPPORTED, "spb_ffn_bwd: all-reduce needs contiguous weight gradients");
      ptrs[n] = g->dw1;
      counts[n++] = (size_t)(2 * H * C);
    } else {
      SPB_CHECK_ARG(g->ldw_13 == cols, SPB_E_UNSUPPORTED, "spb_ffn_bwd: all-reduce needs contiguous weight gradients");
      if (g->dw1) { ptrs[n] = g->dw1; counts[n++] = (size_t)(rows * cols); }
      if (gated && g->dw3) { ptrs[n] = g->dw3; counts[n++] = (size_t)(rows * cols); }
    }
    rc = comm_allreduce_side(comm, ptrs, counts, n, st);
    if (rc) return rc;
  }
  if (g->dw2) {
    if (in_out) rc = spb_gemm(dt, SPB_F32, gimpl, H, C, N, w8 + L.act, H, 1, dy, C, 0, g->dw2, g->ldw_2, nullptr, 0, stream);
    else rc = spb_gemm(dt, SPB_F32, gimpl, C, H, N, dy, C, 1, w8 + L.act, H, 0, g->dw2, g->ldw_2, nullptr, 0, stream);
    if (rc) return rc;
    if (comm != nullptr) {
      SPB_CHECK_ARG(g->ldw_2 == (in_out ? C : H), SPB_E_UNSUPPORTED, "spb_ffn_bwd: all-reduce needs a contiguous dw2");
      float* ptrs[1] = {g->dw2};
      size_t counts[1] = {(size_t)(H * C)};
      rc = comm_allreduce_side(comm, ptrs, counts, 1, st);
      if (rc) return rc;
    }
  }
  // 4. dxhat = dg13 . [w1 | w3]^T  (+ dy through the residual when there is no norm)
  void* dxhat = (d->norm_kind != SPB_NORM_NONE) ? (void*)(w8 + L.dxhat) : g->dx;
  const void* dx_res = (d->residual && d->norm_kind == SPB_NORM_NONE) ? dy : nullptr;
  const int tBx = in_out ? 1 : 0;
  if (dxhat != nullptr) {
    if (packed) {
      rc = gemm_any(dt, dt, gimpl, N, C, 2 * H, dg13, G, 0, p->w1, p->ldw_13, tBx, dxhat, C, nullptr, 0, st, nullptr, 0, 0, 0, dx_res, C);
      if (rc) return rc;
    } else {
      rc = gemm_any(dt, dt, gimpl, N, C, H, dg13, G, 0, p->w1, p->ldw_13, tBx, dxhat, C, nullptr, 0, st, nullptr, 0, 0, 0, dx_res, C);
      if (rc) return rc;
      if (gated) {
        rc = gemm_any(dt, dt, gimpl, N, C, H, dg13 + H * e, G, 0, p->w3, p->ldw_13, tBx, dxhat, C, nullptr, 1, st, nullptr, 0, 0, 0);
        if (rc) return rc;
      }
    }
  }
  // 5. norm VJP (+ dy through the residual)
  if (d->norm_kind != SPB_NORM_NONE) {
    rc = norm_bwd_add(dt, d->norm_kind, d->norm_upcast, x, p->norm_w, (const float*)(w8 + L.stats), w8 + L.dxhat, g->dx, g->dnorm_w,
                      g->dnorm_b, partial, N, d->C, d->residual ? dy : nullptr, st);
    if (rc) return rc;
  }
  if (comm != nullptr) {
    float* ptrs[2];
    size_t counts[2];
    int n = 0;
    if (g->dnorm_w && d->norm_kind != SPB_NORM_NONE) { ptrs[n] = g->dnorm_w; counts[n++] = (size_t)d->C; }
    if (g->dnorm_b && d->norm_kind == SPB_NORM_LAYER) { ptrs[n] = g->dnorm_b; counts[n++] = (size_t)d->C; }
    rc = comm_allreduce_join(comm, ptrs, counts, n, st);
    if (rc) return rc;
    if (g->db1 || g->db2) {  // bias gradients: small, on the main stream
      float* bp[2];
      size_t bc[2];
      int nb = 0;
      if (g->db1) { bp[nb] = g->db1; bc[nb++] = (size_t)H; }
      if (g->db2) { bp[nb] = g->db2; bc[nb++] = (size_t)C; }
      rc = spb_comm_allreduce_f32(comm, bp, bc, nb, stream);
      if (rc) return rc;
    }
  }
  return 0;
}


// ------------------------------------------------------------------------------------------------ KV-cache decode
size_t spb_attn_decode_scratch_bytes(int32_t B, int32_t Tq, int32_t Hq, int32_t Hkv, int32_t hd, int32_t Tk) {
  if (B <= 0 || Tq <= 0 || Hq <= 0 || Hkv <= 0 || hd <= 0 || Tk <= 0 || Hq % Hkv != 0) return 0;
  return attn_decode_scratch_bytes(B, Tq, Hq, Hkv, hd, Tk);
}

int spb_attn_decode(int dtype, int32_t B, int32_t Tq, int32_t Hq, int32_t Hkv, int32_t hd, int32_t Tk, float scale, const void* q, int64_t ldq,
                    const void* k_cache, int64_t ldk, int64_t k_batch_stride, const void* v_cache, int64_t ldv, int64_t v_batch_stride, void* o,
                    int64_t ldo, float* lse, void* scratch, spb_stream_t stream) {
  SPB_CHECK_ARG(q && k_cache && v_cache && o && scratch, SPB_E_INVALID, "spb_attn_decode: null pointer");
  SPB_CHECK_ARG(B > 0 && Tq > 0 && Hq > 0 && Hkv > 0 && hd > 0 && Tk > 0 && Hq % Hkv == 0, SPB_E_INVALID, "spb_attn_decode: bad shape");
  SPB_CHECK_ARG(ldq >= (int64_t)Hq * hd && ldo >= (int64_t)Hq * hd && ldk >= (int64_t)Hkv * hd && ldv >= (int64_t)Hkv * hd, SPB_E_INVALID,
                "spb_attn_decode: leading dimension smaller than heads*hd");
  SPB_CHECK_ARG((uintptr_t)scratch % 16 == 0, SPB_E_INVALID, "spb_attn_decode: scratch must be 16-byte aligned");
  return attn_decode(dtype, B, Tq, Hq, Hkv, hd, Tk, scale, q, ldq, k_cache, ldk, k_batch_stride, v_cache, ldv, v_batch_stride, o, ldo, lse, scratch,
                     (cudaStream_t)stream);
}

int spb_block_decode_ws_layout(const spb_block_desc* d, int32_t Tmax, spb_decode_ws* out) {
  int rc = check_desc(d);
  if (rc) return rc;
  SPB_CHECK_ARG(out != nullptr && Tmax >= d->T && Tmax <= (1 << 24), SPB_E_INVALID, "spb_block_decode_ws_layout: bad argument (Tmax=%d Tq=%d)", Tmax, d->T);
  const size_t M = (size_t)d->B * d->T, e = esize(d->dtype);
  const size_t Nq = (size_t)d->Hq * d->hd, Nkv = (size_t)d->Hkv * d->hd, Ntot = Nq + 2 * Nkv;
  SPB_CHECK_ARG(M <= 64, SPB_E_LIMIT, "decode: B*Tq = %zu rows; at most 64 (longer inputs go through spb_block_fwd)", M);
  size_t off = 0;
  auto take = [&](size_t bytes) {
    size_t o = off;
    off = align_up(off + bytes, 256);
    return o;
  };
  memset(out, 0, sizeof(*out));
  out->xhat = take(M * d->C * e);
  out->qkv = take(M * Ntot * e);
  out->attn_o = take(M * Nq * e);
  size_t part = small_m_gemm_scratch_bytes((int)M, (int)Ntot, d->C, d->dtype);
  const size_t part_o = small_m_gemm_scratch_bytes((int)M, d->C, (int)Nq, d->dtype);
  if (part_o > part) part = part_o;
  {  // unpacked weights project q, k, v one after the other through the same scratch: their split plans differ from the packed one
    const size_t pq = small_m_gemm_scratch_bytes((int)M, (int)Nq, d->C, d->dtype), pk = small_m_gemm_scratch_bytes((int)M, (int)Nkv, d->C, d->dtype);
    if (pq > part) part = pq;
    if (pk > part) part = pk;
  }
  out->gemm_part = take(part);
  // the split plan depends on Tk; the scratch grows with the number of key splits, which is capped - take the worst over Tk <= Tmax
  size_t ap = 0;
  for (int tk = 128; ; tk *= 2) {
    const int t = tk < Tmax ? tk : Tmax;
    const size_t b = attn_decode_scratch_bytes(d->B, d->T, d->Hq, d->Hkv, d->hd, t);
    if (b > ap) ap = b;
    if (t == Tmax) break;
  }
  out->attn_part = take(ap);
  out->total = off;
  return 0;
}

// len_dev == nullptr: the cache length is the host value `cache_len`.  len_dev != nullptr (spb_block_decode_dev): the length is
// read from device memory by the kernels and advanced by Tq at the end of the step, so nothing in the launches depends on it.
static int block_decode_impl(const spb_block_desc* d, const spb_block_params* p, const void* x, void* k_cache, void* v_cache, int32_t cache_len,
                             int32_t* len_dev, int32_t Tmax, void* y, void* ws, spb_stream_t stream) {
  int rc = check_desc(d);
  if (rc) return rc;
  SPB_CHECK_ARG(y && k_cache && v_cache, SPB_E_INVALID, "spb_block_decode: null pointer");
  rc = check_block_params("spb_block_decode", d, p, x, ws);
  if (rc) return rc;
  SPB_CHECK_ARG(Tmax >= d->T, SPB_E_INVALID, "spb_block_decode: cache capacity %d below Tq %d", Tmax, d->T);
  SPB_CHECK_ARG(len_dev != nullptr || (cache_len >= 0 && cache_len + d->T <= Tmax), SPB_E_INVALID,
                "spb_block_decode: cache_len %d + Tq %d exceeds the cache capacity %d", cache_len, d->T, Tmax);
  SPB_CHECK_ARG(d->weight_layout == SPB_W_IN_OUT, SPB_E_UNSUPPORTED, "spb_block_decode: [in,out] weights only (the cache branch is the Llama notebook's)");
  SPB_CHECK_ARG(!(d->causal && d->T > 1), SPB_E_UNSUPPORTED,
                "spb_block_decode: a causal mask over Tq > 1 new tokens has no counterpart in llama3:824-825 (its mask slice does not broadcast)");
  SPB_CHECK_ARG(attn_decode_supported(d->dtype, d->hd), SPB_E_UNSUPPORTED, "spb_block_decode: head dim %d (64, 128, 256)", d->hd);
  spb_decode_ws L;
  rc = spb_block_decode_ws_layout(d, Tmax, &L);
  if (rc) return rc;
  char* w8 = (char*)ws;
  cudaStream_t st = (cudaStream_t)stream;
  const int M = d->B * d->T, dt = d->dtype;
  const int64_t Nq = (int64_t)d->Hq * d->hd, Nkv = (int64_t)d->Hkv * d->hd, Ntot = Nq + 2 * Nkv;
  const size_t e = esize(dt);
  // 1. norm
  const void* xin = x;
  if (d->norm_kind != SPB_NORM_NONE) {
    rc = spb_norm_fwd(dt, d->norm_kind, d->norm_upcast, d->norm_eps, x, p->norm_w, p->norm_b, w8 + L.xhat, nullptr, M, d->C, stream);
    if (rc) return rc;
    xin = w8 + L.xhat;
  }
  // 2. q | k | v projections: weight-read-bound small-M GEMMs
  char* qkv = w8 + L.qkv;
  float* part = (float*)(w8 + L.gemm_part);
  const char *wq = (const char*)p->wq, *wk = (const char*)p->wk, *wv = (const char*)p->wv;
  bool packed = p->ldw_q == p->ldw_k && p->ldw_k == p->ldw_v && wk == wq + Nq * e && wv == wk + Nkv * e;
  if (d->qkv_bias) packed = packed && (const char*)p->bk == (const char*)p->bq + Nq * e && (const char*)p->bv == (const char*)p->bk + Nkv * e;
  int fused_splits = 0;  // > 0: the projection left K-slice partials in `part`; the rotary/append kernel finishes it
  if (packed && small_m_mma_ok(dt, M, (int)Ntot, d->C, xin, d->C, wq, p->ldw_q)) {
    rc = small_m_mma(M, (int)Ntot, d->C, xin, d->C, wq, p->ldw_q, part, &fused_splits, st);
    if (rc) return rc;
  } else if (packed) {
    rc = small_m_gemm(dt, M, (int)Ntot, d->C, xin, d->C, wq, p->ldw_q, qkv, Ntot, d->qkv_bias ? p->bq : nullptr, nullptr, 0, part, st);
    if (rc) return rc;
  } else {
    rc = small_m_gemm(dt, M, (int)Nq, d->C, xin, d->C, wq, p->ldw_q, qkv, Ntot, d->qkv_bias ? p->bq : nullptr, nullptr, 0, part, st);
    if (rc) return rc;
    rc = small_m_gemm(dt, M, (int)Nkv, d->C, xin, d->C, wk, p->ldw_k, qkv + Nq * e, Ntot, d->qkv_bias ? p->bk : nullptr, nullptr, 0, part, st);
    if (rc) return rc;
    rc = small_m_gemm(dt, M, (int)Nkv, d->C, xin, d->C, wv, p->ldw_v, qkv + (Nq + Nkv) * e, Ntot, d->qkv_bias ? p->bv : nullptr, nullptr, 0, part, st);
    if (rc) return rc;
  }
  // 3. rotary at positions cache_len .. and the in-place append of the new k, v rows
  rc = decode_rope_append(dt, qkv, Ntot, d->B, d->T, d->Hq, d->Hkv, d->hd, d->rope_kind == SPB_ROPE_TABLE ? p->rope : nullptr, cache_len, k_cache,
                          v_cache, Nkv, (int64_t)Tmax * Nkv, st, fused_splits > 0 ? part : nullptr, fused_splits, d->qkv_bias ? p->bq : nullptr,
                          len_dev, Tmax);
  if (rc) return rc;
  // 4. attention of the new rows over the whole cache
  rc = attn_decode(dt, d->B, d->T, d->Hq, d->Hkv, d->hd, len_dev ? Tmax : cache_len + d->T, d->scale, qkv, Ntot, k_cache, Nkv, (int64_t)Tmax * Nkv,
                   v_cache, Nkv, (int64_t)Tmax * Nkv, w8 + L.attn_o, Nq, nullptr, w8 + L.attn_part, st, len_dev, d->T);
  if (rc) return rc;
  // 5. output projection (+ bias, + x); its finishing kernel advances the device-side length
  return small_m_gemm(dt, M, d->C, (int)Nq, w8 + L.attn_o, Nq, p->wo, p->ldw_o, y, d->C, d->out_bias ? p->bo : nullptr, d->residual ? x : nullptr,
                      d->C, part, st, len_dev, d->T, Tmax);
}

int spb_block_decode(const spb_block_desc* d, const spb_block_params* p, const void* x, void* k_cache, void* v_cache, int32_t cache_len,
                     int32_t Tmax, void* y, void* ws, spb_stream_t stream) {
  return block_decode_impl(d, p, x, k_cache, v_cache, cache_len, nullptr, Tmax, y, ws, stream);
}

int spb_block_decode_dev(const spb_block_desc* d, const spb_block_params* p, const void* x, void* k_cache, void* v_cache, int32_t* cache_len_dev,
                         int32_t Tmax, void* y, void* ws, spb_stream_t stream) {
  SPB_CHECK_ARG(cache_len_dev != nullptr, SPB_E_INVALID, "spb_block_decode_dev: null cache_len_dev");
  return block_decode_impl(d, p, x, k_cache, v_cache, 0, cache_len_dev, Tmax, y, ws, stream);
}

}  // extern "C"
