// spb_jax_ffi.cc - XLA FFI handlers over the C ABI of libspb200.so: the jax.ffi face of the attention block.
//
// What it replaces in the reference: the body of `attention(...)` behind `rms_norm(...)` as transformer_block calls it
// (llama3/LLaMA-jax.ipynb:881, traced under jit + value_and_grad at :993-995).  ffi/spb_jax.py registers the
// symbols below and wraps them in a jax.custom_vjp with the notebook's own function signatures; SpbBlockDecode is the
// KV-cache step of the same function (:815-829) over fixed-capacity caches.
//
// Build where JAX is installed (this image has neither jax nor the xla/ffi headers, so the file is shipped as source
// and is not compiled by build.py):
//   g++ -O2 -std=c++17 -fPIC -shared ffi/spb_jax_ffi.cc -o libspb_jax_ffi.so \
//       -I$(python -c 'import jax; print(jax.ffi.include_dir())') -Iinclude -I/usr/local/cuda/include \
//       -Lsolvingpapers_b200 -lspb200 -Wl,-rpath,'$ORIGIN/solvingpapers_b200'
//
// Handler rules and how they are met: XLA owns every buffer - the workspace and all gradients are RESULT
// buffers, the handlers never allocate, free or retain; work is enqueued on the stream XLA passes; nothing
// synchronises; errors are returned as ffi::Error, nothing throws; no mutable globals (XLA may call from one host
// thread per device).  Only xla/ffi/api/ffi.h's public C++ API is used - no XLA struct layout is restated here.
#include <cmath>
#include <cstdint>
#include <string>

#include <cuda_runtime_api.h>

#include "xla/ffi/api/ffi.h"

#include "spb200.h"

namespace ffi = xla::ffi;

namespace {

struct Shapes {
  spb_block_desc d;
  int64_t nq, nkv, ntot;
  size_t esize;
};

ffi::Error fail(const std::string& what, int rc) {
  return ffi::Error(rc < 0 ? ffi::ErrorCode::kInvalidArgument : ffi::ErrorCode::kInternal,
                    what + " failed with code " + std::to_string(rc) + ": " + spb_last_error());
}

// static arguments of attention(...) (llama3:809) + rms_norm (llama3:536) -> spb_block_desc
ffi::ErrorOr<Shapes> make_desc(const ffi::AnyBuffer& x, int32_t n_heads, int32_t n_kv_heads, float eps, bool causal, bool norm,
                               bool residual) {
  auto dims = x.dimensions();  // [B, T, C]
  if (dims.size() != 3) return ffi::Unexpected(ffi::Error(ffi::ErrorCode::kInvalidArgument, "spb: x must be [B, T, C]"));
  if (x.element_type() != ffi::F32 && x.element_type() != ffi::BF16)
    return ffi::Unexpected(ffi::Error(ffi::ErrorCode::kInvalidArgument, "spb: x must be float32 or bfloat16"));
  if (n_heads <= 0 || n_kv_heads <= 0 || dims[2] % n_heads != 0 || n_heads % n_kv_heads != 0)
    return ffi::Unexpected(ffi::Error(ffi::ErrorCode::kInvalidArgument, "spb: bad head counts"));
  Shapes s = {};
  spb_block_desc& d = s.d;
  d.B = (int32_t)dims[0];
  d.T = (int32_t)dims[1];
  d.C = (int32_t)dims[2];
  d.Hq = n_heads;
  d.Hkv = n_kv_heads;
  d.hd = d.C / n_heads;  // llama3:811 head_dim = dim // n_heads
  d.dtype = x.element_type() == ffi::BF16 ? SPB_BF16 : SPB_F32;
  d.norm_kind = norm ? SPB_NORM_RMS : SPB_NORM_NONE;
  d.norm_eps = eps;
  d.norm_upcast = 0;
  d.rope_kind = SPB_ROPE_TABLE;
  d.causal = causal ? 1 : 0;  // `mask is not None` (the tracer's values cannot be read under jit)
  d.scale = 1.0f / std::sqrt((float)d.hd);
  d.weight_layout = SPB_W_IN_OUT;  // init_attention_weights stores [in, out] (llama3:684-687)
  d.qkv_bias = 0;
  d.out_bias = 0;
  d.attn_impl = SPB_IMPL_AUTO;
  d.gemm_impl = SPB_IMPL_AUTO;
  d.residual = residual ? 1 : 0;
  s.nq = (int64_t)d.Hq * d.hd;
  s.nkv = (int64_t)d.Hkv * d.hd;
  s.ntot = s.nq + 2 * s.nkv;
  s.esize = d.dtype == SPB_BF16 ? 2 : 4;
  return s;
}

// wq, wk, wv, wo as the notebook stores them: four separate [in, out] arrays (params['attention'], llama3:683-688)
void fill_params(spb_block_params* p, const Shapes& s, const ffi::AnyBuffer& wq, const ffi::AnyBuffer& wk, const ffi::AnyBuffer& wv,
                 const ffi::AnyBuffer& wo, const ffi::AnyBuffer& norm_w, const ffi::Buffer<ffi::F32>& rope) {
  p->norm_w = s.d.norm_kind != SPB_NORM_NONE ? norm_w.untyped_data() : nullptr;
  p->norm_b = nullptr;
  p->wq = wq.untyped_data();
  p->wk = wk.untyped_data();
  p->wv = wv.untyped_data();
  p->ldw_q = s.nq;
  p->ldw_k = s.nkv;
  p->ldw_v = s.nkv;
  p->bq = p->bk = p->bv = nullptr;
  p->wo = wo.untyped_data();
  p->ldw_o = s.d.C;
  p->bo = nullptr;
  p->rope = rope.typed_data();  // [T, hd/2, 4] fp32 (cos, -sin, sin, cos): ffi/spb_jax.py builds it from freqs_cis
}

// ---- forward -------------------------------------------------------------------------------------------------------
// args   : x [B,T,C], wq, wk, wv, wo, norm_w [C], rope [T, hd/2, 4] f32
// results: y [B,T,C], k [B,T,Hkv,hd], v [B,T,Hkv,hd] (the new_cache of llama3:819), ws u8[spb_block_ws_layout().total]
ffi::Error BlockFwdImpl(cudaStream_t stream, ffi::AnyBuffer x, ffi::AnyBuffer wq, ffi::AnyBuffer wk, ffi::AnyBuffer wv, ffi::AnyBuffer wo,
                        ffi::AnyBuffer norm_w, ffi::Buffer<ffi::F32> rope, ffi::Result<ffi::AnyBuffer> y, ffi::Result<ffi::AnyBuffer> k,
                        ffi::Result<ffi::AnyBuffer> v, ffi::Result<ffi::Buffer<ffi::U8>> ws, int32_t n_heads, int32_t n_kv_heads, float eps,
                        bool causal, bool norm, bool residual) {
  auto sh = make_desc(x, n_heads, n_kv_heads, eps, causal, norm, residual);
  if (sh.has_error()) return sh.error();
  const Shapes& s = sh.value();
  spb_ws_layout lay;
  int rc = spb_block_ws_layout(&s.d, &lay);
  if (rc) return fail("spb_block_ws_layout", rc);
  if (ws->size_bytes() < lay.total) return ffi::Error(ffi::ErrorCode::kInvalidArgument, "spb: workspace result is smaller than spb_block_ws_layout().total");
  spb_block_params p = {};
  fill_params(&p, s, wq, wk, wv, wo, norm_w, rope);
  rc = spb_block_fwd(&s.d, &p, x.untyped_data(), y->untyped_data(), ws->typed_data(), stream);
  if (rc) return fail("spb_block_fwd", rc);
  // (k, v): the post-rotary k | v column ranges of the packed qkv residual, copied out as arrays that own their memory
  // (the workspace is consumed - donated - by the backward).  Strided device-to-device copies on the same stream.
  const char* qkv = (const char*)ws->typed_data() + lay.qkv;
  const size_t rows = (size_t)s.d.B * s.d.T, pitch = (size_t)s.ntot * s.esize, width = (size_t)s.nkv * s.esize;
  cudaError_t ce = cudaMemcpy2DAsync(k->untyped_data(), width, qkv + s.nq * s.esize, pitch, width, rows, cudaMemcpyDeviceToDevice, stream);
  if (ce == cudaSuccess)
    ce = cudaMemcpy2DAsync(v->untyped_data(), width, qkv + (s.nq + s.nkv) * s.esize, pitch, width, rows, cudaMemcpyDeviceToDevice, stream);
  if (ce != cudaSuccess) return ffi::Error(ffi::ErrorCode::kInternal, std::string("spb: copying new_cache failed: ") + cudaGetErrorString(ce));
  return ffi::Error::Success();
}

// ---- backward ------------------------------------------------------------------------------------------------------
// args   : x, dy [B,T,C], wq, wk, wv, wo, norm_w, rope, ws (the forward's result; aliased to the ws result: the backward
//          writes its scratch into it)
// results: dx [B,T,C] (x's dtype), dwq, dwk, dwv, dwo (fp32, the weights' shapes), dnorm_w [C] fp32, ws (alias of the input)
// attrs  : comm = 0, or the int64 value of a communicator handle from spb_comm_init_rank: the weight gradients are then
//          summed over ranks by ncclAllReduce inside spb_block_bwd (north star item 4: "allreduced by NCCL from the FFI layer")
ffi::Error BlockBwdImpl(cudaStream_t stream, ffi::AnyBuffer x, ffi::AnyBuffer dy, ffi::AnyBuffer wq, ffi::AnyBuffer wk, ffi::AnyBuffer wv,
                        ffi::AnyBuffer wo, ffi::AnyBuffer norm_w, ffi::Buffer<ffi::F32> rope, ffi::Buffer<ffi::U8> ws_in,
                        ffi::Result<ffi::AnyBuffer> dx, ffi::Result<ffi::Buffer<ffi::F32>> dwq, ffi::Result<ffi::Buffer<ffi::F32>> dwk,
                        ffi::Result<ffi::Buffer<ffi::F32>> dwv, ffi::Result<ffi::Buffer<ffi::F32>> dwo, ffi::Result<ffi::Buffer<ffi::F32>> dnorm_w,
                        ffi::Result<ffi::Buffer<ffi::U8>> ws, int32_t n_heads, int32_t n_kv_heads, float eps, bool causal, bool norm,
                        bool residual, int64_t comm) {
  auto sh = make_desc(x, n_heads, n_kv_heads, eps, causal, norm, residual);
  if (sh.has_error()) return sh.error();
  const Shapes& s = sh.value();
  spb_ws_layout lay;
  int rc = spb_block_ws_layout(&s.d, &lay);
  if (rc) return fail("spb_block_ws_layout", rc);
  if (ws->size_bytes() < lay.total) return ffi::Error(ffi::ErrorCode::kInvalidArgument, "spb: workspace is smaller than spb_block_ws_layout().total");
  if (ws->typed_data() != ws_in.typed_data()) {
    // XLA did not honour input_output_aliases (e.g. the operand is still live elsewhere): bring the residuals over
    cudaError_t ce = cudaMemcpyAsync(ws->typed_data(), ws_in.typed_data(), lay.total, cudaMemcpyDeviceToDevice, stream);
    if (ce != cudaSuccess) return ffi::Error(ffi::ErrorCode::kInternal, std::string("spb: workspace copy failed: ") + cudaGetErrorString(ce));
  }
  spb_block_params p = {};
  fill_params(&p, s, wq, wk, wv, wo, norm_w, rope);
  spb_block_grads g = {};
  g.dx = dx->untyped_data();
  g.dnorm_w = s.d.norm_kind != SPB_NORM_NONE ? dnorm_w->typed_data() : nullptr;
  g.dnorm_b = nullptr;
  g.dwq = dwq->typed_data();
  g.dwk = dwk->typed_data();
  g.dwv = dwv->typed_data();
  g.ldw_q = s.nq;
  g.ldw_k = s.nkv;
  g.ldw_v = s.nkv;
  g.dbq = g.dbk = g.dbv = nullptr;
  g.dwo = dwo->typed_data();
  g.ldw_o = s.d.C;
  g.dbo = nullptr;
  rc = spb_block_bwd(&s.d, &p, x.untyped_data(), dy.untyped_data(), ws->typed_data(), &g, (void*)(intptr_t)comm, stream);
  if (rc) return fail("spb_block_bwd", rc);
  if (s.d.norm_kind == SPB_NORM_NONE) {  // the result buffer still has to hold defined values
    cudaError_t ce = cudaMemsetAsync(dnorm_w->typed_data(), 0, dnorm_w->size_bytes(), stream);
    if (ce != cudaSuccess) return ffi::Error(ffi::ErrorCode::kInternal, std::string("spb: memset failed: ") + cudaGetErrorString(ce));
  }
  return ffi::Error::Success();
}

// ---- KV-cache decode step ---------------------------------------------------------------------------------------------
// attention(params, rms_norm(x, w), mask, freqs_cis, n_heads, n_kv_heads, cache=(k, v), position=p) of llama3:815-829 as
// `generate` (:499-511) calls it, for caches of FIXED capacity [B, Tmax, Hkv, hd] that are appended in place, with the number
// of valid rows in device memory (spb_block_decode_dev): nothing in the call depends on the length, so the step traces once
// under jit and serves every token (the notebook's concatenate-per-step form retraces for every new length).
// args   : x [B,Tq,C], wq, wk, wv, wo, norm_w, rope [>= Tmax, hd/2, 4] f32, k_cache, v_cache [B,Tmax,Hkv,hd], cache_len s32[1]
// results: y [B,Tq,C], k_cache, v_cache, cache_len (aliases of the three inputs: input_output_aliases in ffi/spb_jax.py),
//          ws u8[spb_block_decode_ws_layout().total]
ffi::Error BlockDecodeImpl(cudaStream_t stream, ffi::AnyBuffer x, ffi::AnyBuffer wq, ffi::AnyBuffer wk, ffi::AnyBuffer wv, ffi::AnyBuffer wo,
                           ffi::AnyBuffer norm_w, ffi::Buffer<ffi::F32> rope, ffi::AnyBuffer k_in, ffi::AnyBuffer v_in,
                           ffi::Buffer<ffi::S32> len_in, ffi::Result<ffi::AnyBuffer> y, ffi::Result<ffi::AnyBuffer> k_cache,
                           ffi::Result<ffi::AnyBuffer> v_cache, ffi::Result<ffi::Buffer<ffi::S32>> cache_len,
                           ffi::Result<ffi::Buffer<ffi::U8>> ws, int32_t n_heads, int32_t n_kv_heads, float eps, bool norm, bool residual) {
  auto sh = make_desc(x, n_heads, n_kv_heads, eps, /*causal=*/false, norm, residual);  // the cache branch applies no mask (llama3:824-825)
  if (sh.has_error()) return sh.error();
  const Shapes& s = sh.value();
  auto kd = k_in.dimensions();
  if (kd.size() != 4 || kd[0] != s.d.B || kd[2] != s.d.Hkv || kd[3] != s.d.hd)
    return ffi::Error(ffi::ErrorCode::kInvalidArgument, "spb: k_cache / v_cache must be [B, Tmax, n_kv_heads, head_dim]");
  const int32_t Tmax = (int32_t)kd[1];
  if (rope.dimensions().size() != 3 || rope.dimensions()[0] < Tmax)
    return ffi::Error(ffi::ErrorCode::kInvalidArgument, "spb: the rope table must cover every position of the cache ([>= Tmax, hd/2, 4])");
  spb_decode_ws lay;
  int rc = spb_block_decode_ws_layout(&s.d, Tmax, &lay);
  if (rc) return fail("spb_block_decode_ws_layout", rc);
  if (ws->size_bytes() < lay.total) return ffi::Error(ffi::ErrorCode::kInvalidArgument, "spb: workspace result is smaller than spb_block_decode_ws_layout().total");
  // XLA normally honours the aliases; when it does not (an operand still live elsewhere) the caches are brought over first
  cudaError_t ce = cudaSuccess;
  if (k_cache->untyped_data() != k_in.untyped_data())
    ce = cudaMemcpyAsync(k_cache->untyped_data(), k_in.untyped_data(), k_in.size_bytes(), cudaMemcpyDeviceToDevice, stream);
  if (ce == cudaSuccess && v_cache->untyped_data() != v_in.untyped_data())
    ce = cudaMemcpyAsync(v_cache->untyped_data(), v_in.untyped_data(), v_in.size_bytes(), cudaMemcpyDeviceToDevice, stream);
  if (ce == cudaSuccess && cache_len->typed_data() != len_in.typed_data())
    ce = cudaMemcpyAsync(cache_len->typed_data(), len_in.typed_data(), sizeof(int32_t), cudaMemcpyDeviceToDevice, stream);
  if (ce != cudaSuccess) return ffi::Error(ffi::ErrorCode::kInternal, std::string("spb: cache copy failed: ") + cudaGetErrorString(ce));
  spb_block_params p = {};
  fill_params(&p, s, wq, wk, wv, wo, norm_w, rope);
  rc = spb_block_decode_dev(&s.d, &p, x.untyped_data(), k_cache->untyped_data(), v_cache->untyped_data(), cache_len->typed_data(), Tmax,
                            y->untyped_data(), ws->typed_data(), stream);
  return rc ? fail("spb_block_decode_dev", rc) : ffi::Error::Success();
}

// ---- the two stand-alone functions of the notebook (cells c15, c17) ---------------------------------------------------
// rms_norm(x, weight, eps) llama3:536-538; stats [N, 2] f32 (mean, rstd) is the residual of its VJP
ffi::Error RmsNormImpl(cudaStream_t stream, ffi::AnyBuffer x, ffi::AnyBuffer w, ffi::Result<ffi::AnyBuffer> y, ffi::Result<ffi::Buffer<ffi::F32>> stats,
                       float eps) {
  auto dims = x.dimensions();
  if (dims.size() < 1) return ffi::Error(ffi::ErrorCode::kInvalidArgument, "spb: rms_norm needs at least one axis");
  const int64_t C = dims[dims.size() - 1];
  int64_t N = 1;
  for (size_t i = 0; i + 1 < dims.size(); ++i) N *= dims[i];
  const int dt = x.element_type() == ffi::BF16 ? SPB_BF16 : SPB_F32;
  const int rc = spb_norm_fwd(dt, SPB_NORM_RMS, 0, eps, x.untyped_data(), w.untyped_data(), nullptr, y->untyped_data(), stats->typed_data(), N,
                              (int32_t)C, stream);
  return rc ? fail("spb_norm_fwd", rc) : ffi::Error::Success();
}

// VJP of rms_norm: dx (x's dtype), dw [C] f32; `partial` is a scratch RESULT of spb_norm_bwd_partial_bytes(N, C) bytes
ffi::Error RmsNormBwdImpl(cudaStream_t stream, ffi::AnyBuffer x, ffi::AnyBuffer w, ffi::Buffer<ffi::F32> stats, ffi::AnyBuffer dy,
                          ffi::Result<ffi::AnyBuffer> dx, ffi::Result<ffi::Buffer<ffi::F32>> dw, ffi::Result<ffi::Buffer<ffi::U8>> partial) {
  auto dims = x.dimensions();
  if (dims.size() < 1) return ffi::Error(ffi::ErrorCode::kInvalidArgument, "spb: rms_norm needs at least one axis");
  const int64_t C = dims[dims.size() - 1];
  int64_t N = 1;
  for (size_t i = 0; i + 1 < dims.size(); ++i) N *= dims[i];
  if (partial->size_bytes() < spb_norm_bwd_partial_bytes(N, (int32_t)C))
    return ffi::Error(ffi::ErrorCode::kInvalidArgument, "spb: rms_norm VJP scratch is smaller than spb_norm_bwd_partial_bytes(N, C)");
  const int dt = x.element_type() == ffi::BF16 ? SPB_BF16 : SPB_F32;
  const int rc = spb_norm_bwd(dt, SPB_NORM_RMS, 0, x.untyped_data(), w.untyped_data(), stats.typed_data(), dy.untyped_data(), dx->untyped_data(),
                              dw->typed_data(), nullptr, partial->typed_data(), N, (int32_t)C, stream);
  return rc ? fail("spb_norm_bwd", rc) : ffi::Error::Success();
}

// apply_rotary_emb on ONE of xq / xk (llama3:592-601): x [B,T,H,hd] -> out, table [T, hd/2, 4]; transpose=true is its VJP
ffi::Error RopeImpl(cudaStream_t stream, ffi::AnyBuffer x, ffi::Buffer<ffi::F32> table, ffi::Result<ffi::AnyBuffer> out, bool transpose) {
  auto dims = x.dimensions();
  if (dims.size() != 4) return ffi::Error(ffi::ErrorCode::kInvalidArgument, "spb: apply_rotary_emb takes [B, T, H, hd]");
  const int dt = x.element_type() == ffi::BF16 ? SPB_BF16 : SPB_F32;
  if (out->untyped_data() != x.untyped_data()) {
    cudaError_t ce = cudaMemcpyAsync(out->untyped_data(), x.untyped_data(), x.size_bytes(), cudaMemcpyDeviceToDevice, stream);
    if (ce != cudaSuccess) return ffi::Error(ffi::ErrorCode::kInternal, std::string("spb: copy failed: ") + cudaGetErrorString(ce));
  }
  const int rc = spb_rope_apply(dt, out->untyped_data(), dims[2] * dims[3], dims[0] * dims[1], (int32_t)dims[1], (int32_t)dims[2], (int32_t)dims[3],
                                table.typed_data(), transpose ? 1 : 0, stream);
  return rc ? fail("spb_rope_apply", rc) : ffi::Error::Success();
}

}  // namespace

XLA_FFI_DEFINE_HANDLER_SYMBOL(SpbBlockFwd, BlockFwdImpl,
                              ffi::Ffi::Bind()
                                  .Ctx<ffi::PlatformStream<cudaStream_t>>()
                                  .Arg<ffi::AnyBuffer>()          // x
                                  .Arg<ffi::AnyBuffer>()          // wq
                                  .Arg<ffi::AnyBuffer>()          // wk
                                  .Arg<ffi::AnyBuffer>()          // wv
                                  .Arg<ffi::AnyBuffer>()          // wo
                                  .Arg<ffi::AnyBuffer>()          // norm_w
                                  .Arg<ffi::Buffer<ffi::F32>>()   // rope
                                  .Ret<ffi::AnyBuffer>()          // y
                                  .Ret<ffi::AnyBuffer>()          // k
                                  .Ret<ffi::AnyBuffer>()          // v
                                  .Ret<ffi::Buffer<ffi::U8>>()    // ws
                                  .Attr<int32_t>("n_heads")
                                  .Attr<int32_t>("n_kv_heads")
                                  .Attr<float>("eps")
                                  .Attr<bool>("causal")
                                  .Attr<bool>("norm")
                                  .Attr<bool>("residual"));

XLA_FFI_DEFINE_HANDLER_SYMBOL(SpbBlockBwd, BlockBwdImpl,
                              ffi::Ffi::Bind()
                                  .Ctx<ffi::PlatformStream<cudaStream_t>>()
                                  .Arg<ffi::AnyBuffer>()          // x
                                  .Arg<ffi::AnyBuffer>()          // dy
                                  .Arg<ffi::AnyBuffer>()          // wq
                                  .Arg<ffi::AnyBuffer>()          // wk
                                  .Arg<ffi::AnyBuffer>()          // wv
                                  .Arg<ffi::AnyBuffer>()          // wo
                                  .Arg<ffi::AnyBuffer>()          // norm_w
                                  .Arg<ffi::Buffer<ffi::F32>>()   // rope
                                  .Arg<ffi::Buffer<ffi::U8>>()    // ws (forward's result)
                                  .Ret<ffi::AnyBuffer>()          // dx
                                  .Ret<ffi::Buffer<ffi::F32>>()   // dwq
                                  .Ret<ffi::Buffer<ffi::F32>>()   // dwk
                                  .Ret<ffi::Buffer<ffi::F32>>()   // dwv
                                  .Ret<ffi::Buffer<ffi::F32>>()   // dwo
                                  .Ret<ffi::Buffer<ffi::F32>>()   // dnorm_w
                                  .Ret<ffi::Buffer<ffi::U8>>()    // ws (aliased)
                                  .Attr<int32_t>("n_heads")
                                  .Attr<int32_t>("n_kv_heads")
                                  .Attr<float>("eps")
                                  .Attr<bool>("causal")
                                  .Attr<bool>("norm")
                                  .Attr<bool>("residual")
                                  .Attr<int64_t>("comm"));

XLA_FFI_DEFINE_HANDLER_SYMBOL(SpbBlockDecode, BlockDecodeImpl,
                              ffi::Ffi::Bind()
                                  .Ctx<ffi::PlatformStream<cudaStream_t>>()
                                  .Arg<ffi::AnyBuffer>()          // x
                                  .Arg<ffi::AnyBuffer>()          // wq
                                  .Arg<ffi::AnyBuffer>()          // wk
                                  .Arg<ffi::AnyBuffer>()          // wv
                                  .Arg<ffi::AnyBuffer>()          // wo
                                  .Arg<ffi::AnyBuffer>()          // norm_w
                                  .Arg<ffi::Buffer<ffi::F32>>()   // rope
                                  .Arg<ffi::AnyBuffer>()          // k_cache
                                  .Arg<ffi::AnyBuffer>()          // v_cache
                                  .Arg<ffi::Buffer<ffi::S32>>()   // cache_len
                                  .Ret<ffi::AnyBuffer>()          // y
                                  .Ret<ffi::AnyBuffer>()          // k_cache (alias)
                                  .Ret<ffi::AnyBuffer>()          // v_cache (alias)
                                  .Ret<ffi::Buffer<ffi::S32>>()   // cache_len (alias, advanced by Tq)
                                  .Ret<ffi::Buffer<ffi::U8>>()    // ws
                                  .Attr<int32_t>("n_heads")
                                  .Attr<int32_t>("n_kv_heads")
                                  .Attr<float>("eps")
                                  .Attr<bool>("norm")
                                  .Attr<bool>("residual"));

XLA_FFI_DEFINE_HANDLER_SYMBOL(SpbRmsNorm, RmsNormImpl,
                              ffi::Ffi::Bind()
                                  .Ctx<ffi::PlatformStream<cudaStream_t>>()
                                  .Arg<ffi::AnyBuffer>()          // x
                                  .Arg<ffi::AnyBuffer>()          // weight
                                  .Ret<ffi::AnyBuffer>()          // y
                                  .Ret<ffi::Buffer<ffi::F32>>()   // stats
                                  .Attr<float>("eps"));

XLA_FFI_DEFINE_HANDLER_SYMBOL(SpbRmsNormBwd, RmsNormBwdImpl,
                              ffi::Ffi::Bind()
                                  .Ctx<ffi::PlatformStream<cudaStream_t>>()
                                  .Arg<ffi::AnyBuffer>()          // x
                                  .Arg<ffi::AnyBuffer>()          // weight
                                  .Arg<ffi::Buffer<ffi::F32>>()   // stats
                                  .Arg<ffi::AnyBuffer>()          // dy
                                  .Ret<ffi::AnyBuffer>()          // dx
                                  .Ret<ffi::Buffer<ffi::F32>>()   // dweight
                                  .Ret<ffi::Buffer<ffi::U8>>());  // scratch

XLA_FFI_DEFINE_HANDLER_SYMBOL(SpbRope, RopeImpl,
                              ffi::Ffi::Bind().Ctx<ffi::PlatformStream<cudaStream_t>>().Arg<ffi::AnyBuffer>().Arg<ffi::Buffer<ffi::F32>>().Ret<ffi::AnyBuffer>().Attr<bool>("transpose"));
