// comm.cu - batch-sharded data parallelism: the block's weight gradients are summed over ranks with ncclAllReduce
// (fp32, ncclSum) issued from this library.  One process per GPU; NVLink 5 / NVSwitch, NCCL picks NVLS or
// ring.  libnccl is resolved with dlopen at communicator creation, so (a) the library loads on hosts without NCCL or a
// GPU and (b) inside a torch process the already-loaded libnccl.so.2 (torch's bundled 2.28) is the one that is used -
// two NCCL copies in one process are avoided.
#include <dlfcn.h>
#include <nccl.h>  // types only; nothing is linked

#include "common.cuh"
#include "kernels.h"

namespace spb {

struct NcclApi {
  void* handle = nullptr;
  ncclResult_t (*GetUniqueId)(ncclUniqueId*) = nullptr;
  ncclResult_t (*CommInitRank)(ncclComm_t*, int, ncclUniqueId, int) = nullptr;
  ncclResult_t (*CommDestroy)(ncclComm_t) = nullptr;
  ncclResult_t (*AllReduce)(const void*, void*, size_t, ncclDataType_t, ncclRedOp_t, ncclComm_t, cudaStream_t) = nullptr;
  ncclResult_t (*CommInitRankConfig)(ncclComm_t*, int, ncclUniqueId, int, ncclConfig_t*) = nullptr;
  ncclResult_t (*CommInitAll)(ncclComm_t*, int, const int*) = nullptr;
  ncclResult_t (*MemAlloc)(void**, size_t) = nullptr;
  ncclResult_t (*MemFree)(void*) = nullptr;
  ncclResult_t (*CommRegister)(const ncclComm_t, void*, size_t, void**) = nullptr;
  ncclResult_t (*CommDeregister)(const ncclComm_t, void*) = nullptr;
  ncclResult_t (*CommWindowRegister)(ncclComm_t, void*, size_t, ncclWindow_t*, int) = nullptr;
  ncclResult_t (*CommWindowDeregister)(ncclComm_t, ncclWindow_t) = nullptr;
  ncclResult_t (*GroupStart)() = nullptr;
  ncclResult_t (*GroupEnd)() = nullptr;
  const char* (*GetErrorString)(ncclResult_t) = nullptr;
};

static NcclApi* nccl_api() {
  static NcclApi api;
  static bool tried = false;
  if (!tried) {
    tried = true;
    const char* names[] = {"libnccl.so.2", "libnccl.so"};
    for (const char* nm : names) {
      api.handle = dlopen(nm, RTLD_NOW | RTLD_GLOBAL);
      if (api.handle) break;
    }
    if (api.handle) {
      api.GetUniqueId = (decltype(api.GetUniqueId))dlsym(api.handle, "ncclGetUniqueId");
      api.CommInitRank = (decltype(api.CommInitRank))dlsym(api.handle, "ncclCommInitRank");
      api.CommDestroy = (decltype(api.CommDestroy))dlsym(api.handle, "ncclCommDestroy");
      api.AllReduce = (decltype(api.AllReduce))dlsym(api.handle, "ncclAllReduce");
      api.CommInitRankConfig = (decltype(api.CommInitRankConfig))dlsym(api.handle, "ncclCommInitRankConfig");
      api.CommInitAll = (decltype(api.CommInitAll))dlsym(api.handle, "ncclCommInitAll");
      api.MemAlloc = (decltype(api.MemAlloc))dlsym(api.handle, "ncclMemAlloc");
      api.MemFree = (decltype(api.MemFree))dlsym(api.handle, "ncclMemFree");
      api.CommRegister = (decltype(api.CommRegister))dlsym(api.handle, "ncclCommRegister");
      api.CommDeregister = (decltype(api.CommDeregister))dlsym(api.handle, "ncclCommDeregister");
      api.CommWindowRegister = (decltype(api.CommWindowRegister))dlsym(api.handle, "ncclCommWindowRegister");
      api.CommWindowDeregister = (decltype(api.CommWindowDeregister))dlsym(api.handle, "ncclCommWindowDeregister");
      api.GroupStart = (decltype(api.GroupStart))dlsym(api.handle, "ncclGroupStart");
      api.GroupEnd = (decltype(api.GroupEnd))dlsym(api.handle, "ncclGroupEnd");
      api.GetErrorString = (decltype(api.GetErrorString))dlsym(api.handle, "ncclGetErrorString");
    }
  }
  if (!api.handle || !api.GetUniqueId || !api.CommInitRank || !api.AllReduce || !api.GroupStart || !api.GroupEnd) return nullptr;
  return &api;
}

#define SPB_NCCL(api, call)                                                                               \
  do {                                                                                                    \
    ncclResult_t r__ = (call);                                                                            \
    if (r__ != ncclSuccess) {                                                                             \
      set_error("%s failed: %s", #call, (api)->GetErrorString ? (api)->GetErrorString(r__) : "nccl error"); \
      return 1000 + (int)r__;                                                                             \
    }                                                                                                     \
  } while (0)

// A gradient buffer obtained through spb_comm_alloc: NCCL's own allocator (cuMem-backed, multicast-capable) and registered
// with the communicator, so that the NVLS all-reduce works on it in place - no staging copies through NCCL's internal
// buffers, and with the "efficiency" CTA policy a handful of CTAs instead of ~32.
struct CommBuf {
  void* ptr = nullptr;
  size_t bytes = 0;
  ncclWindow_t win = nullptr;  // symmetric window (preferred)
  void* reg = nullptr;         // plain registration handle (fallback)
};
constexpr int COMM_MAX_BUFS = 32;

struct Comm {
  ncclComm_t comm = nullptr;
  CommBuf bufs[COMM_MAX_BUFS];
  int nbufs = 0;
  cudaStream_t side = nullptr;
  cudaEvent_t ev_main = nullptr, ev_side = nullptr;
  int nranks = 1, rank = 0;
  bool side_pending = false;
};

static int allreduce_on(Comm* c, float* const* ptrs, const size_t* counts, int n, cudaStream_t st) {
  NcclApi* api = nccl_api();
  SPB_CHECK_ARG(api != nullptr, SPB_E_STATE, "libnccl.so.2 could not be loaded");
  if (n <= 0) return 0;
  SPB_NCCL(api, api->GroupStart());
  for (int i = 0; i < n; ++i) {
    ncclResult_t r = api->AllReduce(ptrs[i], ptrs[i], counts[i], ncclFloat32, ncclSum, c->comm, st);
    if (r != ncclSuccess) {
      api->GroupEnd();
      set_error("ncclAllReduce failed: %s", api->GetErrorString ? api->GetErrorString(r) : "nccl error");
      return 1000 + (int)r;
    }
  }
  SPB_NCCL(api, api->GroupEnd());
  return 0;
}

int comm_allreduce_side(void* comm, float* const* ptrs, const size_t* counts, int n, cudaStream_t main) {
  Comm* c = (Comm*)comm;
  SPB_CUDA(cudaEventRecord(c->ev_main, main));
  SPB_CUDA(cudaStreamWaitEvent(c->side, c->ev_main, 0));
  int rc = allreduce_on(c, ptrs, counts, n, c->side);
  if (rc) return rc;
  SPB_CUDA(cudaEventRecord(c->ev_side, c->side));
  c->side_pending = true;
  return 0;
}

int comm_allreduce_join(void* comm, float* const* ptrs, const size_t* counts, int n, cudaStream_t main) {
  Comm* c = (Comm*)comm;
  int rc = allreduce_on(c, ptrs, counts, n, main);
  if (rc) return rc;
  if (c->side_pending) {
    SPB_CUDA(cudaStreamWaitEvent(main, c->ev_side, 0));
    c->side_pending = false;
  }
  return 0;
}

void comm_side_cleanup(void* comm, cudaStream_t main) {
  Comm* c = (Comm*)comm;
  if (c != nullptr && c->side_pending) {
    cudaStreamWaitEvent(main, c->ev_side, 0);
    c->side_pending = false;
  }
}

}  // namespace spb

using namespace spb;

extern "C" {

int spb_comm_unique_id(void* out128) {
  SPB_CHECK_ARG(out128 != nullptr, SPB_E_INVALID, "spb_comm_unique_id: null pointer");
  static_assert(sizeof(ncclUniqueId) == 128, "ncclUniqueId is 128 bytes");
  NcclApi* api = nccl_api();
  SPB_CHECK_ARG(api != nullptr, SPB_E_STATE, "libnccl.so.2 could not be loaded");
  ncclUniqueId id;
  SPB_NCCL(api, api->GetUniqueId(&id));
  memcpy(out128, &id, 128);
  return 0;
}

int spb_comm_init_rank(const void* id128, int nranks, int rank, void** comm_out) {
  SPB_CHECK_ARG(id128 && comm_out, SPB_E_INVALID, "spb_comm_init_rank: null pointer");
  SPB_CHECK_ARG(nranks >= 1 && rank >= 0 && rank < nranks, SPB_E_INVALID, "spb_comm_init_rank: rank %d of %d", rank, nranks);
  NcclApi* api = nccl_api();
  SPB_CHECK_ARG(api != nullptr, SPB_E_STATE, "libnccl.so.2 could not be loaded");
  Comm* c = new Comm();
  c->nranks = nranks;
  c->rank = rank;
  ncclUniqueId id;
  memcpy(&id, id128, 128);
  // Communicator configuration.  SPB_NCCL_CTA_POLICY (0 default, 1 efficiency: fewer CTAs where the algorithm allows it - the
  // all-reduce runs beside a persistent GEMM that wants every SM), SPB_NCCL_MAX_CTAS, SPB_NCCL_NVLS_CTAS: tuning knobs.
  auto env_int = [](const char* name, int dflt) {
    const char* e = getenv(name);
    return e ? atoi(e) : dflt;
  };
  const int policy = env_int("SPB_NCCL_CTA_POLICY", -1), max_ctas = env_int("SPB_NCCL_MAX_CTAS", -1), nvls_ctas = env_int("SPB_NCCL_NVLS_CTAS", -1);
  ncclResult_t r;
  if (api->CommInitRankConfig != nullptr && (policy >= 0 || max_ctas > 0 || nvls_ctas > 0)) {
    ncclConfig_t cfg = NCCL_CONFIG_INITIALIZER;
    if (policy >= 0) cfg.CTAPolicy = policy;
    if (max_ctas > 0) cfg.maxCTAs = max_ctas;
    if (nvls_ctas > 0) cfg.nvlsCTAs = nvls_ctas;
    r = api->CommInitRankConfig(&c->comm, nranks, id, rank, &cfg);
  } else {
    r = api->CommInitRank(&c->comm, nranks, id, rank);
  }
  if (r != ncclSuccess) {
    set_error("ncclCommInitRank failed: %s", api->GetErrorString ? api->GetErrorString(r) : "nccl error");
    delete c;
    return 1000 + (int)r;
  }
  {  // the collective's few CTAs should be placed as soon as SMs free up, ahead of the compute grids: highest priority
    int lo = 0, hi = 0;
    SPB_CUDA(cudaDeviceGetStreamPriorityRange(&lo, &hi));
    static const bool prio = [] { const char* e = getenv("SPB_COMM_PRIORITY"); return e == nullptr || atoi(e) != 0; }();
    SPB_CUDA(cudaStreamCreateWithPriority(&c->side, cudaStreamNonBlocking, prio ? hi : lo));
  }
  SPB_CUDA(cudaEventCreateWithFlags(&c->ev_main, cudaEventDisableTiming));
  SPB_CUDA(cudaEventCreateWithFlags(&c->ev_side, cudaEventDisableTiming));
  *comm_out = c;
  return 0;
}

int spb_comm_allreduce_f32(void* comm, float* const* ptrs, const size_t* counts, int n, spb_stream_t stream) {
  SPB_CHECK_ARG(comm && ptrs && counts && n >= 0, SPB_E_INVALID, "spb_comm_allreduce_f32: bad argument");
  return allreduce_on((Comm*)comm, ptrs, counts, n, (cudaStream_t)stream);
}

int spb_comm_alloc(void* comm, size_t bytes, void** ptr_out) {
  SPB_CHECK_ARG(comm && ptr_out && bytes > 0, SPB_E_INVALID, "spb_comm_alloc: bad argument");
  Comm* c = (Comm*)comm;
  NcclApi* api = nccl_api();
  SPB_CHECK_ARG(api != nullptr && api->MemAlloc && api->MemFree, SPB_E_STATE, "this libnccl has no ncclMemAlloc");
  SPB_CHECK_ARG(c->nbufs < COMM_MAX_BUFS, SPB_E_LIMIT, "spb_comm_alloc: at most %d registered buffers per communicator", COMM_MAX_BUFS);
  CommBuf b;
  b.bytes = (bytes + 4095) / 4096 * 4096;  // NCCL_WIN_REQUIRED_ALIGNMENT
  SPB_NCCL(api, api->MemAlloc(&b.ptr, b.bytes));
  // symmetric window first (collective: every rank calls spb_comm_alloc in the same order with the same size); plain
  // registration if this NCCL or this topology refuses it; unregistered (still a valid buffer) if both do
  ncclResult_t r = ncclInvalidUsage;
  // measured (r02g, C3 gradients, 2 and 8 GPUs): NCCL 2.28's symmetric-window all-reduce kernel runs these 64-100 MB buffers on 3-10
  // channels and is 2-4x slower than NVLS over a plainly registered buffer, so the window is opt-in (SPB_NCCL_WINDOW=1)
  static const bool want_window = [] { const char* e = getenv("SPB_NCCL_WINDOW"); return e != nullptr && atoi(e) != 0; }();
  if (api->CommWindowRegister && want_window) r = api->CommWindowRegister(c->comm, b.ptr, b.bytes, &b.win, NCCL_WIN_COLL_SYMMETRIC);
  if (r != ncclSuccess) {
    b.win = nullptr;
    if (api->CommRegister && api->CommRegister(c->comm, b.ptr, b.bytes, &b.reg) != ncclSuccess) b.reg = nullptr;
  }
  c->bufs[c->nbufs++] = b;
  *ptr_out = b.ptr;
  return 0;
}

// 0 = not from spb_comm_alloc, 1 = registered (ncclCommRegister), 2 = symmetric window
int spb_comm_buffer_kind(void* comm, const void* ptr) {
  if (comm == nullptr) return 0;
  Comm* c = (Comm*)comm;
  for (int i = 0; i < c->nbufs; ++i)
    if (c->bufs[i].ptr == ptr) return c->bufs[i].win ? 2 : (c->bufs[i].reg ? 1 : 0);
  return 0;
}

static void comm_release_buf(Comm* c, NcclApi* api, int i) {
  CommBuf& b = c->bufs[i];
  if (b.win && api->CommWindowDeregister) api->CommWindowDeregister(c->comm, b.win);
  if (b.reg && api->CommDeregister) api->CommDeregister(c->comm, b.reg);
  if (b.ptr && api->MemFree) api->MemFree(b.ptr);
  b = CommBuf();
}

int spb_comm_free(void* comm, void* ptr) {
  SPB_CHECK_ARG(comm && ptr, SPB_E_INVALID, "spb_comm_free: null pointer");
  Comm* c = (Comm*)comm;
  NcclApi* api = nccl_api();
  SPB_CHECK_ARG(api != nullptr, SPB_E_STATE, "libnccl.so.2 could not be loaded");
  for (int i = 0; i < c->nbufs; ++i) {
    if (c->bufs[i].ptr == ptr) {
      if (c->side) cudaStreamSynchronize(c->side);
      comm_release_buf(c, api, i);
      c->bufs[i] = c->bufs[c->nbufs - 1];
      c->bufs[--c->nbufs] = CommBuf();
      return 0;
    }
  }
  SPB_CHECK_ARG(false, SPB_E_INVALID, "spb_comm_free: pointer was not allocated by spb_comm_alloc on this communicator");
}

// single-process harness: one communicator per visible device 0..ndev-1 (ncclCommInitAll); comms_out[ndev]
int spb_comm_init_all(int ndev, void** comms_out) {
  SPB_CHECK_ARG(ndev >= 1 && ndev <= 64 && comms_out, SPB_E_INVALID, "spb_comm_init_all: bad argument");
  NcclApi* api = nccl_api();
  SPB_CHECK_ARG(api != nullptr && api->CommInitAll, SPB_E_STATE, "libnccl.so.2 could not be loaded");
  ncclComm_t raw[64];
  SPB_NCCL(api, api->CommInitAll(raw, ndev, nullptr));
  int prev = 0;
  cudaGetDevice(&prev);
  for (int i = 0; i < ndev; ++i) {
    Comm* c = new Comm();
    c->comm = raw[i];
    c->nranks = ndev;
    c->rank = i;
    cudaSetDevice(i);
    SPB_CUDA(cudaStreamCreateWithFlags(&c->side, cudaStreamNonBlocking));
    SPB_CUDA(cudaEventCreateWithFlags(&c->ev_main, cudaEventDisableTiming));
    SPB_CUDA(cudaEventCreateWithFlags(&c->ev_side, cudaEventDisableTiming));
    comms_out[i] = c;
  }
  cudaSetDevice(prev);
  return 0;
}

int spb_comm_destroy(void* comm) {
  if (comm == nullptr) return 0;
  Comm* c = (Comm*)comm;
  NcclApi* api = nccl_api();
  if (c->side) cudaStreamSynchronize(c->side);
  if (api) for (int i = c->nbufs - 1; i >= 0; --i) comm_release_buf(c, api, i);
  c->nbufs = 0;
  if (api && api->CommDestroy && c->comm) api->CommDestroy(c->comm);
  if (c->ev_main) cudaEventDestroy(c->ev_main);
  if (c->ev_side) cudaEventDestroy(c->ev_side);
  if (c->side) cudaStreamDestroy(c->side);
  delete c;
  return 0;
}

}  // extern "C"
