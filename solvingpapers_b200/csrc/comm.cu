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

struct Comm {
  ncclComm_t comm = nullptr;
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
  ncclResult_t r = api->CommInitRank(&c->comm, nranks, id, rank);
  if (r != ncclSuccess) {
    set_error("ncclCommInitRank failed: %s", api->GetErrorString ? api->GetErrorString(r) : "nccl error");
    delete c;
    return 1000 + (int)r;
  }
  SPB_CUDA(cudaStreamCreateWithFlags(&c->side, cudaStreamNonBlocking));
  SPB_CUDA(cudaEventCreateWithFlags(&c->ev_main, cudaEventDisableTiming));
  SPB_CUDA(cudaEventCreateWithFlags(&c->ev_side, cudaEventDisableTiming));
  *comm_out = c;
  return 0;
}

int spb_comm_allreduce_f32(void* comm, float* const* ptrs, const size_t* counts, int n, spb_stream_t stream) {
  SPB_CHECK_ARG(comm && ptrs && counts && n >= 0, SPB_E_INVALID, "spb_comm_allreduce_f32: bad argument");
  return allreduce_on((Comm*)comm, ptrs, counts, n, (cudaStream_t)stream);
}

int spb_comm_destroy(void* comm) {
  if (comm == nullptr) return 0;
  Comm* c = (Comm*)comm;
  NcclApi* api = nccl_api();
  if (c->side) cudaStreamSynchronize(c->side);
  if (api && api->CommDestroy && c->comm) api->CommDestroy(c->comm);
  if (c->ev_main) cudaEventDestroy(c->ev_main);
  if (c->ev_side) cudaEventDestroy(c->ev_side);
  if (c->side) cudaStreamDestroy(c->side);
  delete c;
  return 0;
}

}  // extern "C"
