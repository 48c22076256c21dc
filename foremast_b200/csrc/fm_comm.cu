// fm_comm.cu -- the one exchange step of the path behind the C ABI: merging the dense per-service
// counters of the shards of a multi-GPU sweep.
//
// The reference scales with shared-nothing brain pods that meet only in Elasticsearch
// (docs/guides/design.md:37-41); here one process (or thread) per GPU scores its shard of the series
// and the per-service counters -- u32[4*S] summed, u64[S] worst-unit keys maxed -- are all-reduced
// over NVLink by NCCL, on the context's own stream, so the host never waits between scoring and
// fm_result_verdicts.  NCCL is opened at run time (dlopen "libnccl.so.2"): a single-GPU host needs
// no NCCL at all, and a process that already carries one (torch's) keeps using that copy.
#include <dlfcn.h>

#include <cstdlib>
#include <cstring>
#include <mutex>
#include <string>

#include "fm_common.cuh"
#include "fm_internal.h"

namespace {

// the handful of NCCL entry points used, with the types of nccl.h 2.x spelled out (stable ABI)
typedef struct ncclComm *ncclComm_t;
typedef struct { char internal[128]; } ncclUniqueId;
enum { kNcclSuccess = 0, kNcclSum = 0, kNcclMax = 2, kNcclUint32 = 3, kNcclUint64 = 5 };
struct NcclApi {
  void *handle = nullptr;
  int (*GetUniqueId)(ncclUniqueId *) = nullptr;
  int (*CommInitRank)(ncclComm_t *, int, ncclUniqueId, int) = nullptr;
  int (*CommDestroy)(ncclComm_t) = nullptr;
  int (*AllReduce)(const void *, void *, size_t, int, int, ncclComm_t, cudaStream_t) = nullptr;
  int (*GroupStart)() = nullptr;
  int (*GroupEnd)() = nullptr;
  const char *(*GetErrorString)(int) = nullptr;
  std::string error;
};

NcclApi *nccl_api() {
  static NcclApi api;
  static std::once_flag once;
  std::call_once(once, [] {
    // FOREMAST_GPU_NCCL_LIB names a specific copy; otherwise the soname -- which the loader resolves to a copy
    // the process already carries (a host framework's), else to the system one
    const char *forced = getenv("FOREMAST_GPU_NCCL_LIB");
    if (forced && *forced) api.handle = dlopen(forced, RTLD_NOW | RTLD_GLOBAL);
    for (const char *name : {"libnccl.so.2", "libnccl.so"}) {
      if (api.handle) break;
      api.handle = dlopen(name, RTLD_NOW | RTLD_GLOBAL);
    }
    if (!api.handle) {
      api.error = std::string("libnccl.so.2 not found: ") + (dlerror() ? dlerror() : "");
      return;
    }
    auto sym = [&](const char *n) {
      void *p = dlsym(api.handle, n);
      if (!p && api.error.empty()) api.error = std::string("NCCL symbol missing: ") + n;
      return p;
    };
    api.GetUniqueId = reinterpret_cast<decltype(api.GetUniqueId)>(sym("ncclGetUniqueId"));
    api.CommInitRank = reinterpret_cast<decltype(api.CommInitRank)>(sym("ncclCommInitRank"));
    api.CommDestroy = reinterpret_cast<decltype(api.CommDestroy)>(sym("ncclCommDestroy"));
    api.AllReduce = reinterpret_cast<decltype(api.AllReduce)>(sym("ncclAllReduce"));
    api.GroupStart = reinterpret_cast<decltype(api.GroupStart)>(sym("ncclGroupStart"));
    api.GroupEnd = reinterpret_cast<decltype(api.GroupEnd)>(sym("ncclGroupEnd"));
    api.GetErrorString = reinterpret_cast<decltype(api.GetErrorString)>(sym("ncclGetErrorString"));
  });
  return &api;
}

int comm_err(fm_ctx *ctx, const NcclApi *api, const char *what, int rc) {
  return fm_set_error(ctx, FM_E_COMM, std::string(what) + ": " + (api && api->GetErrorString ? api->GetErrorString(rc) : "NCCL unavailable"));
}

}  // namespace

extern "C" {

int fm_comm_unique_id(void *id_out, uint32_t cap) {
  if (!id_out || cap < FM_COMM_ID_BYTES) return FM_E_INVALID;
  NcclApi *api = nccl_api();
  if (!api->error.empty()) return FM_E_COMM;
  ncclUniqueId id;
  if (api->GetUniqueId(&id) != kNcclSuccess) return FM_E_COMM;
  memcpy(id_out, &id, sizeof(id));
  return FM_OK;
}

int fm_comm_init(fm_ctx *ctx, const void *unique_id, int32_t rank, int32_t world) {
  if (!ctx || !unique_id || world < 1 || rank < 0 || rank >= world) return FM_E_INVALID;
  std::lock_guard<std::mutex> g(fm_ctx_mutex(ctx));
  if (fm_ctx_comm(ctx)->comm) return fm_set_error(ctx, FM_E_STATE, "this context already has a communicator");
  NcclApi *api = nccl_api();
  if (!api->error.empty()) return fm_set_error(ctx, FM_E_COMM, api->error);
  if (cudaSetDevice(fm_ctx_device(ctx)) != cudaSuccess) return fm_set_error(ctx, FM_E_CUDA, "cudaSetDevice failed");
  ncclUniqueId id;
  memcpy(&id, unique_id, sizeof(id));
  ncclComm_t comm = nullptr;
  const int rc = api->CommInitRank(&comm, world, id, rank);
  if (rc != kNcclSuccess) return comm_err(ctx, api, "ncclCommInitRank", rc);
  fm_comm_state *cs = fm_ctx_comm(ctx);
  cs->comm = comm;
  cs->rank = rank;
  cs->world = world;
  return FM_OK;
}

int fm_comm_info(const fm_ctx *ctx, int32_t *rank, int32_t *world) {
  if (!ctx) return FM_E_INVALID;
  const fm_comm_state *cs = fm_ctx_comm(const_cast<fm_ctx *>(ctx));
  if (rank) *rank = cs->comm ? cs->rank : 0;
  if (world) *world = cs->comm ? cs->world : 1;
  return FM_OK;
}

void fm_comm_destroy(fm_ctx *ctx) {
  if (!ctx) return;
  std::lock_guard<std::mutex> g(fm_ctx_mutex(ctx));
  fm_comm_state *cs = fm_ctx_comm(ctx);
  if (!cs->comm) return;
  cudaSetDevice(fm_ctx_device(ctx));
  cudaStreamSynchronize(fm_ctx_stream(ctx));
  nccl_api()->CommDestroy(static_cast<ncclComm_t>(cs->comm));
  cs->comm = nullptr;
  cs->world = 1;
  cs->rank = 0;
}

}  // extern "C"

// The merge itself, called by fm_result_merge (fm_api.cu) with the context locked and its device current:
// both all-reduces in one NCCL group (one launch), in place, on the library's stream.
int fm_comm_merge_counters(fm_ctx *ctx, uint32_t *d_counts, unsigned long long *d_worst, uint32_t n_services) {
  fm_comm_state *cs = fm_ctx_comm(ctx);
  if (!cs->comm || cs->world == 1 || n_services == 0) return FM_OK;  // a single shard is its own merge
  NcclApi *api = nccl_api();
  ncclComm_t comm = static_cast<ncclComm_t>(cs->comm);
  cudaStream_t st = fm_ctx_stream(ctx);
  int rc = api->GroupStart();
  if (rc == kNcclSuccess) rc = api->AllReduce(d_counts, d_counts, 4ull * n_services, kNcclUint32, kNcclSum, comm, st);
  if (rc == kNcclSuccess) rc = api->AllReduce(d_worst, d_worst, n_services, kNcclUint64, kNcclMax, comm, st);
  const int rc_end = api->GroupEnd();
  if (rc == kNcclSuccess) rc = rc_end;
  if (rc != kNcclSuccess) return comm_err(ctx, api, "ncclAllReduce of the per-service counters", rc);
  return FM_OK;
}
