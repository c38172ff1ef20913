// Thin NCCL wrapper for the two latency-bound collectives of the sharded path:
//   plan selection : one ncclAllReduce(ncclUint64, ncclMin) over packed (cost,index) keys
//   ndiv row sums  : ncclAllGather of [sx | sz] row blocks, ncclSum of the scalar loss
// libnccl is dlopen'ed lazily so the library loads on machines without NCCL (no link-time
// dependency); symbols resolve against the copy torch already loaded when hosted by PyTorch.
#include <dlfcn.h>
#include <string.h>
#include "common.cuh"
#include "peer.cuh"

namespace ndp {

typedef struct ncclComm* ncclComm_t;
typedef struct { char internal[128]; } ncclUniqueId;
enum { ncclSuccess_ = 0 };
// values from nccl.h (stable across NCCL 2.x)
enum { kNcclInt8 = 0, kNcclUint64 = 5, kNcclFloat32 = 7 };
enum { kNcclSum = 0, kNcclMin = 3 };

struct NcclApi {
  void* h = nullptr;
  int (*GetUniqueId)(ncclUniqueId*) = nullptr;
  int (*CommInitRank)(ncclComm_t*, int, ncclUniqueId, int) = nullptr;
  int (*CommDestroy)(ncclComm_t) = nullptr;
  int (*AllReduce)(const void*, void*, size_t, int, int, ncclComm_t, cudaStream_t) = nullptr;
  int (*AllGather)(const void*, void*, size_t, int, ncclComm_t, cudaStream_t) = nullptr;
  const char* (*GetErrorString)(int) = nullptr;
};

static NcclApi* nccl() {
  static NcclApi api;
  static bool tried = false;
  if (tried) return api.h ? &api : nullptr;
  tried = true;
  const char* names[] = {"libnccl.so.2", "libnccl.so"};
  for (const char* n : names) {
    api.h = dlopen(n, RTLD_NOW | RTLD_GLOBAL);
    if (api.h) break;
  }
  if (!api.h) { set_error("cannot dlopen libnccl.so.2: %s", dlerror()); return nullptr; }
#define NDP_SYM(field, name)                                                  \
  *(void**)(&api.field) = dlsym(api.h, name);                                 \
  if (!api.field) { set_error("libnccl lacks %s", name); api.h = nullptr; return nullptr; }
  NDP_SYM(GetUniqueId, "ncclGetUniqueId")
  NDP_SYM(CommInitRank, "ncclCommInitRank")
  NDP_SYM(CommDestroy, "ncclCommDestroy")
  NDP_SYM(AllReduce, "ncclAllReduce")
  NDP_SYM(AllGather, "ncclAllGather")
  NDP_SYM(GetErrorString, "ncclGetErrorString")
#undef NDP_SYM
  return &api;
}

}  // namespace ndp

using namespace ndp;

#define NDP_NCCL(call)                                                          \
  do {                                                                          \
    int r__ = (call);                                                           \
    if (r__ != 0) {                                                             \
      set_error("%s failed: %s", #call, api->GetErrorString(r__));              \
      return NDP_ERR_NCCL;                                                      \
    }                                                                           \
  } while (0)

extern "C" int ndp_comm_unique_id(void* id128_host) {
  NDP_CHECK_ARG(id128_host, "comm_unique_id: null");
  NcclApi* api = nccl();
  if (!api) return NDP_ERR_NCCL;
  ncclUniqueId id;
  NDP_NCCL(api->GetUniqueId(&id));
  memcpy(id128_host, &id, sizeof(id));
  return NDP_OK;
}

extern "C" int ndp_comm_create(ndp_comm** out, const void* id128_host, int rank, int world) {
  NDP_CHECK_ARG(out && id128_host && world >= 1 && rank >= 0 && rank < world, "comm_create: bad arguments");
  if (int rc = device_check()) return rc;
  NcclApi* api = nccl();
  if (!api) return NDP_ERR_NCCL;
  ncclUniqueId id;
  memcpy(&id, id128_host, sizeof(id));
  ndp_comm* c = new ndp_comm{};
  c->rank = rank; c->world = world;
  int r = api->CommInitRank(&c->comm, world, id, rank);
  if (r != 0) {
    set_error("ncclCommInitRank failed: %s", api->GetErrorString(r));
    delete c;
    return NDP_ERR_NCCL;
  }
  *out = c;
  return NDP_OK;
}

extern "C" int ndp_comm_allreduce_min_u64(ndp_comm* c, uint64_t* buf, int count, void* stream) {
  NDP_CHECK_ARG(c && buf && count >= 0, "allreduce_min_u64: bad arguments");
  NcclApi* api = nccl();
  if (!api) return NDP_ERR_NCCL;
  NDP_NCCL(api->AllReduce(buf, buf, (size_t)count, kNcclUint64, kNcclMin, c->comm, (cudaStream_t)stream));
  count_launch();
  return NDP_OK;
}

extern "C" int ndp_comm_allreduce_sum_f32(ndp_comm* c, float* buf, int count, void* stream) {
  NDP_CHECK_ARG(c && buf && count >= 0, "allreduce_sum_f32: bad arguments");
  NcclApi* api = nccl();
  if (!api) return NDP_ERR_NCCL;
  NDP_NCCL(api->AllReduce(buf, buf, (size_t)count, kNcclFloat32, kNcclSum, c->comm, (cudaStream_t)stream));
  count_launch();
  return NDP_OK;
}

extern "C" int ndp_comm_allgather_f32(ndp_comm* c, const float* send, float* recv, int count_per_rank,
                                      void* stream) {
  NDP_CHECK_ARG(c && send && recv && count_per_rank >= 0, "allgather_f32: bad arguments");
  NcclApi* api = nccl();
  if (!api) return NDP_ERR_NCCL;
  NDP_NCCL(api->AllGather(send, recv, (size_t)count_per_rank, kNcclFloat32, c->comm, (cudaStream_t)stream));
  count_launch();
  return NDP_OK;
}

extern "C" int ndp_comm_allgather_bytes(ndp_comm* c, const void* send, void* recv, size_t bytes_per_rank,
                                        void* stream) {
  NDP_CHECK_ARG(c && send && recv, "allgather_bytes: bad arguments");
  NcclApi* api = nccl();
  if (!api) return NDP_ERR_NCCL;
  NDP_NCCL(api->AllGather(send, recv, bytes_per_rank, kNcclInt8, c->comm, (cudaStream_t)stream));
  count_launch();
  return NDP_OK;
}

extern "C" int ndp_comm_rank(const ndp_comm* c) { return c ? c->rank : -1; }
extern "C" int ndp_comm_world(const ndp_comm* c) { return c ? c->world : -1; }

// ---------------------------------------------------------------- peer window (peer.cuh)
static void window_release(ndp_comm* c) {
  for (int r = 0; r < c->world && r < kMaxPeers; ++r)
    if (c->peer[r] && r != c->rank) cudaIpcCloseMemHandle(c->peer[r]);
  if (c->win) cudaFree(c->win);
  c->win = nullptr; c->win_bytes = 0;
  for (int r = 0; r < kMaxPeers; ++r) c->peer[r] = nullptr;
}

namespace ndp {
int comm_window(ndp_comm* c, size_t payload, cudaStream_t s, PeerWin* w) {
  memset(w, 0, sizeof(*w));
  if (c->win_state < 0 || c->world > kMaxPeers) { c->win_state = -1; return NDP_OK; }
  static const bool off = getenv("NDP_NO_PEER_WINDOW") != nullptr;      // diagnostics: force the NCCL collectives
  if (off) { c->win_state = -1; return NDP_OK; }
  const size_t need = kWinHeader + payload;
  if (!c->win || c->win_bytes < need) {
    // collective re-allocation: every rank takes this branch in the same call (same arguments on every rank)
    cudaStreamCaptureStatus cap = cudaStreamCaptureStatusNone;
    NDP_CUDA(cudaStreamIsCapturing(s, &cap));
    if (cap != cudaStreamCaptureStatusNone) { set_error("peer window: cannot (re)allocate under stream capture; run the call once eagerly first"); return NDP_ERR_ARG; }
    NcclApi* api = nccl();
    if (!api) return NDP_ERR_NCCL;
    NDP_CUDA(cudaDeviceSynchronize());                       // nobody may still read the old window after the exchange below
    size_t bytes = (need + (4u << 20) - 1) / (4u << 20) * (4u << 20);
    char* nw = nullptr;
    NDP_CUDA(cudaMalloc(&nw, bytes));
    NDP_CUDA(cudaMemset(nw, 0, bytes));
    struct Ex { cudaIpcMemHandle_t h; unsigned long long bytes; int ok; int pad; };
    Ex mine{};
    mine.bytes = bytes;
    mine.ok = 1;
    if (c->world > 1 && cudaIpcGetMemHandle(&mine.h, nw) != cudaSuccess) { mine.ok = 0; cudaGetLastError(); }
    Ex* dex = nullptr;
    Ex hex[kMaxPeers];
    NDP_CUDA(cudaMalloc(&dex, sizeof(Ex) * c->world));
    NDP_CUDA(cudaMemcpy(dex + c->rank, &mine, sizeof(Ex), cudaMemcpyHostToDevice));
    NDP_NCCL(api->AllGather(dex + c->rank, dex, sizeof(Ex), kNcclInt8, c->comm, s));
    NDP_CUDA(cudaStreamSynchronize(s));
    NDP_CUDA(cudaMemcpy(hex, dex, sizeof(Ex) * c->world, cudaMemcpyDeviceToHost));
    // every peer has synchronised its device before contributing: the old windows are idle everywhere
    window_release(c);
    int ok = 1;
    for (int r = 0; r < c->world; ++r) ok &= hex[r].ok && hex[r].bytes == bytes;
    c->win = nw; c->win_bytes = bytes; c->peer[c->rank] = nw;
    for (int r = 0; r < c->world && ok; ++r) {
      if (r == c->rank) continue;
      void* pp = nullptr;
      if (cudaIpcOpenMemHandle(&pp, hex[r].h, cudaIpcMemLazyEnablePeerAccess) != cudaSuccess) { ok = 0; cudaGetLastError(); break; }
      c->peer[r] = reinterpret_cast<char*>(pp);
    }
    // agree on the outcome (a rank that could not map a peer takes everyone to the NCCL path)
    unsigned long long vote = ok ? 1ull : 0ull;
    NDP_CUDA(cudaMemcpy(dex, &vote, 8, cudaMemcpyHostToDevice));
    NDP_NCCL(api->AllReduce(dex, dex, 1, kNcclUint64, kNcclMin, c->comm, s));
    NDP_CUDA(cudaStreamSynchronize(s));
    NDP_CUDA(cudaMemcpy(&vote, dex, 8, cudaMemcpyDeviceToHost));
    cudaFree(dex);
    if (!vote) { window_release(c); c->win_state = -1; return NDP_OK; }
    c->win_state = 1;
  }
  for (int r = 0; r < c->world; ++r) w->peer[r] = c->peer[r];
  w->rank = c->rank; w->world = c->world;
  return NDP_OK;
}

int comm_window_check(ndp_comm* c, cudaStream_t s) {
  if (!c->win) return NDP_OK;
  unsigned err = 0;
  NDP_CUDA(cudaMemcpyAsync(&err, c->win + kWinErrOff, 4, cudaMemcpyDeviceToHost, s));
  NDP_CUDA(cudaStreamSynchronize(s));
  if (err) {
    NDP_CUDA(cudaMemsetAsync(c->win + kWinErrOff, 0, 4, s));
    set_error("peer window: rank %d timed out waiting for rank %u (a peer did not make the matching call)", c->rank, err - 1);
    return NDP_ERR_NCCL;
  }
  return NDP_OK;
}
}  // namespace ndp

extern "C" int ndp_comm_peer_status(ndp_comm* c, int* mapped, void* stream) {
  NDP_CHECK_ARG(c, "comm_peer_status: null communicator");
  if (mapped) *mapped = c->win_state == 1 ? 1 : 0;
  return comm_window_check(c, (cudaStream_t)stream);
}

extern "C" void ndp_comm_destroy(ndp_comm* c) {
  if (!c) return;
  cudaDeviceSynchronize();
  window_release(c);
  NcclApi* api = nccl();
  if (api && c->comm) api->CommDestroy(c->comm);
  delete c;
}
