// fcm_comm.cu — the all-reduce of the sharded recount, inside libfcm.so (include/fcm.h section 3b).
//
// Replaces nothing in the reference (it has no distributed path); it completes the sharding of
// State::new's `graph.flagser_count()` (src/lib.rs:51): every rank counts the simplices whose source
// vertex is in its shard, the u64 per-dimension vectors are summed over the ranks of the box.
//   FCM_COMM_PEER  fused: the count kernel's last CTA pushes the vector into every rank's window over
//                  NVLink (comm_last_cta, fcm_internal.cuh).  This file only sets the windows up.
//   FCM_COMM_NCCL  ncclAllReduce behind the count kernel, libnccl.so.2 bound at run time with dlopen.
#include "fcm_internal.cuh"

#include <dlfcn.h>
#include <unistd.h>

#include <cstdlib>
#include <cstring>

#define REQUIRE(cond, msg)                                 \
  do {                                                     \
    if (!(cond)) { fcm_set_error("%s", msg); return FCM_EINVAL; } \
  } while (0)

// the blob the ranks exchange
struct CommHandle {
  u32 magic, kind, rank, nranks;
  u32 pid, device;
  unsigned long long window;              // the rank's own device pointer (valid inside its process)
  cudaIpcMemHandle_t ipc;                 // 64 bytes
  unsigned char nccl_id[128];             // rank 0 only (ncclUniqueId)
  unsigned char hostid[16];               // boot id prefix: ranks must share one box
};
static_assert(sizeof(CommHandle) <= FCM_COMM_HANDLE_BYTES, "handle blob too small");
#define COMM_MAGIC 0x434d4346u            // "FCMC"

// the NCCL entry points used, bound by name (types as in nccl.h 2.27: opaque comm, 128-byte id, int enums)
typedef struct { char internal[128]; } nccl_id_t;
struct NcclApi {
  void* lib = nullptr;
  int (*GetUniqueId)(nccl_id_t*) = nullptr;
  int (*CommInitRank)(void**, int, nccl_id_t, int) = nullptr;
  int (*CommDestroy)(void*) = nullptr;
  int (*AllReduce)(const void*, void*, size_t, int, int, void*, cudaStream_t) = nullptr;
  const char* (*GetErrorString)(int) = nullptr;
};
enum { NCCL_UINT64 = 5, NCCL_SUM = 0 };   // ncclUint64, ncclSum

struct fcm_comm {
  fcm_ctx* ctx = nullptr;
  int device = 0;
  int kind = FCM_COMM_PEER;
  u32 rank = 0, nranks = 1;
  bool connected = false, broken = false;
  CommSlot* window = nullptr;             // 2 slots, cudaMalloc (IPC-exportable)
  unsigned long long peer[FCM_COMM_MAXRANKS];
  void* ipc_open[FCM_COMM_MAXRANKS];      // mappings to close
  u64 epoch = 0;
  unsigned long long timeout_ns = 10000ULL * 1000000ULL;
  NcclApi nccl;
  void* nccl_comm = nullptr;
};

static int nccl_bind(NcclApi& a) {
  if (a.lib) return FCM_OK;
  // a host that already carries NCCL (torch, an MPI launcher) keeps using that copy
  const char* names[] = {"libnccl.so.2", "libnccl.so"};
  for (const char* nm : names) {
    a.lib = dlopen(nm, RTLD_NOW | RTLD_NOLOAD | RTLD_GLOBAL);
    if (a.lib) break;
  }
  for (int i = 0; i < 2 && !a.lib; i++) a.lib = dlopen(names[i], RTLD_NOW | RTLD_GLOBAL);
  if (!a.lib) { fcm_set_error("FCM_COMM_NCCL: libnccl.so.2 not found (%s)", dlerror()); return FCM_EUNSUPPORTED; }
  a.GetUniqueId = (int (*)(nccl_id_t*))dlsym(a.lib, "ncclGetUniqueId");
  a.CommInitRank = (int (*)(void**, int, nccl_id_t, int))dlsym(a.lib, "ncclCommInitRank");
  a.CommDestroy = (int (*)(void*))dlsym(a.lib, "ncclCommDestroy");
  a.AllReduce = (int (*)(const void*, void*, size_t, int, int, void*, cudaStream_t))dlsym(a.lib, "ncclAllReduce");
  a.GetErrorString = (const char* (*)(int))dlsym(a.lib, "ncclGetErrorString");
  if (!a.GetUniqueId || !a.CommInitRank || !a.CommDestroy || !a.AllReduce || !a.GetErrorString) {
    fcm_set_error("FCM_COMM_NCCL: libnccl lacks an entry point");
    return FCM_EUNSUPPORTED;
  }
  return FCM_OK;
}
#define NCCL_CALL(c, call)                                                                                   \
  do {                                                                                                       \
    int r__ = (call);                                                                                        \
    if (r__ != 0) { fcm_set_error("%s -> %s", #call, (c)->nccl.GetErrorString(r__)); return FCM_ECUDA; }     \
  } while (0)

static void host_id(unsigned char out[16]) {
  memset(out, 0, 16);
  FILE* f = fopen("/proc/sys/kernel/random/boot_id", "r");
  if (f) { size_t got = fread(out, 1, 16, f); (void)got; fclose(f); }
}

extern "C" int fcm_comm_create(fcm_ctx* ctx, int kind, uint32_t rank, uint32_t nranks, fcm_comm** out,
                               uint8_t handle_out[FCM_COMM_HANDLE_BYTES]) {
  REQUIRE(ctx && out && handle_out, "ctx/out/handle_out is null");
  *out = nullptr;
  REQUIRE(kind == FCM_COMM_PEER || kind == FCM_COMM_NCCL, "comm kind must be FCM_COMM_PEER or FCM_COMM_NCCL");
  REQUIRE(nranks >= 1 && nranks <= FCM_COMM_MAXRANKS && rank < nranks, "need rank < nranks <= FCM_COMM_MAXRANKS");
  FCM_CUDA(cudaSetDevice(ctx->device));
  fcm_comm* c = new fcm_comm();
  c->ctx = ctx; c->device = ctx->device; c->kind = kind; c->rank = rank; c->nranks = nranks;
  memset(c->peer, 0, sizeof(c->peer));
  memset(c->ipc_open, 0, sizeof(c->ipc_open));
  if (const char* t = getenv("FCM_COMM_TIMEOUT_MS")) { const double ms = atof(t); if (ms > 0) c->timeout_ns = (unsigned long long)(ms * 1e6); }
  CommHandle h;
  memset(&h, 0, sizeof(h));
  h.magic = COMM_MAGIC; h.kind = (u32)kind; h.rank = rank; h.nranks = nranks;
  h.pid = (u32)getpid(); h.device = (u32)ctx->device;
  host_id(h.hostid);
  if (kind == FCM_COMM_PEER) {
    cudaError_t e = cudaMalloc(&c->window, 2 * sizeof(CommSlot));
    if (e == cudaSuccess) e = cudaMemset(c->window, 0, 2 * sizeof(CommSlot));
    if (e == cudaSuccess) e = cudaDeviceSynchronize();
    if (e != cudaSuccess) { fcm_set_error("fcm_comm_create: %s", cudaGetErrorString(e)); fcm_comm_destroy(c); return FCM_ECUDA; }
    h.window = (unsigned long long)c->window;
    if (nranks > 1) {
      e = cudaIpcGetMemHandle(&h.ipc, c->window);
      // no IPC on this platform: the ranks can still share one process (connect checks)
      if (e != cudaSuccess) { cudaGetLastError(); memset(&h.ipc, 0, sizeof(h.ipc)); }
    }
  } else {
    int rc = nccl_bind(c->nccl);
    if (rc) { fcm_comm_destroy(c); return rc; }
    if (rank == 0) {
      nccl_id_t id;
      int r = c->nccl.GetUniqueId(&id);
      if (r != 0) { fcm_set_error("ncclGetUniqueId -> %s", c->nccl.GetErrorString(r)); fcm_comm_destroy(c); return FCM_ECUDA; }
      memcpy(h.nccl_id, &id, 128);
    }
  }
  memset(handle_out, 0, FCM_COMM_HANDLE_BYTES);
  memcpy(handle_out, &h, sizeof(h));
  *out = c;
  return FCM_OK;
}

extern "C" int fcm_comm_connect(fcm_comm* c, const uint8_t* handles) {
  REQUIRE(c && handles, "comm/handles is null");
  REQUIRE(!c->connected, "comm is already connected");
  FCM_CUDA(cudaSetDevice(c->ctx->device));
  CommHandle mine;
  memcpy(&mine, handles + (size_t)c->rank * FCM_COMM_HANDLE_BYTES, sizeof(mine));
  for (u32 r = 0; r < c->nranks; r++) {
    CommHandle h;
    memcpy(&h, handles + (size_t)r * FCM_COMM_HANDLE_BYTES, sizeof(h));
    if (h.magic != COMM_MAGIC || h.rank != r || h.nranks != c->nranks || (int)h.kind != c->kind) {
      fcm_set_error("fcm_comm_connect: blob %u is not rank %u's handle of a %u-rank comm of this kind", r, r, c->nranks);
      return FCM_EINVAL;
    }
    if (memcmp(h.hostid, mine.hostid, 16) != 0) { fcm_set_error("fcm_comm_connect: rank %u runs on another host (one box only)", r); return FCM_EUNSUPPORTED; }
    if (c->kind != FCM_COMM_PEER) continue;
    if (r == c->rank) { c->peer[r] = (unsigned long long)c->window; continue; }
    if (h.pid == (u32)getpid()) {
      // same process: the pointer is valid here; another device needs peer access
      if ((int)h.device != c->ctx->device) {
        int can = 0;
        FCM_CUDA(cudaDeviceCanAccessPeer(&can, c->ctx->device, (int)h.device));
        if (!can) { fcm_set_error("fcm_comm_connect: device %d cannot access device %u (no NVLink/P2P path)", c->ctx->device, h.device); return FCM_EUNSUPPORTED; }
        cudaError_t e = cudaDeviceEnablePeerAccess((int)h.device, 0);
        if (e != cudaSuccess && e != cudaErrorPeerAccessAlreadyEnabled) { fcm_set_error("cudaDeviceEnablePeerAccess(%u) -> %s", h.device, cudaGetErrorString(e)); return FCM_ECUDA; }
        cudaGetLastError();
      }
      c->peer[r] = h.window;
    } else {
      void* p = nullptr;
      cudaError_t e = cudaIpcOpenMemHandle(&p, h.ipc, cudaIpcMemLazyEnablePeerAccess);
      if (e != cudaSuccess) { cudaGetLastError(); fcm_set_error("cudaIpcOpenMemHandle(rank %u) -> %s", r, cudaGetErrorString(e)); return FCM_ECUDA; }
      c->ipc_open[r] = p;
      c->peer[r] = (unsigned long long)p;
    }
  }
  if (c->kind == FCM_COMM_NCCL) {
    CommHandle h0;
    memcpy(&h0, handles, sizeof(h0));
    nccl_id_t id;
    memcpy(&id, h0.nccl_id, 128);
    NCCL_CALL(c, c->nccl.CommInitRank(&c->nccl_comm, (int)c->nranks, id, (int)c->rank));
  }
  c->connected = true;
  return FCM_OK;
}

extern "C" int fcm_comm_destroy(fcm_comm* c) {
  if (!c) return FCM_OK;
  cudaSetDevice(c->device);
  cudaDeviceSynchronize();                 // (the context may be gone already: only the device ordinal is used)
  for (u32 r = 0; r < FCM_COMM_MAXRANKS; r++) if (c->ipc_open[r]) cudaIpcCloseMemHandle(c->ipc_open[r]);
  if (c->nccl_comm) c->nccl.CommDestroy(c->nccl_comm);
  if (c->window) cudaFree(c->window);
  cudaGetLastError();
  delete c;
  return FCM_OK;
}

extern "C" int fcm_graph_count_allreduce_async(fcm_graph* g, fcm_comm* c, uint64_t** counts_dev) {
  REQUIRE(g && c, "graph/comm is null");
  REQUIRE(g->ctx == c->ctx, "graph and comm must live on the same context");
  REQUIRE(c->connected || c->nranks == 1, "comm is not connected");
  if (c->broken) { fcm_set_error("comm is broken (an earlier all-reduce timed out): destroy it"); return FCM_ECUDA; }
  FCM_CUDA(cudaSetDevice(g->ctx->device));
  CommArgs ca;
  memset(&ca, 0, sizeof(ca));
  ca.nranks = c->nranks; ca.rank = c->rank; ca.timeout_ns = c->timeout_ns;
  if (c->kind == FCM_COMM_PEER) {
    if (!c->connected) c->peer[0] = (unsigned long long)c->window;   // a single rank needs no exchange
    ca.mode = COMM_PEER_EPILOGUE;
    ca.slot = (u32)(c->epoch & 1);
    memcpy(ca.peer, c->peer, sizeof(ca.peer));
  } else ca.mode = COMM_ERR_WORD;
  int rc = fcm_count_launch(g, c->rank, c->nranks, &ca);
  if (rc) return rc;
  c->epoch++;
  if (c->kind == FCM_COMM_NCCL && c->nccl_comm)
    NCCL_CALL(c, c->nccl.AllReduce(g->counts, g->counts, FCM_MAXD + 1, NCCL_UINT64, NCCL_SUM, c->nccl_comm, g->ctx->stream));
  if (counts_dev) *counts_dev = g->counts;
  return FCM_OK;
}

extern "C" int fcm_graph_count_allreduce(fcm_graph* g, fcm_comm* c, uint64_t* counts, int* len) {
  REQUIRE(counts && len, "counts/len is null");
  int rc = fcm_graph_count_allreduce_async(g, c, nullptr);
  if (rc) return rc;
  // pinned staging: a pageable read-back would sit in the driver until the kernel - which waits for the other
  // ranks - is done, and keep ranks driven by other threads of this process from launching theirs
  u64* const all = g->ctx->pin;
  const u32* const ctrl = (const u32*)(all + FCM_MAXD + 1);
  FCM_CUDA(cudaMemcpyAsync(all, g->counts, (FCM_MAXD + 1) * 8, cudaMemcpyDeviceToHost, g->ctx->stream));
  FCM_CUDA(cudaMemcpyAsync((void*)ctrl, g->ctrl, CT_WORDS * 4, cudaMemcpyDeviceToHost, g->ctx->stream));
  FCM_CUDA(cudaStreamSynchronize(g->ctx->stream));
  if (ctrl[CT_ERR] == 3) { c->broken = true; fcm_set_error("the all-reduce of the counts timed out: a rank never arrived"); return FCM_ECUDA; }
  if (ctrl[CT_ERR] || all[FCM_MAXD]) { fcm_set_error("a simplex of dimension >= %d was met (on %llu rank(s))", FCM_MAXD, (unsigned long long)all[FCM_MAXD]); return FCM_EDIM; }
  memcpy(counts, all, FCM_MAXD * 8);
  int l = FCM_MAXD;
  while (l > 0 && counts[l - 1] == 0) l--;
  *len = l;
  return FCM_OK;
}
