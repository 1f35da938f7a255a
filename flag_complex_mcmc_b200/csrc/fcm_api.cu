// fcm_api.cu — the extern "C" surface of libfcm.so (include/fcm.h) and the host-side setup.
#include "fcm_internal.cuh"
#include <vector>

#include <algorithm>
#include <chrono>
#include <cmath>
#include <cstdarg>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <mutex>
#include <thread>

// ---------------------------------------------------------------------------------------------------
// errors
// ---------------------------------------------------------------------------------------------------
static thread_local char g_err[512] = "";
void fcm_set_error(const char* fmt, ...) {
  va_list ap;
  va_start(ap, fmt);
  vsnprintf(g_err, sizeof(g_err), fmt, ap);
  va_end(ap);
}
extern "C" const char* fcm_last_error(void) { return g_err; }
extern "C" int fcm_abi_version(void) { return FCM_ABI_VERSION; }
extern "C" int fcm_device_count(void) {
  int n = 0;
  if (cudaGetDeviceCount(&n) != cudaSuccess) { cudaGetLastError(); return 0; }
  return n;
}
extern "C" void fcm_free(void* p) { free(p); }

#define REQUIRE(cond, msg)                                 \
  do {                                                     \
    if (!(cond)) { fcm_set_error("%s", msg); return FCM_EINVAL; } \
  } while (0)

static int use_ctx(const fcm_ctx* c) {
  FCM_CUDA(cudaSetDevice(c->device));
  return FCM_OK;
}
#define USE(c)                          \
  do {                                  \
    int rc__ = use_ctx(c);              \
    if (rc__) return rc__;              \
  } while (0)

// ---------------------------------------------------------------------------------------------------
// context
// ---------------------------------------------------------------------------------------------------
extern "C" int fcm_ctx_create(int device, fcm_ctx** out) {
  REQUIRE(out, "fcm_ctx_create: out is null");
  *out = nullptr;
  int ndev = fcm_device_count();
  if (ndev <= 0) { fcm_set_error("no CUDA device available (libfcm has no CPU fallback)"); return FCM_ENODEVICE; }
  if (device < 0 || device >= ndev) { fcm_set_error("device ordinal %d out of range (0..%d)", device, ndev - 1); return FCM_EINVAL; }
  FCM_CUDA(cudaSetDevice(device));
  fcm_ctx* c = new fcm_ctx();
  c->device = device;
  cudaDeviceProp prop;
  FCM_CUDA(cudaGetDeviceProperties(&prop, device));
  c->sm_count = prop.multiProcessorCount;
  c->smem_optin = prop.sharedMemPerBlockOptin;
  {
    cudaMemPool_t pool;
    FCM_CUDA(cudaDeviceGetDefaultMemPool(&pool, device));
    unsigned long long keep = ~0ULL;                    // freed graph memory stays in the pool for the next graph
    FCM_CUDA(cudaMemPoolSetAttribute(pool, cudaMemPoolAttrReleaseThreshold, &keep));
  }
  if (const char* fg = getenv("FCM_L2_FETCH")) {
    // L2 fetch-granularity hint (32 / 64 / 128 bytes), measurements only: on B200 the DRAM sectors moved per missing sector
    // stay at two whatever the hint says (profiles/README.md r02_m), so the library leaves the device limit alone
    if (atoi(fg) > 0 && cudaDeviceSetLimit(cudaLimitMaxL2FetchGranularity, (size_t)atoi(fg)) != cudaSuccess) cudaGetLastError();
  }
  FCM_CUDA(cudaStreamCreateWithFlags(&c->own, cudaStreamNonBlocking));
  c->stream = c->own;
  FCM_CUDA(cudaEventCreate(&c->ev0));
  FCM_CUDA(cudaEventCreate(&c->ev1));
  FCM_CUDA(cudaHostAlloc((void**)&c->pin, FCM_PIN_WORDS * 8, cudaHostAllocDefault));
  *out = c;
  return FCM_OK;
}
extern "C" int fcm_ctx_destroy(fcm_ctx* c) {
  if (!c) return FCM_OK;
  USE(c);
  cudaStreamSynchronize(c->stream);
  cudaEventDestroy(c->ev0);
  cudaEventDestroy(c->ev1);
  cudaStreamDestroy(c->own);
  if (c->pin) cudaFreeHost(c->pin);
  for (int i = 0; i < 2; i++) if (c->stage_host[i]) cudaFreeHost(c->stage_host[i]);
  for (int i = 0; i < 16; i++) if (c->stage_evs[i]) cudaEventDestroy(c->stage_evs[i]);
  delete c;
  return FCM_OK;
}
extern "C" int fcm_ctx_set_stream(fcm_ctx* c, void* s) {
  REQUIRE(c, "ctx is null");
  c->stream = s ? (cudaStream_t)s : c->own;
  return FCM_OK;
}
extern "C" int fcm_ctx_synchronize(fcm_ctx* c) {
  REQUIRE(c, "ctx is null");
  USE(c);
  FCM_CUDA(cudaStreamSynchronize(c->stream));
  return FCM_OK;
}
extern "C" int fcm_ctx_timer_begin(fcm_ctx* c) {
  REQUIRE(c, "ctx is null");
  USE(c);
  FCM_CUDA(cudaEventRecord(c->ev0, c->stream));
  return FCM_OK;
}
extern "C" int fcm_ctx_timer_end(fcm_ctx* c, float* ms) {
  REQUIRE(c && ms, "ctx/ms is null");
  USE(c);
  FCM_CUDA(cudaEventRecord(c->ev1, c->stream));
  FCM_CUDA(cudaEventSynchronize(c->ev1));
  FCM_CUDA(cudaEventElapsedTime(ms, c->ev0, c->ev1));
  return FCM_OK;
}
extern "C" uint64_t fcm_ctx_launches(const fcm_ctx* c) { return c ? c->launches : 0; }

// Big uploads from pageable host memory (an edge list is 8 bytes per edge: 66 MB at C4).  The driver's own pageable path
// copies through one thread.  Here FCM_STAGE_THREADS host threads (spawned once per call) each take one stripe of the list
// and push it through their own two slices of the context's pinned staging buffers: fill a slice with memcpy, enqueue
// its cudaMemcpyAsync on the context stream, fill the other slice while that one is on the wire.  The copies of all
// threads land on the same stream (they write disjoint ranges, their order does not matter).  Small lists take the plain call.
#define FCM_STAGE_THREADS 8
int fcm_upload_edges(fcm_ctx* c, void* dst_dev, const void* src_host, size_t bytes) {
  cudaStream_t s = c->stream;
  if (bytes < (4u << 20)) { FCM_CUDA(cudaMemcpyAsync(dst_dev, src_host, bytes, cudaMemcpyHostToDevice, s)); return FCM_OK; }
  for (int i = 0; i < 2; i++)
    if (!c->stage_host[i]) FCM_CUDA(cudaHostAlloc((void**)&c->stage_host[i], FCM_STAGE_BYTES, cudaHostAllocDefault));
  for (int i = 0; i < 2 * FCM_STAGE_THREADS; i++)
    if (!c->stage_evs[i]) FCM_CUDA(cudaEventCreateWithFlags(&c->stage_evs[i], cudaEventDisableTiming));
  const int nth = (int)std::max(1u, std::min(std::thread::hardware_concurrency(), (unsigned)FCM_STAGE_THREADS));
  const size_t slice = (FCM_STAGE_BYTES / FCM_STAGE_THREADS) & ~(size_t)4095;      // per thread and buffer: 1 MB
  const size_t stripe = ((bytes + nth - 1) / nth + 4095) & ~(size_t)4095;
  const int device = c->device;
  std::vector<cudaError_t> err(nth, cudaSuccess);
  std::vector<std::thread> th;
  for (int t = 0; t < nth; t++) th.emplace_back([&, t]() {
    cudaError_t e = cudaSetDevice(device);
    const size_t lo = std::min(bytes, stripe * t), hi = std::min(bytes, stripe * (t + 1));
    int buf = 0;
    for (size_t done = lo; done < hi && e == cudaSuccess; buf ^= 1) {
      const size_t now = std::min(slice, hi - done);
      unsigned char* const d = c->stage_host[buf] + slice * t;
      cudaEvent_t ev = c->stage_evs[2 * t + buf];
      e = cudaEventSynchronize(ev);                          // the copy that last used this slice has left it
      if (e != cudaSuccess) break;
      memcpy(d, (const unsigned char*)src_host + done, now);
      e = cudaMemcpyAsync((unsigned char*)dst_dev + done, d, now, cudaMemcpyHostToDevice, s);
      if (e == cudaSuccess) e = cudaEventRecord(ev, s);
      done += now;
    }
    err[t] = e;
  });
  for (auto& t : th) t.join();
  for (int t = 0; t < nth; t++)
    if (err[t] != cudaSuccess) { fcm_set_error("edge-list upload: %s", cudaGetErrorString(err[t])); return FCM_ECUDA; }
  return FCM_OK;
}

// ---------------------------------------------------------------------------------------------------
// device graph
// ---------------------------------------------------------------------------------------------------
static int graph_create(fcm_ctx* c, u32 n, u64 ne, const uint32_t (*edges)[2], int normalise, fcm_graph** out) {
  REQUIRE(c && out, "ctx/out is null");
  REQUIRE(ne == 0 || edges, "edges is null");
  *out = nullptr;
  USE(c);
  fcm_graph* g = new fcm_graph();
  g->ctx = c;
  g->n = n;
  g->W = ((n + 63) / 64 + 1) & ~1u;                  // even number of words: 16-byte row stride
  if (g->W == 0) g->W = 2;
  int rc = FCM_OK;
  uint2* edges_dev = nullptr;
  auto fail = [&](int code) { fcm_pool_free(edges_dev, c->stream); fcm_graph_destroy(g); return code; };
#define G_CUDA(call) do { cudaError_t e__ = (call); if (e__ != cudaSuccess) { fcm_set_error("%s:%d: %s -> %s", __FILE__, __LINE__, #call, cudaGetErrorString(e__)); return fail(FCM_ECUDA); } } while (0)
  const bool timing = getenv("FCM_TIMING") != nullptr;
  auto now = []() { return std::chrono::steady_clock::now(); };
  auto ms_since = [](std::chrono::steady_clock::time_point t0) { return std::chrono::duration<double, std::milli>(std::chrono::steady_clock::now() - t0).count(); };
  auto t0 = now();
  G_CUDA(fcm_pool_alloc(&g->adj, std::max<size_t>((size_t)n * g->W, 2) * 8, c->stream));
  G_CUDA(fcm_pool_alloc(&g->csr_off, ((size_t)n + 1) * 8, c->stream));
  G_CUDA(fcm_pool_alloc(&g->order, std::max<u32>(n, 1) * 4, c->stream));
  G_CUDA(fcm_pool_alloc(&g->overflow, std::max<u32>(n, 1) * 4, c->stream));
  G_CUDA(fcm_pool_alloc(&g->counts, (FCM_MAXD + 1) * 8, c->stream));
  G_CUDA(fcm_pool_alloc(&g->ctrl, CT_WORDS * 4, c->stream));
  if (ne) {
    G_CUDA(fcm_pool_alloc(&edges_dev, ne * 8, c->stream));
    if ((rc = fcm_upload_edges(c, edges_dev, edges, ne * 8))) return fail(rc);
  }
  if (timing) { cudaStreamSynchronize(c->stream); fprintf(stderr, "[fcm] graph_create: alloc + upload of %.1f MB: %.2f ms\n", ne * 8 / 1e6, ms_since(t0)); t0 = now(); }
  // rows, CSR and the source order (vertices by descending out-degree: the heavy ones start first and shards
  // (part, nparts) taken round-robin from this order are degree-balanced) are built on the device
  rc = fcm_graph_build_device(g, edges_dev, ne, normalise);
  if (rc) return fail(rc);
  fcm_pool_free(edges_dev, c->stream);
  if (timing) { cudaStreamSynchronize(c->stream); fprintf(stderr, "[fcm] graph_create: device build (rows, scan, sort, csr): %.2f ms\n", ms_since(t0)); }
#undef G_CUDA
  *out = g;
  return FCM_OK;
}
extern "C" int fcm_graph_create(fcm_ctx* c, uint32_t n, uint64_t ne, const uint32_t (*edges)[2], fcm_graph** out) {
  return graph_create(c, n, ne, edges, 0, out);
}
extern "C" int fcm_graph_create_normalised(fcm_ctx* c, uint32_t n, uint64_t ne, const uint32_t (*edges)[2], fcm_graph** out) {
  return graph_create(c, n, ne, edges, 1, out);
}
extern "C" int fcm_graph_destroy(fcm_graph* g) {
  if (!g) return FCM_OK;
  if (g->ctx) cudaSetDevice(g->ctx->device);
  cudaStream_t s = g->ctx ? g->ctx->stream : nullptr;
  fcm_pool_free(g->adj, s); fcm_pool_free(g->csr_off, s); fcm_pool_free(g->csr, s); fcm_pool_free(g->order, s); fcm_pool_free(g->oldid, s);
  fcm_pool_free(g->overflow, s); fcm_pool_free(g->counts, s); fcm_pool_free(g->ctrl, s); fcm_pool_free(g->scratch, s);
  delete g;
  return FCM_OK;
}
extern "C" int fcm_graph_set_kernel(fcm_graph* g, int kind) {
  REQUIRE(g, "graph is null");
  REQUIRE(kind == 0 || kind == 1, "kernel kind must be 0 (auto) or 1 (dense-row)");
  g->kernel_kind = kind;
  return FCM_OK;
}
extern "C" int fcm_graph_info(const fcm_graph* g, uint32_t* n, uint64_t* ne, uint32_t* W, uint32_t* maxout) {
  REQUIRE(g, "graph is null");
  if (n) *n = g->n;
  if (ne) *ne = g->ne;
  if (W) *W = g->W;
  if (maxout) *maxout = g->maxout;
  return FCM_OK;
}
extern "C" int fcm_graph_count_async(fcm_graph* g, uint32_t part, uint32_t nparts, uint64_t** counts_dev) {
  REQUIRE(g, "graph is null");
  REQUIRE(nparts >= 1 && part < nparts, "need part < nparts");
  USE(g->ctx);
  int rc = fcm_count_launch(g, part, nparts);
  if (rc) return rc;
  if (counts_dev) *counts_dev = g->counts;
  return FCM_OK;
}
extern "C" int fcm_graph_status(fcm_graph* g) {
  REQUIRE(g, "graph is null");
  USE(g->ctx);
  u32* const ctrl = (u32*)g->ctx->pin;
  FCM_CUDA(cudaMemcpyAsync(ctrl, g->ctrl, CT_WORDS * 4, cudaMemcpyDeviceToHost, g->ctx->stream));
  FCM_CUDA(cudaStreamSynchronize(g->ctx->stream));
  if (ctrl[CT_ERR] == 3) { fcm_set_error("the all-reduce of the counts timed out: a rank never arrived"); return FCM_ECUDA; }
  if (ctrl[CT_ERR]) { fcm_set_error("a simplex of dimension >= %d was met", FCM_MAXD); return FCM_EDIM; }
  return FCM_OK;
}
extern "C" int fcm_graph_count(fcm_graph* g, uint32_t part, uint32_t nparts, uint64_t* counts, int* len) {
  REQUIRE(counts && len, "counts/len is null");
  int rc = fcm_graph_count_async(g, part, nparts, nullptr);
  if (rc) return rc;
  u64* const pin = g->ctx->pin;
  const u32* const ctrl = (const u32*)(pin + FCM_MAXD + 1);
  FCM_CUDA(cudaMemcpyAsync(pin, g->counts, FCM_MAXD * 8, cudaMemcpyDeviceToHost, g->ctx->stream));
  FCM_CUDA(cudaMemcpyAsync((void*)ctrl, g->ctrl, CT_WORDS * 4, cudaMemcpyDeviceToHost, g->ctx->stream));
  FCM_CUDA(cudaStreamSynchronize(g->ctx->stream));
  memcpy(counts, pin, FCM_MAXD * 8);
  if (ctrl[CT_ERR]) { fcm_set_error("a simplex of dimension >= %d was met", FCM_MAXD); return FCM_EDIM; }
  int l = FCM_MAXD;
  while (l > 0 && counts[l - 1] == 0) l--;
  *len = l;
  return FCM_OK;
}

// flag_complex::for_each_cell (src/bin/all_cxs.rs:60-64) as a bulk call: every cell of dimension
// dim_min..dim_max, one row of dim_max + 1 vertex ids (padded with 0xffffffff), rows in the order
// the recursion meets them when candidates are walked in ascending order (lexicographic, a prefix
// before its extensions).
extern "C" int fcm_graph_cells(fcm_graph* g, int dim_min, int dim_max, uint32_t* cells_out, uint64_t cap, uint64_t* ncells) {
  REQUIRE(g && ncells, "graph/ncells is null");
  REQUIRE(dim_min >= 0 && dim_max >= dim_min, "need 0 <= dim_min <= dim_max");
  REQUIRE(cap == 0 || cells_out, "cells_out is null");
  USE(g->ctx);
  cudaStream_t s = g->ctx->stream;
  const size_t stride = (size_t)dim_max + 1;
  u32* cells_dev = nullptr;
  unsigned long long* n_dev = nullptr;
  FCM_CUDA(cudaMalloc(&n_dev, 8));
  if (cap) { cudaError_t ce = cudaMalloc(&cells_dev, cap * stride * 4); if (ce != cudaSuccess) { cudaFree(n_dev); fcm_set_error("cudaMalloc of the cell buffer failed"); return FCM_ENOMEM; } }
  int rc = fcm_cells_launch(g, (u32)dim_min, (u32)dim_max, cells_dev, cap, n_dev);
  unsigned long long total = 0;
  u32 ctrl[CT_WORDS] = {0};
  if (!rc) {
    cudaError_t ce = cudaMemcpyAsync(&total, n_dev, 8, cudaMemcpyDeviceToHost, s);
    if (ce == cudaSuccess) ce = cudaMemcpyAsync(ctrl, g->ctrl, sizeof(ctrl), cudaMemcpyDeviceToHost, s);
    if (ce == cudaSuccess) ce = cudaStreamSynchronize(s);
    if (ce != cudaSuccess) { fcm_set_error("for_each_cell: %s", cudaGetErrorString(ce)); rc = FCM_ECUDA; }
  }
  if (!rc && total <= cap && total) {
    cudaError_t ce = cudaMemcpy(cells_out, cells_dev, (size_t)total * stride * 4, cudaMemcpyDeviceToHost);
    if (ce != cudaSuccess) { fcm_set_error("for_each_cell: %s", cudaGetErrorString(ce)); rc = FCM_ECUDA; }
    else {
      // the device meets the cells in no particular order: put them in the recursion's order.
      // FCM_NONE padding sorts a prefix AFTER its extensions, so compare lengths first on ties.
      std::vector<size_t> idx((size_t)total);
      for (size_t i = 0; i < idx.size(); i++) idx[i] = i;
      auto less = [&](size_t a, size_t b) {
        const u32 *x = cells_out + a * stride, *y = cells_out + b * stride;
        for (size_t t = 0; t < stride; t++) {
          if (x[t] == y[t]) continue;
          if (x[t] == FCM_NONE) return true;              // x is a proper prefix of y: first
          if (y[t] == FCM_NONE) return false;
          return x[t] < y[t];
        }
        return false;
      };
      std::sort(idx.begin(), idx.end(), less);
      std::vector<u32> tmp((size_t)total * stride);
      for (size_t i = 0; i < idx.size(); i++) memcpy(&tmp[i * stride], cells_out + idx[i] * stride, stride * 4);
      memcpy(cells_out, tmp.data(), tmp.size() * 4);
    }
  }
  cudaFree(n_dev);
  if (cells_dev) cudaFree(cells_dev);
  if (rc) return rc;
  *ncells = total;
  return FCM_OK;
}

// flagser_count_unweighted precedent (src/flagser.rs:7-21): malloc'd result, NULL on error
static std::mutex g_default_mu;
static fcm_ctx* g_default_ctx = nullptr;
extern "C" size_t* fcm_count_unweighted(size_t nvertices, size_t nedges, const uint32_t (*edges)[2], size_t* res_size) {
  if (!res_size) { fcm_set_error("res_size is null"); return nullptr; }
  *res_size = 0;
  if (nvertices > 0xffffffffull) { fcm_set_error("too many vertices"); return nullptr; }
  std::lock_guard<std::mutex> lk(g_default_mu);
  if (!g_default_ctx && fcm_ctx_create(0, &g_default_ctx)) return nullptr;
  fcm_graph* g = nullptr;
  if (fcm_graph_create(g_default_ctx, (u32)nvertices, nedges, edges, &g)) return nullptr;
  u64 counts[FCM_MAXD];
  int len = 0;
  int rc = fcm_graph_count(g, 0, 1, counts, &len);
  fcm_graph_destroy(g);
  if (rc) return nullptr;
  size_t* res = (size_t*)malloc(sizeof(size_t) * (len ? len : 1));
  if (!res) { fcm_set_error("out of host memory"); return nullptr; }
  for (int i = 0; i < len; i++) res[i] = (size_t)counts[i];
  *res_size = (size_t)len;
  return res;
}
