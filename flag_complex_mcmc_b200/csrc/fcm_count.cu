// fcm_count.cu — directed-flag-complex simplex counting on sm_100a.
//
// Replaces `graph.flagser_count()` of the external flag-complex crate as called from
// src/lib.rs:51 (State::new), :130 (Bounds::calculate, normalised graph) and
// flag_complex::count_cells (src/bin/all_cxs.rs:123).  Algorithm = the flagser directed-clique
// recursion: every simplex (v0..vd) owns the candidate set
// P = out(v0) ∩ ... ∩ out(vd); each member of P closes a (d+1)-simplex.
//
// Two kernels, identical results:
//   count_local_kernel<WK, EMIT>  one CTA per source vertex v0.  out(v0) is relabelled 0..k-1 and the
//       k x k induced adjacency is gathered ONCE from the packed HBM rows into shared memory
//       (k*WK 64-bit words).  Every deeper intersection is then a WK-word AND in shared memory
//       instead of a W-word row read (C4: WK <= 8 against W = 490); the simplices under v0 go through
//       a per-warp stack of pending tuples, one simplex per lane per step (see the kernel).
//       EMIT = true additionally writes the cells of a dimension range out (for_each_cell).
//   count_dense_kernel      one warp per source vertex, candidate sets of full row width staged
//       in a per-warp stack, rows read with coalesced 128-bit loads, popc + warp reduce.  This is
//       the "one adjacency row per simplex" kernel of the roofline model; it also
//       serves sources whose out-degree exceeds the shared-memory tile.
#include "fcm_internal.cuh"

#include <algorithm>
#include <cstring>
#include <mutex>

#include <cooperative_groups.h>
#include <cub/device/device_radix_sort.cuh>
#include <cub/device/device_scan.cuh>

namespace cg = cooperative_groups;

#define DENSE_WARPS 4
#define MAXL_DENSE (FCM_MAXD - 1)   // level l holds candidates for vertex index l+1

// ---------------------------------------------------------------------------------------------------
// graph construction
// ---------------------------------------------------------------------------------------------------
__global__ void set_bits_kernel(const uint2* __restrict__ edges, u64 ne, u32 n, u32 W, int normalise,
                                const u32* __restrict__ newid, u64* __restrict__ adj, u32* __restrict__ ctrl) {
  u64 i = (u64)blockIdx.x * blockDim.x + threadIdx.x;
  u64 stride = (u64)gridDim.x * blockDim.x;
  for (; i < ne; i += stride) {
    uint2 e = edges[i];
    u32 a = e.x, b = e.y;
    if (a >= n || b >= n || a == b) { atomicExch(&ctrl[CT_ERR], 1u); continue; }
    if (normalise && a < b) { u32 t = a; a = b; b = t; }   // src/lib.rs:126-129: edge big -> small (the caller's ids decide)
    if (newid) { a = newid[a]; b = newid[b]; }             // the graph's internal numbering (locality order below)
    atomicOr((unsigned long long*)&adj[(u64)a * W + (b >> 6)], 1ULL << (b & 63));
  }
}

// ---------------------------------------------------------------------------------------------------
// locality order of the vertices (large graphs)
// ---------------------------------------------------------------------------------------------------
// The count is invariant under renumbering the vertices, its speed is not: count_local_kernel gathers the k x k tile
// of a source by bit probes into the rows of its out-neighbours, and with ids scattered over the graph (BASELINE
// config 4: ids are hashes of the position) every probe is its own sector.  Graphs beyond LOC_MIN_N vertices are
// therefore built under an internal numbering in which neighbours get nearby ids; fcm_graph_cells maps ids back
// (EmitArgs::oldid), nothing else of the API sees vertex ids.  The order, all on the device in ONE cooperative launch
// over the uploaded edge list (no CSR exists yet):
//   1. breadth-first levels of the undirected graph from vertex 0, level-synchronous sweeps over the edge list (at most
//      LOC_MAXLEV of them: what lies deeper is treated as one last level);
//   2. the same from the farthest vertex found (largest level, smallest id): levels along the graph's long axis;
//   3. two Jacobi sweeps x[a] <- (x[a] + sum of x[b] over the edges a -> b) / (outdeg(a) + 1) on the level as a fixed-
//      point number, which turns the handful of levels into a smooth coordinate (C4: 6 levels -> median |i - j| over
//      the edges 3 829 -> 1 234 of 31 346; the positions' own long axis gives 1 228; tools/relabel_probe.py);
//   4. stable radix sort of the vertices by x.
// Integer arithmetic throughout: every rank of a sharded count builds the same numbering (the shards are dealt from
// the degree order of the renumbered graph).  FCM_RELABEL=0 switches it off, =1 forces it for every graph (tests).
#define LOC_MIN_N 4096u
#define LOC_INF 0xffffffffu
#define LOC_FRAC 20
#define LOC_MAXLEV 63u
__global__ void __launch_bounds__(256)
locality_kernel(const uint2* __restrict__ edges, u64 ne, u32 n, u32* lev, u32* x, unsigned long long* acc, u32* deg,
                u32* ctl, unsigned long long* far, u32* iota) {
  cg::grid_group grid = cg::this_grid();
  const u64 tid = (u64)blockIdx.x * blockDim.x + threadIdx.x, nt = (u64)gridDim.x * blockDim.x;
  const int lane = threadIdx.x & 31;
  u32 stamp = 0, root = 0;
  for (int pass = 0; pass < 2; pass++) {
    for (u64 i = tid; i < n; i += nt) __stcg(&lev[i], i == root ? 0u : LOC_INF);
    grid.sync();
    for (u32 L = 0; L < LOC_MAXLEV; L++) {               // deeper levels would be clamped below anyway: a graph of large diameter costs 63 sweeps, not its diameter
      stamp++;
      bool any = false;
      for (u64 i = tid; i < ne; i += nt) {
        const uint2 e = __ldg(&edges[i]);
        if (e.x >= n || e.y >= n) continue;              // reported by set_bits_kernel
        const u32 la = __ldcg(&lev[e.x]), lb = __ldcg(&lev[e.y]);
        if (la == L && lb == LOC_INF) { __stcg(&lev[e.y], L + 1); any = true; }
        else if (lb == L && la == LOC_INF) { __stcg(&lev[e.x], L + 1); any = true; }
      }
      if (any) __stcg(&ctl[0], stamp);                   // this sweep reached a new vertex
      grid.sync();
      if (__ldcg(&ctl[0]) < stamp) break;                // (a thread already in the next sweep may have written stamp + 1)
    }
    for (u64 i = tid; i < n; i += nt) {
      const u32 l = __ldcg(&lev[i]);
      if (l != LOC_INF) atomicMax(&far[pass], ((unsigned long long)l << 32) | (LOC_INF - (u32)i));
    }
    grid.sync();
    root = LOC_INF - (u32)__ldcg(&far[pass]);
  }
  const u32 maxlev = (u32)(__ldcg(&far[1]) >> 32);
  for (u64 i = tid; i < n; i += nt) {
    u32 l = __ldcg(&lev[i]);
    if (l == LOC_INF) l = maxlev + 1;                    // other components: behind the reached ones
    x[i] = min(l, LOC_MAXLEV) << LOC_FRAC;
    deg[i] = 0; acc[i] = 0; iota[i] = (u32)i;
  }
  grid.sync();
  for (int sweep = 0; sweep < 2; sweep++) {
    for (u64 i0 = tid - lane; i0 < ne; i0 += nt) {       // warp-uniform trip count
      const u64 i = i0 + lane;
      uint2 e = make_uint2(LOC_INF, LOC_INF);
      if (i < ne) e = __ldg(&edges[i]);
      const bool ok = e.x < n && e.y < n;
      const u32 a = ok ? e.x : LOC_INF;
      const u32 v = ok ? __ldcg(&x[e.y]) : 0u;
      // edge lists come sorted by source (src/io.rs reads them in file order): one atomic per run of equal sources
      const u32 m = __match_any_sync(0xffffffffu, a);
      const u32 sum = __reduce_add_sync(m, v);           // <= 32 * 63 * 2^20 < 2^32
      if (ok && lane == __ffs((int)m) - 1) {
        atomicAdd(&acc[a], (unsigned long long)sum);
        if (sweep == 0) atomicAdd(&deg[a], (u32)__popc(m));
      }
    }
    grid.sync();
    for (u64 i = tid; i < n; i += nt) {
      const unsigned long long t = __ldcg(&acc[i]) + x[i];
      x[i] = (u32)(t / (__ldcg(&deg[i]) + 1u));
      acc[i] = 0;
    }
    grid.sync();
  }
}
__global__ void invert_order_kernel(const u32* __restrict__ oldid, u32 n, u32* __restrict__ newid) {
  const u32 i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i < n) newid[oldid[i]] = i;
}

static bool want_locality_order(u32 n, u64 ne) {
  const char* env = getenv("FCM_RELABEL");
  if (env && env[0] == '0') return false;
  if (env && env[0] == '1') return n >= 2 && ne > 0;
  return n > LOC_MIN_N && ne > 0;
}

// g->oldid (kept with the graph: internal id -> caller's id) and *newid_out (caller's id -> internal id, for the build only)
static int locality_order(fcm_graph* g, const uint2* edges_dev, u64 ne, u32** newid_out) {
  fcm_ctx* c = g->ctx;
  cudaStream_t s = c->stream;
  const u32 n = g->n;
  int coop = 0;
  FCM_CUDA(cudaDeviceGetAttribute(&coop, cudaDevAttrCooperativeLaunch, c->device));
  if (!coop) return FCM_OK;                               // the build then keeps the caller's ids
  u32 *lev = nullptr, *x = nullptr, *deg = nullptr, *ctl = nullptr, *iota = nullptr, *xs = nullptr, *newid = nullptr;
  unsigned long long *acc = nullptr, *far = nullptr;
  void* tmp = nullptr;
  auto release = [&]() { fcm_pool_free(lev, s); fcm_pool_free(x, s); fcm_pool_free(deg, s); fcm_pool_free(ctl, s); fcm_pool_free(iota, s);
                         fcm_pool_free(xs, s); fcm_pool_free(acc, s); fcm_pool_free(far, s); fcm_pool_free(tmp, s); };
#define L_CUDA(call) do { cudaError_t e__ = (call); if (e__ != cudaSuccess) { fcm_set_error("%s:%d: %s -> %s", __FILE__, __LINE__, #call, cudaGetErrorString(e__)); release(); fcm_pool_free(newid, s); return FCM_ECUDA; } } while (0)
  L_CUDA(fcm_pool_alloc(&lev, (size_t)n * 4, s));
  L_CUDA(fcm_pool_alloc(&x, (size_t)n * 4, s));
  L_CUDA(fcm_pool_alloc(&deg, (size_t)n * 4, s));
  L_CUDA(fcm_pool_alloc(&iota, (size_t)n * 4, s));
  L_CUDA(fcm_pool_alloc(&xs, (size_t)n * 4, s));
  L_CUDA(fcm_pool_alloc(&newid, (size_t)n * 4, s));
  L_CUDA(fcm_pool_alloc(&acc, (size_t)n * 8, s));
  L_CUDA(fcm_pool_alloc(&ctl, 16, s));
  L_CUDA(fcm_pool_alloc(&far, 16, s));
  L_CUDA(fcm_pool_alloc(&g->oldid, (size_t)n * 4, s));
  L_CUDA(cudaMemsetAsync(ctl, 0, 16, s));
  L_CUDA(cudaMemsetAsync(far, 0, 16, s));
  int per_sm = 0;
  L_CUDA(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, locality_kernel, 256, 0));
  // all blocks of a cooperative launch must be resident together: at most half of an SM's thread slots, so that the launch
  // also fits beside a kernel that does not drain (another rank's count waiting in its all-reduce on the same device)
  per_sm = std::max(1, std::min(per_sm, 4));
  const u32 blocks = (u32)std::max<u64>(1, std::min<u64>((u64)c->sm_count * per_sm, (ne + 255) / 256));
  u64 ne_arg = ne;
  u32 n_arg = n;
  void* args[] = {(void*)&edges_dev, (void*)&ne_arg, (void*)&n_arg, (void*)&lev, (void*)&x, (void*)&acc, (void*)&deg, (void*)&ctl, (void*)&far, (void*)&iota};
  L_CUDA(cudaLaunchCooperativeKernel((const void*)locality_kernel, dim3(blocks), dim3(256), args, 0, s));
  c->launches++;
  size_t tbytes = 0;
  L_CUDA(cub::DeviceRadixSort::SortPairs(nullptr, tbytes, x, xs, iota, g->oldid, (int)n, 0, 32, s));
  L_CUDA(fcm_pool_alloc((unsigned char**)&tmp, std::max<size_t>(tbytes, 16), s));
  L_CUDA(cub::DeviceRadixSort::SortPairs(tmp, tbytes, x, xs, iota, g->oldid, (int)n, 0, 32, s));
  invert_order_kernel<<<(n + 255) / 256, 256, 0, s>>>(g->oldid, n, newid);
  c->launches += 2;
  L_CUDA(cudaGetLastError());
#undef L_CUDA
  release();
  *newid_out = newid;
  return FCM_OK;
}

// one warp per row: out-degree -> off[v + 1] (scanned in place afterwards) and the sort key of the source order
__global__ void row_degree_kernel(const u64* __restrict__ adj, u32 n, u32 W, u64* __restrict__ off,
                                  u32* __restrict__ key, u32* __restrict__ val) {
  u32 warp = (blockIdx.x * blockDim.x + threadIdx.x) >> 5, lane = threadIdx.x & 31;
  if (warp >= n) return;
  const u64* row = adj + (u64)warp * W;
  u32 c = 0;
  for (u32 w = lane; w < W; w += 32) c += __popcll(row[w]);
  c = __reduce_add_sync(0xffffffffu, c);
  if (lane == 0) { off[warp + 1] = c; key[warp] = ~c; val[warp] = warp; }   // ascending ~degree = descending degree
}

// one warp per row: ascending out-list into csr[off[v]..]
__global__ void row_to_csr_kernel(const u64* __restrict__ adj, u32 n, u32 W,
                                  const u64* __restrict__ off, u32* __restrict__ csr) {
  u32 warp = (blockIdx.x * blockDim.x + threadIdx.x) >> 5, lane = threadIdx.x & 31;
  if (warp >= n) return;
  const u64* row = adj + (u64)warp * W;
  u64 base = off[warp];
  for (u32 w0 = 0; w0 < W; w0 += 32) {
    u32 w = w0 + lane;
    u64 bits = w < W ? row[w] : 0;
    u32 c = __popcll(bits);
    u32 incl = c;                        // inclusive warp scan
    for (int d = 1; d < 32; d <<= 1) { u32 t = __shfl_up_sync(0xffffffffu, incl, d); if (lane >= d) incl += t; }
    u64 o = base + incl - c;
    while (bits) { csr[o++] = w * 64 + (__ffsll((long long)bits) - 1); bits &= bits - 1; }
    base += __shfl_sync(0xffffffffu, incl, 31);
  }
}
// [0] = number of distinct edges (last scanned offset), [1] = largest out-degree (~ smallest sort key)
__global__ void build_summary_kernel(const u64* __restrict__ off, const u32* __restrict__ key_sorted, u32 n, u64* __restrict__ out) {
  out[0] = off[n];
  out[1] = n ? (u64)(~key_sorted[0]) : 0ULL;
}

// Bit rows -> CSR offsets (device scan), out-lists, and the source order (vertices by descending out-degree, ties by
// ascending id: a stable device radix sort) - all on the device.  One small read-back (edge count, largest out-degree,
// the bad-edge flag) is the only host hop: the host needs them to size the CSR and to choose the tile.
int fcm_graph_build_device(fcm_graph* g, const uint2* edges_dev, u64 ne_in, int normalise) {
  fcm_ctx* c = g->ctx;
  cudaStream_t s = c->stream;
  const u32 n = g->n;
  size_t words = (size_t)n * g->W;
  FCM_CUDA(cudaMemsetAsync(g->adj, 0, words * 8, s));
  FCM_CUDA(cudaMemsetAsync(g->ctrl, 0, CT_WORDS * 4, s));
  u32* newid = nullptr;
  if (want_locality_order(n, ne_in)) { const int rc = locality_order(g, edges_dev, ne_in, &newid); if (rc) return rc; }
  if (ne_in) {
    int blocks = (int)std::min<u64>((ne_in + 255) / 256, 148 * 16);
    set_bits_kernel<<<blocks, 256, 0, s>>>(edges_dev, ne_in, n, g->W, normalise, newid, g->adj, g->ctrl);
    c->launches++;
  }
  fcm_pool_free(newid, s);                               // stream-ordered: behind set_bits_kernel
  FCM_CUDA(cudaMemsetAsync(g->csr_off, 0, 8, s));
  u32 *key = nullptr, *val = nullptr, *key2 = nullptr;
  void* tmp = nullptr;
  u64* summary = nullptr;
  auto release = [&]() { fcm_pool_free(key, s); fcm_pool_free(val, s); fcm_pool_free(key2, s); fcm_pool_free(tmp, s); fcm_pool_free(summary, s); };
#define B_CUDA(call) do { cudaError_t e__ = (call); if (e__ != cudaSuccess) { fcm_set_error("%s:%d: %s -> %s", __FILE__, __LINE__, #call, cudaGetErrorString(e__)); release(); return FCM_ECUDA; } } while (0)
  const u32 nn = std::max<u32>(n, 1);
  B_CUDA(fcm_pool_alloc(&key, (size_t)nn * 4, s));
  B_CUDA(fcm_pool_alloc(&val, (size_t)nn * 4, s));
  B_CUDA(fcm_pool_alloc(&key2, (size_t)nn * 4, s));
  B_CUDA(fcm_pool_alloc(&summary, 16, s));
  if (n) {
    int blocks = (n * 32 + 255) / 256;
    row_degree_kernel<<<blocks, 256, 0, s>>>(g->adj, n, g->W, g->csr_off, key, val);
    c->launches++;
    size_t t_scan = 0, t_sort = 0;
    B_CUDA(cub::DeviceScan::InclusiveSum(nullptr, t_scan, g->csr_off + 1, g->csr_off + 1, (int)n, s));
    B_CUDA(cub::DeviceRadixSort::SortPairs(nullptr, t_sort, key, key2, val, g->order, (int)n, 0, 32, s));
    const size_t tbytes = std::max(t_scan, t_sort);
    B_CUDA(fcm_pool_alloc((unsigned char**)&tmp, std::max<size_t>(tbytes, 16), s));
    size_t tb = tbytes;
    B_CUDA(cub::DeviceScan::InclusiveSum(tmp, tb, g->csr_off + 1, g->csr_off + 1, (int)n, s));
    tb = tbytes;
    B_CUDA(cub::DeviceRadixSort::SortPairs(tmp, tb, key, key2, val, g->order, (int)n, 0, 32, s));
    c->launches += 2;
  }
  build_summary_kernel<<<1, 1, 0, s>>>(g->csr_off, key2, n, summary);
  c->launches++;
  u64* const pin = c->pin;
  B_CUDA(cudaMemcpyAsync(pin, summary, 16, cudaMemcpyDeviceToHost, s));
  B_CUDA(cudaMemcpyAsync(pin + 2, g->ctrl, CT_WORDS * 4, cudaMemcpyDeviceToHost, s));
  B_CUDA(cudaStreamSynchronize(s));
  release();
#undef B_CUDA
  if (((const u32*)(pin + 2))[CT_ERR]) { fcm_set_error("edge list holds a vertex id >= n or a self loop"); return FCM_EINVAL; }
  g->ne = pin[0];
  g->maxout = (u32)pin[1];
  FCM_CUDA(fcm_pool_alloc(&g->csr, std::max<u64>(g->ne, 1) * 4, s));
  if (n) {
    int blocks = (n * 32 + 255) / 256;
    row_to_csr_kernel<<<blocks, 256, 0, s>>>(g->adj, n, g->W, g->csr_off, g->csr);
    c->launches++;
  }
  FCM_CUDA(cudaGetLastError());
  return FCM_OK;
}

// ---------------------------------------------------------------------------------------------------
// block-local relabelled count
// ---------------------------------------------------------------------------------------------------
// One CTA per source vertex v0 (persistent CTAs, sources handed out by a ticket in degree-descending
// order).  out(v0) is relabelled 0..k-1 and its k x k induced adjacency L is gathered once into shared
// memory (row stride LS = WK|1 64-bit words: odd, so rows fall on different banks).  The simplices
// under v0 are then enumerated through a per-warp STACK OF PENDING SIMPLICES in shared memory: an
// entry is the tuple (v1..vd) of local ids, 10 bits each, d <= 6.
//   fill   when fewer than 32 entries are pending the warp takes the next row i (= v1) and pushes one
//          entry (i, j) per member j of L[i] (the triangles v0,v1,v2); lane l extracts 2*WK bits;
//   step   the warp pops the top 32 entries, one per lane: P = L[v1] & .. & L[vd] is rebuilt from the
//          rows (d*WK shared-memory loads, no per-thread stack), popc(P) simplices of d+2 vertices are
//          counted, and every member of P is pushed back as a child entry (warp scan for the offsets).
// Every lane thus runs the same straight-line AND/popc code on a different simplex whatever the shape
// of the tree (a per-thread DFS leaves most lanes idle: a triangle of the cortical graph has 1.26
// children on average, spread over WK words).  Tuples deeper than 6 ids, or a full stack, fall back to
// a generic per-thread DFS (deep_count), which is rare.
template <int WK> struct LocalCfg {
  static constexpr int LS = WK | 1;
  static constexpr int THREADS = WK >= 16 ? 128 : 256;
  static constexpr int MINBLOCKS = WK >= 16 ? 1 : 3;
  static constexpr int MAXIDS = 6;                                 // ids per stack entry (10 bits each + depth)
  static constexpr int BPL = 2 * WK;                               // bits of a row each lane extracts (32 lanes x BPL = 64*WK)
};
__host__ __device__ constexpr u32 stack_cap(u32 kcap) { return kcap + 128; }   // entries per warp (>= kcap + 32; a full stack falls back to deep_count)

// for_each_cell (flag_complex::for_each_cell, src/bin/all_cxs.rs:60-64): where the cells of dimension
// emit_min..emit_max go.  Row r of `cells` = the vertices (v0..vd) of one cell, padded with FCM_NONE.
struct EmitArgs { u32 emit_min, emit_max, stride; u32* cells; u64 cap; unsigned long long* ncells; const u32* oldid; };   // oldid: internal -> caller's ids (null = same)

__device__ __forceinline__ void emit_cell(const EmitArgs& ea, u32 v0, const u32* verts, u64 e, u32 d) {
  if (d < ea.emit_min || d > ea.emit_max) return;
  const u64 row = atomicAdd(ea.ncells, 1ULL);
  if (row >= ea.cap) return;
  u32* out = ea.cells + row * ea.stride;
  out[0] = ea.oldid ? ea.oldid[v0] : v0;
  for (u32 t = 0; t < ea.stride - 1; t++) {
    u32 v = FCM_NONE;
    if (t < d) { v = verts[(e >> (10 * t)) & 1023u]; if (ea.oldid) v = ea.oldid[v]; }
    out[1 + t] = v;
  }
}

// P = candidates for the vertex of index `lvl` (non-empty, already counted).  Generic levels: explicit
// stack in local memory, counts through shared atomics (rare: dimension > MAXIDS + 1, or a full stack).
// EMIT: the walk also writes out the cells it meets (for_each_cell never drops a subtree): `e` = the
// tuple of local ids (10 bits each) of the simplex P belongs to; the walk stops at dimension emit_max.
template <int WK, bool EMIT>
__device__ __noinline__ void deep_count(const u64 (&P0)[WK], int lvl, const u64* __restrict__ L,
                                        unsigned long long* s_cnt, u32* ctrl, const EmitArgs& ea, u32 v0,
                                        const u32* verts, u64 e) {
  constexpr int LS = LocalCfg<WK>::LS;
  u64 P[FCM_MAXD][WK];
  int cw[FCM_MAXD];
  u64 cb[FCM_MAXD];
#pragma unroll
  for (int x = 0; x < WK; x++) P[0][x] = P0[x];
  int l = 0, w = 0;
  u64 bits = P[0][0];
  for (;;) {
    while (bits == 0) { if (++w == WK) break; bits = P[l][w]; }
    if (w == WK) {
      if (l == 0) break;
      --l; w = cw[l]; bits = cb[l];
      continue;
    }
    const int j = w * 64 + (__ffsll((long long)bits) - 1);
    bits &= bits - 1;
    const u64* Lj = L + (size_t)j * LS;
    u32 cpc = 0;
    const int d = lvl + l + 1;                          // flag_count index the children of this node close
    if constexpr (EMIT) {
      // the simplex (v0, tuple.., j) has dimension d - 1 = lvl + l; its ids sit at positions 0 .. lvl + l - 1
      const u32 npos = (u32)(lvl + l - 1);              // position of j in the tuple
      e = (e & ((1ULL << (10 * npos)) - 1ULL)) | ((u64)j << (10 * npos));
      emit_cell(ea, v0, verts, e, (u32)(lvl + l));
      if ((u32)(lvl + l) >= ea.emit_max) continue;      // nothing deeper is asked for (MAXIDS bounds emit_max)
    }
    if (d < FCM_MAXD) {
#pragma unroll
      for (int x = 0; x < WK; x++) { u64 q = P[l][x] & Lj[x]; P[l + 1][x] = q; cpc += __popcll(q); }
      if (cpc) {
        atomicAdd(&s_cnt[d], (unsigned long long)cpc);
        cw[l] = w; cb[l] = bits; ++l; w = 0; bits = P[l][0];
      }
    } else {
#pragma unroll
      for (int x = 0; x < WK; x++) cpc += __popcll(P[l][x] & Lj[x]);
      if (cpc) atomicExch(&ctrl[CT_ERR], 2u);           // a simplex of dimension >= FCM_MAXD
    }
  }
}

template <int WK, bool EMIT>
__global__ void __launch_bounds__(LocalCfg<WK>::THREADS, LocalCfg<WK>::MINBLOCKS)
count_local_kernel(const u64* __restrict__ adj, const u64* __restrict__ csr_off,
                   const u32* __restrict__ csr, const u32* __restrict__ order, u32 W, u32 nv_total,
                   u32 part, u32 nparts, u32 kcap, u32* __restrict__ ctrl,
                   u32* __restrict__ overflow, u64* __restrict__ counts, const EmitArgs ea,
                   const __grid_constant__ CommArgs ca) {
  using Cfg = LocalCfg<WK>;
  constexpr int LS = Cfg::LS, NT = Cfg::THREADS, NW = NT / 32, BPL = Cfg::BPL, MAXIDS = Cfg::MAXIDS;
  constexpr int RL = MAXIDS - 1;                        // cnt[r]: flag_count index 3 + r (tuples of 2 + r ids)
  extern __shared__ __align__(16) unsigned char smem_raw[];
  u64* L = (u64*)smem_raw;                              // kcap * LS
  u32* verts = (u32*)(L + (size_t)kcap * LS);           // kcap
  u64* stacks = (u64*)(verts + kcap + (kcap & 1));      // NW * stack_cap(kcap)
  __shared__ u32 s_task, s_next;
  __shared__ unsigned long long s_cnt[FCM_MAXD];

  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  const u32 scap = stack_cap(kcap);
  u64* const stk = stacks + (size_t)warp * scap;
  u64 c0 = 0, c01 = 0, c2 = 0;                          // c0, c01: thread 0 only (sources, edges)
  u64 cnt[RL];
#pragma unroll
  for (int r = 0; r < RL; r++) cnt[r] = 0;
  if (tid < FCM_MAXD) s_cnt[tid] = 0;

  // number of source vertices of this shard: indices part, part+nparts, ... < nv_total
  const u32 ntasks = nv_total > part ? (nv_total - part + nparts - 1) / nparts : 0;

  for (;;) {
    __syncthreads();
    if (tid == 0) { s_task = atomicAdd(&ctrl[CT_TICKET], 1u); s_next = 0; }
    __syncthreads();
    const u32 t = s_task;
    if (t >= ntasks) break;
    const u32 v0 = order[part + t * nparts];
    const u64 o0 = csr_off[v0];
    const u32 k = (u32)(csr_off[v0 + 1] - o0);
    if (k > kcap) {                                     // tile too small: hand to the dense-row kernel
      if (tid == 0) overflow[atomicAdd(&ctrl[CT_NOVER], 1u)] = v0;
      continue;
    }
    if (tid == 0) { c0 += 1; c01 += k; }
    if constexpr (EMIT) {                               // the vertex and its out-edges
      if (tid == 0) emit_cell(ea, v0, nullptr, 0, 0);
      if (ea.emit_min <= 1 && ea.emit_max >= 1)
        for (u32 i = tid; i < k; i += NT) { const u32 v1 = csr[o0 + i]; emit_cell(ea, v0, &v1, 0, 1); }
      if (ea.emit_max < 2) continue;
    }
    if (k < 2) continue;                                // no 2-simplex possible
    for (u32 i = tid; i < k; i += NT) verts[i] = csr[o0 + i];
    __syncthreads();
    // ---- gather the k x k induced adjacency.  Tasks = (32 rows) x (32 columns); a lane owns a column,
    //      the loads of 8 rows are in flight together, lane r keeps the ballot of row r ----
    {
      u32* L32 = (u32*)L;
      const u32 nch = (k + 31) >> 5;
      for (u32 task = warp; task < nch * nch; task += NW) {
        const u32 rb = task / nch, c = task - rb * nch;
        const u32 j = c * 32 + lane;
        const u32 gv = j < k ? verts[j] : 0u;
        const u64* col = adj + (gv >> 6);
        const u32 gs = gv & 63;
        const u32 r0 = rb * 32, nr = min(32u, k - r0);
        u32 mine = 0;
#pragma unroll 1
        for (u32 r = 0; r < nr; r += 8) {
          u64 wv[8];
#pragma unroll
          for (int q = 0; q < 8; q++) wv[q] = r + q < nr ? __ldg(&col[(u64)verts[r0 + r + q] * W]) : 0ULL;
#pragma unroll
          for (int q = 0; q < 8; q++) {
            const u32 b = __ballot_sync(0xffffffffu, j < k && ((wv[q] >> gs) & 1ULL));
            if (lane == (int)(r + q)) mine = b;
          }
        }
        if ((u32)lane < nr) L32[(size_t)(r0 + lane) * (2 * LS) + c] = mine;
      }
      // words past the last column chunk
      if (nch < 2 * WK)
        for (u32 i = tid; i < k * (2 * WK - nch); i += NT) {
          const u32 row = i / (2 * WK - nch), c = nch + i % (2 * WK - nch);
          L32[(size_t)row * (2 * LS) + c] = 0;
        }
    }
    __syncthreads();
    // ---- enumerate ----
    {
      const u32* L32 = (const u32*)L;
      u32 top = 0;                                      // pending entries: stk[0..top)
      bool rows_left = true;
      for (;;) {
        if (top < 32 && rows_left) {
          u32 i = 0;
          if (lane == 0) i = atomicAdd(&s_next, 1u);
          i = __shfl_sync(0xffffffffu, i, 0);
          if (i >= k) { rows_left = false; continue; }
          // this lane's BPL bits of row i
          u32 bits;
          if constexpr (BPL >= 32) bits = L32[(size_t)i * (2 * LS) + lane];
          else bits = (L32[(size_t)i * (2 * LS) + ((lane * BPL) >> 5)] >> ((lane * BPL) & 31)) & ((1u << BPL) - 1u);
          const u32 pc = __popc(bits);
          c2 += pc;
          u32 incl = pc;
#pragma unroll
          for (int d = 1; d < 32; d <<= 1) { const u32 tt = __shfl_up_sync(0xffffffffu, incl, d); if (lane >= d) incl += tt; }
          const u32 c1 = __shfl_sync(0xffffffffu, incl, 31);
          if (!EMIT && c1 < 2) continue;                // a lone triangle has no 3-simplex above it
          u32 o = top + incl - pc;
          while (bits) {
            const u32 j = lane * BPL + (__ffs((int)bits) - 1);
            bits &= bits - 1;
            stk[o++] = (2ULL << 60) | ((u64)j << 10) | i;
          }
          top += c1;
          __syncwarp();
          continue;
        }
        if (top == 0) break;
        const u32 take = min(32u, top);
        const bool have = (u32)lane < take;
        u64 e = 0;
        u64 P[WK];
        u32 pc = 0, d = 0;
        if (have) {
          e = stk[top - take + lane];
          d = (u32)(e >> 60);
          const u64* R0 = L + (size_t)(e & 1023u) * LS;
#pragma unroll
          for (int y = 0; y < WK; y++) P[y] = R0[y];
#pragma unroll 1
          for (u32 t = 1; t < d; t++) {
            const u64* Rt = L + (size_t)((e >> (10 * t)) & 1023u) * LS;
#pragma unroll
            for (int y = 0; y < WK; y++) P[y] &= Rt[y];
          }
#pragma unroll
          for (int y = 0; y < WK; y++) pc += __popcll(P[y]);
#pragma unroll
          for (int r = 0; r < RL; r++) cnt[r] += (d == (u32)(2 + r)) ? pc : 0u;
          if constexpr (EMIT) {
            emit_cell(ea, v0, verts, e, d);
            if (d >= ea.emit_max) pc = 0;               // nothing deeper is asked for
          }
        }
        top -= take;
        // children: one entry per member of P.  A lone candidate has no child simplex of its own (the
        // graph has no self loops: P & L[j] = 0 for P = {j}), so it is not pushed.
        if (!EMIT && pc < 2) pc = 0;
        const bool deep = pc && d >= (u32)MAXIDS;
        const u32 want = deep ? 0u : pc;
        u32 incl = want;
#pragma unroll
        for (int s2 = 1; s2 < 32; s2 <<= 1) { const u32 tt = __shfl_up_sync(0xffffffffu, incl, s2); if (lane >= s2) incl += tt; }
        const u32 total = __shfl_sync(0xffffffffu, incl, 31);
        __syncwarp();                                   // every lane has read its entry
        if (top + total > scap) {                       // stack full: this batch's subtrees by per-thread DFS
          if (pc) deep_count<WK, EMIT>(P, (int)d + 1, L, s_cnt, ctrl, ea, v0, verts, e);
        } else {
          if (deep) deep_count<WK, EMIT>(P, (int)d + 1, L, s_cnt, ctrl, ea, v0, verts, e);
          else if (pc) {
            u32 o = top + incl - want;
            const u64 base = (e & 0x0fffffffffffffffULL) | ((u64)(d + 1) << 60);
            const u32 sh = 10 * d;
#pragma unroll
            for (int y = 0; y < WK; y++) {
              u64 b = P[y];
              while (b) {
                const u32 j = y * 64 + (__ffsll((long long)b) - 1);
                b &= b - 1;
                stk[o++] = base | ((u64)j << sh);
              }
            }
          }
          top += total;
        }
        __syncwarp();
      }
    }
  }
  // flush: thread -> block (shared atomics) -> global
  __syncthreads();
  if (c0) atomicAdd(&s_cnt[0], (unsigned long long)c0);
  if (c01) atomicAdd(&s_cnt[1], (unsigned long long)c01);
  if (c2) atomicAdd(&s_cnt[2], (unsigned long long)c2);
#pragma unroll
  for (int r = 0; r < RL; r++)
    if (cnt[r]) atomicAdd(&s_cnt[3 + r], (unsigned long long)cnt[r]);
  __syncthreads();
  if (tid < FCM_MAXD && s_cnt[tid]) atomicAdd((unsigned long long*)&counts[tid], s_cnt[tid]);
  // the all-reduce over the ranks of the box, fused: the CTA that finishes last owns the device's totals
  if (ca.mode != COMM_NONE) {
    __threadfence();
    __syncthreads();
    if (tid == 0) s_task = atomicAdd(&ctrl[CT_DONE], 1u) == gridDim.x - 1 ? 1u : 0u;
    __syncthreads();
    if (s_task) comm_last_cta(ca, counts, ctrl);
  }
}

// ---------------------------------------------------------------------------------------------------
// dense-row count: one warp per source vertex, W-word candidate sets, 128-bit row loads
// ---------------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(DENSE_WARPS * 32)
count_dense_kernel(const u64* __restrict__ adj, const u32* __restrict__ tasks, u32 ntasks_total,
                   const u32* __restrict__ ntasks_dev, u32 part, u32 nparts, u32 W, u32* __restrict__ ticket,
                   u32* __restrict__ ctrl, u64* __restrict__ scratch, u64* __restrict__ counts,
                   const __grid_constant__ CommArgs ca) {
  if (ntasks_dev) ntasks_total = *ntasks_dev;             // the overflow list's length is only known on the device
  const int lane = threadIdx.x & 31, wib = threadIdx.x >> 5;
  const u32 gwarp = blockIdx.x * DENSE_WARPS + wib;
  u64* P = scratch + (size_t)gwarp * MAXL_DENSE * W;     // P[l*W + x]
  __shared__ int s_cw[DENSE_WARPS][MAXL_DENSE];
  __shared__ u64 s_cb[DENSE_WARPS][MAXL_DENSE];
  __shared__ unsigned long long s_cnt[FCM_MAXD];
  if (threadIdx.x < FCM_MAXD) s_cnt[threadIdx.x] = 0;
  __syncthreads();
  u64 cnt[FCM_MAXD];                                     // warp-uniform; lane 0's copy is flushed
#pragma unroll
  for (int d = 0; d < FCM_MAXD; d++) cnt[d] = 0;
  bool dim_overflow = false;
  const u32 ntasks = ntasks_total > part ? (ntasks_total - part + nparts - 1) / nparts : 0;
  const u32 W2 = W >> 1;                                 // W is even: rows are 16-byte aligned

  for (;;) {
    u32 t = 0;
    if (lane == 0) t = atomicAdd(ticket, 1u);
    t = __shfl_sync(0xffffffffu, t, 0);
    if (t >= ntasks) break;
    const u32 v0 = tasks[part + t * nparts];
    cnt[0] += 1;
    {   // level 0 = out(v0)
      const ulonglong2* r = (const ulonglong2*)(adj + (u64)v0 * W);
      ulonglong2* p0 = (ulonglong2*)P;
      u32 pc = 0;
      for (u32 x = lane; x < W2; x += 32) { ulonglong2 q = __ldg(&r[x]); p0[x] = q; pc += __popcll(q.x) + __popcll(q.y); }
      pc = __reduce_add_sync(0xffffffffu, pc);
      cnt[1] += pc;
      if (!pc) continue;
    }
    __syncwarp();
    int l = 0;
    u32 w = 0;
    u64 bits = P[0];
    for (;;) {
      while (bits == 0) { if (++w == W) break; bits = P[(size_t)l * W + w]; }
      if (w == W) {
        if (l == 0) break;
        --l; w = s_cw[wib][l]; bits = s_cb[wib][l];
        continue;
      }
      const u32 j = w * 64 + (__ffsll((long long)bits) - 1);
      bits &= bits - 1;
      const ulonglong2* rj = (const ulonglong2*)(adj + (u64)j * W);
      const ulonglong2* pl = (const ulonglong2*)(P + (size_t)l * W);
      u32 cpc = 0;
      __syncwarp();                                      // every lane is done reading level l+1
      if (l + 1 < MAXL_DENSE) {
        ulonglong2* pn = (ulonglong2*)(P + (size_t)(l + 1) * W);
        for (u32 x = lane; x < W2; x += 32) {
          ulonglong2 a = pl[x], b = __ldg(&rj[x]);
          a.x &= b.x; a.y &= b.y;
          pn[x] = a;
          cpc += __popcll(a.x) + __popcll(a.y);
        }
        cpc = __reduce_add_sync(0xffffffffu, cpc);
        cnt[l + 2] += cpc;
        if (cpc) {
          if (lane == 0) { s_cw[wib][l] = w; s_cb[wib][l] = bits; }
          __syncwarp();
          ++l; w = 0; bits = P[(size_t)l * W];
        }
      } else {
        for (u32 x = lane; x < W2; x += 32) {
          ulonglong2 a = pl[x], b = __ldg(&rj[x]);
          cpc += __popcll(a.x & b.x) + __popcll(a.y & b.y);
        }
        cpc = __reduce_add_sync(0xffffffffu, cpc);
        if (cpc) dim_overflow = true;
      }
    }
  }
  if (lane == 0) {
#pragma unroll
    for (int d = 0; d < FCM_MAXD; d++)
      if (cnt[d]) atomicAdd(&s_cnt[d], (unsigned long long)cnt[d]);
    if (dim_overflow) atomicExch(&ctrl[CT_ERR], 2u);
  }
  __syncthreads();
  if (threadIdx.x < FCM_MAXD && s_cnt[threadIdx.x])
    atomicAdd((unsigned long long*)&counts[threadIdx.x], s_cnt[threadIdx.x]);
  if (ca.mode != COMM_NONE) {
    __shared__ u32 s_last;
    __threadfence();
    __syncthreads();
    if (threadIdx.x == 0) s_last = atomicAdd(&ctrl[CT_DONE], 1u) == gridDim.x - 1 ? 1u : 0u;
    __syncthreads();
    if (s_last) comm_last_cta(ca, counts, ctrl);
  }
}

// ---------------------------------------------------------------------------------------------------
// launcher
// ---------------------------------------------------------------------------------------------------
template <int WK, bool EMIT>
static int launch_local(fcm_graph* g, u32 part, u32 nparts, u32 kcap, const EmitArgs& ea, const CommArgs& ca) {
  fcm_ctx* c = g->ctx;
  constexpr int LS = LocalCfg<WK>::LS;
  constexpr int LOCAL_THREADS = LocalCfg<WK>::THREADS;
  size_t smem = (size_t)kcap * LS * 8 + (size_t)(kcap + (kcap & 1)) * 4 + (size_t)(LOCAL_THREADS / 32) * stack_cap(kcap) * 8;
  auto kern = count_local_kernel<WK, EMIT>;
  static std::mutex launch_mu;                          // the attribute is state of the function: set-and-launch is one step (host threads)
  std::lock_guard<std::mutex> launch_lock(launch_mu);
  FCM_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
  int per_sm = 0;
  FCM_CUDA(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, kern, LOCAL_THREADS, smem));
  if (per_sm < 1) per_sm = 1;
  u32 ntasks = g->n > part ? (g->n - part + nparts - 1) / nparts : 0;
  u32 blocks = (u32)std::min<u64>((u64)c->sm_count * per_sm, std::max<u32>(ntasks, 1));
  kern<<<blocks, LOCAL_THREADS, smem, c->stream>>>(g->adj, g->csr_off, g->csr, g->order, g->W, g->n,
                                                   part, nparts, kcap, g->ctrl, g->overflow, g->counts, ea, ca);
  c->launches++;
  FCM_CUDA(cudaGetLastError());
  return FCM_OK;
}

static int launch_dense(fcm_graph* g, const u32* tasks, u32 ntasks_total, const u32* ntasks_dev, u32 part, u32 nparts,
                        u32* ticket, const CommArgs& ca) {
  fcm_ctx* c = g->ctx;
  u32 blocks = (u32)c->sm_count * 4;
  size_t need = (size_t)blocks * DENSE_WARPS * MAXL_DENSE * g->W;
  if (need > g->scratch_words) {
    fcm_pool_free(g->scratch, c->stream);
    g->scratch = nullptr; g->scratch_words = 0;
    FCM_CUDA(fcm_pool_alloc(&g->scratch, need * 8, c->stream));
    g->scratch_words = need;
  }
  count_dense_kernel<<<blocks, DENSE_WARPS * 32, 0, c->stream>>>(g->adj, tasks, ntasks_total, ntasks_dev, part, nparts,
                                                                g->W, ticket, g->ctrl, g->scratch, g->counts, ca);
  c->launches++;
  FCM_CUDA(cudaGetLastError());
  return FCM_OK;
}

// Enqueue the count of shard (part, nparts) on the context stream; result in g->counts (device).  With comm the
// last CTA of the last kernel also runs the all-reduce over the ranks (fcm_comm.cu).  Nothing here waits for
// the device.
int fcm_count_launch(fcm_graph* g, u32 part, u32 nparts, const CommArgs* comm) {
  fcm_ctx* c = g->ctx;
  CommArgs none;
  memset(&none, 0, sizeof(none));
  const CommArgs& ca = comm ? *comm : none;
  FCM_CUDA(cudaMemsetAsync(g->counts, 0, (FCM_MAXD + 1) * 8, c->stream));
  FCM_CUDA(cudaMemsetAsync(g->ctrl, 0, CT_WORDS * 4, c->stream));
  if (g->n == 0 && ca.mode == COMM_NONE) return FCM_OK;
  if (g->kernel_kind == 1 || g->n == 0) return launch_dense(g, g->order, g->n, nullptr, part, nparts, &g->ctrl[CT_TICKET], ca);
  // tile: the smallest WK in {1,2,4,8,16} covering the largest out-degree (capped at 1024 columns)
  u32 kmax = g->maxout;
  int rc;
  u32 kcap;
  // (rows: the actual largest out-degree rounded up to 32, so that more CTAs fit one SM)
  kcap = std::min<u32>(1024u, std::max<u32>(32u, (kmax + 31u) & ~31u));
  const EmitArgs noemit = {0, 0, 0, nullptr, 0, nullptr, nullptr};
  // sources that do not fit the tile are appended to g->overflow and counted by the dense-row kernel behind
  // this one (their number stays on the device: ctrl[CT_NOVER]); the all-reduce then rides on that kernel
  const bool spill = kmax > kcap;
  const CommArgs& cl = spill ? none : ca;
  if (kmax <= 64) rc = launch_local<1, false>(g, part, nparts, kcap, noemit, cl);
  else if (kmax <= 128) rc = launch_local<2, false>(g, part, nparts, kcap, noemit, cl);
  else if (kmax <= 256) rc = launch_local<4, false>(g, part, nparts, kcap, noemit, cl);
  else if (kmax <= 512) rc = launch_local<8, false>(g, part, nparts, kcap, noemit, cl);
  else rc = launch_local<16, false>(g, part, nparts, kcap, noemit, cl);
  if (rc) return rc;
  if (spill) return launch_dense(g, g->overflow, 0, &g->ctrl[CT_NOVER], 0, 1, &g->ctrl[CT_TICKET2], ca);
  return FCM_OK;
}

// for_each_cell: the cells of dimension dim_min..dim_max of the whole graph into cells_dev (rows of
// dim_max + 1 vertices); *ncells_dev counts every cell found, also those beyond cap.
int fcm_cells_launch(fcm_graph* g, u32 dim_min, u32 dim_max, u32* cells_dev, u64 cap, unsigned long long* ncells_dev) {
  fcm_ctx* c = g->ctx;
  FCM_CUDA(cudaMemsetAsync(g->counts, 0, (FCM_MAXD + 1) * 8, c->stream));
  FCM_CUDA(cudaMemsetAsync(g->ctrl, 0, CT_WORDS * 4, c->stream));
  FCM_CUDA(cudaMemsetAsync(ncells_dev, 0, 8, c->stream));
  if (g->n == 0) return FCM_OK;
  const u32 kmax = g->maxout;
  if (kmax > 1024) { fcm_set_error("for_each_cell: out-degree %u exceeds the 1024-column tile", kmax); return FCM_ELIMIT; }
  if (dim_max > (u32)LocalCfg<1>::MAXIDS) { fcm_set_error("for_each_cell: dim_max > %d", LocalCfg<1>::MAXIDS); return FCM_ELIMIT; }
  const u32 kcap = std::min<u32>(1024u, std::max<u32>(32u, (kmax + 31u) & ~31u));
  const EmitArgs ea = {dim_min, dim_max, dim_max + 1, cells_dev, cap, ncells_dev, g->oldid};
  CommArgs none;
  memset(&none, 0, sizeof(none));
  if (kmax <= 64) return launch_local<1, true>(g, 0, 1, kcap, ea, none);
  if (kmax <= 128) return launch_local<2, true>(g, 0, 1, kcap, ea, none);
  if (kmax <= 256) return launch_local<4, true>(g, 0, 1, kcap, ea, none);
  if (kmax <= 512) return launch_local<8, true>(g, 0, 1, kcap, ea, none);
  return launch_local<16, true>(g, 0, 1, kcap, ea, none);
}

// ---------------------------------------------------------------------------------------------------
// fcm_graph_count_noalloc: the dense-row count with its scratch in a file-scope device array.
// No counterpart in the reference; added for callers that cannot call
// cudaMalloc (code under CUDA graph capture, tight serving loops, harnesses that forbid it).  The entry point
// calls no cudaMalloc / cudaFree / pool routine: count_dense_kernel's candidate-set stacks, its ticket, its error
// flag and the count vector all live in g_noalloc_scratch below.  Every GPU that loads libfcm.so holds the array
// (about 4.1 MB) until the process exits.
//   limit   rows of at most FCM_NOALLOC_MAXW 64-bit words, i.e. n <= 64 * FCM_NOALLOC_MAXW = 4096 vertices (the row
//           stride is ceil(n/64) + 1 rounded down to even); larger graphs are refused with FCM_ELIMIT.
//   sharing every call on a device uses the same array and finds the previous call's contents in it: the words that
//           must start at zero (counts, ticket, flags) are cleared by the call; the stacks are written before they are
//           read.  Calls that overlap on the GPU would race - inside one process the call holds a mutex from its
//           first enqueue to its synchronize, so they cannot.
// In host code `g_noalloc_scratch` is a host-side stand-in for the symbol, not a device address: the address the
// kernel gets comes from cudaGetSymbolAddress, asked once per device with that device current (noalloc_scratch()).
// ---------------------------------------------------------------------------------------------------
#define FCM_NOALLOC_MAXW_ 64                               // == FCM_NOALLOC_MAXW (include/fcm.h)
#define NOALLOC_BLOCKS 64                                  // 64 blocks x DENSE_WARPS warps = 256 stacks
#define NOALLOC_HDR_WORDS 64                               // [0, 33) counts, [40, 44) = u32 ctrl[CT_WORDS], rest unused
#define NOALLOC_WORDS (NOALLOC_HDR_WORDS + (size_t)NOALLOC_BLOCKS * DENSE_WARPS * MAXL_DENSE * FCM_NOALLOC_MAXW_)
static_assert(FCM_NOALLOC_MAXW_ == FCM_NOALLOC_MAXW, "fcm.h and fcm_count.cu disagree on the limit");
static_assert(NOALLOC_WORDS * 8 <= (5u << 20), "the static scratch is meant to stay modest");
static __device__ __align__(256) unsigned long long g_noalloc_scratch[NOALLOC_WORDS];

#define NOALLOC_MAXDEV 64
static std::mutex g_noalloc_mu;
static u64* g_noalloc_ptr[NOALLOC_MAXDEV];                 // device address of the array on each device (null: not asked yet)

// device address of g_noalloc_scratch on `device` (which must be current); nullptr + error text on failure
static u64* noalloc_scratch(int device) {
  if (device < 0 || device >= NOALLOC_MAXDEV) { fcm_set_error("device ordinal %d beyond %d", device, NOALLOC_MAXDEV - 1); return nullptr; }
  if (!g_noalloc_ptr[device]) {
    void* p = nullptr;
    const cudaError_t e = cudaGetSymbolAddress(&p, g_noalloc_scratch);
    if (e != cudaSuccess || !p) { fcm_set_error("cudaGetSymbolAddress(g_noalloc_scratch) -> %s", cudaGetErrorString(e)); cudaGetLastError(); return nullptr; }
    g_noalloc_ptr[device] = (u64*)p;
  }
  return g_noalloc_ptr[device];
}

extern "C" int fcm_graph_count_noalloc(fcm_graph* g, uint32_t part, uint32_t nparts, uint64_t* counts, int* len) {
  if (!g || !counts || !len) { fcm_set_error("fcm_graph_count_noalloc: graph/counts/len is null"); return FCM_EINVAL; }
  if (nparts < 1 || part >= nparts) { fcm_set_error("fcm_graph_count_noalloc: need part < nparts"); return FCM_EINVAL; }
  if (g->W > FCM_NOALLOC_MAXW_) {
    fcm_set_error("fcm_graph_count_noalloc: rows of %u words exceed the static scratch (FCM_NOALLOC_MAXW = %d words, n <= %d vertices); "
                  "use fcm_graph_count", g->W, FCM_NOALLOC_MAXW_, 64 * FCM_NOALLOC_MAXW_);
    return FCM_ELIMIT;
  }
  fcm_ctx* c = g->ctx;
  FCM_CUDA(cudaSetDevice(c->device));
  std::lock_guard<std::mutex> lk(g_noalloc_mu);            // one user of the array at a time (see above)
  u64* const base = noalloc_scratch(c->device);
  if (!base) return FCM_ECUDA;
  u64* const cnt = base;
  u32* const ctrl = (u32*)(base + 40);
  u64* const stacks = base + NOALLOC_HDR_WORDS;
  cudaStream_t s = c->stream;
  FCM_CUDA(cudaMemsetAsync(base, 0, NOALLOC_HDR_WORDS * 8, s));   // counts, ticket, flags start at zero; the stacks need not
  if (g->n) {
    CommArgs none;
    memset(&none, 0, sizeof(none));
    count_dense_kernel<<<NOALLOC_BLOCKS, DENSE_WARPS * 32, 0, s>>>(g->adj, g->order, g->n, nullptr, part, nparts, g->W,
                                                                   &ctrl[CT_TICKET], ctrl, stacks, cnt, none);
    c->launches++;
    FCM_CUDA(cudaGetLastError());
  }
  u64* const pin = c->pin;                                 // pinned words of the context (allocated at fcm_ctx_create)
  FCM_CUDA(cudaMemcpyAsync(pin, base, (FCM_MAXD + 1 + 7 + CT_WORDS / 2) * 8, cudaMemcpyDeviceToHost, s));
  FCM_CUDA(cudaStreamSynchronize(s));
  const u32* const hctrl = (const u32*)(pin + 40);
  if (hctrl[CT_ERR]) { fcm_set_error("a simplex of dimension >= %d was met", FCM_MAXD); return FCM_EDIM; }
  memcpy(counts, pin, FCM_MAXD * 8);
  int l = FCM_MAXD;
  while (l > 0 && counts[l - 1] == 0) l--;
  *len = l;
  return FCM_OK;
}
