// fcm_count.cu — directed-flag-complex simplex counting on sm_100a.
//
// Replaces `graph.flagser_count()` of the external flag-complex crate as called from
// src/lib.rs:51 (State::new), :130 (Bounds::calculate, normalised graph) and
// flag_complex::count_cells (src/bin/all_cxs.rs:123).  Algorithm = the flagser directed-clique
// recursion: every simplex (v0..vd) owns the candidate set
// P = out(v0) ∩ ... ∩ out(vd); each member of P closes a (d+1)-simplex.
//
// Two kernels, identical results:
//   count_local_kernel<WK>  one CTA per source vertex v0.  out(v0) is relabelled 0..k-1 and the
//       k x k induced adjacency is gathered ONCE from the packed HBM rows into shared memory
//       (k*WK 64-bit words).  Every deeper intersection is then a WK-word AND in shared memory
//       instead of a W-word row read (C4: WK <= 8 against W = 490), one DFS per thread.
//   count_dense_kernel      one warp per source vertex, candidate sets of full row width staged
//       in a per-warp stack, rows read with coalesced 128-bit loads, popc + warp reduce.  This is
//       the "one adjacency row per simplex" kernel of the roofline model; it also
//       serves sources whose out-degree exceeds the shared-memory tile.
#include "fcm_internal.cuh"

#include <algorithm>

#define LOCAL_THREADS 128
#define DENSE_WARPS 4
#define MAXL_LOCAL (FCM_MAXD - 2)   // level l holds candidates for vertex index l+2
#define MAXL_DENSE (FCM_MAXD - 1)   // level l holds candidates for vertex index l+1

// ---------------------------------------------------------------------------------------------------
// graph construction
// ---------------------------------------------------------------------------------------------------
__global__ void set_bits_kernel(const uint2* __restrict__ edges, u64 ne, u32 n, u32 W, int normalise,
                                u64* __restrict__ adj, u32* __restrict__ ctrl) {
  u64 i = (u64)blockIdx.x * blockDim.x + threadIdx.x;
  u64 stride = (u64)gridDim.x * blockDim.x;
  for (; i < ne; i += stride) {
    uint2 e = edges[i];
    u32 a = e.x, b = e.y;
    if (a >= n || b >= n || a == b) { atomicExch(&ctrl[CT_ERR], 1u); continue; }
    if (normalise && a < b) { u32 t = a; a = b; b = t; }   // src/lib.rs:126-129: edge big -> small
    atomicOr((unsigned long long*)&adj[(u64)a * W + (b >> 6)], 1ULL << (b & 63));
  }
}

// one warp per row: out-degree
__global__ void row_degree_kernel(const u64* __restrict__ adj, u32 n, u32 W, u64* __restrict__ off) {
  u32 warp = (blockIdx.x * blockDim.x + threadIdx.x) >> 5, lane = threadIdx.x & 31;
  if (warp >= n) return;
  const u64* row = adj + (u64)warp * W;
  u32 c = 0;
  for (u32 w = lane; w < W; w += 32) c += __popcll(row[w]);
  c = __reduce_add_sync(0xffffffffu, c);
  if (lane == 0) off[warp + 1] = c;     // exclusive scan done afterwards (host, n values)
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

int fcm_graph_build_device(fcm_graph* g, const uint2* edges_dev, u64 ne_in, int normalise,
                           std::vector<u64>& off_host) {
  fcm_ctx* c = g->ctx;
  cudaStream_t s = c->stream;
  size_t words = (size_t)g->n * g->W;
  FCM_CUDA(cudaMemsetAsync(g->adj, 0, words * 8, s));
  FCM_CUDA(cudaMemsetAsync(g->ctrl, 0, CT_WORDS * 4, s));
  if (ne_in) {
    int blocks = (int)std::min<u64>((ne_in + 255) / 256, 148 * 16);
    set_bits_kernel<<<blocks, 256, 0, s>>>(edges_dev, ne_in, g->n, g->W, normalise, g->adj, g->ctrl);
    c->launches++;
  }
  FCM_CUDA(cudaMemsetAsync(g->csr_off, 0, 8, s));
  if (g->n) {
    int blocks = (g->n * 32 + 255) / 256;
    row_degree_kernel<<<blocks, 256, 0, s>>>(g->adj, g->n, g->W, g->csr_off);
    c->launches++;
  }
  off_host.assign(g->n + 1, 0);
  u32 ctrl_h[CT_WORDS];
  FCM_CUDA(cudaMemcpyAsync(off_host.data(), g->csr_off, (g->n + 1) * 8, cudaMemcpyDeviceToHost, s));
  FCM_CUDA(cudaMemcpyAsync(ctrl_h, g->ctrl, sizeof(ctrl_h), cudaMemcpyDeviceToHost, s));
  FCM_CUDA(cudaStreamSynchronize(s));
  if (ctrl_h[CT_ERR]) { fcm_set_error("edge list holds a vertex id >= n or a self loop"); return FCM_EINVAL; }
  u32 maxout = 0;
  for (u32 v = 0; v < g->n; v++) { u64 d = off_host[v + 1]; if (d > maxout) maxout = (u32)d; off_host[v + 1] = off_host[v] + d; }
  g->maxout = maxout;
  g->ne = off_host[g->n];
  FCM_CUDA(cudaMemcpyAsync(g->csr_off, off_host.data(), (g->n + 1) * 8, cudaMemcpyHostToDevice, s));
  FCM_CUDA(cudaMalloc(&g->csr, std::max<u64>(g->ne, 1) * 4));
  if (g->n) {
    int blocks = (g->n * 32 + 255) / 256;
    row_to_csr_kernel<<<blocks, 256, 0, s>>>(g->adj, g->n, g->W, g->csr_off, g->csr);
    c->launches++;
  }
  FCM_CUDA(cudaGetLastError());
  return FCM_OK;
}

// ---------------------------------------------------------------------------------------------------
// block-local relabelled count
// ---------------------------------------------------------------------------------------------------
template <int WK> struct LocalCfg { static constexpr int LS = WK | 1; };   // odd row stride: no bank conflicts

template <int WK>
__global__ void __launch_bounds__(LOCAL_THREADS)
count_local_kernel(const u64* __restrict__ adj, const u64* __restrict__ csr_off,
                   const u32* __restrict__ csr, const u32* __restrict__ order, u32 W, u32 nv_total,
                   u32 part, u32 nparts, u32 kcap, u32* __restrict__ ctrl,
                   u32* __restrict__ overflow, u64* __restrict__ counts) {
  constexpr int LS = LocalCfg<WK>::LS;
  extern __shared__ __align__(16) unsigned char smem_raw[];
  u64* L = (u64*)smem_raw;                              // kcap * LS
  u32* verts = (u32*)(L + (size_t)kcap * LS);           // kcap
  __shared__ u32 s_task, s_next;
  __shared__ unsigned long long s_cnt[FCM_MAXD];

  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  constexpr int NW = LOCAL_THREADS / 32;
  u64 cnt[FCM_MAXD];
#pragma unroll
  for (int d = 0; d < FCM_MAXD; d++) cnt[d] = 0;
  if (tid < FCM_MAXD) s_cnt[tid] = 0;
  bool dim_overflow = false;

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
    if (tid == 0) { cnt[0] += 1; cnt[1] += k; }
    if (k > kcap) {                                     // tile too small: hand to the dense-row kernel
      if (tid == 0) { cnt[0] -= 1; cnt[1] -= k; overflow[atomicAdd(&ctrl[CT_NOVER], 1u)] = v0; }
      continue;
    }
    if (k < 2) continue;                                // no 2-simplex possible
    for (u32 i = tid; i < k; i += LOCAL_THREADS) verts[i] = csr[o0 + i];
    __syncthreads();
    // gather the k x k induced adjacency: warp per row, lane per column, ballot -> 32 columns/instr
    u32* L32 = (u32*)L;
    const u32 nch = (k + 31) >> 5;
    for (u32 i = warp; i < k; i += NW) {
      const u64* row = adj + (u64)verts[i] * W;
      u32* dst = L32 + (size_t)i * (2 * LS);
      for (u32 c = 0; c < 2 * WK; c++) {
        u32 b = 0;
        if (c < nch) {
          u32 j = c * 32 + lane;
          bool bit = false;
          if (j < k) { u32 gv = verts[j]; bit = (__ldg(&row[gv >> 6]) >> (gv & 63)) & 1ULL; }
          b = __ballot_sync(0xffffffffu, bit);
        }
        if (lane == 0) dst[c] = b;
      }
    }
    __syncthreads();
    // one DFS per thread, roots = local rows (v1), fetched dynamically
    for (;;) {
      u32 i = atomicAdd(&s_next, 1u);
      if (i >= k) break;
      u64 P[MAXL_LOCAL][WK];
      int cw[MAXL_LOCAL];
      u64 cb[MAXL_LOCAL];
      u32 pc = 0;
#pragma unroll
      for (int x = 0; x < WK; x++) { u64 q = L[(size_t)i * LS + x]; P[0][x] = q; pc += __popcll(q); }
      cnt[2] += pc;
      if (!pc) continue;
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
        if (l + 1 < MAXL_LOCAL) {
#pragma unroll
          for (int x = 0; x < WK; x++) { u64 q = P[l][x] & Lj[x]; P[l + 1][x] = q; cpc += __popcll(q); }
          cnt[l + 3] += cpc;
          if (cpc) { cw[l] = w; cb[l] = bits; ++l; w = 0; bits = P[l][0]; }
        } else {
#pragma unroll
          for (int x = 0; x < WK; x++) cpc += __popcll(P[l][x] & Lj[x]);
          if (cpc) dim_overflow = true;                 // a simplex of dimension >= FCM_MAXD
        }
      }
    }
  }
  // flush: thread -> block (shared atomics) -> global
  __syncthreads();
#pragma unroll
  for (int d = 0; d < FCM_MAXD; d++)
    if (cnt[d]) atomicAdd(&s_cnt[d], (unsigned long long)cnt[d]);
  if (dim_overflow) atomicExch(&ctrl[CT_ERR], 2u);
  __syncthreads();
  if (tid < FCM_MAXD && s_cnt[tid]) atomicAdd((unsigned long long*)&counts[tid], s_cnt[tid]);
}

// ---------------------------------------------------------------------------------------------------
// dense-row count: one warp per source vertex, W-word candidate sets, 128-bit row loads
// ---------------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(DENSE_WARPS * 32)
count_dense_kernel(const u64* __restrict__ adj, const u32* __restrict__ tasks, u32 ntasks_total,
                   u32 part, u32 nparts, u32 W, u32* __restrict__ ticket, u32* __restrict__ ctrl,
                   u64* __restrict__ scratch, u64* __restrict__ counts) {
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
}

// ---------------------------------------------------------------------------------------------------
// launcher
// ---------------------------------------------------------------------------------------------------
template <int WK>
static int launch_local(fcm_graph* g, u32 part, u32 nparts, u32 kcap) {
  fcm_ctx* c = g->ctx;
  constexpr int LS = LocalCfg<WK>::LS;
  size_t smem = (size_t)kcap * LS * 8 + (size_t)kcap * 4;
  auto kern = count_local_kernel<WK>;
  FCM_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
  int per_sm = 0;
  FCM_CUDA(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, kern, LOCAL_THREADS, smem));
  if (per_sm < 1) per_sm = 1;
  u32 ntasks = g->n > part ? (g->n - part + nparts - 1) / nparts : 0;
  u32 blocks = (u32)std::min<u64>((u64)c->sm_count * per_sm, std::max<u32>(ntasks, 1));
  kern<<<blocks, LOCAL_THREADS, smem, c->stream>>>(g->adj, g->csr_off, g->csr, g->order, g->W, g->n,
                                                   part, nparts, kcap, g->ctrl, g->overflow, g->counts);
  c->launches++;
  FCM_CUDA(cudaGetLastError());
  return FCM_OK;
}

static int launch_dense(fcm_graph* g, const u32* tasks, u32 ntasks_total, u32 part, u32 nparts,
                        u32* ticket, bool tasks_count_on_device) {
  fcm_ctx* c = g->ctx;
  (void)tasks_count_on_device;
  u32 blocks = (u32)c->sm_count * 4;
  size_t need = (size_t)blocks * DENSE_WARPS * MAXL_DENSE * g->W;
  if (need > g->scratch_words) {
    if (g->scratch) FCM_CUDA(cudaFree(g->scratch));
    g->scratch = nullptr; g->scratch_words = 0;
    FCM_CUDA(cudaMalloc(&g->scratch, need * 8));
    g->scratch_words = need;
  }
  count_dense_kernel<<<blocks, DENSE_WARPS * 32, 0, c->stream>>>(g->adj, tasks, ntasks_total, part, nparts,
                                                                g->W, ticket, g->ctrl, g->scratch, g->counts);
  c->launches++;
  FCM_CUDA(cudaGetLastError());
  return FCM_OK;
}

// Enqueue the count of shard (part, nparts) on the context stream; result in g->counts (device).
int fcm_count_launch(fcm_graph* g, u32 part, u32 nparts) {
  fcm_ctx* c = g->ctx;
  FCM_CUDA(cudaMemsetAsync(g->counts, 0, FCM_MAXD * 8, c->stream));
  FCM_CUDA(cudaMemsetAsync(g->ctrl, 0, CT_WORDS * 4, c->stream));
  if (g->n == 0) return FCM_OK;
  if (g->kernel_kind == 1) return launch_dense(g, g->order, g->n, part, nparts, &g->ctrl[CT_TICKET], false);
  // tile: the smallest WK in {1,2,4,8,16} covering the largest out-degree (capped at 1024 columns)
  u32 kmax = g->maxout;
  int rc;
  u32 kcap;
  if (kmax <= 64) { kcap = 64; rc = launch_local<1>(g, part, nparts, kcap); }
  else if (kmax <= 128) { kcap = 128; rc = launch_local<2>(g, part, nparts, kcap); }
  else if (kmax <= 256) { kcap = 256; rc = launch_local<4>(g, part, nparts, kcap); }
  else if (kmax <= 512) { kcap = 512; rc = launch_local<8>(g, part, nparts, kcap); }
  else { kcap = 1024; rc = launch_local<16>(g, part, nparts, kcap); }
  if (rc) return rc;
  if (kmax > kcap) {
    // sources that did not fit the tile were appended to g->overflow; their number is only known
    // on the device, so the dense kernel takes it from ctrl[CT_NOVER] via a tiny host hop.
    u32 nover = 0;
    FCM_CUDA(cudaMemcpyAsync(&nover, &g->ctrl[CT_NOVER], 4, cudaMemcpyDeviceToHost, c->stream));
    FCM_CUDA(cudaStreamSynchronize(c->stream));
    if (nover) return launch_dense(g, g->overflow, nover, 0, 1, &g->ctrl[CT_TICKET2], true);
  }
  return FCM_OK;
}
