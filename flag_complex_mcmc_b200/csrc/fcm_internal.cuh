// fcm_internal.cuh — shared declarations of libfcm.so (not part of the ABI).
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>
#include <string>
#include <unordered_map>
#include <vector>

#include "../../include/fcm.h"

typedef uint64_t u64;
typedef uint32_t u32;

#define FCM_NONE 0xffffffffu

// ---- error plumbing ------------------------------------------------------------------------------
void fcm_set_error(const char* fmt, ...);
#define FCM_CUDA(call)                                                                       \
  do {                                                                                       \
    cudaError_t e__ = (call);                                                                \
    if (e__ != cudaSuccess) {                                                                \
      fcm_set_error("%s:%d: %s -> %s", __FILE__, __LINE__, #call, cudaGetErrorString(e__));  \
      return FCM_ECUDA;                                                                      \
    }                                                                                        \
  } while (0)

struct fcm_ctx {
  int device = 0;
  int sm_count = 0;
  size_t smem_optin = 0;
  cudaStream_t own = nullptr, stream = nullptr;
  cudaEvent_t ev0 = nullptr, ev1 = nullptr;
  u64 launches = 0;
  unsigned char* stage_host[2] = {nullptr, nullptr};   // pinned staging of big host -> device uploads (edge lists), FCM_STAGE_BYTES each
  cudaEvent_t stage_evs[16] = {};                      // per staging thread and slice: the copy out of it has completed
  u64* pin = nullptr;               // FCM_PIN_WORDS of pinned host memory: small read-backs that must not take the
                                    // driver's pageable-copy path (it blocks the other host threads' launches)
};
#define FCM_PIN_WORDS 128
#define FCM_STAGE_BYTES (8u << 20)

// ---- device graph for recounts ---------------------------------------------------------------------
struct fcm_graph {
  fcm_ctx* ctx = nullptr;
  u32 n = 0, W = 0, maxout = 0;     // W = u64 words per adjacency row (even: 16-byte row stride)
  u64 ne = 0;                       // number of distinct directed edges
  u64* adj = nullptr;               // n * W
  u64* csr_off = nullptr;           // n + 1
  u32* csr = nullptr;               // ne, each out-list ascending
  u32* order = nullptr;             // vertices by descending out-degree
  u32* oldid = nullptr;             // n, or null: internal vertex id -> caller's id when the graph was built under the locality order (fcm_count.cu)
  u64* counts = nullptr;            // FCM_MAXD + 1 ([FCM_MAXD]: ranks in error, all-reduce paths only)
  u32* ctrl = nullptr;              // [0] ticket, [1] overflow count, [2] error flag, [3] ticket2
  u32* overflow = nullptr;          // n
  u64* scratch = nullptr;           // dense-row kernel P stacks
  size_t scratch_words = 0;
  int kernel_kind = 0;
};

// Device memory of graphs comes from the device's stream-ordered pool (its release threshold is
// raised at fcm_ctx_create, so the memory of a destroyed graph serves the next one: flagser_count is
// called over and over on small graphs by the research binaries, src/bin/all_cxs.rs:57-83).
template <typename T>
static inline cudaError_t fcm_pool_alloc(T** p, size_t bytes, cudaStream_t s) { return cudaMallocAsync((void**)p, bytes, s); }
static inline void fcm_pool_free(void* p, cudaStream_t s) { if (p) cudaFreeAsync(p, s); }

// device control words
enum { CT_TICKET = 0, CT_NOVER = 1, CT_ERR = 2, CT_TICKET2 = 3, CT_DONE = 4, CT_WORDS = 8 };
// CT_ERR values: 1 bad edge list (build), 2 a simplex of dimension >= FCM_MAXD, 3 all-reduce timed out

// ---- all-reduce of the per-dimension counts across the ranks of one box (fcm_comm.cu) ----------------
// A rank's window: two slots used by alternating all-reduce calls.  Every rank adds its vector into slot s
// of EVERY rank's window (system-scope atomics: NVLink for the peers), then bumps `arrived` there; a rank
// reads its own slot once `arrived` == nranks and clears it for the call after the next.  (Slot s is only
// written again by a peer that has finished the call in between, which needs this rank's contribution to
// that call, which this rank issues after the clear: two slots suffice.)
struct CommSlot {
  unsigned long long sum[FCM_MAXD + 1];   // [FCM_MAXD]: number of ranks whose shard met a simplex of dimension >= FCM_MAXD
  u32 arrived;
  u32 pad[61];                            // 512 bytes per slot
};
static_assert(sizeof(CommSlot) == 512, "CommSlot layout");
enum { COMM_NONE = 0, COMM_PEER_EPILOGUE = 1, COMM_ERR_WORD = 2 };
struct CommArgs {                         // kernel parameter of the count kernels
  u32 mode;                               // COMM_*: nothing / fused all-reduce by the last CTA / last CTA copies the error flag behind the counts (NCCL follows)
  u32 nranks, rank, slot;
  unsigned long long timeout_ns;
  unsigned long long peer[FCM_COMM_MAXRANKS];   // rank r's window as mapped into this process
};

#ifdef __CUDACC__
__device__ __forceinline__ u32 comm_ld_acquire_sys(const u32* p) {
  u32 v;
  asm volatile("ld.acquire.sys.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
  return v;
}
__device__ __forceinline__ unsigned long long comm_ld_relaxed_sys(const unsigned long long* p) {
  unsigned long long v;
  asm volatile("ld.relaxed.sys.global.u64 %0, [%1];" : "=l"(v) : "l"(p) : "memory");
  return v;
}
__device__ __forceinline__ unsigned long long comm_globaltimer() {
  unsigned long long t;
  asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t));
  return t;
}
// Run by every thread of the LAST CTA of a count (>= 64 threads), after all CTAs' atomics on counts[] have been
// fenced: counts[0..FCM_MAXD) becomes the sum over all ranks, counts[FCM_MAXD] the number of ranks in error.
__device__ __forceinline__ void comm_last_cta(const CommArgs& ca, u64* counts, u32* ctrl) {
  const int tid = threadIdx.x;
  __threadfence();
  unsigned long long v = 0;
  if (tid < FCM_MAXD) v = __ldcg((const unsigned long long*)&counts[tid]);
  else if (tid == FCM_MAXD) v = __ldcg(&ctrl[CT_ERR]) ? 1ULL : 0ULL;
  if (ca.mode == COMM_ERR_WORD) {
    if (tid == FCM_MAXD) counts[FCM_MAXD] = v;
    return;
  }
  if (tid <= FCM_MAXD) {
    if (v)
      for (u32 k = 0; k < ca.nranks; k++) {              // start with the own window: the ranks do not all hit rank 0 first
        u32 q = ca.rank + k; if (q >= ca.nranks) q -= ca.nranks;
        atomicAdd_system(&((CommSlot*)ca.peer[q] + ca.slot)->sum[tid], v);
      }
    __threadfence_system();
  }
  __syncthreads();
  CommSlot* const mine = (CommSlot*)ca.peer[ca.rank] + ca.slot;
  if (tid == 0) {
    __threadfence_system();
    for (u32 k = 0; k < ca.nranks; k++) {
      u32 q = ca.rank + k; if (q >= ca.nranks) q -= ca.nranks;
      atomicAdd_system(&((CommSlot*)ca.peer[q] + ca.slot)->arrived, 1u);
    }
    const unsigned long long t0 = comm_globaltimer();
    u32 spins = 0;
    while (comm_ld_acquire_sys(&mine->arrived) < ca.nranks) {
      __nanosleep(64);
      if ((++spins & 255u) == 0 && comm_globaltimer() - t0 > ca.timeout_ns) { atomicExch(&ctrl[CT_ERR], 3u); break; }
    }
  }
  __syncthreads();
  if (tid <= FCM_MAXD) {
    __threadfence_system();
    const unsigned long long tot = comm_ld_relaxed_sys(&mine->sum[tid]);
    counts[tid] = tot;
    if (tid == FCM_MAXD && tot) atomicMax(&ctrl[CT_ERR], 2u);
    mine->sum[tid] = 0;
  }
  __syncthreads();
  if (tid == 0) { mine->arrived = 0; __threadfence_system(); }
}
#endif

// fcm_count.cu
int fcm_count_launch(fcm_graph* g, u32 part, u32 nparts, const CommArgs* comm = nullptr);
int fcm_cells_launch(fcm_graph* g, u32 dim_min, u32 dim_max, u32* cells_dev, u64 cap, unsigned long long* ncells_dev);
int fcm_graph_build_device(fcm_graph* g, const uint2* edges_dev, u64 ne_in, int normalise);
// host edge list -> device, through the context's pinned staging buffers (fcm_api.cu)
int fcm_upload_edges(fcm_ctx* c, void* dst_dev, const void* src_host, size_t bytes);

// ---- ensemble of chains ----------------------------------------------------------------------------
#define FCM_MAXB (FCM_MAXD + 1)     // bounds vectors may be one longer (src/lib.rs:152 push(10))
#define FCM_MAXEDIT 128                 // change_edges of one transition (clique_swap of two 8-cliques: 112)
#define FCM_MAXCQ 32                    // vertices of d = c ++ a ++ b in clique_swap (src/lib.rs:240): cliques of order <= 16

struct ChainState {                 // one per chain, in HBM
  u64 rng[4];
  u64 sampled, accepted, sumK, nonempty;
  u64 flag_count[FCM_MAXD];
  u32 E, ND;
  int fc_len;
  int err;
};

struct EnsParams {
  u32 n, W, nchains;
  u64 npairs;
  u32 maxc, Kcap, KWcap;
  // static, shared by all chains
  const uint2* pair_ab;             // pair id -> (big, small)
  const u32* nb_off;                // npairs + 1 (total list length < 2^32 is checked at create)
  const u32* nb_list;               // common undirected neighbours, ascending
  // static clique index (null when not built: graph too large for it, see fcm_ensemble.cu).  For the
  // pair (big,small) with common neighbours w_0..w_{c-1}:
  //   pinfo[2*pair]   = (big, small, nb_off, k4_off)      pinfo[2*pair+1] = (c, m, 0, 0)
  //   tri[nb_off + i] = (slot of w_i -> big, slot of w_i -> small)
  //   k4[k4_off + e]  = slot of w_i -> w_j << 10 | j << 5 | i   for the m adjacent i < j (lists of <= 32 only)
  const uint4* pinfo;
  const uint2* tri;
  const u32* k4;
  // per chain: one block of stride32 32-bit words = [adj | items | pos | dbl | dblpos | ori]; the pointers
  // below are chain 0's, chain k's are stride32 words further (one base register per chain)
  u64 stride32;
  u32 off_items, off_pos, off_dbl, off_dblpos, off_ori;   // word offsets inside a block (adj is at 0)
  u64* adj;                         // n * W        packed 64-bit out-neighbour rows
  u32* items;                       // 2*npairs     directed-edge slots in container order
  u32* pos;                         // 2*npairs     slot -> index in items | NONE
  u32* dbl;                         // npairs       pair ids in container order
  u32* dblpos;                      // npairs
  u32* ori;                         // ceil(2*npairs/32)  orientation bits: bit `slot` = that directed edge exists
  ChainState* st;                   // nchains
  // settings
  u64 bmin[FCM_MAXB], bmax[FCM_MAXB];
  int nmin, nmax;
  u32 sample_policy;                // edge-sampler policy (fcm_ensemble_set_policy): 0 = element r, 1 = element len-1-r
  double cum[3], scale;             // WeightedIndex<f64>
  // clique moves (src/lib.rs:214-290); cq_norders = 0 when the ensemble was created without them
  int cq_norders;                   // cliques_by_order.len()
  double cq_cum[FCM_MAXD], cq_scale; // clique_order_distribution (src/bin/sample.rs:87-88)
  u32 cq_count[FCM_MAXD + 1];       // maximal cliques of order o + 1
  u32 cq_base[FCM_MAXD + 1];        // first vertex of the first clique of order o + 1 in cq_verts
  const u32* cq_verts;              // cliques of one order back to back, vertices ascending
  const u64* pkeys;                 // (big << 32 | small) of every adjacent pair, ascending
  const u32* pkey_pair;             // ... and its pair id
  const u32* pkey_begin;            // n + 1: first key whose bigger vertex is >= v (the binary search of a pair stays in one vertex's run)
};

struct fcm_ensemble {
  fcm_ctx* ctx = nullptr;
  EnsParams p;
  u64 ne0 = 0, nd0 = 0;
  std::unordered_map<u64, u32> pairmap;   // (big<<32|small) -> pair id   (host side, for fcm_apply)
  std::vector<uint2> pair_ab_host;
  std::vector<void*> dev_allocs;
  u64 bytes_per_chain = 0;
  size_t smem_per_warp = 0;
  int warps_per_block = 4;
  int force_full = 0;                     // 1: every proposal through the full pre/post count path
  int force_rows = 0;                     // 1: hot path gathers from the adjacency rows even when the clique index exists
  int have_index = 0;                     // the static clique index was built
  int cliques = 0;                        // clique moves enabled (bigger per-warp shared memory)
  int deep_hint = 0;                      // the start graph has simplices beyond the hot loop's register levels (dimension > 5)
  int stats = 1;                          // 1: the kernels keep the neighbourhood-size statistic (sumK / nonempty)
  u64 index_bytes = 0;
  cudaAccessPolicyWindow l2_window = {};    // the clique index slab, given to every chain-kernel launch (num_bytes = 0: none)
  fcm_trace_rec* trace_dev = nullptr;
  u64 trace_cap = 0;
  u32* edit_dev = nullptr;                // FCM_MAXEDIT * 2 (slot, add)
  u64* io_dev = nullptr;                  // pre/post exchange: 2*FCM_MAXD + 4
  // BitOutput (src/io.rs:128-212), built on first use
  std::vector<u32> eb_rank2pair;          // rank in (big, small) order -> pair id (host copy)
  u32* eb_rank2pair_dev = nullptr;
  u32* eb_out_dev = nullptr;              // packed samples, eb_out_cap chains x eb_words words
  u32 eb_out_cap = 0;
};

// fcm_chain.cu
size_t fcm_chain_smem_per_warp(u32 Kcap, u32 KWcap, bool cliques);
int fcm_chain_launch_step(fcm_ensemble* e, u64 nsteps, fcm_trace_rec* trace_dev, u32 trace_chain);
int fcm_chain_launch_apply(fcm_ensemble* e, u32 chain, int nedits, int revert);
int fcm_chain_launch_replicate(fcm_ensemble* e);
