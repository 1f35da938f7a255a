/* fcm.h — C ABI of libfcm.so: the B200-native replacement of flag-complex-mcmc's data-parallel
 * hot path (directed-flag-complex simplex counting + per-move local delta + bounds test).
 *
 * Every entry point replaces one reference interface; the reference file:line is cited beside it
 * (paths relative to the reference repository, TheJonny/flag-complex-mcmc).  The Rust side binds
 * these with a ~60-line `extern "C"` block (INTEGRATION.md shows it); nothing here mentions torch.
 *
 * Conventions
 *   - plain pointers + sizes; Node = uint32_t, Edge = uint32_t[2] = (from, to)   src/flagser.rs:9
 *   - counts are uint64_t (Rust usize on 64-bit)                                src/lib.rs:34
 *   - every function returns 0 on success, a negative FCM_E* code otherwise; fcm_last_error()
 *     gives the message (thread-local).  The reference panics/aborts in the same situations
 *     (src/lib.rs:65,86,104,309); a Rust shim turns non-zero into panic!().
 *   - no CPU fallback: every compute entry point fails with FCM_ENODEVICE when no CUDA device
 *     is usable.
 *   - process-global side effects, all on the device of the context: fcm_ctx_create raises the release threshold of the
 *     device's default memory pool (memory of destroyed graphs is kept for the next one: flagser_count is called over and
 *     over by the research binaries); fcm_ensemble_create raises cudaLimitPersistingL2CacheSize to cover the clique index
 *     (the window itself is a per-launch attribute, no stream is modified; fcm_ensemble_destroy resets the persisting
 *     lines); loading the library reserves ~4.1 MB per GPU for fcm_graph_count_noalloc's static scratch.
 *   - a handle (ctx / graph / ensemble) is owned by one host thread at a time (the reference's
 *     `&mut self`, src/lib.rs:61,81,181); different handles may be used from different threads.
 */
#ifndef FCM_H
#define FCM_H
#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define FCM_MAXD 32            /* longest flag_count vector handled (dimension <= 31) */
#define FCM_ABI_VERSION 2

enum {
  FCM_OK = 0,
  FCM_EINVAL = -1,             /* bad argument (null pointer, vertex id >= n, self loop ...) */
  FCM_ENODEVICE = -2,          /* no usable CUDA device / driver */
  FCM_ECUDA = -3,              /* a CUDA call failed; message holds cudaGetErrorString */
  FCM_ENOMEM = -4,
  FCM_EDIM = -5,               /* a simplex of dimension >= FCM_MAXD was met */
  FCM_ENOTADJ = -6,            /* an edit names a pair that is not adjacent in the underlying
                                  undirected graph (reference: HashMap index panic, src/lib.rs:104) */
  FCM_EASSERT = -7,            /* reference assert!(s >= p) would fail (src/lib.rs:65,86) */
  FCM_EUNSUPPORTED = -8,       /* a clique move was drawn on an ensemble created without clique moves */
  FCM_ELIMIT = -9              /* neighbourhood larger than the kernel's local-matrix capacity, or a clique move
                                  beyond the limits stated at fcm_ensemble_create */
};

typedef struct fcm_ctx fcm_ctx;
typedef struct fcm_graph fcm_graph;
typedef struct fcm_ensemble fcm_ensemble;

const char* fcm_last_error(void);
int fcm_abi_version(void);
/* number of CUDA devices visible (0 when none; never fails) */
int fcm_device_count(void);

/* ------------------------------------------------------------------------------------------------
 * 1. Stateless count — signature and ownership of the reference's C ABI precedent
 *    `flagser_count_unweighted` (src/flagser.rs:7-21): callee mallocs the result, caller copies
 *    and free()s it (fcm_free == free).  Returns NULL on error.  Uses device 0.
 *    Replaces `graph.flagser_count()`  (src/lib.rs:51,130) / `flag_complex::count_cells`
 *    (src/bin/all_cxs.rs:123).  Result has no trailing zeros; result[0] = nvertices.
 * ---------------------------------------------------------------------------------------------- */
size_t* fcm_count_unweighted(size_t nvertices, size_t nedges, const uint32_t (*edges)[2],
                             size_t* res_size);
void fcm_free(void* p);

/* ------------------------------------------------------------------------------------------------
 * 2. Context = one CUDA device + one stream.  device = ordinal (>= 0).
 *    fcm_ctx_set_stream lets the host run the kernels on its own stream (cudaStream_t passed as
 *    void*; NULL restores the context's private stream).
 * ---------------------------------------------------------------------------------------------- */
int fcm_ctx_create(int device, fcm_ctx** out);
int fcm_ctx_destroy(fcm_ctx* ctx);
int fcm_ctx_set_stream(fcm_ctx* ctx, void* cuda_stream);
int fcm_ctx_synchronize(fcm_ctx* ctx);
/* device-side timer on the context's stream (CUDA events): begin ... end -> milliseconds */
int fcm_ctx_timer_begin(fcm_ctx* ctx);
int fcm_ctx_timer_end(fcm_ctx* ctx, float* ms_out);
/* number of kernels this library has launched on this context since creation */
uint64_t fcm_ctx_launches(const fcm_ctx* ctx);

/* ------------------------------------------------------------------------------------------------
 * 3. Device-resident graph for (sharded) full recounts: State::new's `graph.flagser_count()`
 *    (src/lib.rs:51) and Bounds::calculate's count of the normalised graph (src/lib.rs:125-130).
 *    Adjacency is kept in HBM as packed 64-bit out-neighbour rows (row stride padded to 16 B).
 *    Graphs of more than 4096 vertices are stored under an internal numbering of the vertices (a locality
 *    order computed on the device at create: neighbours get nearby ids, which is what the count's gather
 *    wants).  Counts do not depend on it, fcm_graph_cells reports the caller's ids; which source vertices a
 *    shard (part, nparts) holds is the library's choice - only the sum over the parts is defined.
 *    Environment: FCM_RELABEL=0 keeps the caller's ids, FCM_RELABEL=1 renumbers every graph (tests).
 * ---------------------------------------------------------------------------------------------- */
int fcm_graph_create(fcm_ctx* ctx, uint32_t n, uint64_t nedges, const uint32_t (*edges)[2],
                     fcm_graph** out);
/* graph of the normalised orientation (one edge big->small per adjacent pair, src/lib.rs:126-129) */
int fcm_graph_create_normalised(fcm_ctx* ctx, uint32_t n, uint64_t nedges,
                                const uint32_t (*edges)[2], fcm_graph** out);
int fcm_graph_destroy(fcm_graph* g);
/* Count the simplices whose source (first) vertex v0 is in shard `part` of `nparts`
 * (degree-balanced round-robin over source vertices; part 0 of 1 = whole graph).
 * counts[FCM_MAXD] is filled (zero padded); *len = top non-empty dimension + 1 OF THIS SHARD.
 * Summing the shards' vectors (fcm_graph_count_allreduce below, or any all-reduce of FCM_MAXD x u64 of the
 * caller's own) gives flagser_count.  Synchronous. */
int fcm_graph_count(fcm_graph* g, uint32_t part, uint32_t nparts, uint64_t* counts, int* len);
/* Same, asynchronous on the context stream: result stays in a device buffer of FCM_MAXD u64
 * (pointer returned through counts_dev) so that a collective can consume it without a host hop. */
int fcm_graph_count_async(fcm_graph* g, uint32_t part, uint32_t nparts, uint64_t** counts_dev);
/* Synchronise and report what the last asynchronous count on this graph met: FCM_OK, FCM_EDIM (a simplex of
 * dimension >= FCM_MAXD: the counts are truncated) or FCM_ECUDA (all-reduce timed out / CUDA error). */
int fcm_graph_status(fcm_graph* g);

/* ------------------------------------------------------------------------------------------------
 * 3b. The sharded recount's all-reduce, inside the library: State::new's `graph.flagser_count()`
 *     (src/lib.rs:51, and :130) split by source vertex over the GPUs of one box.  One fcm_comm per rank
 *     (= per GPU); the ranks may live in one process or in one process each.
 *       FCM_COMM_PEER  every rank owns a small window in its HBM that all ranks map (cudaIpc across
 *                      processes, peer access inside one).  The LAST CTA of the count kernel adds the
 *                      rank's u64 vector into every rank's window with system-scope atomics over
 *                      NVLink/NVSwitch and waits for the other ranks' contributions: count and all-reduce
 *                      are ONE kernel launch, no host hop, no extra launch.
 *       FCM_COMM_NCCL  ncclAllReduce(sum, uint64, FCM_MAXD + 1) enqueued on the context stream right behind
 *                      the count kernel (libnccl.so.2 is dlopen'ed; libfcm.so has no link-time NCCL need).
 *     Setup, both kinds: every rank calls fcm_comm_create and gets a FCM_COMM_HANDLE_BYTES blob; the host
 *     exchanges the blobs by any means (MPI, torch.distributed all_gather, a file; memcpy when the ranks
 *     share a process) and gives all nranks of them, in rank order, to fcm_comm_connect.  With
 *     FCM_COMM_NCCL fcm_comm_connect is ncclCommInitRank, i.e. itself a collective: every rank must be
 *     inside it at the same time (one thread per rank when they share a process).
 *     Like any collective, every rank must issue the same sequence of all-reduce calls on a comm; a rank
 *     that never arrives makes the others give up after FCM_COMM_TIMEOUT_MS (default 10000, environment)
 *     with FCM_ECUDA, after which the comm must be destroyed.
 * ---------------------------------------------------------------------------------------------- */
typedef struct fcm_comm fcm_comm;
#define FCM_COMM_HANDLE_BYTES 256
#define FCM_COMM_MAXRANKS 16
enum { FCM_COMM_PEER = 0, FCM_COMM_NCCL = 1 };
int fcm_comm_create(fcm_ctx* ctx, int kind, uint32_t rank, uint32_t nranks, fcm_comm** out,
                    uint8_t handle_out[FCM_COMM_HANDLE_BYTES]);
int fcm_comm_connect(fcm_comm* c, const uint8_t* handles /* nranks x FCM_COMM_HANDLE_BYTES, rank order */);
int fcm_comm_destroy(fcm_comm* c);
/* Count shard `rank of nranks` of g (g and c on the same context) and all-reduce: afterwards EVERY rank holds
 * the whole graph's flagser_count.  Asynchronous on the context stream; *counts_dev = device buffer of
 * FCM_MAXD + 1 u64 (entry FCM_MAXD: number of ranks that met a simplex of dimension >= FCM_MAXD). */
int fcm_graph_count_allreduce_async(fcm_graph* g, fcm_comm* c, uint64_t** counts_dev);
/* Same, synchronous: counts[FCM_MAXD] zero padded, *len = top non-empty dimension + 1 of the WHOLE graph. */
int fcm_graph_count_allreduce(fcm_graph* g, fcm_comm* c, uint64_t* counts, int* len);

/* fcm_graph_count without any allocation.  NO COUNTERPART IN THE REFERENCE: added
 * for callers that cannot call cudaMalloc (code under CUDA graph capture, tight serving loops, harnesses that
 * forbid it).  Same arguments and result as fcm_graph_count (the dense-row kernel counts); the kernel's scratch - its
 * candidate-set stacks, ticket, error flag and count vector - lives in one file-scope device array of about 4.1 MB that
 * every GPU loading libfcm.so holds until exit.  The call uses no cudaMalloc / cudaFree / memory-pool routine.
 * Limit: adjacency rows of at most FCM_NOALLOC_MAXW 64-bit words (n <= 4096 vertices); beyond that FCM_ELIMIT.
 * All calls on a device share the array: the call clears what must start at zero, and calls that overlap on the GPU
 * would race (within one process the library serialises them).  Synchronous. */
#define FCM_NOALLOC_MAXW 64
int fcm_graph_count_noalloc(fcm_graph* g, uint32_t part, uint32_t nparts, uint64_t* counts, int* len);

/* flag_complex::for_each_cell(&G, &mut F, min, max) (src/bin/all_cxs.rs:60-64,79-83) as a bulk call:
 * every cell of dimension dim_min..dim_max (dim_max <= 6) as one row of dim_max + 1 vertex ids
 * (v0..vd, padded with 0xffffffff), rows in the order the recursion meets them with candidates in
 * ascending order.  *ncells = number of such cells; rows are written only when *ncells <= cap
 * (call with cap = 0 first to size the buffer).  Synchronous. */
int fcm_graph_cells(fcm_graph* g, int dim_min, int dim_max, uint32_t* cells_out, uint64_t cap,
                    uint64_t* ncells);
/* which kernel to use: 0 = auto (block-local relabelled kernel, dense-row kernel for sources whose
 * out-degree exceeds the shared-memory tile), 1 = force dense-row kernel (the roofline model's
 * "one adjacency row per simplex" kernel).  Results are identical. */
int fcm_graph_set_kernel(fcm_graph* g, int kind);
/* bytes of adjacency rows resident in HBM, words per row */
int fcm_graph_info(const fcm_graph* g, uint32_t* n, uint64_t* nedges, uint32_t* words_per_row,
                   uint32_t* max_out_degree);

/* State::compute_edge_neighborhoods (src/lib.rs:331-356; printed by src/bin/edgeset_nbhd.rs): for every
 * adjacent pair, key [big, small], the ascending common undirected neighbours - computed on the
 * device (bit-row AND + compaction).  Pairs come in order of first appearance in the edge list.
 * As in the reference, whose adjacency lists are built from the directed edges without deduplication
 * (src/lib.rs:333-336), a neighbour joined to BOTH endpoints by double edges is listed twice.
 * Call with the three buffers NULL to get *npairs and *list_len, then with pairs_out[2*npairs],
 * off_out[npairs+1] (exclusive prefix), list_out[list_len]. */
int fcm_edge_neighborhoods(fcm_ctx* ctx, uint32_t n, uint64_t nedges, const uint32_t (*edges)[2],
                           uint64_t* npairs, uint64_t* list_len, uint32_t* pairs_out,
                           uint64_t* off_out, uint32_t* list_out);

/* Graph::compute_maximal_cliques of the external flag-complex crate (called by State::new, src/lib.rs:41-49, which
 * files the cliques by order into State::cliques_by_order): the maximal cliques of the underlying undirected graph,
 * each with its vertices ascending, the cliques in lexicographic order (the order the oracle assumes; parity
 * unpinned).  HOST code - it is host code in the reference too, fcm_ensemble_create runs the same
 * enumeration on host threads and needs no device for it; this entry point exposes that step so that a port of
 * State::new can fill cliques_by_order, and so that it is tested without a GPU.  Call with sizes_out = verts_out = NULL
 * to get *ncliques and *nverts (= sum of the orders), then with sizes_out[ncliques], verts_out[nverts].
 * FCM_EDIM for a clique of more than FCM_MAXD vertices (the chain kernels' limit). */
int fcm_maximal_cliques(uint32_t n, uint64_t nedges, const uint32_t (*edges)[2], uint64_t* ncliques, uint64_t* nverts,
                        uint32_t* sizes_out, uint32_t* verts_out);

/* ------------------------------------------------------------------------------------------------
 * 4. Ensemble of independent chains on one device: MCMCSampler (src/lib.rs:164-198), one warp per
 *    chain.  All chains start from the same graph (edge order = insertion order of the edge list,
 *    as read_flag_file inserts them, src/io.rs:26-33) and share bounds / move mix; chain k uses
 *    Xoshiro256StarStar::seed_from_u64(seeds[k]) (src/bin/sample.rs:86).
 *    flag_count0/len0 = the initial flagser_count of the graph (State::new, src/lib.rs:51).
 *    move_weights[4] = MOVE_DISTRIBUTION (src/bin/sample.rs:16-17): flip, double-edge move,
 *    clique permute, clique swap (src/lib.rs:214-290).  When one of the last two is non-zero the
 *    maximal cliques of the underlying undirected graph are enumerated at creation (State::new,
 *    src/lib.rs:41-49; order handed to `choose`: every clique ascending, cliques lexicographic) and
 *    clique_order_distribution = WeightedIndex(len^0.2) (src/bin/sample.rs:87-88) is built.
 *    Trace records of clique moves: a, b = index of the chosen clique(s) within their order,
 *    dm0 = order index, dm1 = number of change_edges, K = |edgeset_neighborhood|.
 *    Limits of a clique move: the cliques of a swap hold <= 32 vertices together and a transition
 *    <= 128 change_edges; beyond that the chain stops with the sticky error FCM_ELIMIT.
 * ---------------------------------------------------------------------------------------------- */
int fcm_ensemble_create(fcm_ctx* ctx, uint32_t n, uint64_t nedges, const uint32_t (*edges)[2],
                        uint32_t nchains, const uint64_t* seeds,
                        const uint64_t* flag_count0, int len0,
                        const uint64_t* bounds_min, int nmin, const uint64_t* bounds_max, int nmax,
                        const double move_weights[4], fcm_ensemble** out);
int fcm_ensemble_destroy(fcm_ensemble* e);
int fcm_ensemble_set_bounds(fcm_ensemble* e, const uint64_t* bounds_min, int nmin,
                            const uint64_t* bounds_max, int nmax);

/* how a proposal's count change is computed: 0 = auto (delta path: only the simplices through the
 * changed edges are enumerated; the hot loop looks the orientation of the pair's triangles and
 * 4-cliques up in a static clique index when the graph is small enough for it, else gathers them
 * from the adjacency rows; falls back per proposal when a common-neighbour list exceeds 64),
 * 1 = always the full pre/post counts of the neighbourhood graph as the reference does them
 * (src/lib.rs:63,71), 2 = delta path gathered from the adjacency rows even when the index exists.
 * flag_count, accept/reject and the trace are identical. */
int fcm_ensemble_set_kernel(fcm_ensemble* e, int kind);

/* The edge-sampler policy, i.e. which element of the insertion-ordered edge container Graph::sample_edge /
 * sample_double_edge (external flag-complex crate; src/lib.rs:293,306,309) return for a draw r = gen_range(0..len):
 * 0 (default) = element r, the documented assumption of oracle/fcm_oracle.cpp; 1 = element len-1-r.  Policy 1 is not a
 * claim about the crate: the crate's source is not available here (parity unpinned), and this switch is the single place
 * where its real rule goes once known - the tests show the CUDA path and the oracle change together under it. */
int fcm_ensemble_set_policy(fcm_ensemble* e, int policy);

/* Instrumentation switch.  on (default): every non-empty proposal adds its neighbourhood size
 * |State::edgeset_neighborhood| (src/lib.rs:99-111) to the statistic fcm_ensemble_read_all returns
 * in stats[2].  off: the statistic stands still (a double-edge move then skips the size of the union
 * of its two neighbourhoods, which nothing else needs); flag_count, sampled, accepted, the RNG and
 * the edge containers are unaffected.  Traced steps always carry K. */
int fcm_ensemble_set_stats(fcm_ensemble* e, int on);
/* whether the static clique index was built for this graph and its size in HBM (shared by all chains) */
int fcm_ensemble_index_info(const fcm_ensemble* e, int* have_index, uint64_t* index_bytes);

/* one record per proposal (optional trace; identical layout in oracle/fcm_oracle.cpp) */
typedef struct fcm_trace_rec {
  uint32_t flags;              /* bits 0-1 move kind, bit 4 accepted, bit 5 empty transition */
  uint32_t a, b;               /* flip: sampled edge (from,to); double move: the single edge [a,b] */
  uint32_t dm0, dm1;           /* double move: the deleted direction `delme` */
  uint32_t K;                  /* |neighbourhood| (0 for an empty transition) */
  uint32_t draws, pad;
} fcm_trace_rec;
#define FCM_TR_KIND_MASK 3u
#define FCM_TR_ACCEPT (1u << 4)
#define FCM_TR_EMPTY (1u << 5)

/* `nproposals` iterations of the loop body of MCMCSampler::next (src/lib.rs:182-192) on every
 * chain (next() itself = nproposals = sample_distance).  Asynchronous on the context stream. */
int fcm_ensemble_step(fcm_ensemble* e, uint64_t nproposals);
/* same, synchronous, writing chain `chain`'s per-proposal trace to trace_out[nproposals] (host) */
int fcm_ensemble_step_trace(fcm_ensemble* e, uint64_t nproposals, uint32_t chain,
                            fcm_trace_rec* trace_out);
/* Synchronise and read one chain.  Any output pointer may be NULL.  flag_count_out[FCM_MAXD];
 * edges_out[2*nedges] in container order (State.graph.edges()); rng_out[4] = xoshiro state. */
int fcm_ensemble_read(fcm_ensemble* e, uint32_t chain, uint64_t* flag_count_out, int* len_out,
                      uint64_t* sampled_out, uint64_t* accepted_out, uint32_t* edges_out,
                      uint64_t* nedges_out, uint64_t* ndouble_out, uint64_t* rng_out);
/* Synchronise and read all chains at once: flag_counts[nchains*FCM_MAXD], lens[nchains],
 * sampled[nchains], accepted[nchains]; any may be NULL.  stats[2] (may be NULL): [0] = total
 * neighbourhood size K over all non-empty proposals of all chains since creation, [1] = number of
 * non-empty proposals (both feed the algorithmic-bytes model of DESIGN.md). */
int fcm_ensemble_read_all(fcm_ensemble* e, uint64_t* flag_counts, int* lens, uint64_t* sampled,
                          uint64_t* accepted, uint64_t* stats);
/* sticky device-side error of the ensemble (FCM_OK, FCM_EDIM, FCM_EASSERT, FCM_ELIMIT) */
int fcm_ensemble_status(fcm_ensemble* e);
int fcm_ensemble_info(const fcm_ensemble* e, uint32_t* n, uint64_t* npairs, uint32_t* nchains,
                      uint32_t* max_common_nbhd, uint64_t* bytes_per_chain);

/* State::apply_transition / revert_transition (src/lib.rs:61-95) on one chain with explicit
 * edits (a, b, add) x nedits; pre/post[FCM_MAXD] are the API-visible local counts (src/bin/
 * seo_bt_flip_only_once.rs:51-74 reads post[2]-pre[2]).  Synchronous. */
int fcm_apply(fcm_ensemble* e, uint32_t chain, const uint32_t* edits3, int nedits,
              uint64_t* pre, int* prelen, uint64_t* post, int* postlen);
int fcm_revert(fcm_ensemble* e, uint32_t chain, const uint32_t* edits3, int nedits,
               const uint64_t* pre, int prelen, const uint64_t* post, int postlen);
/* Bounds::check (src/lib.rs:157-160) of one chain's current state: *ok = 0/1 */
int fcm_bounds_check(fcm_ensemble* e, uint32_t chain, int* ok);

/* ------------------------------------------------------------------------------------------------
 * 5. io::BitOutput (src/io.rs:128-212): the `.edgebits` sample format, packed on the device.
 *    Slots = both directions of every adjacent pair, sorted by (max(a,b), min(a,b), a < b)
 *    (src/io.rs:152-159); a sample = one bit per slot (set iff the chain's graph has that directed
 *    edge), LSB-first in bytes, ceil(nslots/8) bytes (src/io.rs:180-194).
 * ---------------------------------------------------------------------------------------------- */
int fcm_ensemble_edgebits_info(const fcm_ensemble* e, uint64_t* nslots, uint64_t* nbytes);
/* BitOutput.edges: slots_out[2*nslots] = (a, b) of every slot in file order */
int fcm_ensemble_edgebits_slots(fcm_ensemble* e, uint32_t* slots_out);
/* BitOutput::save for chains [chain_first, chain_first + nchains): synchronises, packs on the device
 * and writes nchains rows of nbytes bytes to out (host). */
int fcm_ensemble_edgebits(fcm_ensemble* e, uint32_t chain_first, uint32_t nchains, uint8_t* out);

/* ------------------------------------------------------------------------------------------------
 * 6. Checkpoint / resume of a chain: the device analogue of io::save_state / load_state
 *    (src/io.rs:51-62).  A chain is its edge container in order (fcm_ensemble_read: edges_out), its
 *    double-edge container in order (fcm_ensemble_read_doubles: [min,max] pairs), the RNG state, the
 *    counters and flag_count.  fcm_ensemble_restore puts such a state into chain `chain` of an
 *    ensemble created from a graph with the same underlying undirected graph; the chain then
 *    continues exactly as the checkpointed one would have.
 * ---------------------------------------------------------------------------------------------- */
int fcm_ensemble_read_doubles(fcm_ensemble* e, uint32_t chain, uint32_t* pairs_out, uint64_t* ndouble_out);
int fcm_ensemble_restore(fcm_ensemble* e, uint32_t chain, const uint32_t* edges, uint64_t nedges,
                         const uint32_t* doubles, uint64_t ndouble, const uint64_t* rng4,
                         const uint64_t* flag_count, int len, uint64_t sampled, uint64_t accepted);

#ifdef __cplusplus
}
#endif
#endif /* FCM_H */
