"""Host-side mirror of the reference crate's interface for the hot path (crate `directed-scm`,
src/lib.rs) over libfcm.so.  Names, argument meaning and error behaviour follow the reference:

  read_flag_file                      src/io.rs:18-35
  flagser_count / count_cells         external flag-complex crate; call sites src/lib.rs:51,130
  Bounds.calculate / Bounds.check     src/lib.rs:119-160
  util: all_le, factorial (the (x-1)! quirk), binomial, calc_relax_de, OEIS_A058298   src/util.rs:53-105
  initialize_new_sampler              src/bin/sample.rs:80-105
  MCMCSampler.next / acceptance_ratio src/lib.rs:181-197   (here: an ensemble of chains, one per seed)
  State.apply_transition / revert_transition   src/lib.rs:61-95

All counting and all chain steps run in CUDA (libfcm.so); nothing here falls back to the CPU.
"""
import ctypes as C
import sys
import math

import numpy as np

from . import _native as N
from ._native import FCM_MAXD, FcmError, check, edges_array, lib, ptr

U64 = (1 << 64) - 1
USIZE_MAX = U64

MOVE_DISTRIBUTION = [0.1, 0.1, 0.6, 0.2]          # src/bin/sample.rs:16
MOVE_DISTRIBUTION_SIMPLE = [0.5, 0.5, 0.0, 0.0]   # src/bin/sample.rs:17

OEIS_A058298 = [2, 3, 6, 8, 12, 24, 30, 40, 60, 120, 144, 180, 240, 360, 720,
                840, 1008, 1260, 1680, 2520, 5040, 5760, 6720, 8064, 10080,
                13440, 20160, 40320, 45360, 51840, 60480, 72576, 90720,
                120960, 181440, 362880, 403200, 453600, 518400, 604800,
                725760, 907200, 1209600, 1814400, 3628800, 3991680, 4435200, 4989600,
                5702400, 6652800, 7983360, 9979200, 13305600, 19958400, 39916800,
                43545600, 47900160, 53222400, 59875200, 68428800,
                79833600, 95800320, 119750400, 159667200]     # src/util.rs:98-105


def f64_as_usize(x):
    """Rust `f64 as usize` (src/bin/sample.rs:90-91): saturating, NaN -> 0."""
    if not (x > 0.0):
        return 0
    if x >= 18446744073709551616.0:
        return U64
    return int(x)


# ---- util.rs -------------------------------------------------------------------------------------
def all_le(a, b, z=0):
    """src/util.rs:53-63: element-wise a <= b with the shorter side padded by z."""
    m = max(len(a), len(b))
    for i in range(m):
        l = a[i] if i < len(a) else z
        r = b[i] if i < len(b) else z
        if l > r:
            return False
    return True


def factorial(x):
    """src/util.rs:65-71: `for i in 1..x` => (x-1)!  — the quirk is part of the bounds; usize wraps."""
    res = 1
    for i in range(1, x):
        res = (res * i) & U64
    return res


def binomial(n, k):
    """src/util.rs:73-77 on the quirky factorial, integer division (panics on n < k / zero divisor)."""
    if k > n:
        raise OverflowError("binomial: n - k underflows (reference panics)")
    den = (factorial(k) * factorial(n - k)) & U64
    if den == 0:
        raise ZeroDivisionError("binomial: divisor wrapped to zero (reference panics)")
    return factorial(n) // den


def calc_relax_de(sc):
    """src/util.rs:79-93."""
    out = []
    for d in range(len(sc)):
        ind = 1
        lost = []
        while True:
            if ind >= len(OEIS_A058298):
                raise IndexError("calc_relax_de: OEIS_A058298 index out of range (reference panics)")
            if not (OEIS_A058298[ind] < sc[d]):
                break
            lost.append(OEIS_A058298[ind] - OEIS_A058298[ind - 1])
            ind += 1
        a = max(lost) if lost else 1
        out.append(min(a, factorial(d + 1)))
    return out


# ---- io.rs ---------------------------------------------------------------------------------------
def _parse_node(tok):
    """Rust `u32::from_str` (src/io.rs:31 `i.parse().unwrap()`): optional '+', then decimal digits only."""
    t = tok[1:] if tok.startswith("+") else tok
    if not t or not all("0" <= ch <= "9" for ch in t) or int(t) > 0xFFFFFFFF:
        raise ValueError("not a vertex id: %r" % tok)
    return int(t)


def read_flag_file(fname):
    """src/io.rs:18-35 -> (n, edges uint32 (E,2) in file order).  Lines as str::lines() yields them, tokens split
    on ' ' only; a token that is not a u32 raises ValueError where the reference panics."""
    with open(fname, "rb") as f:
        text = f.read().decode("utf-8")
    lines = text.split("\n")
    if lines and lines[-1] == "":
        lines.pop()                                                  # str::lines(): no empty last line after a final '\n'
    lines = [ln[:-1] if ln.endswith("\r") else ln for ln in lines]
    n = len([t for t in lines[1].split(" ") if t != ""])
    edges = []
    for ln in lines[3:]:
        t = [x for x in ln.split(" ") if x != ""]
        if len(t) >= 2:
            edges.append((_parse_node(t[0]), _parse_node(t[1])))
    return n, np.array(edges, dtype=np.uint32).reshape(-1, 2)


# ---- device plumbing -------------------------------------------------------------------------------
def save_flag_file(fname, n, edges):
    """io::save_flag_file (src/io.rs:37-49): "dim 0:", n ones, "dim 1:", the edges sorted, weight 1."""
    e = np.asarray(edges, dtype=np.uint32).reshape(-1, 2)
    e = e[np.lexsort((e[:, 1], e[:, 0]))]
    with open(fname, "w") as f:
        f.write("dim 0:\n" + " ".join(["1"] * int(n)) + "\ndim 1:\n")
        f.write("".join("%d %d 1\n" % (a, b) for a, b in e))


class Context:
    """One CUDA device + stream (fcm_ctx)."""

    def __init__(self, device=0, stream=None):
        self.h = C.c_void_p()
        check(lib().fcm_ctx_create(int(device), C.byref(self.h)))
        self.device = int(device)
        if stream is not None:
            self.set_stream(stream)

    def set_stream(self, stream):
        check(lib().fcm_ctx_set_stream(self.h, C.c_void_p(int(stream) if stream else 0)))

    def synchronize(self):
        check(lib().fcm_ctx_synchronize(self.h))

    def timer_begin(self):
        check(lib().fcm_ctx_timer_begin(self.h))

    def timer_end(self):
        ms = C.c_float()
        check(lib().fcm_ctx_timer_end(self.h, C.byref(ms)))
        return ms.value

    @property
    def launches(self):
        return int(lib().fcm_ctx_launches(self.h))

    def close(self):
        if self.h:
            lib().fcm_ctx_destroy(self.h)
            self.h = C.c_void_p()

    def __del__(self):
        try:
            if not sys.is_finalizing():                 # at interpreter exit the CUDA runtime may be gone already
                self.close()
        except Exception:
            pass


class DeviceGraph:
    """Adjacency resident in HBM as packed 64-bit rows (fcm_graph)."""

    def __init__(self, ctx, n, edges, normalised=False):
        self.ctx = ctx
        e = edges_array(edges)
        self.h = C.c_void_p()
        f = lib().fcm_graph_create_normalised if normalised else lib().fcm_graph_create
        check(f(ctx.h, int(n), int(e.shape[0]), ptr(e), C.byref(self.h)))
        n_, ne, w, mo = C.c_uint32(), C.c_uint64(), C.c_uint32(), C.c_uint32()
        check(lib().fcm_graph_info(self.h, C.byref(n_), C.byref(ne), C.byref(w), C.byref(mo)))
        self.n, self.nedges, self.words_per_row, self.max_out_degree = n_.value, ne.value, w.value, mo.value

    def set_kernel(self, kind):
        check(lib().fcm_graph_set_kernel(self.h, int(kind)))

    def count_shard(self, part=0, nparts=1):
        """uint64[FCM_MAXD] of the simplices whose source vertex lies in the shard (zero padded)."""
        out = np.zeros(FCM_MAXD, dtype=np.uint64)
        ln = C.c_int()
        check(lib().fcm_graph_count(self.h, int(part), int(nparts), ptr(out), C.byref(ln)))
        return out

    def count_async(self, part=0, nparts=1):
        """Enqueue the count; returns the device address of the uint64[FCM_MAXD] result."""
        p = C.c_void_p()
        check(lib().fcm_graph_count_async(self.h, int(part), int(nparts), C.byref(p)))
        return p.value

    def flagser_count(self):
        """`graph.flagser_count()`: list without trailing zeros."""
        return trim(self.count_shard(0, 1))

    def count_allreduce(self, comm):
        """Shard `comm.rank of comm.nranks` counted and summed over the ranks inside the library
        (fused last-CTA epilogue over NVLink, or NCCL): every rank gets the whole graph's uint64[FCM_MAXD]."""
        out = np.zeros(FCM_MAXD, dtype=np.uint64)
        ln = C.c_int()
        check(lib().fcm_graph_count_allreduce(self.h, comm.h, ptr(out), C.byref(ln)))
        return out

    def count_allreduce_async(self, comm):
        """Enqueue count + all-reduce on the context stream; returns the device address of uint64[FCM_MAXD + 1]."""
        p = C.c_void_p()
        check(lib().fcm_graph_count_allreduce_async(self.h, comm.h, C.byref(p)))
        return p.value

    def status(self):
        check(lib().fcm_graph_status(self.h))

    def cells(self, dim_min, dim_max):
        """All cells of dimension dim_min..dim_max: (ncells, dim_max + 1) uint32 rows (v0..vd, padded
        with 0xffffffff) in the recursion's order (flag_complex::for_each_cell)."""
        cnt = C.c_uint64()
        check(lib().fcm_graph_cells(self.h, int(dim_min), int(dim_max), None, 0, C.byref(cnt)))
        out = np.zeros((cnt.value, dim_max + 1), dtype=np.uint32)
        if cnt.value:
            check(lib().fcm_graph_cells(self.h, int(dim_min), int(dim_max), ptr(out), cnt.value, C.byref(cnt)))
        return out

    def close(self):
        if self.h:
            lib().fcm_graph_destroy(self.h)
            self.h = C.c_void_p()

    def __del__(self):
        try:
            if not sys.is_finalizing():                 # at interpreter exit the CUDA runtime may be gone already
                self.close()
        except Exception:
            pass


COMM_PEER, COMM_NCCL = 0, 1
COMM_HANDLE_BYTES = 256


class Comm:
    """One rank of the recount's all-reduce (fcm_comm).  `exchange(blob) -> list of all ranks' blobs in rank
    order` is the host's transport (torch.distributed all_gather, MPI, ...); ranks sharing a process pass
    `Comm.connect_local([...])` instead."""

    def __init__(self, ctx, rank, nranks, kind=COMM_PEER, exchange=None):
        self.ctx, self.rank, self.nranks, self.kind = ctx, int(rank), int(nranks), int(kind)
        self.h = C.c_void_p()
        self.handle = np.zeros(COMM_HANDLE_BYTES, dtype=np.uint8)
        check(lib().fcm_comm_create(ctx.h, self.kind, self.rank, self.nranks, C.byref(self.h), ptr(self.handle)))
        if exchange is not None:
            self.connect(exchange(self.handle))

    def connect(self, handles):
        allh = np.ascontiguousarray(np.concatenate([np.asarray(h, dtype=np.uint8).reshape(-1) for h in handles]))
        assert allh.size == self.nranks * COMM_HANDLE_BYTES, "need one handle per rank"
        check(lib().fcm_comm_connect(self.h, ptr(allh)))

    @staticmethod
    def connect_local(comms):
        """Ranks living in this process (FCM_COMM_PEER): hand every comm all the handles."""
        hs = [c.handle for c in sorted(comms, key=lambda c: c.rank)]
        for c in comms:
            c.connect(hs)

    def close(self):
        if self.h:
            lib().fcm_comm_destroy(self.h)
            self.h = C.c_void_p()

    def __del__(self):
        try:
            if not sys.is_finalizing():                 # at interpreter exit the CUDA runtime may be gone already
                self.close()
        except Exception:
            pass


def trim(v):
    v = [int(x) for x in v]
    while v and v[-1] == 0:
        v.pop()
    return v


def flagser_count(n, edges, ctx=None):
    """Stateless count.  Without a context this is fcm_count_unweighted, the drop-in for the
    reference's `flagser_count_unweighted` C ABI (src/flagser.rs:7-21)."""
    e = edges_array(edges)
    if ctx is not None:
        g = DeviceGraph(ctx, n, e)
        try:
            return g.flagser_count()
        finally:
            g.close()
    sz = C.c_size_t()
    res = lib().fcm_count_unweighted(int(n), int(e.shape[0]), ptr(e), C.byref(sz))
    if not res:
        raise FcmError(-1, lib().fcm_last_error().decode())
    try:
        arr = (C.c_size_t * sz.value).from_address(res)
        return [int(x) for x in arr]
    finally:
        lib().fcm_free(res)


count_cells = flagser_count      # flag_complex::count_cells (src/bin/all_cxs.rs:123)


def for_each_cell(n, edges, f, dim_min, dim_max, ctx=None):
    """flag_complex::for_each_cell(&G, &mut F, min, max) (src/bin/all_cxs.rs:60-64): F(cell) for every
    cell of dimension dim_min..dim_max, cell = list of its vertices (v0..vd); enumerated on the GPU."""
    own = ctx is None
    ctx = Context(0) if own else ctx
    g = DeviceGraph(ctx, n, edges)
    try:
        for row in g.cells(dim_min, dim_max):
            f([int(x) for x in row if x != 0xFFFFFFFF])
    finally:
        g.close()
        if own:
            ctx.close()


def compute_edge_neighborhoods(ctx, n, edges):
    """State::compute_edge_neighborhoods (src/lib.rs:331-356) on the GPU -> {(big, small): [common
    undirected neighbours, ascending]} in order of first appearance of the pair."""
    e = edges_array(edges)
    npairs, ll = C.c_uint64(), C.c_uint64()
    check(lib().fcm_edge_neighborhoods(ctx.h, int(n), int(e.shape[0]), ptr(e), C.byref(npairs), C.byref(ll), None, None, None))
    pairs = np.zeros((npairs.value, 2), dtype=np.uint32)
    off = np.zeros(npairs.value + 1, dtype=np.uint64)
    lst = np.zeros(max(ll.value, 1), dtype=np.uint32)
    check(lib().fcm_edge_neighborhoods(ctx.h, int(n), int(e.shape[0]), ptr(e), C.byref(npairs), C.byref(ll),
                                       ptr(pairs), ptr(off), ptr(lst)))
    return {(int(a), int(b)): [int(x) for x in lst[int(off[q]):int(off[q + 1])]] for q, (a, b) in enumerate(pairs)}


def compute_maximal_cliques(n, edges):
    """Graph::compute_maximal_cliques (external crate, called by State::new, src/lib.rs:41): the maximal cliques of the
    underlying undirected graph, vertices ascending, cliques in lexicographic order.  Host code in the reference and here
    (libfcm.so runs the same enumeration inside fcm_ensemble_create); needs no device."""
    e = edges_array(edges)
    nc, nv = C.c_uint64(), C.c_uint64()
    check(lib().fcm_maximal_cliques(int(n), int(e.shape[0]), ptr(e), C.byref(nc), C.byref(nv), None, None))
    sizes = np.zeros(max(nc.value, 1), dtype=np.uint32)
    verts = np.zeros(max(nv.value, 1), dtype=np.uint32)
    check(lib().fcm_maximal_cliques(int(n), int(e.shape[0]), ptr(e), C.byref(nc), C.byref(nv), ptr(sizes), ptr(verts)))
    ends = np.cumsum(sizes[:nc.value], dtype=np.int64)
    return [verts[int(b) - int(k):int(b)].tolist() for k, b in zip(sizes[:nc.value], ends)]


def cliques_by_order(n, edges):
    """State::cliques_by_order (src/lib.rs:41-49): cliques_by_order[k] = the maximal cliques of k + 1 vertices"""
    cl = compute_maximal_cliques(n, edges)
    out = [[] for _ in range(max((len(c) for c in cl), default=0))]
    for c in cl:
        out[len(c) - 1].append(c)
    return out


def undirected_edges(edges):
    """[a,b] with a > b once per adjacent pair (src/lib.rs:344 asserts a > b)."""
    e = edges_array(edges)
    big = np.maximum(e[:, 0], e[:, 1]).astype(np.uint64)
    small = np.minimum(e[:, 0], e[:, 1]).astype(np.uint64)
    key = np.unique((big << np.uint64(32)) | small)
    return np.stack([(key >> np.uint64(32)).astype(np.uint32), (key & np.uint64(0xFFFFFFFF)).astype(np.uint32)], axis=1)


# ---- Bounds ----------------------------------------------------------------------------------------
class Bounds:
    """src/lib.rs:113-161."""

    def __init__(self, flag_count_min, flag_count_max):
        self.flag_count_min = [int(x) for x in flag_count_min]
        self.flag_count_max = [int(x) for x in flag_count_max]

    @staticmethod
    def target(flag_count, target_relaxation):
        """src/bin/sample.rs:89-92: exact for d < 2, floor(s*(1-+t)) in f64 for d >= 2."""
        mn = [s if d < 2 else f64_as_usize(math.floor(float(s) * (1.0 - target_relaxation))) for d, s in enumerate(flag_count)]
        mx = [s if d < 2 else f64_as_usize(math.floor(float(s) * (1.0 + target_relaxation))) for d, s in enumerate(flag_count)]
        return Bounds(mn, mx)

    @staticmethod
    def calculate(flag_count, ncliques, n_undirected_edges, target_bounds):
        """src/lib.rs:119-156.  `ncliques` = flagser_count of the normalised graph (line 130)."""
        if n_undirected_edges == flag_count[1]:                      # SEO case, :135-137
            return Bounds(target_bounds.flag_count_min, ncliques)
        mn = list(target_bounds.flag_count_min)
        mx = list(target_bounds.flag_count_max)
        relax_de = calc_relax_de(flag_count)
        ln = len(flag_count)
        for d in range(2, ln):
            dense = binomial(ln - 2, d - 1)
            relax = (relax_de[d] * dense) & U64
            mx[d] = max((mn[d] + relax) & U64, mx[d])
            if mx[d] < relax:
                raise OverflowError("flag_count_max[d] - relax underflows (reference panics in debug)")
            mn[d] = min(mx[d] - relax, mn[d])
        mx[2] = USIZE_MAX                                            # :151
        mx.append(10)                                                # :152
        return Bounds(mn, mx)

    def check_counts(self, flag_count):
        """src/lib.rs:157-160 on a host copy of flag_count."""
        return all_le(self.flag_count_min, flag_count) and all_le(flag_count, self.flag_count_max)


# ---- the sampler: an ensemble of chains --------------------------------------------------------------
class MCMCSampler:
    """`MCMCSampler` (src/lib.rs:164-198) for `len(seeds)` independent chains on one GPU.

    next() advances every chain by sample_distance proposals.  Per-chain results are read with
    flag_count(k) / edges(k) / sampled / accepted."""

    def __init__(self, ctx, n, edges, seeds, bounds, flag_count0, move_weights=MOVE_DISTRIBUTION_SIMPLE,
                 sample_distance=0):
        self.ctx = ctx
        e = edges_array(edges)
        self.n = int(n)
        self.seeds = np.ascontiguousarray(np.asarray(seeds, dtype=np.uint64).reshape(-1))
        self.nchains = int(self.seeds.size)
        self.bounds = bounds
        fc0 = np.asarray(flag_count0, dtype=np.uint64)
        mn = np.asarray(bounds.flag_count_min, dtype=np.uint64)
        mx = np.asarray(bounds.flag_count_max, dtype=np.uint64)
        mw = np.asarray(move_weights, dtype=np.float64)
        if mw.size != 4:
            raise ValueError("move_weights must have 4 entries")
        self.move_weights = [float(x) for x in mw]
        self.h = C.c_void_p()
        check(lib().fcm_ensemble_create(ctx.h, self.n, int(e.shape[0]), ptr(e), self.nchains, ptr(self.seeds),
                                        ptr(fc0), int(fc0.size), ptr(mn), int(mn.size), ptr(mx), int(mx.size),
                                        ptr(mw), C.byref(self.h)))
        if sample_distance == 0:                                     # src/bin/sample.rs:102
            E = float(int(fc0[1])) if fc0.size > 1 else 0.0
            sd = math.ceil(2.0 * E * math.log2(E)) if E > 0 else 0
            sample_distance = int(sd)
        self.sample_distance = int(sample_distance)

    def set_kernel(self, kind):
        """0 = delta path, from the static clique index + orientation bits when it was built (default),
        1 = full pre/post neighbourhood counts, 2 = delta path gathered from the adjacency rows;
        identical results."""
        check(lib().fcm_ensemble_set_kernel(self.h, int(kind)))

    # -- stepping --
    def step(self, nproposals):
        """nproposals iterations of the loop body of next() on every chain (asynchronous)."""
        check(lib().fcm_ensemble_step(self.h, int(nproposals)))

    def next(self):
        self.step(self.sample_distance)
        return self

    def step_trace(self, nproposals, chain=0):
        rec = np.zeros(int(nproposals), dtype=N.TRACE_DTYPE)
        check(lib().fcm_ensemble_step_trace(self.h, int(nproposals), int(chain), ptr(rec)))
        return rec

    def status(self):
        check(lib().fcm_ensemble_status(self.h))

    # -- reading --
    def read(self, chain=0, want_edges=True):
        fc = np.zeros(FCM_MAXD, dtype=np.uint64)
        ln, s, a, ne, nd = C.c_int(), C.c_uint64(), C.c_uint64(), C.c_uint64(), C.c_uint64()
        rng = np.zeros(4, dtype=np.uint64)
        info = self.info()
        ebuf = np.zeros((2 * info["npairs"], 2), dtype=np.uint32) if want_edges else None
        check(lib().fcm_ensemble_read(self.h, int(chain), ptr(fc), C.byref(ln), C.byref(s), C.byref(a),
                                      ptr(ebuf) if want_edges else None, C.byref(ne), C.byref(nd), ptr(rng)))
        return {"flag_count": [int(x) for x in fc[:ln.value]], "sampled": s.value, "accepted": a.value,
                "nedges": ne.value, "ndouble": nd.value, "rng": [int(x) for x in rng],
                "edges": ebuf[:ne.value].copy() if want_edges else None}

    def read_all(self):
        fcs = np.zeros((self.nchains, FCM_MAXD), dtype=np.uint64)
        lens = np.zeros(self.nchains, dtype=np.int32)
        s = np.zeros(self.nchains, dtype=np.uint64)
        a = np.zeros(self.nchains, dtype=np.uint64)
        st = np.zeros(2, dtype=np.uint64)
        check(lib().fcm_ensemble_read_all(self.h, ptr(fcs), ptr(lens), ptr(s), ptr(a), ptr(st)))
        return {"flag_counts": fcs, "lens": lens, "sampled": s, "accepted": a, "sumK": int(st[0]),
                "nonempty": int(st[1])}

    def acceptance_ratio(self, chain=0):
        r = self.read(chain, want_edges=False)
        return r["accepted"] / r["sampled"] if r["sampled"] else float("nan")

    def info(self):
        n, npairs, nch, maxc, bpc = C.c_uint32(), C.c_uint64(), C.c_uint32(), C.c_uint32(), C.c_uint64()
        check(lib().fcm_ensemble_info(self.h, C.byref(n), C.byref(npairs), C.byref(nch), C.byref(maxc), C.byref(bpc)))
        return {"n": n.value, "npairs": npairs.value, "nchains": nch.value, "max_common_nbhd": maxc.value,
                "bytes_per_chain": bpc.value}

    def set_policy(self, policy):
        """edge-sampler policy seam (include/fcm.h): 0 = element r of the container (default), 1 = element len-1-r"""
        check(lib().fcm_ensemble_set_policy(self.h, int(policy)))

    def set_stats(self, on):
        """Switch the neighbourhood-size statistic (sumK / nonempty of read_all) on (default) or off."""
        check(lib().fcm_ensemble_set_stats(self.h, 1 if on else 0))

    def index_info(self):
        have, nbytes = C.c_int32(), C.c_uint64()
        check(lib().fcm_ensemble_index_info(self.h, C.byref(have), C.byref(nbytes)))
        return {"have_index": bool(have.value), "index_bytes": nbytes.value}

    def set_bounds(self, bounds):
        mn = np.asarray(bounds.flag_count_min, dtype=np.uint64)
        mx = np.asarray(bounds.flag_count_max, dtype=np.uint64)
        check(lib().fcm_ensemble_set_bounds(self.h, ptr(mn), int(mn.size), ptr(mx), int(mx.size)))
        self.bounds = bounds

    # -- State::apply_transition / revert_transition (src/lib.rs:61-95) --
    def apply_transition(self, change_edges, chain=0):
        """change_edges = [((a, b), add), ...] -> (pre, post) lists."""
        ed = np.array([[a, b, 1 if add else 0] for (a, b), add in change_edges], dtype=np.uint32).reshape(-1, 3)
        pre = np.zeros(FCM_MAXD, dtype=np.uint64)
        post = np.zeros(FCM_MAXD, dtype=np.uint64)
        pl, ql = C.c_int(), C.c_int()
        check(lib().fcm_apply(self.h, int(chain), ptr(ed), int(ed.shape[0]), ptr(pre), C.byref(pl), ptr(post), C.byref(ql)))
        return [int(x) for x in pre[:pl.value]], [int(x) for x in post[:ql.value]]

    def revert_transition(self, change_edges, counters, chain=0):
        pre, post = counters
        ed = np.array([[a, b, 1 if add else 0] for (a, b), add in change_edges], dtype=np.uint32).reshape(-1, 3)
        p = np.asarray(pre, dtype=np.uint64)
        q = np.asarray(post, dtype=np.uint64)
        check(lib().fcm_revert(self.h, int(chain), ptr(ed), int(ed.shape[0]), ptr(p), int(p.size), ptr(q), int(q.size)))

    def bounds_check(self, chain=0):
        ok = C.c_int()
        check(lib().fcm_bounds_check(self.h, int(chain), C.byref(ok)))
        return bool(ok.value)

    # -- io::save_state / load_state (src/io.rs:51-62): checkpoint / resume of one chain --
    def checkpoint(self, chain=0):
        r = self.read(chain, want_edges=True)
        nd = C.c_uint64()
        dbl = np.zeros((max(r["ndouble"], 1), 2), dtype=np.uint32)
        check(lib().fcm_ensemble_read_doubles(self.h, int(chain), ptr(dbl), C.byref(nd)))
        return {"edges": r["edges"], "doubles": dbl[:nd.value].copy(), "rng": np.asarray(r["rng"], dtype=np.uint64),
                "flag_count": np.asarray(r["flag_count"], dtype=np.uint64), "sampled": r["sampled"], "accepted": r["accepted"]}

    def restore(self, chain, ck):
        e = edges_array(ck["edges"])
        d = edges_array(ck["doubles"])
        rng = np.ascontiguousarray(np.asarray(ck["rng"], dtype=np.uint64))
        fc = np.ascontiguousarray(np.asarray(ck["flag_count"], dtype=np.uint64))
        check(lib().fcm_ensemble_restore(self.h, int(chain), ptr(e), int(e.shape[0]), ptr(d), int(d.shape[0]), ptr(rng),
                                         ptr(fc), int(fc.size), int(ck["sampled"]), int(ck["accepted"])))

    # -- io::BitOutput (src/io.rs:128-212): samples packed on the device --
    def edgebits_info(self):
        ns, nb = C.c_uint64(), C.c_uint64()
        check(lib().fcm_ensemble_edgebits_info(self.h, C.byref(ns), C.byref(nb)))
        return {"nslots": ns.value, "nbytes": nb.value}

    def edgebits_slots(self):
        """BitOutput.edges: (nslots, 2) directed slots in file order (src/io.rs:152-159)."""
        out = np.zeros((self.edgebits_info()["nslots"], 2), dtype=np.uint32)
        check(lib().fcm_ensemble_edgebits_slots(self.h, ptr(out)))
        return out

    def edgebits(self, chain_first=0, nchains=None):
        """One BitOutput::save row per chain: (nchains, nbytes) uint8 (src/io.rs:180-194)."""
        nchains = self.nchains - chain_first if nchains is None else nchains
        out = np.zeros((nchains, self.edgebits_info()["nbytes"]), dtype=np.uint8)
        if out.size:
            check(lib().fcm_ensemble_edgebits(self.h, int(chain_first), int(nchains), ptr(out)))
        return out

    def close(self):
        if self.h:
            lib().fcm_ensemble_destroy(self.h)
            self.h = C.c_void_p()

    def __del__(self):
        try:
            if not sys.is_finalizing():                 # at interpreter exit the CUDA runtime may be gone already
                self.close()
        except Exception:
            pass


def initialize_new_sampler(ctx, n, edges, seeds, target_relaxation=0.01, simple=True, sample_distance=0):
    """src/bin/sample.rs:80-105 for one or many seeds: State::new's full count, target bounds,
    Bounds::calculate (second full count on the normalised graph), move mix, sample distance."""
    e = edges_array(edges)
    g = DeviceGraph(ctx, n, e)
    flag_count = g.flagser_count()                                   # State::new, src/lib.rs:51
    g.close()
    gn = DeviceGraph(ctx, n, e, normalised=True)
    ncliques = gn.flagser_count()                                    # src/lib.rs:125-130
    n_und = gn.nedges
    gn.close()
    target = Bounds.target(flag_count, target_relaxation)
    bounds = Bounds.calculate(flag_count, ncliques, n_und, target)
    mw = MOVE_DISTRIBUTION_SIMPLE if simple else MOVE_DISTRIBUTION
    s = MCMCSampler(ctx, n, e, seeds, bounds, flag_count, mw, sample_distance)
    s.initial_flag_count = flag_count
    s.ncliques = ncliques
    return s


class BitOutput:
    """io::BitOutput (src/io.rs:128-212) for one chain of a device ensemble: `dir/graph.flag` +
    `dir/<k>.edgebits` files of `chunk_size` samples each; every save() is one device-side pack and
    a D2H copy of ceil(nslots/8) bytes."""

    def __init__(self, sampler, directory, chain=0, n=None, edges=None):
        import os
        os.makedirs(directory, exist_ok=True)
        self.sampler, self.dir, self.chain = sampler, directory, chain
        self.edges = sampler.edgebits_slots()
        if edges is not None:
            save_flag_file(os.path.join(directory, "graph.flag"), n, edges)
        # src/io.rs:161: max(2e9 / (len/8), 1); the reference divides by zero below 8 slots
        self.chunk_size = max(2_000_000_000 // max(len(self.edges) // 8, 1), 1)
        self.index_in_file = 0
        self.index_in_dir = 0
        self.current_file = None

    def save(self):
        import os
        if self.index_in_file == 0:
            assert self.current_file is None
            self.current_file = open(os.path.join(self.dir, "%d.edgebits" % self.index_in_dir), "wb")
        self.current_file.write(self.sampler.edgebits(self.chain, 1)[0].tobytes())
        self.index_in_file += 1
        if self.index_in_file == self.chunk_size:
            self.current_file.close()
            self.current_file = None
            self.index_in_file = 0
            self.index_in_dir += 1

    def flush(self):
        if self.current_file is not None:
            self.current_file.flush()


def _undirected_hash(edges):
    """FNV-1a over the sorted list of adjacent pairs: identifies the underlying undirected graph, which no move changes."""
    e = np.asarray(edges, dtype=np.uint64).reshape(-1, 2)
    key = np.unique(np.maximum(e[:, 0], e[:, 1]) << np.uint64(32) | np.minimum(e[:, 0], e[:, 1]))
    h = 1469598103934665603
    for x in key.tolist():
        h = ((h ^ x) * 1099511628211) & 0xFFFFFFFFFFFFFFFF
    return h


def save_state(fname, sample_number, sampler, chain=0):
    """io::save_state (src/io.rs:51-56): (sample_number, chain state) to `fname` via a temp file + rename.
    Beside the chain (edge containers in order, RNG, counters) the file holds what the reference's MCMCSampler
    carries (src/lib.rs:164-178): bounds, move distribution, sample_distance - and a hash of the undirected graph."""
    import os
    ck = sampler.checkpoint(chain)
    tmp = fname + ".tmp"
    with open(tmp, "wb") as f:
        np.savez(f, sample_number=np.int64(sample_number), sample_distance=np.int64(sampler.sample_distance),
                 bounds_min=np.asarray(sampler.bounds.flag_count_min, dtype=np.uint64),
                 bounds_max=np.asarray(sampler.bounds.flag_count_max, dtype=np.uint64),
                 move_weights=np.asarray(sampler.move_weights, dtype=np.float64),
                 undirected_hash=np.uint64(_undirected_hash(ck["edges"])),
                 sampled=np.uint64(ck["sampled"]), accepted=np.uint64(ck["accepted"]),
                 **{k: ck[k] for k in ("edges", "doubles", "rng", "flag_count")})
    os.replace(tmp, fname)


def load_state(fname, sampler, chain=0, adopt_settings=False):
    """io::load_state (src/io.rs:58-62) into chain `chain` of `sampler` -> sample_number.

    The reference rebuilds a whole one-chain MCMCSampler from the file.  Here the target is one chain of an
    ensemble whose bounds, move distribution and sample_distance are shared by all chains, so a checkpoint taken
    under different settings is REFUSED (ValueError) rather than silently changing the other chains - unless
    adopt_settings=True, which takes bounds and sample_distance from the file for the whole ensemble (what the
    reference does for its single chain).  The underlying undirected graph must be the ensemble's."""
    z = np.load(fname)
    mn, mx = [int(x) for x in z["bounds_min"]], [int(x) for x in z["bounds_max"]]
    sd = int(z["sample_distance"])
    mw = [float(x) for x in z["move_weights"]] if "move_weights" in z.files else None
    if mw is not None and list(mw) != [float(x) for x in sampler.move_weights]:
        raise ValueError("checkpoint was taken with move distribution %s, the ensemble runs %s" % (mw, list(sampler.move_weights)))
    same = (mn == list(sampler.bounds.flag_count_min) and mx == list(sampler.bounds.flag_count_max) and sd == sampler.sample_distance)
    if not same and not adopt_settings:
        raise ValueError("checkpoint bounds / sample_distance differ from the ensemble's (pass adopt_settings=True to "
                         "take them over for every chain)")
    if "undirected_hash" in z.files and sampler.nchains:
        here = _undirected_hash(sampler.read(0)["edges"])
        if int(z["undirected_hash"]) != here:
            raise ValueError("checkpoint belongs to a different underlying undirected graph")
    sampler.restore(chain, {k: z[k] for k in ("edges", "doubles", "rng", "flag_count")} |
                    {"sampled": int(z["sampled"]), "accepted": int(z["accepted"])})
    if not same:
        sampler.set_bounds(Bounds(mn, mx))
        sampler.sample_distance = sd
    return int(z["sample_number"])
