"""ctypes binding of the CPU oracle (oracle/libfcm_oracle.so).  TEST INFRASTRUCTURE: imported only
by tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs."""
import ctypes as C
import os
import subprocess

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
ORC_DIR = os.path.join(ROOT, "oracle")
ORC_MAXD = 32
TRACE_DTYPE = np.dtype([(k, np.uint32) for k in ("flags", "a", "b", "dm0", "dm1", "K", "draws", "pad")])

_lib = None


def build():
    subprocess.check_call(["make", "-s", "-C", ORC_DIR], stdout=subprocess.DEVNULL, stderr=subprocess.DEVNULL)


def lib():
    global _lib
    if _lib is not None:
        return _lib
    path = os.path.join(ORC_DIR, "libfcm_oracle.so")
    if not os.path.exists(path):
        build()
    L = C.CDLL(path)
    vp, u32, u64, i32, dbl = C.c_void_p, C.c_uint32, C.c_uint64, C.c_int, C.c_double
    P = C.POINTER
    L.orc_read_flag_header.argtypes = [C.c_char_p, P(u32), P(u64)]
    L.orc_read_flag.argtypes = [C.c_char_p, vp]
    L.orc_count.argtypes = [u32, u64, vp, i32, vp, P(i32)]
    L.orc_count_sources.argtypes = [u32, u64, vp, vp, u32, vp]
    L.orc_count_sources_mt.argtypes = [u32, u64, vp, vp, u32, i32, vp]
    L.orc_ncliques.argtypes = [u32, u64, vp, i32, vp, P(i32)]
    L.orc_maximal_cliques.argtypes = [u32, u64, vp, vp, vp, P(u64)]
    L.orc_maximal_cliques.restype = u64
    L.orc_rng_seed.argtypes = [u64, vp]
    L.orc_rng_seed.restype = None
    L.orc_rng_next.argtypes = [vp]
    L.orc_rng_next.restype = u64
    L.orc_rng_gen_range.argtypes = [vp, u64]
    L.orc_rng_gen_range.restype = u64
    L.orc_rng_gen_bool.argtypes = [vp, dbl]
    L.orc_weighted_index.argtypes = [vp, i32, vp, vp, P(dbl)]
    L.orc_factorial.argtypes = [u64]
    L.orc_factorial.restype = u64
    L.orc_binomial.argtypes = [u64, u64]
    L.orc_binomial.restype = u64
    L.orc_calc_relax_de.argtypes = [vp, i32, vp]
    L.orc_all_le.argtypes = [vp, i32, vp, i32]
    L.orc_intersect_sorted.argtypes = [vp, i32, vp, i32, vp]
    L.orc_sampler_new.argtypes = [u32, u64, vp, u64, dbl, vp, u64, i32, P(i32)]
    L.orc_sampler_new.restype = vp
    L.orc_sampler_free.argtypes = [vp]
    L.orc_sampler_free.restype = None
    L.orc_sampler_set_bounds.argtypes = [vp, vp, i32, vp, i32]
    L.orc_sampler_set_bounds.restype = None
    L.orc_set_policy.argtypes = [i32]
    L.orc_set_policy.restype = None
    L.orc_sampler_steps.argtypes = [vp, u64, vp]
    L.orc_sampler_next.argtypes = [vp]
    L.orc_sampler_info.argtypes = [vp, P(u64), P(u64), P(u64), P(u64), P(u64), P(i32), P(i32), P(i32)]
    L.orc_sampler_info.restype = None
    L.orc_sampler_read.argtypes = [vp, vp, vp, vp, vp, vp, vp]
    L.orc_sampler_read.restype = None
    L.orc_apply.argtypes = [vp, vp, i32, vp, P(i32), vp, P(i32)]
    L.orc_revert.argtypes = [vp, vp, i32, vp, i32, vp, i32]
    L.orc_bounds_check.argtypes = [vp]
    L.orc_nbhd_stats.argtypes = [vp, P(u64), P(u32)]
    L.orc_nbhd_stats.restype = u64
    L.orc_edgebits_slots.argtypes = [u64, vp, vp]
    L.orc_edgebits_slots.restype = u64
    L.orc_edgebits_save.argtypes = [u64, vp, u64, vp, vp]
    L.orc_edgebits_save.restype = None
    L.orc_cells.argtypes = [u32, u64, vp, i32, i32, vp, u64]
    L.orc_cells.restype = u64
    L.orc_edge_neighborhoods.argtypes = [u32, u64, vp, vp, vp, vp, P(u64)]
    L.orc_edge_neighborhoods.restype = u64
    _lib = L
    return L


def _e(edges):
    return np.ascontiguousarray(np.asarray(edges, dtype=np.uint32).reshape(-1, 2))


def _p(a):
    return a.ctypes.data_as(C.c_void_p) if a is not None and a.size else None


def count(n, edges, nthreads=1):
    e = _e(edges)
    out = np.zeros(ORC_MAXD, dtype=np.uint64)
    ln = C.c_int()
    rc = lib().orc_count(int(n), e.shape[0], _p(e), int(nthreads), _p(out), C.byref(ln))
    assert rc == 0
    return [int(x) for x in out[:ln.value]]


def count_sources(n, edges, sources):
    e = _e(edges)
    s = np.ascontiguousarray(np.asarray(sources, dtype=np.uint32))
    out = np.zeros(ORC_MAXD, dtype=np.uint64)
    rc = lib().orc_count_sources(int(n), e.shape[0], _p(e), _p(s), int(s.size), _p(out))
    assert rc == 0
    return out


def count_sources_mt(n, edges, sources, nthreads):
    e = _e(edges)
    s = np.ascontiguousarray(np.asarray(sources, dtype=np.uint32))
    out = np.zeros(ORC_MAXD, dtype=np.uint64)
    rc = lib().orc_count_sources_mt(int(n), e.shape[0], _p(e), _p(s), int(s.size), int(nthreads), _p(out))
    assert rc == 0
    return out


def maximal_cliques(n, edges):
    """compute_maximal_cliques of the restatement: list of ascending vertex lists, lexicographic"""
    e = _e(edges)
    tot = C.c_uint64()
    nc = lib().orc_maximal_cliques(int(n), e.shape[0], _p(e), None, None, C.byref(tot))
    sizes = np.zeros(max(nc, 1), dtype=np.uint32)
    verts = np.zeros(max(tot.value, 1), dtype=np.uint32)
    lib().orc_maximal_cliques(int(n), e.shape[0], _p(e), _p(sizes), _p(verts), C.byref(tot))
    out, at = [], 0
    for k in sizes[:nc]:
        out.append([int(v) for v in verts[at:at + int(k)]])
        at += int(k)
    return out


def ncliques(n, edges, nthreads=1):
    e = _e(edges)
    out = np.zeros(ORC_MAXD, dtype=np.uint64)
    ln = C.c_int()
    lib().orc_ncliques(int(n), e.shape[0], _p(e), int(nthreads), _p(out), C.byref(ln))
    return [int(x) for x in out[:ln.value]]


def edgebits_slots(edges0):
    """BitOutput::new's slot list (src/io.rs:152-159) -> (nslots, 2) uint32"""
    e = _e(edges0)
    ns = lib().orc_edgebits_slots(e.shape[0], _p(e), None)
    out = np.zeros((ns, 2), dtype=np.uint32)
    lib().orc_edgebits_slots(e.shape[0], _p(e), _p(out))
    return out


def edgebits_save(slots, edges):
    """BitOutput::save (src/io.rs:180-194) of the graph with directed edges `edges` -> uint8[ceil(nslots/8)]"""
    sl, e = _e(slots), _e(edges)
    out = np.zeros((sl.shape[0] + 7) // 8, dtype=np.uint8)
    lib().orc_edgebits_save(sl.shape[0], _p(sl), e.shape[0], _p(e), _p(out))
    return out


def cells(n, edges, dmin, dmax):
    """for_each_cell in the recursion's order -> (ncells, dmax + 1) uint32, padded with 0xffffffff"""
    e = _e(edges)
    cnt = lib().orc_cells(int(n), e.shape[0], _p(e), int(dmin), int(dmax), None, 0)
    out = np.zeros((cnt, dmax + 1), dtype=np.uint32)
    lib().orc_cells(int(n), e.shape[0], _p(e), int(dmin), int(dmax), _p(out), cnt)
    return out


def edge_neighborhoods(n, edges):
    """compute_edge_neighborhoods -> {(big, small): [common undirected neighbours]}"""
    e = _e(edges)
    tot = C.c_uint64()
    npairs = lib().orc_edge_neighborhoods(int(n), e.shape[0], _p(e), None, None, None, C.byref(tot))
    pairs = np.zeros((npairs, 2), dtype=np.uint32)
    off = np.zeros(npairs + 1, dtype=np.uint64)
    lst = np.zeros(max(tot.value, 1), dtype=np.uint32)
    lib().orc_edge_neighborhoods(int(n), e.shape[0], _p(e), _p(pairs), _p(off), _p(lst), C.byref(tot))
    return {(int(a), int(b)): [int(x) for x in lst[int(off[q]):int(off[q + 1])]] for q, (a, b) in enumerate(pairs)}


def read_flag_file(path):
    n, ne = C.c_uint32(), C.c_uint64()
    rc = lib().orc_read_flag_header(path.encode(), C.byref(n), C.byref(ne))
    assert rc == 0, path
    e = np.zeros((ne.value, 2), dtype=np.uint32)
    lib().orc_read_flag(path.encode(), _p(e))
    return n.value, e


class Rng:
    def __init__(self, seed=None, state=None):
        self.s = np.zeros(4, dtype=np.uint64)
        if state is not None:
            self.s[:] = state
        else:
            lib().orc_rng_seed(int(seed), _p(self.s))

    def next_u64(self):
        return int(lib().orc_rng_next(_p(self.s)))

    def gen_range(self, n):
        return int(lib().orc_rng_gen_range(_p(self.s), int(n)))

    def gen_bool(self, p):
        return bool(lib().orc_rng_gen_bool(_p(self.s), float(p)))

    def weighted_index(self, weights):
        w = np.asarray(weights, dtype=np.float64)
        return int(lib().orc_weighted_index(_p(w), int(w.size), _p(self.s), None, None))


def set_policy(policy):
    """edge-sampler policy of the oracle (process-wide): 0 = element r, 1 = element len-1-r"""
    lib().orc_set_policy(int(policy))


class Sampler:
    """oracle MCMCSampler (initialize_new_sampler + next), one chain."""

    def __init__(self, n, edges, seed, target_relaxation=0.01, move_weights=(0.5, 0.5, 0.0, 0.0),
                 sample_distance=0, nthreads=1):
        e = _e(edges)
        mw = np.asarray(move_weights, dtype=np.float64)
        err = C.c_int()
        self.h = lib().orc_sampler_new(int(n), e.shape[0], _p(e), int(seed), float(target_relaxation), _p(mw),
                                       int(sample_distance), int(nthreads), C.byref(err))
        if not self.h:
            raise RuntimeError("oracle sampler init failed: %d" % err.value)

    def info(self):
        sd, s, a, ne, nd = (C.c_uint64() for _ in range(5))
        fl, mnl, mxl = C.c_int(), C.c_int(), C.c_int()
        lib().orc_sampler_info(self.h, C.byref(sd), C.byref(s), C.byref(a), C.byref(ne), C.byref(nd),
                               C.byref(fl), C.byref(mnl), C.byref(mxl))
        return {"sample_distance": sd.value, "sampled": s.value, "accepted": a.value, "nedges": ne.value,
                "ndouble": nd.value, "fc_len": fl.value, "min_len": mnl.value, "max_len": mxl.value}

    def read(self):
        i = self.info()
        fc = np.zeros(max(i["fc_len"], 1), dtype=np.uint64)
        mn = np.zeros(max(i["min_len"], 1), dtype=np.uint64)
        mx = np.zeros(max(i["max_len"], 1), dtype=np.uint64)
        ed = np.zeros((max(i["nedges"], 1), 2), dtype=np.uint32)
        db = np.zeros((max(i["ndouble"], 1), 2), dtype=np.uint32)
        rng = np.zeros(4, dtype=np.uint64)
        lib().orc_sampler_read(self.h, _p(fc), _p(mn), _p(mx), _p(ed), _p(db), _p(rng))
        i.update({"flag_count": [int(x) for x in fc[:i["fc_len"]]], "bounds_min": [int(x) for x in mn[:i["min_len"]]],
                  "bounds_max": [int(x) for x in mx[:i["max_len"]]], "edges": ed[:i["nedges"]],
                  "dbl": db[:i["ndouble"]], "rng": [int(x) for x in rng]})
        return i

    def set_bounds(self, mn, mx):
        a = np.asarray(mn, dtype=np.uint64)
        b = np.asarray(mx, dtype=np.uint64)
        lib().orc_sampler_set_bounds(self.h, _p(a), int(a.size), _p(b), int(b.size))

    def steps(self, nsteps, trace=True):
        rec = np.zeros(int(nsteps), dtype=TRACE_DTYPE) if trace else None
        rc = lib().orc_sampler_steps(self.h, int(nsteps), _p(rec) if trace else None)
        assert rc == 0, rc
        return rec

    def next(self):
        assert lib().orc_sampler_next(self.h) == 0

    def apply(self, change_edges):
        ed = np.array([[a, b, 1 if add else 0] for (a, b), add in change_edges], dtype=np.uint32).reshape(-1, 3)
        pre = np.zeros(ORC_MAXD, dtype=np.uint64)
        post = np.zeros(ORC_MAXD, dtype=np.uint64)
        pl, ql = C.c_int(), C.c_int()
        lib().orc_apply(self.h, _p(ed), int(ed.shape[0]), _p(pre), C.byref(pl), _p(post), C.byref(ql))
        return [int(x) for x in pre[:pl.value]], [int(x) for x in post[:ql.value]]

    def revert(self, change_edges, counters):
        pre, post = counters
        ed = np.array([[a, b, 1 if add else 0] for (a, b), add in change_edges], dtype=np.uint32).reshape(-1, 3)
        p = np.asarray(pre, dtype=np.uint64)
        q = np.asarray(post, dtype=np.uint64)
        lib().orc_revert(self.h, _p(ed), int(ed.shape[0]), _p(p), int(p.size), _p(q), int(q.size))

    def bounds_check(self):
        return bool(lib().orc_bounds_check(self.h))

    def nbhd_stats(self):
        s, m = C.c_uint64(), C.c_uint32()
        n = lib().orc_nbhd_stats(self.h, C.byref(s), C.byref(m))
        return {"pairs": n, "sum_len": s.value, "max_len": m.value}

    def close(self):
        if self.h:
            lib().orc_sampler_free(self.h)
            self.h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass
