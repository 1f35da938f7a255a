"""GPU: simplex counting through the C ABI, bit-exact against the golden vectors and the oracle."""
import ctypes as C

import numpy as np
import pytest

import oracle_binding as orc
import flag_complex_mcmc_b200 as fcm
from flag_complex_mcmc_b200 import _native as N
from flag_complex_mcmc_b200 import directed_scm as ds
from tools import synth

pytestmark = pytest.mark.gpu


def rand_graph(n, p, seed, pdouble=0.0):
    r = np.random.default_rng(seed)
    m = r.random((n, n)) < p
    if pdouble:
        m |= (m.T & (r.random((n, n)) < pdouble))
    np.fill_diagonal(m, False)
    a, b = np.nonzero(m)
    e = np.stack([a, b], 1).astype(np.uint32)
    return e[r.permutation(e.shape[0])]


def test_golden_counts_stateless_abi(expected, graphs):
    for name, (n, e) in graphs.items():
        assert fcm.flagser_count(n, e) == expected[name]["flag_count"], name


@pytest.mark.parametrize("kind", [0, 1])
def test_golden_counts_both_kernels(kind, ctx, expected, graphs):
    for name, (n, e) in graphs.items():
        g = fcm.DeviceGraph(ctx, n, e)
        g.set_kernel(kind)
        assert g.flagser_count() == expected[name]["flag_count"], (name, kind)
        g.close()
        gn = fcm.DeviceGraph(ctx, n, e, normalised=True)
        gn.set_kernel(kind)
        assert gn.flagser_count() == expected[name]["ncliques"], (name, kind)
        assert gn.nedges == len(fcm.undirected_edges(e))
        gn.close()


@pytest.mark.parametrize("kind", [0, 1])
def test_shards_sum_to_whole_and_match_oracle_shards(kind, ctx, graphs):
    n, e = graphs["c3_er_2000"]
    g = fcm.DeviceGraph(ctx, n, e)
    g.set_kernel(kind)
    whole = g.count_shard(0, 1)
    for nparts in (2, 3, 8):
        tot = np.zeros(N.FCM_MAXD, dtype=np.uint64)
        for part in range(nparts):
            tot += g.count_shard(part, nparts)
        assert np.array_equal(tot, whole), nparts
    # shard = degree-descending order taken round-robin; recompute the same shard with the oracle
    deg = np.bincount(e[:, 0], minlength=n)
    order = np.argsort(-deg, kind="stable")
    assert np.array_equal(g.count_shard(1, 4), orc.count_sources(n, e, order[1::4]))
    g.close()


@pytest.mark.parametrize("n,p,pd,seed", [(1, 0, 0, 0), (2, 1.0, 1.0, 1), (33, 0.3, 0.2, 2), (64, 0.5, 0.5, 3),
                                         (65, 0.2, 0.0, 4), (130, 0.25, 0.1, 5), (300, 0.1, 0.05, 6),
                                         (700, 0.3, 0.0, 7), (1500, 0.02, 0.02, 8)])
def test_random_graphs_vs_oracle(n, p, pd, seed, ctx):
    e = rand_graph(n, p, seed, pd)
    want = orc.count(n, e, nthreads=8)
    for kind in (0, 1):
        g = fcm.DeviceGraph(ctx, n, e)
        g.set_kernel(kind)
        assert g.flagser_count() == want, (n, kind)
        g.close()


def test_bidirected_cliques_closed_form(ctx):
    # complete bidirected K_m: every ordered (d+1)-tuple is a simplex: m!/(m-d-1)!
    for m in (4, 7, 9):
        e = np.array([(a, b) for a in range(m) for b in range(m) if a != b], dtype=np.uint32)
        want = [int(np.prod(np.arange(m - d, m + 1, dtype=np.float64))) for d in range(m)]
        for kind in (0, 1):
            g = fcm.DeviceGraph(ctx, m, e)
            g.set_kernel(kind)
            assert g.flagser_count() == want, (m, kind)
            g.close()


def test_out_degree_beyond_tile_goes_to_dense_kernel(ctx):
    # hub with 1300 out-neighbours (> 1024-column tile) over a sparse random graph
    n = 1400
    e = rand_graph(n, 0.01, 11)
    hub = np.stack([np.zeros(1300, dtype=np.uint32), np.arange(1, 1301, dtype=np.uint32)], 1)
    e = np.concatenate([e[e[:, 0] != 0], hub])
    want = orc.count(n, e, nthreads=8)
    g = fcm.DeviceGraph(ctx, n, e)
    assert g.max_out_degree == 1300
    assert g.flagser_count() == want
    g.close()


def test_edge_cases(ctx):
    assert fcm.flagser_count(5, np.zeros((0, 2), dtype=np.uint32)) == [5]
    assert fcm.flagser_count(0, np.zeros((0, 2), dtype=np.uint32)) == []
    # duplicate lines collapse (edge *set*), order irrelevant
    e = np.array([[0, 1], [0, 1], [1, 2], [0, 2], [0, 2]], dtype=np.uint32)
    assert fcm.flagser_count(3, e) == [3, 3, 1]
    # vertex id out of range / self loop -> loud error, not a wrong count
    with pytest.raises(fcm.FcmError):
        fcm.flagser_count(3, np.array([[0, 3]], dtype=np.uint32))
    with pytest.raises(fcm.FcmError):
        fcm.DeviceGraph(ctx, 3, np.array([[1, 1]], dtype=np.uint32))
    L = N.lib()
    sz = C.c_size_t()
    assert not L.fcm_count_unweighted(3, 1, None, C.byref(sz))


def test_deep_dimensions_transitive_tournament(ctx):
    # transitive tournament on m vertices: every subset is a simplex -> counts are binomials
    import math
    m = 22
    e = np.array([(a, b) for a in range(m) for b in range(a + 1, m)], dtype=np.uint32)
    want = [math.comb(m, d + 1) for d in range(m)]
    for kind in (0, 1):
        g = fcm.DeviceGraph(ctx, m, e)
        g.set_kernel(kind)
        assert g.flagser_count() == want, kind
        g.close()


def test_c4_full_size_known_answer(ctx, expected):
    n, e = synth.config("c4_geo_31346")
    assert e.shape[0] == expected["c4_geo_31346"]["nedges"]
    g = fcm.DeviceGraph(ctx, n, e)
    got = g.flagser_count()
    assert got == expected["c4_geo_31346"]["flag_count"]
    # size-independent property: shards sum to the whole
    tot = np.zeros(N.FCM_MAXD, dtype=np.uint64)
    for part in range(8):
        tot += g.count_shard(part, 8)
    assert fcm.directed_scm.trim(tot) == got
    g.close()


def test_c4_known_answer_under_the_callers_numbering(ctx, expected, monkeypatch):
    # the same count with the library's locality order of the vertices switched off
    monkeypatch.setenv("FCM_RELABEL", "0")
    n, e = synth.config("c4_geo_31346")
    g = fcm.DeviceGraph(ctx, n, e)
    assert g.flagser_count() == expected["c4_geo_31346"]["flag_count"]
    g.close()


def test_locality_order_is_invisible(ctx, graphs, expected, monkeypatch):
    """Graphs beyond 4096 vertices are built under an internal numbering of the vertices (breadth-first levels from a far
    vertex, smoothed, sorted: fcm_count.cu); FCM_RELABEL=1 forces it for every graph.  Counts, shard sums, the normalised
    graph's counts and both kernels are unchanged, and for_each_cell still speaks the caller's ids in the recursion's order."""
    monkeypatch.setenv("FCM_RELABEL", "1")
    for name in ("counterexample_any_order", "c2_chunglu_279", "c3_er_2000"):
        n, e = graphs[name]
        g = fcm.DeviceGraph(ctx, n, e)
        assert g.flagser_count() == expected[name]["flag_count"], name
        whole = g.count_shard(0, 1)
        tot = np.zeros(N.FCM_MAXD, dtype=np.uint64)
        for part in range(3):
            tot += g.count_shard(part, 3)
        assert np.array_equal(tot, whole), name
        g.set_kernel(1)
        assert g.flagser_count() == expected[name]["flag_count"], name
        g.close()
        gn = fcm.DeviceGraph(ctx, n, e, normalised=True)
        assert gn.flagser_count() == expected[name]["ncliques"], name
        gn.close()
    for name, dmin, dmax in (("counterexample_any_order", 0, 3), ("c2_chunglu_279", 2, 2)):
        n, e = graphs[name]
        g = ds.DeviceGraph(ctx, n, e)
        assert np.array_equal(g.cells(dmin, dmax), orc.cells(n, e, dmin, dmax)), name
        g.close()
    # several components, vertex 0 and others isolated, a two-vertex graph, an edge list in shuffled order
    e = np.concatenate([rand_graph(40, 0.3, 5, pdouble=0.3) + 1, rand_graph(30, 0.4, 6) + 50]).astype(np.uint32)
    for n, ee in ((85, e), (2, np.array([[1, 0]], dtype=np.uint32)), (64, rand_graph(64, 0.5, 3, pdouble=0.5))):
        g = ds.DeviceGraph(ctx, n, ee)
        assert g.flagser_count() == orc.count(n, ee), n
        assert np.array_equal(g.cells(0, 2), orc.cells(n, ee, 0, 2)), n
        g.close()
    monkeypatch.setenv("FCM_RELABEL", "0")
    n, e = graphs["c2_chunglu_279"]
    g = fcm.DeviceGraph(ctx, n, e)
    assert g.flagser_count() == expected["c2_chunglu_279"]["flag_count"]
    g.close()
    # beyond 4096 vertices the order is on by default; a graph whose diameter (5999) is far beyond the 63 levels the
    # sweeps go: a directed path with a transitive triangle at its far end
    monkeypatch.delenv("FCM_RELABEL")
    n = 6000
    path = np.stack([np.arange(n - 1), np.arange(1, n)], 1).astype(np.uint32)
    ee = np.concatenate([path, np.array([[n - 3, n - 1]], dtype=np.uint32)])
    g = ds.DeviceGraph(ctx, n, ee)
    assert g.flagser_count() == [n, n, 1]
    tri = g.cells(2, 2)
    assert tri.tolist() == [[n - 3, n - 2, n - 1]]
    assert np.array_equal(g.cells(1, 1), ee[np.lexsort((ee[:, 1], ee[:, 0]))])
    g.close()


@pytest.mark.parametrize("name,dmin,dmax", [("counterexample_any_order", 0, 3), ("c2_chunglu_279", 2, 2),
                                            ("c2_chunglu_279", 0, 6), ("counterexample_seo_greedy_5_start", 1, 4)])
def test_for_each_cell_matches_oracle_rows(name, dmin, dmax, ctx, graphs):
    # flag_complex::for_each_cell (src/bin/all_cxs.rs:60-64): same cells, same order as the recursion
    n, e = graphs[name]
    g = ds.DeviceGraph(ctx, n, e)
    got = g.cells(dmin, dmax)
    want = orc.cells(n, e, dmin, dmax)
    assert got.shape == want.shape
    assert np.array_equal(got, want)
    fc = g.flagser_count()                                # the plain count still works on the same handle
    assert fc == orc.count(n, e)
    g.close()
    seen = []
    ds.for_each_cell(n, e, seen.append, 2, 2, ctx)
    assert seen == [[int(x) for x in r if x != 0xFFFFFFFF] for r in orc.cells(n, e, 2, 2)]


def bidirected_clique(m):
    return np.array([(a, b) for a in range(m) for b in range(m) if a != b], dtype=np.uint32)


def transitive_tournament(m):
    return np.array([(a, b) for a in range(m) for b in range(a + 1, m)], dtype=np.uint32)


@pytest.mark.parametrize("case", ["k9_bidirected", "tournament_40", "dense_random_40"])
def test_for_each_cell_dense_graphs_never_drop_a_subtree(case, ctx):
    """A popped batch whose children overflow the per-warp stack must still emit every cell below it
    (the batch goes through the per-thread walk, which emits too).  Dense graphs: ~k children per entry."""
    import math
    if case == "k9_bidirected":
        n, e, dmax = 9, bidirected_clique(9), 4
        fc = [math.perm(9, d + 1) for d in range(dmax + 1)]           # every ordered (d+1)-tuple is a simplex
    elif case == "tournament_40":
        n, e, dmax = 40, transitive_tournament(40), 4
        fc = [math.comb(40, d + 1) for d in range(dmax + 1)]          # (its full count has 2^40 simplices: not asked for)
    else:
        n, e, dmax = 40, rand_graph(40, 0.45, 11, pdouble=0.5), 4
        fc = orc.count(n, e)
    g = ds.DeviceGraph(ctx, n, e)
    got = g.cells(0, dmax)
    want = orc.cells(n, e, 0, dmax)
    assert got.shape == want.shape
    assert np.array_equal(got, want)
    dims = (got != 0xFFFFFFFF).sum(1) - 1
    for d in range(dmax + 1):
        assert int((dims == d).sum()) == (fc[d] if d < len(fc) else 0), d
    if case != "tournament_40":
        assert g.flagser_count()[:dmax + 1] == fc[:dmax + 1]
    g.close()


def test_for_each_cell_local_subgraph_like_all_cxs(ctx):
    # the use in src/bin/all_cxs.rs:57-64: 2-cells of a small induced neighbourhood graph, as sorted sets
    e = rand_graph(24, 0.4, 5, pdouble=0.2)
    g = ds.DeviceGraph(ctx, 24, e)
    got = sorted(tuple(sorted(int(x) for x in r)) for r in g.cells(2, 2))
    want = sorted(tuple(sorted(int(x) for x in r)) for r in orc.cells(24, e, 2, 2))
    assert got == want and len(got) == g.flagser_count()[2]
    g.close()


@pytest.mark.parametrize("name", ["counterexample_any_order", "c2_chunglu_279", "counterexample_seo_greedy_5_start", "c3_er_2000"])
def test_edge_neighborhoods_match_oracle(name, ctx, graphs):
    # State::compute_edge_neighborhoods (src/lib.rs:331-356): same keys [big, small], same sorted lists
    n, e = graphs[name]
    got = ds.compute_edge_neighborhoods(ctx, n, e)
    want = orc.edge_neighborhoods(n, e)
    assert got == want
    # order of the device result: first appearance of the pair in the edge list
    first = []
    seen = set()
    for a, b in e.tolist():
        k = (max(a, b), min(a, b))
        if k not in seen:
            seen.add(k)
            first.append(k)
    assert list(got.keys()) == first
