"""GPU: (1) the limits of the clique-move path (include/fcm.h: cliques of a swap hold <= 32 vertices together, a
transition <= 128 change_edges) stop the chain with the sticky FCM_ELIMIT *before* anything of the chain is
modified - the reference (src/lib.rs:214-290) has no such limits, so the error must fire rather than corrupt;
(2) the C ABI from several host threads at once, the way src/bin/all_cxs.rs:33-41 runs 100 OS threads with a
private State each (thread-safe per handle); (3) chains on the C4-sized graph (22 % of
its pairs have K > 64, multi-word local rows are required)."""
import threading

import numpy as np
import pytest

import oracle_binding as orc
import flag_complex_mcmc_b200 as fcm
from flag_complex_mcmc_b200 import _native as N
from flag_complex_mcmc_b200 import directed_scm as ds
from tools import synth

pytestmark = pytest.mark.gpu


def tournament(vs):
    return [(vs[i], vs[j]) for i in range(len(vs)) for j in range(i + 1, len(vs))]


def check_untouched(s, chain, n, e0, fc0):
    """the chain's graph, containers and flag_count are those of the start"""
    L = N.lib()
    import ctypes as C
    fc = np.zeros(N.FCM_MAXD, dtype=np.uint64)
    ln, smp, acc, ne, nd = C.c_int(), C.c_uint64(), C.c_uint64(), C.c_uint64(), C.c_uint64()
    edges = np.zeros((e0.shape[0], 2), dtype=np.uint32)
    rc = L.fcm_ensemble_read(s.h, chain, N.ptr(fc), C.byref(ln), C.byref(smp), C.byref(acc), N.ptr(edges), C.byref(ne), C.byref(nd), None)
    assert rc == N.FCM_ELIMIT                       # sticky error is reported, the data is still delivered
    assert ne.value == e0.shape[0]
    assert sorted(map(tuple, edges.tolist())) == sorted(map(tuple, e0.tolist()))
    assert [int(x) for x in fc[:ln.value]] == fc0
    assert acc.value == smp.value                   # every proposal before the failing one was counted; the failing one was not
    return smp.value


@pytest.mark.parametrize("case", ["swap_two_disjoint_17_cliques", "permute_24_clique_more_than_128_edits"])
def test_clique_move_limits_fire_without_corrupting_the_chain(case, ctx):
    if case == "swap_two_disjoint_17_cliques":
        n = 34
        e = np.array(tournament(list(range(17))) + tournament(list(range(17, 34))), dtype=np.uint32)
        weights = (0.0, 0.0, 0.0, 1.0)
    else:
        n = 24
        e = np.array(tournament(list(range(24))), dtype=np.uint32)
        weights = (0.0, 0.0, 1.0, 0.0)
    fc0 = fcm.flagser_count(n, e, ctx)
    big = [2 ** 62] * (len(fc0) + 1)
    s = fcm.MCMCSampler(ctx, n, e, [1, 2, 3, 4], ds.Bounds([0] * len(big), big), fc0, weights, 10)
    s.step(3)
    with pytest.raises(fcm.FcmError) as ei:
        s.status()
    assert ei.value.code == N.FCM_ELIMIT
    stopped = 0
    for k in range(4):
        try:
            r = s.read(k)
        except fcm.FcmError as ex:
            assert ex.code == N.FCM_ELIMIT
            stopped += 1
            check_untouched(s, k, n, e, fc0)
            continue
        # a chain that drew only moves inside the limits (a swap of a clique with itself ...) ran on: still consistent
        assert fcm.flagser_count(n, r["edges"], ctx) == r["flag_count"]
    assert stopped >= 1
    s.step(5)                                       # a stopped chain stays stopped, nothing moves
    for k in range(4):
        try:
            s.read(k)
        except fcm.FcmError:
            check_untouched(s, k, n, e, fc0)
    s.close()


def test_stateless_count_from_eight_host_threads(graphs, expected):
    """src/bin/all_cxs.rs:33-41 pattern: many OS threads call the count at once (fcm_count_unweighted, the
    src/flagser.rs:9 precedent ABI) - every call returns its own graph's vector."""
    names = ["counterexample_any_order", "c2_chunglu_279", "counterexample_seo_greedy_5_start", "bug_calc_relax_de"]
    out, errs = {}, []

    def work(t):
        try:
            for rep in range(6):
                nm = names[(t + rep) % len(names)]
                n, e = graphs[nm]
                got = fcm.flagser_count(n, e)
                if got != expected[nm]["flag_count"]:
                    errs.append((t, nm, got))
            out[t] = True
        except Exception as ex:                      # noqa: BLE001
            errs.append((t, repr(ex)))
    th = [threading.Thread(target=work, args=(t,)) for t in range(8)]
    for t in th:
        t.start()
    for t in th:
        t.join()
    assert not errs and len(out) == 8


def test_two_threads_each_with_its_own_ensemble(graphs):
    """one context + ensemble per host thread (the reference's `&mut self` ownership): both chains equal the oracle's"""
    n, e = graphs["c2_chunglu_279"]
    res, errs = {}, []

    def work(t):
        try:
            c = fcm.Context(0)
            s = fcm.initialize_new_sampler(c, n, e, [20 + t], target_relaxation=0.01, simple=(t == 0))
            for _ in range(5):
                s.step(300)
            res[t] = s.read(0)
            s.close()
            c.close()
        except Exception as ex:                      # noqa: BLE001
            errs.append((t, repr(ex)))
    th = [threading.Thread(target=work, args=(t,)) for t in range(2)]
    for t in th:
        t.start()
    for t in th:
        t.join()
    assert not errs, errs
    for t in range(2):
        o = orc.Sampler(n, e, 20 + t, 0.01, move_weights=ds.MOVE_DISTRIBUTION_SIMPLE if t == 0 else ds.MOVE_DISTRIBUTION)
        o.steps(1500, trace=False)
        w = o.read()
        assert res[t]["flag_count"] == w["flag_count"] and res[t]["rng"] == w["rng"]
        assert np.array_equal(res[t]["edges"], w["edges"])


def test_chains_on_the_c4_sized_graph_match_the_oracle(ctx):
    """src/lib.rs:61-79 on the 31 346-vertex cortical-shaped graph (BASELINE config 4's graph): no clique index fits,
    the hot path gathers from the adjacency rows, lists beyond 64 take the full pre/post counts with multi-word local
    rows.  300-proposal trace of one chain and the state after 600 are the oracle's."""
    n, e = synth.config("c4_geo_31346")
    s = fcm.initialize_new_sampler(ctx, n, e, [0, 1, 2, 3], target_relaxation=0.01, simple=True)
    assert s.index_info()["have_index"] == 0
    tr = s.step_trace(300, 1)
    s.status()
    assert int(tr["K"].max()) > 64                    # multi-word local rows were exercised
    o = orc.Sampler(n, e, 1, 0.01, nthreads=16)
    w = o.steps(300)
    for f in ("flags", "a", "b", "dm0", "dm1", "K"):
        assert np.array_equal(tr[f], w[f]), f
    s.step(300)
    o.steps(300, trace=False)
    r, ww = s.read(1), o.read()
    assert r["flag_count"] == ww["flag_count"] and r["rng"] == ww["rng"] and np.array_equal(r["edges"], ww["edges"])
    s.close()
