"""CPU: Graph::compute_maximal_cliques (external crate, called by State::new, src/lib.rs:41-49) as libfcm.so runs it on
host threads - bit-set Bron-Kerbosch per neighbourhood, flat group-sorted output (csrc/fcm_ensemble.cu) - against the
oracle's plain Bron-Kerbosch + sort and against closed forms.  No device involved: this is host code in the reference too."""
import itertools

import numpy as np
import pytest

import oracle_binding as orc
import flag_complex_mcmc_b200 as fcm
from flag_complex_mcmc_b200 import _native as N


def rand_graph(n, p, seed):
    r = np.random.default_rng(seed)
    m = r.random((n, n)) < p
    np.fill_diagonal(m, False)
    a, b = np.nonzero(m)
    e = np.stack([a, b], 1).astype(np.uint32)
    return e[r.permutation(e.shape[0])]


def test_golden_graphs_match_the_oracle(graphs):
    for name, (n, e) in graphs.items():
        if name == "c3_er_2000":
            continue                                           # 700 k cliques: the oracle's list form takes a while; covered below by counts
        assert fcm.compute_maximal_cliques(n, e) == orc.maximal_cliques(n, e), name


@pytest.mark.parametrize("n,p,seed", [(1, 0.0, 0), (2, 1.0, 1), (7, 0.0, 2), (40, 0.1, 3), (60, 0.3, 4), (90, 0.12, 5),
                                      (150, 0.05, 6), (33, 0.6, 7), (200, 0.03, 8)])
def test_random_graphs_match_the_oracle(n, p, seed):
    e = rand_graph(n, p, seed)
    got = fcm.compute_maximal_cliques(n, e)
    assert got == orc.maximal_cliques(n, e)
    assert got == sorted(got) and all(c == sorted(c) for c in got)


def brute_force_maximal_cliques(n, e):
    """every vertex subset that is a clique and has no common neighbour outside it (n <= 12)"""
    adj = [[False] * n for _ in range(n)]
    for a, b in e:
        adj[a][b] = adj[b][a] = True
    out = []
    for mask in range(1, 1 << n):
        vs = [v for v in range(n) if (mask >> v) & 1]
        if all(adj[a][b] for a, b in itertools.combinations(vs, 2)) and \
                not any(all(adj[w][v] for v in vs) for w in range(n) if not (mask >> w) & 1):
            out.append(vs)
    return sorted(out)


@pytest.mark.parametrize("seed", range(12))
def test_small_graphs_against_brute_force(seed):
    """pins the oracle's enumeration (and the library's) to the definition itself"""
    r = np.random.default_rng(100 + seed)
    n = int(r.integers(1, 12))
    e = rand_graph(n, float(r.uniform(0.1, 0.8)), seed)
    want = brute_force_maximal_cliques(n, e.tolist())
    assert orc.maximal_cliques(n, e) == want
    assert fcm.compute_maximal_cliques(n, e) == want


def test_word_widths_and_the_general_form():
    """neighbourhoods of <= 64 / 128 / 256 vertices use 1 / 2 / 4-word sets, larger ones the list form: a hub joined to
    everything on top of a sparse graph walks through all of them"""
    for n in (70, 140, 270, 400):
        base = rand_graph(n, 0.04, n)
        hub = np.array([(0, v) for v in range(1, n)], dtype=np.uint32)
        e = np.concatenate([base, hub])
        assert fcm.compute_maximal_cliques(n, e) == orc.maximal_cliques(n, e), n


def test_closed_forms_and_cliques_by_order():
    # complete graph: one clique; two triangles sharing an edge plus an isolated vertex and a pendant edge
    k6 = np.array([(a, b) for a in range(6) for b in range(6) if a != b], dtype=np.uint32)
    assert fcm.compute_maximal_cliques(6, k6) == [[0, 1, 2, 3, 4, 5]]
    e = np.array([(0, 1), (1, 2), (2, 0), (1, 3), (3, 2), (5, 6), (6, 5)], dtype=np.uint32)
    assert fcm.compute_maximal_cliques(7, e) == [[0, 1, 2], [1, 2, 3], [4], [5, 6]]
    by = fcm.cliques_by_order(7, e)                            # State::cliques_by_order (src/lib.rs:42-49)
    assert by == [[[4]], [[5, 6]], [[0, 1, 2], [1, 2, 3]]]
    # Moon-Moser graph K_{3,3,3}: 27 maximal triangles
    parts = [range(0, 3), range(3, 6), range(6, 9)]
    e = np.array([(a, b) for pa, pb in itertools.combinations(parts, 2) for a in pa for b in pb], dtype=np.uint32)
    got = fcm.compute_maximal_cliques(9, e)
    assert len(got) == 27 and got == [list(t) for t in itertools.product(*parts)]


def test_c3_clique_census(graphs):
    """the C3 graph (697 k maximal cliques): per-order census against the oracle's list, and the lexicographic order"""
    n, e = graphs["c3_er_2000"]
    got = fcm.compute_maximal_cliques(n, e)
    want = orc.maximal_cliques(n, e)
    assert len(got) == len(want)
    assert got == want


def test_errors_and_limits():
    with pytest.raises(fcm.FcmError) as ei:
        fcm.compute_maximal_cliques(3, np.array([(0, 3)], dtype=np.uint32))
    assert ei.value.code == N.FCM_EINVAL
    with pytest.raises(fcm.FcmError):
        fcm.compute_maximal_cliques(3, np.array([(1, 1)], dtype=np.uint32))
    # a clique of FCM_MAXD + 1 vertices is refused like fcm_ensemble_create refuses it
    m = N.FCM_MAXD + 1
    km = np.array([(a, b) for a in range(m) for b in range(a + 1, m)], dtype=np.uint32)
    with pytest.raises(fcm.FcmError) as ei:
        fcm.compute_maximal_cliques(m, km)
    assert ei.value.code == N.FCM_EDIM
    km = np.array([(a, b) for a in range(m - 1) for b in range(a + 1, m - 1)], dtype=np.uint32)
    assert fcm.compute_maximal_cliques(m - 1, km) == [list(range(m - 1))]
