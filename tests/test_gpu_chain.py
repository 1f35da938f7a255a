"""GPU: MCMC chains through the C ABI against the oracle: per-proposal traces (move kind, sampled
edge, neighbourhood size, accept/reject), final flag_count, edge-container order, RNG state."""
import numpy as np
import pytest

import oracle_binding as orc
import flag_complex_mcmc_b200 as fcm
from flag_complex_mcmc_b200 import directed_scm as ds
from test_gpu_count import rand_graph

pytestmark = pytest.mark.gpu

FIELDS = ("flags", "a", "b", "dm0", "dm1", "K")


def assert_trace_equal(got, want):
    for f in FIELDS:
        if not np.array_equal(got[f], want[f]):
            i = int(np.nonzero(got[f] != want[f])[0][0])
            raise AssertionError("trace field %s differs first at proposal %d: got %s want %s" % (f, i, got[i], want[i]))


def compare_chain(s, chain, o):
    r, w = s.read(chain), o.read()
    assert r["flag_count"] == w["flag_count"]
    assert (r["sampled"], r["accepted"]) == (w["sampled"], w["accepted"])
    assert r["rng"] == w["rng"]
    assert (r["nedges"], r["ndouble"]) == (w["nedges"], w["ndouble"])
    assert np.array_equal(r["edges"], w["edges"])        # container order, not just the set


@pytest.mark.parametrize("kernel", [0, 1, 2])
@pytest.mark.parametrize("name,seed,steps", [("counterexample_seo_greedy_5_start", 0, 20000),
                                             ("counterexample_any_order", 3, 3000),
                                             ("bug_calc_relax_de", 1, 8000),
                                             ("c2_chunglu_279", 0, 20000)])
def test_trace_parity_single_chain(name, seed, steps, kernel, ctx, graphs):
    # kernel 0 = delta path (simplices through the changed edges only), 1 = full pre/post counts
    n, e = graphs[name]
    s = fcm.initialize_new_sampler(ctx, n, e, [seed], target_relaxation=0.01, simple=True)
    s.set_kernel(kernel)
    o = orc.Sampler(n, e, seed, 0.01)
    w = o.read()
    assert s.bounds.flag_count_min == w["bounds_min"] and s.bounds.flag_count_max == w["bounds_max"]
    assert s.sample_distance == w["sample_distance"]
    for chunk in (1, 7, steps):
        assert_trace_equal(s.step_trace(chunk, 0), o.steps(chunk))
    compare_chain(s, 0, o)
    acc = s.read(0)["accepted"]
    assert 0 < acc <= steps + 8
    # incremental flag_count == full recount of the chain's final graph (GPU count)
    r = s.read(0)
    assert ds.trim(r["flag_count"]) == fcm.flagser_count(n, r["edges"], ctx)
    s.close()


@pytest.mark.parametrize("kernel", [0, 1, 2])
def test_many_chains_match_per_seed_oracle(kernel, ctx):
    n = 120
    e = rand_graph(n, 0.12, 21, pdouble=0.3)
    seeds = list(range(40, 40 + 70))                      # 70 chains: partial last block
    s = fcm.initialize_new_sampler(ctx, n, e, seeds, target_relaxation=0.05, simple=True)
    s.set_kernel(kernel)
    s.step(1500)
    s.step(500)
    allr = s.read_all()
    assert (allr["sampled"] == 2000).all()
    for k in (0, 1, 31, 32, 69):
        o = orc.Sampler(n, e, seeds[k], 0.05)
        o.steps(2000, trace=False)
        compare_chain(s, k, o)
        assert [int(x) for x in allr["flag_counts"][k][:allr["lens"][k]]] == o.read()["flag_count"]
    s.status()
    s.close()


def test_next_is_sample_distance_proposals(ctx, graphs):
    n, e = graphs["counterexample_any_order"]
    s = fcm.initialize_new_sampler(ctx, n, e, [5, 6], simple=True)
    o = orc.Sampler(n, e, 6)
    s.next()
    o.next()
    compare_chain(s, 1, o)
    assert s.read(1)["sampled"] == s.sample_distance == o.info()["sample_distance"]
    assert abs(s.acceptance_ratio(1) - o.info()["accepted"] / o.info()["sampled"]) < 1e-15
    s.close()


def hub_graph(H, S, ps, seed, pdouble=0.15):
    """H mutually adjacent hubs, S satellites adjacent to every hub and sparsely to each other:
    hub-hub pairs have ~S+H common neighbours but the neighbourhood holds few simplices."""
    r = np.random.default_rng(seed)
    n = H + S
    edges = []
    for a in range(n):
        for b in range(a):
            if not (a < H or b < H or r.random() < ps):
                continue
            u = r.random()
            if u < pdouble:
                edges += [(a, b), (b, a)]
            elif u < 0.5 + pdouble / 2:
                edges.append((a, b))
            else:
                edges.append((b, a))
    e = np.array(edges, dtype=np.uint32)
    return n, e[r.permutation(e.shape[0])]


@pytest.mark.parametrize("H,S,kmin,steps", [(8, 75, 64, 1200), (8, 150, 128, 1200), (8, 290, 256, 1500)])
def test_large_neighbourhoods_multiword_rows(H, S, kmin, steps, ctx):
    # K > 64 (2-word local rows), > 128 (4-word) and > 256 (8-word)
    n, e = hub_graph(H, S, 0.03, 100 + S)
    s = fcm.initialize_new_sampler(ctx, n, e, [S], target_relaxation=0.5, simple=True)
    o = orc.Sampler(n, e, S, 0.5, nthreads=8)
    got, want = s.step_trace(steps, 0), o.steps(steps)
    assert want["K"].max() > kmin
    assert_trace_equal(got, want)
    compare_chain(s, 0, o)
    s.close()


def test_apply_revert_api(ctx, graphs):
    n, e = graphs["c2_chunglu_279"]
    s = fcm.initialize_new_sampler(ctx, n, e, [0], simple=True)
    o = orc.Sampler(n, e, 0)
    es = set(map(tuple, e.tolist()))
    singles = [(a, b) for (a, b) in e.tolist() if (b, a) not in es]
    doubles = [(a, b) for (a, b) in e.tolist() if (b, a) in es and a < b]
    fc0 = s.read(0)["flag_count"]
    for (a, b) in singles[:40:4]:
        ch = [((a, b), False), ((b, a), True)]             # a flip, src/lib.rs:295
        pp, qq = s.apply_transition(ch), o.apply(ch)
        assert pp == qq
        assert s.read(0)["flag_count"] == o.read()["flag_count"]
        assert s.bounds_check(0) == o.bounds_check()
        s.revert_transition(ch, pp)
        o.revert(ch, qq)
        assert ds.trim(s.read(0)["flag_count"]) == ds.trim(fc0)
        assert np.array_equal(s.read(0)["edges"], o.read()["edges"])
    # a double-edge move written by hand (src/lib.rs:322): make a single edge double, delete one side of a double
    (a, b), (x, y) = singles[1], doubles[0]
    ch = [((b, a), True), ((x, y), False)]
    assert s.apply_transition(ch) == o.apply(ch)
    r, w = s.read(0), o.read()
    assert r["flag_count"] == w["flag_count"] and np.array_equal(r["edges"], w["edges"]) and r["ndouble"] == w["ndouble"]
    # empty transition
    assert s.apply_transition([]) == ([], [])
    # a pair that is not adjacent: the reference panics on the HashMap index (src/lib.rs:104)
    non = next((a, b) for a in range(n) for b in range(a) if (a, b) not in es and (b, a) not in es)
    with pytest.raises(fcm.FcmError) as ei:
        s.apply_transition([(non, True)])
    assert ei.value.code == fcm._native.FCM_ENOTADJ
    # chains continue identically after the explicit edits
    assert_trace_equal(s.step_trace(500, 0), o.steps(500))
    s.close()


@pytest.mark.parametrize("kernel", [0, 1, 2])
def test_rejections_happen_and_revert_restores(kernel, ctx, graphs):
    # tight hand-set bounds: most flips are rejected; state must stay consistent
    n, e = graphs["c2_chunglu_279"]
    s = fcm.initialize_new_sampler(ctx, n, e, [9], simple=True)
    s.set_kernel(kernel)
    o = orc.Sampler(n, e, 9)
    fc = s.initial_flag_count
    mn = [fc[0], fc[1]] + [c - 1 for c in fc[2:]]
    mx = [fc[0], fc[1]] + [c + 1 for c in fc[2:]]
    s.set_bounds(ds.Bounds(mn, mx))
    o.set_bounds(mn, mx)
    got, want = s.step_trace(4000, 0), o.steps(4000)
    assert_trace_equal(got, want)
    rej = ((want["flags"] >> 4) & 1) == 0
    assert rej.sum() > 500
    compare_chain(s, 0, o)
    r = s.read(0)
    assert ds.trim(r["flag_count"]) == fcm.flagser_count(n, r["edges"], ctx)
    s.close()


@pytest.mark.parametrize("kernel", [0, 2])
@pytest.mark.parametrize("name,seed,steps,weights", [
    ("counterexample_any_order", 5, 4000, ds.MOVE_DISTRIBUTION),
    ("counterexample_seo_greedy_5_start", 1, 3000, ds.MOVE_DISTRIBUTION),
    ("c2_chunglu_279", 0, 3000, ds.MOVE_DISTRIBUTION),
    ("c2_chunglu_279", 7, 1500, (0.0, 0.0, 1.0, 0.0)),
    ("c2_chunglu_279", 8, 1500, (0.0, 0.0, 0.0, 1.0)),
    ("bug_calc_relax_de", 2, 3000, (0.25, 0.25, 0.25, 0.25))])
def test_clique_moves_trace_parity(name, seed, steps, weights, kernel, ctx, graphs):
    # Transition::clique_permute / clique_swap (src/lib.rs:214-290) mixed with the simple moves: trace
    # (move kind, chosen cliques, order, number of edits, |nbhd|, accept), counts, containers, RNG
    n, e = graphs[name]
    s = fcm.initialize_new_sampler(ctx, n, e, [seed], target_relaxation=0.01, simple=False)
    s.close()
    fc0 = fcm.flagser_count(n, e, ctx)
    o = orc.Sampler(n, e, seed, 0.01, move_weights=weights)
    w = o.read()
    b = ds.Bounds(w["bounds_min"], w["bounds_max"])
    s = fcm.MCMCSampler(ctx, n, e, [seed], b, fc0, weights, w["sample_distance"])
    s.set_kernel(kernel)
    for chunk in (1, 5, steps):
        got, want = s.step_trace(chunk, 0), o.steps(chunk)
        for f in FIELDS:
            if not np.array_equal(got[f], want[f]):
                i = int(np.nonzero(got[f] != want[f])[0][0])
                raise AssertionError("trace field %s differs first at proposal %d: got %s want %s" % (f, i, got[i], want[i]))
    compare_chain(s, 0, o)
    kinds = np.bincount(want["flags"] & 3, minlength=4)
    assert kinds[2] + kinds[3] > 0
    r = s.read(0)
    assert ds.trim(r["flag_count"]) == fcm.flagser_count(n, r["edges"], ctx)
    s.close()


def test_clique_moves_many_chains_untraced(ctx, graphs):
    n, e = graphs["c2_chunglu_279"]
    seeds = list(range(40, 52))
    s = fcm.initialize_new_sampler(ctx, n, e, seeds, target_relaxation=0.01, simple=False)
    s.step(1200)
    s.status()
    for k in (0, 5, 11):
        o = orc.Sampler(n, e, seeds[k], 0.01, move_weights=ds.MOVE_DISTRIBUTION)
        o.steps(1200, trace=False)
        compare_chain(s, k, o)
    s.close()


def test_c3_full_size_chains(ctx, graphs, expected):
    # BASELINE config C3: ER n=2000 p=0.05, many chains; parity on sampled chains + invariants on all
    n, e = graphs["c3_er_2000"]
    nch = 256
    s = fcm.initialize_new_sampler(ctx, n, e, list(range(nch)), simple=True)
    assert s.initial_flag_count == expected["c3_er_2000"]["flag_count"]
    steps = 3000
    s.step(steps)
    allr = s.read_all()
    assert (allr["sampled"] == steps).all()
    assert (allr["flag_counts"][:, 0] == 2000).all() and (allr["flag_counts"][:, 1] == 199938).all()
    for k in (0, 100, 255):
        o = orc.Sampler(n, e, k, nthreads=8)
        o.steps(steps, trace=False)
        compare_chain(s, k, o)
    r = s.read(17)
    assert ds.trim(r["flag_count"]) == fcm.flagser_count(n, r["edges"], ctx)
    s.close()


def test_c3_full_size_default_move_mix(ctx, graphs):
    # BASELINE config C3 with the CLI's default move mix (clique moves included): many chains run
    # untraced; sampled chains are compared with the oracle (counts, containers, RNG state), and every
    # chain keeps |V|, |E| and the incremental == full recount invariant
    n, e = graphs["c3_er_2000"]
    nch, steps = 64, 600
    s = fcm.initialize_new_sampler(ctx, n, e, list(range(nch)), simple=False)
    s.step(steps)
    s.status()
    allr = s.read_all()
    assert (allr["sampled"] == steps).all()
    assert (allr["flag_counts"][:, 0] == 2000).all() and (allr["flag_counts"][:, 1] == 199938).all()
    for k in (3, 63):
        o = orc.Sampler(n, e, k, move_weights=ds.MOVE_DISTRIBUTION, nthreads=8)
        o.steps(steps, trace=False)
        compare_chain(s, k, o)
    r = s.read(31)
    assert ds.trim(r["flag_count"]) == fcm.flagser_count(n, r["edges"], ctx)
    # the three ways to the same chain: delta path from the clique index, from the rows, full counts
    for kernel in (1, 2):
        s2 = fcm.initialize_new_sampler(ctx, n, e, [3], simple=False)
        s2.set_kernel(kernel)
        got = s2.step_trace(150, 0)
        o = orc.Sampler(n, e, 3, move_weights=ds.MOVE_DISTRIBUTION, nthreads=8)
        assert_trace_equal(got, o.steps(150))
        compare_chain(s2, 0, o)
        s2.close()
    s.close()


@pytest.mark.parametrize("n,p,pd,steps,minlen", [(30, 0.3, 0.5, 1500, 6), (50, 0.25, 0.8, 1000, 7), (12, 0.7, 0.2, 600, 8)])
def test_delta_path_deep_and_double_heavy(n, p, pd, steps, minlen, ctx):
    # dense graphs with many double edges: common neighbours carry several tags and the simplices
    # through an edge go beyond the delta path's register levels.  Loose bounds so that chains move;
    # delta path, full path and oracle are compared proposal by proposal.
    e = rand_graph(n, p, 1000 + n, pdouble=pd)
    o = orc.Sampler(n, e, 7, 0.9, nthreads=8)
    want = o.steps(steps)
    assert len(o.read()["flag_count"]) >= minlen         # >= 7: beyond the register levels (index > 5)
    for kernel in (0, 1, 2):
        s = fcm.initialize_new_sampler(ctx, n, e, [7], target_relaxation=0.9, simple=True)
        s.set_kernel(kernel)
        assert_trace_equal(s.step_trace(steps, 0), want)
        compare_chain(s, 0, o)
        r = s.read(0)
        assert ds.trim(r["flag_count"]) == fcm.flagser_count(n, r["edges"], ctx)
        s.close()


def test_delta_path_flag_count_length_growth(ctx):
    # transitive tournament with one edge reversed: flipping it back creates a simplex of a dimension
    # the complex did not have, so flag_count must grow in length exactly as src/lib.rs:72-74 does
    m = 6
    e = [(a, b) for a in range(m) for b in range(a + 1, m)]
    e[e.index((0, m - 1))] = (m - 1, 0)
    e = np.array(e, dtype=np.uint32)
    for seed in range(6):
        o = orc.Sampler(m, e, seed, 5.0)
        want = o.steps(400)
        for kernel in (0, 1, 2):
            s = fcm.initialize_new_sampler(ctx, m, e, [seed], target_relaxation=5.0, simple=True)
            s.set_kernel(kernel)
            assert_trace_equal(s.step_trace(400, 0), want)
            compare_chain(s, 0, o)                         # flag_count incl. its length (trailing zeros)
            s.close()


@pytest.mark.parametrize("name", ["counterexample_seo_greedy_5_start", "c2_chunglu_279"])
def test_switching_kernels_mid_run_keeps_rows_and_orientation_bits_in_step(name, ctx, graphs):
    # the index hot path (0) works from the orientation bits, the row-gather hot path (2) and the full
    # path (1) from the adjacency rows: both views are updated by every accepted move, so a chain may
    # change kernels at any proposal and must still follow the oracle.
    n, e = graphs[name]
    o = orc.Sampler(n, e, 11, 0.05)
    s = fcm.initialize_new_sampler(ctx, n, e, [3, 11], target_relaxation=0.05, simple=True)
    for kernel, steps in ((0, 700), (2, 500), (0, 300), (1, 200), (2, 300), (0, 500)):
        s.set_kernel(kernel)
        assert_trace_equal(s.step_trace(steps, 1), o.steps(steps))
    compare_chain(s, 1, o)
    r = s.read(1)
    assert ds.trim(r["flag_count"]) == fcm.flagser_count(n, r["edges"], ctx)
    s.close()


def test_statistic_switch_does_not_change_the_chain(ctx, graphs):
    # fcm_ensemble_set_stats(off) only freezes sumK / nonempty: trajectory, counts, containers, RNG equal
    n, e = graphs["c2_chunglu_279"]
    o = orc.Sampler(n, e, 5, 0.01)
    o.steps(6000, trace=False)
    s = fcm.initialize_new_sampler(ctx, n, e, [4, 5], target_relaxation=0.01, simple=True)
    s.step(1000)
    a = s.read_all()
    s.set_stats(False)
    s.step(4000)
    b = s.read_all()
    assert (a["sumK"], a["nonempty"]) == (b["sumK"], b["nonempty"]) and a["nonempty"] > 0
    s.set_stats(True)
    s.step(1000)
    c = s.read_all()
    assert c["nonempty"] > b["nonempty"]
    compare_chain(s, 1, o)
    s.close()


@pytest.mark.parametrize("name,kernel", [("c2_chunglu_279", 0), ("counterexample_any_order", 2), ("bug_calc_relax_de", 1)])
def test_edgebits_match_bitoutput_of_the_chain_graph(name, kernel, ctx, graphs, tmp_path):
    # io::BitOutput (src/io.rs:128-212): device-packed sample == oracle's save() of the chain's edge list
    n, e = graphs[name]
    seeds = [0, 1, 2, 3, 4]
    s = fcm.initialize_new_sampler(ctx, n, e, seeds, target_relaxation=0.01, simple=True)
    s.set_kernel(kernel)
    slots = s.edgebits_slots()
    assert np.array_equal(slots, orc.edgebits_slots(e))
    info = s.edgebits_info()
    assert info["nslots"] == slots.shape[0] and info["nbytes"] == (slots.shape[0] + 7) // 8
    bo = fcm.BitOutput(s, str(tmp_path / "out"), chain=3, n=n, edges=e)
    want_file = b""
    for rnd in range(3):
        bits = s.edgebits()
        assert bits.shape == (len(seeds), info["nbytes"])
        for k in range(len(seeds)):
            assert np.array_equal(bits[k], orc.edgebits_save(slots, s.read(k)["edges"])), (rnd, k)
        assert np.array_equal(s.edgebits(2, 2), bits[2:4])
        bo.save()
        want_file += bits[3].tobytes()
        s.step(1500)
    bo.flush()
    assert open(str(tmp_path / "out" / "0.edgebits"), "rb").read() == want_file
    n2, e2 = ds.read_flag_file(str(tmp_path / "out" / "graph.flag"))
    assert n2 == n and sorted(map(tuple, e2.tolist())) == sorted(map(tuple, e.tolist()))
    s.close()


@pytest.mark.parametrize("simple", [True, False])
def test_checkpoint_restore_resumes_exactly(simple, ctx, graphs, tmp_path):
    # io::save_state / load_state (src/io.rs:51-62): a restored chain continues as the original would
    n, e = graphs["c2_chunglu_279"]
    mw = ds.MOVE_DISTRIBUTION_SIMPLE if simple else ds.MOVE_DISTRIBUTION
    s = fcm.initialize_new_sampler(ctx, n, e, [3, 4], simple=simple)
    s.step(800)
    f = str(tmp_path / "state.npz")
    fcm.save_state(f, 7, s, chain=1)
    ck = s.checkpoint(1)
    assert ck["doubles"].shape[0] == s.read(1)["ndouble"]
    s.step(500)
    want = s.read(1)
    s2 = fcm.initialize_new_sampler(ctx, n, e, [99, 100, 101], simple=simple)
    assert fcm.load_state(f, s2, chain=2) == 7
    s2.step(500)
    got = s2.read(2)
    for k in ("flag_count", "rng", "sampled", "accepted", "nedges", "ndouble"):
        assert got[k] == want[k], k
    assert np.array_equal(got["edges"], want["edges"])
    assert np.array_equal(s2.checkpoint(2)["doubles"], s.checkpoint(1)["doubles"])
    o = orc.Sampler(n, e, 4, 0.01, move_weights=mw)
    o.steps(1300, trace=False)
    compare_chain(s2, 2, o)
    o0 = orc.Sampler(n, e, 99, 0.01, move_weights=mw)     # the neighbours of the restored chain are untouched
    o0.steps(500, trace=False)
    compare_chain(s2, 0, o0)
    # a checkpoint taken under other settings is refused instead of changing the neighbours' bounds
    s3 = fcm.initialize_new_sampler(ctx, n, e, [5, 6], target_relaxation=0.05, simple=simple)
    b3 = (list(s3.bounds.flag_count_min), list(s3.bounds.flag_count_max))
    with pytest.raises(ValueError):
        fcm.load_state(f, s3, chain=0)
    assert (list(s3.bounds.flag_count_min), list(s3.bounds.flag_count_max)) == b3
    assert fcm.load_state(f, s3, chain=0, adopt_settings=True) == 7
    assert list(s3.bounds.flag_count_min) == list(s.bounds.flag_count_min) and s3.sample_distance == s.sample_distance
    s3.close()
    if simple:
        s4 = fcm.initialize_new_sampler(ctx, n, e, [5], simple=False)
        with pytest.raises(ValueError):
            fcm.load_state(f, s4, chain=0)              # other move distribution
        s4.close()
    bad = dict(ck)
    bad["edges"] = ck["edges"].copy()
    bad["edges"][0] = [0, 0]
    with pytest.raises(fcm.FcmError):
        s2.restore(1, bad)
    bad["edges"] = ck["edges"][:-1]
    with pytest.raises(fcm.FcmError):
        s2.restore(1, bad)
    s.close()
    s2.close()


@pytest.mark.parametrize("block", range(4))
def test_all_moves_fuzz_random_graphs(block, ctx):
    # many small random graphs (sparse to dense, with and without double edges, overlapping maximal cliques),
    # random move mixes and bounds: per-proposal trace, counts, containers and RNG against the oracle
    r = np.random.default_rng(1000 + block)
    mixes = [ds.MOVE_DISTRIBUTION, (0.0, 0.0, 1.0, 0.0), (0.0, 0.0, 0.0, 1.0), (0.25, 0.25, 0.25, 0.25), (0.0, 0.2, 0.4, 0.4)]
    ran = rejected = 0
    for case in range(12):
        n = int(r.integers(5, 44))
        e = rand_graph(n, float(r.uniform(0.08, 0.45)), int(r.integers(1 << 30)), pdouble=float(r.choice([0.0, 0.2, 0.6])))
        if e.shape[0] == 0 or sum(orc.count(n, e)) > 60000:
            continue                                      # one warp per chain: keep the single traced chain's counts small
        weights = mixes[int(r.integers(len(mixes)))]
        seed = int(r.integers(1 << 40))
        relax = float(r.choice([0.01, 0.05, 0.5]))
        try:
            o = orc.Sampler(n, e, seed, relax, move_weights=weights)
        except RuntimeError:
            continue                                      # the reference panics on this graph (Bounds::calculate)
        w = o.read()
        fc0 = fcm.flagser_count(n, e, ctx)
        s = fcm.MCMCSampler(ctx, n, e, [seed], ds.Bounds(w["bounds_min"], w["bounds_max"]), fc0, weights, w["sample_distance"])
        s.set_kernel(int(r.choice([0, 0, 2])))
        got, want = s.step_trace(500, 0), o.steps(500)
        try:
            assert_trace_equal(got, want)
            compare_chain(s, 0, o)
        except AssertionError as ex:
            raise AssertionError("block %d case %d (n=%d, m=%d, mix=%s, relax=%s): %s" % (block, case, n, e.shape[0], weights, relax, ex))
        rr = s.read(0)
        assert ds.trim(rr["flag_count"]) == fcm.flagser_count(n, rr["edges"], ctx)
        rejected += 500 - int(((want["flags"] >> 4) & 1).sum())
        ran += 1
        s.close()
    assert ran >= 5
    assert rejected > 0


@pytest.mark.parametrize("kernel", [0, 1, 2])
def test_sampler_policy_seam(kernel, ctx, graphs):
    """the edge-sampler policy of the external crate (which element a draw names) is an ASSUMPTION isolated
    behind one switch.  Under a second policy (element len-1-r) the CUDA path and the oracle must still agree proposal by
    proposal, and both must differ from the default policy's trajectory: swapping in the crate's real rule is one place
    on each side (sample_at in fcm_chain.cu, Graph::at in oracle/fcm_oracle.cpp)."""
    n, e = graphs["c2_chunglu_279"]                       # has double edges: both containers are sampled
    traces = {}
    try:
        for policy in (0, 1):
            orc.set_policy(policy)
            s = fcm.initialize_new_sampler(ctx, n, e, [7, 8], target_relaxation=0.01, simple=True)
            s.set_kernel(kernel)
            s.set_policy(policy)
            o = orc.Sampler(n, e, 8, 0.01)
            got, want = s.step_trace(3000, 1), o.steps(3000)
            assert_trace_equal(got, want)
            compare_chain(s, 1, o)
            traces[policy] = got
            s.close()
    finally:
        orc.set_policy(0)
    assert not np.array_equal(traces[0]["a"], traces[1]["a"])          # the policy does decide the trajectory
    with pytest.raises(fcm.FcmError):
        s2 = fcm.initialize_new_sampler(ctx, n, e, [1], simple=True)
        try:
            s2.set_policy(2)
        finally:
            s2.close()


@pytest.mark.parametrize("simple", [True, False])
def test_large_ensemble_runs_the_seven_blocks_per_sm_instance(simple, ctx, graphs):
    """Ensembles of more than 16 chains per SM run the kernel instance compiled for 7 blocks per SM (72 registers), smaller
    ones the 128-register instance (fcm_chain_launch_step): 2500 chains here, three of them against the oracle, and the same
    seeds in a small ensemble must give the same chains."""
    n, e = graphs["c2_chunglu_279"]
    mw = ds.MOVE_DISTRIBUTION_SIMPLE if simple else ds.MOVE_DISTRIBUTION
    seeds = list(range(100, 2600))
    s = fcm.initialize_new_sampler(ctx, n, e, seeds, target_relaxation=0.01, simple=simple)
    steps = 600 if simple else 150
    s.step(steps)
    s.status()
    small = fcm.initialize_new_sampler(ctx, n, e, [seeds[0], seeds[1234], seeds[2499]], target_relaxation=0.01, simple=simple)
    small.step(steps)
    for k, j in ((0, 0), (1234, 1), (2499, 2)):
        o = orc.Sampler(n, e, seeds[k], 0.01, move_weights=mw)
        o.steps(steps, trace=False)
        compare_chain(s, k, o)
        compare_chain(small, j, o)
    s.close()
    small.close()


def test_pair_numbering_is_internal(ctx, graphs, monkeypatch):
    """Pair ids are an internal numbering ((big, small) order by default, first appearance with FCM_PAIR_ORDER=0, the
    round-1 layout kept for A/B measurements): trajectories, containers and samples do not depend on it."""
    n, e = graphs["c2_chunglu_279"]
    out = {}
    for order in ("1", "0"):
        monkeypatch.setenv("FCM_PAIR_ORDER", order)
        s = fcm.initialize_new_sampler(ctx, n, e, [5, 6], target_relaxation=0.01, simple=False)
        o = orc.Sampler(n, e, 6, 0.01, move_weights=ds.MOVE_DISTRIBUTION)
        assert_trace_equal(s.step_trace(400, 1), o.steps(400))
        compare_chain(s, 1, o)
        out[order] = (s.edgebits(1, 1)[0].tobytes(), s.read(1)["flag_count"])
        s.close()
    assert out["0"] == out["1"]
