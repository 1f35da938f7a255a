"""CPU: pin the oracle (and the synthetic-graph generator) against the golden vectors, and the
host-side mirror (flag_complex_mcmc_b200.directed_scm) against the oracle."""
import hashlib

import numpy as np
import pytest

import oracle_binding as orc
from flag_complex_mcmc_b200 import directed_scm as ds
from tools import synth


def test_counts_match_golden(expected, graphs):
    for name, (n, e) in graphs.items():
        assert orc.count(n, e) == expected[name]["flag_count"], name
        assert orc.count(n, e, nthreads=4) == expected[name]["flag_count"], name
        assert orc.ncliques(n, e) == expected[name]["ncliques"], name


def test_generator_matches_numpy_restatement(expected, fixtures):
    n, e = synth.config("c2_chunglu_279")
    assert np.array_equal(e, fixtures["c2_chunglu_279"])
    n, e = synth.config("c3_er_2000")
    assert e.shape[0] == expected["c3_er_2000"]["nedges"]
    assert hashlib.sha256(e.tobytes()).hexdigest() == expected["c3_er_2000"]["sha256_edges_u32le"]


def test_shards_sum_to_whole(graphs):
    n, e = graphs["counterexample_seo_greedy_5_start"]
    whole = orc.count(n, e)
    tot = np.zeros(orc.ORC_MAXD, dtype=np.uint64)
    for part in range(3):
        tot += orc.count_sources(n, e, np.arange(part, n, 3))
    assert ds.trim(tot) == whole


def test_xoshiro_known_answers():
    # reference vector of xoshiro256** (rand_xoshiro docs/tests): state [1,2,3,4]
    r = orc.Rng(state=[1, 2, 3, 4])
    assert [r.next_u64() for _ in range(4)] == [11520, 0, 1509978240, 1215971899390074240]
    # SplitMix64 seeding: seed 0 gives the published SplitMix64 stream
    r = orc.Rng(seed=0)
    assert [int(x) for x in r.s] == [0xe220a8397b1dcdaf, 0x6e789e6aa1b965f4, 0x06c45d188009454f, 0xf88bb8a8724c81ec]


def test_gen_range_and_bool_properties():
    r = orc.Rng(seed=7)
    xs = [r.gen_range(10) for _ in range(2000)]
    assert min(xs) == 0 and max(xs) == 9
    # rand 0.8 UniformInt::sample_single restated in Python ints: widening multiply + rejection zone
    a, b = orc.Rng(seed=9), orc.Rng(seed=9)
    for rng_ in (1 << 20, 3937, 199938, 3, (1 << 63) + 5):
        zone = ((rng_ << (64 - rng_.bit_length())) & ((1 << 64) - 1)) - 1
        for _ in range(50):
            while True:
                m = b.next_u64() * rng_
                if (m & ((1 << 64) - 1)) <= zone:
                    break
            assert a.gen_range(rng_) == m >> 64
    a, b = orc.Rng(seed=11), orc.Rng(seed=11)
    for _ in range(100):
        assert a.gen_bool(0.5) == (b.next_u64() < (1 << 63))


def test_weighted_index_simple_mix_is_top_bit():
    # weights [0.5,0.5,0,0]: index 1 iff the 52-bit mantissa draw >= 0.5  <=> top bit of next_u64
    a, b = orc.Rng(seed=3), orc.Rng(seed=3)
    for _ in range(500):
        assert a.weighted_index([0.5, 0.5, 0.0, 0.0]) == (b.next_u64() >> 63)
    a = orc.Rng(seed=5)
    ks = [a.weighted_index([0.1, 0.1, 0.6, 0.2]) for _ in range(4000)]
    frac = [ks.count(i) / 4000 for i in range(4)]
    assert abs(frac[0] - 0.1) < 0.03 and abs(frac[2] - 0.6) < 0.04 and abs(frac[3] - 0.2) < 0.03


def test_util_quirks_match_reference_text():
    L = orc.lib()
    # src/util.rs:65-71: loop 1..x  =>  (x-1)!
    assert [L.orc_factorial(x) for x in range(7)] == [1, 1, 1, 2, 6, 24, 120]
    assert [ds.factorial(x) for x in range(7)] == [1, 1, 1, 2, 6, 24, 120]
    for n in range(0, 12):
        for k in range(0, n + 1):
            assert L.orc_binomial(n, k) == ds.binomial(n, k)
    assert ds.binomial(5, 2) == 24 // (1 * 2)       # f(5)/(f(2) f(3)) = 24/(1*2)
    for sc in ([200, 3937, 7703, 1545, 34], [279, 2200, 2743, 2431, 1839, 1054, 379, 64], [1, 2, 3], [0]):
        a = np.asarray(sc, dtype=np.uint64)
        out = np.zeros(len(sc), dtype=np.uint64)
        assert L.orc_calc_relax_de(orc._p(a), len(sc), orc._p(out)) == 0
        assert [int(x) for x in out] == ds.calc_relax_de(sc)
    big = np.asarray([159667201], dtype=np.uint64)
    assert L.orc_calc_relax_de(orc._p(big), 1, orc._p(np.zeros(1, dtype=np.uint64))) == -1   # reference panics
    with pytest.raises(IndexError):
        ds.calc_relax_de([159667201])


def test_all_le_zero_padding():
    cases = [([1, 2], [1, 2, 0], True), ([1, 2, 1], [1, 2], False), ([], [0], True), ([0, 0, 1], [5], False),
             ([1], [1, 7], True)]
    for a, b, want in cases:
        aa, bb = np.asarray(a, dtype=np.uint64), np.asarray(b, dtype=np.uint64)
        assert bool(orc.lib().orc_all_le(orc._p(aa), len(a), orc._p(bb), len(b))) == want
        assert ds.all_le(a, b) == want


# the reference's only unit test, `test_intersect` (src/util.rs:107-156), literally: its 8 (a, b, expected) cases in order
REFERENCE_TEST_INTERSECT = [
    ([1, 2, 5, 8, 9], [0, 2, 2, 2, 3, 4, 5, 9, 9, 9], [2, 5, 9]),       # :109-112
    ([0, 2, 2, 2, 3, 4, 5, 9, 9, 9], [1, 2, 5, 8, 9], [2, 5, 9]),       # :114-117
    ([], [0, 2, 2, 2, 3, 4, 5, 9, 9, 9], []),                           # :119-122
    ([2], [3], []),                                                     # :124-127
    ([2], [0, 1, 2, 3, 4, 5], [2]),                                     # :129-134
    ([0], [0, 1, 2, 3, 4, 5], [0]),                                     # :136-141
    ([0], [0, 1, 2, 3, 4, 5], [0]),                                     # :143-148 (the reference repeats the case)
    ([5], [0, 1, 2, 3, 4, 5], [5]),                                     # :150-155
]


def oracle_intersect(x, y):
    xa, ya = np.asarray(x, dtype=np.uint32), np.asarray(y, dtype=np.uint32)
    out = np.zeros(16, dtype=np.uint32)
    k = orc.lib().orc_intersect_sorted(orc._p(xa), len(x), orc._p(ya), len(y), orc._p(out))
    return [int(v) for v in out[:k]]


def test_intersect_sorted_reference_vectors():
    """Golden vectors the reference itself holds (src/util.rs:107-156): the oracle's intersect_sorted gives the
    reference's expected output on each, twice in a row where the reference calls it twice (:131-133)."""
    assert len(REFERENCE_TEST_INTERSECT) == 8
    for a, b, want in REFERENCE_TEST_INTERSECT:
        assert oracle_intersect(a, b) == want
        assert oracle_intersect(a, b) == oracle_intersect(a, b)


def test_intersect_sorted_more_cases():
    a = [1, 2, 5, 8, 9]
    b = [0, 2, 2, 2, 3, 4, 5, 9, 9, 9]
    cases = [(a, [], []), ([], a, []), ([1], a, [1]), (a, [1], [1]), ([9], a, [9]), (a, [9], [9]), ([2, 5], b, [2, 5])]
    for x, y, want in cases:
        assert oracle_intersect(x, y) == want


@pytest.mark.parametrize("name", ["counterexample_seo_greedy_5_start", "bug_calc_relax_de", "c2_chunglu_279",
                                  "counterexample_any_order"])
def test_host_bounds_mirror_equals_oracle(name, expected, graphs):
    n, e = graphs[name]
    fc, ncl = expected[name]["flag_count"], expected[name]["ncliques"]
    s = orc.Sampler(n, e, seed=0, target_relaxation=0.01)
    r = s.read()
    tgt = ds.Bounds.target(fc, 0.01)
    b = ds.Bounds.calculate(fc, ncl, len(ds.undirected_edges(e)), tgt)
    assert b.flag_count_min == r["bounds_min"] and b.flag_count_max == r["bounds_max"]
    assert r["flag_count"] == fc
    import math
    assert r["sample_distance"] == math.ceil(2.0 * fc[1] * math.log2(fc[1]))
    assert s.bounds_check()


def test_oracle_chain_invariants(graphs):
    n, e = graphs["c2_chunglu_279"]
    s = orc.Sampler(n, e, seed=1)
    tr = s.steps(3000)
    r = s.read()
    assert r["sampled"] == 3000 and r["accepted"] == int(((tr["flags"] >> 4) & 1).sum())
    assert r["flag_count"][0] == 279 and r["flag_count"][1] == 2200          # moves keep n and E
    # incremental flag_count == full recount of the final graph
    assert ds.trim(r["flag_count"]) == orc.count(n, r["edges"])
    # apply then revert restores flag_count
    e0 = tuple(int(x) for x in r["edges"][0])
    if not any((r["edges"] == [e0[1], e0[0]]).all(axis=1)):
        ch = [((e0[0], e0[1]), False), ((e0[1], e0[0]), True)]
        pp = s.apply(ch)
        s.revert(ch, pp)
        assert ds.trim(s.read()["flag_count"]) == ds.trim(r["flag_count"])


def test_seo_fixture_double_move_always_empty(graphs):
    n, e = graphs["counterexample_seo_greedy_5_start"]
    s = orc.Sampler(n, e, seed=0)
    tr = s.steps(2000)
    dm = (tr["flags"] & 3) == 1
    assert dm.any() and ((tr["flags"][dm] >> 5) & 1).all() and ((tr["flags"][dm] >> 4) & 1).all()


def test_read_flag_file_mirror(tmp_path):
    p = tmp_path / "g.flag"
    p.write_text("dim 0:\n1 1 1  1\ndim 1:\n0 1 1\n\n2 3\n3  0 0.5\n")
    n, e = ds.read_flag_file(str(p))
    n2, e2 = orc.read_flag_file(str(p))
    assert n == n2 == 4 and np.array_equal(e, e2) and e.tolist() == [[0, 1], [2, 3], [3, 0]]


def _bitoutput_literal(edges0, edges):
    """src/io.rs:152-159 + :180-194 as literal Python (small inputs only)."""
    ed = [tuple(x) for x in np.asarray(edges0).tolist()]
    ed += [(b, a) for a, b in ed]
    ed.sort(key=lambda ab: (max(ab), min(ab), ab[0] < ab[1]))
    slots = [ed[i] for i in range(len(ed)) if i == 0 or ed[i] != ed[i - 1]]
    have = set(tuple(x) for x in np.asarray(edges).tolist())
    out, bit, byte = [], 1, 0
    for s in slots:
        if s in have:
            byte |= bit
        bit = ((bit << 1) | (bit >> 7)) & 0xFF
        if bit == 1:
            out.append(byte)
            byte = 0
    if bit != 1:
        out.append(byte)
    return np.array(slots, dtype=np.uint32).reshape(-1, 2), np.array(out, dtype=np.uint8)


def test_edgebits_oracle_matches_literal_bitoutput(graphs):
    # toy: 0->1, 1->0, 2->0, 1->2 : pairs (1,0),(2,0),(2,1); slots (1,0),(0,1),(2,0),(0,2),(2,1),(1,2)
    e0 = np.array([[0, 1], [1, 0], [2, 0], [1, 2]], dtype=np.uint32)
    sl = orc.edgebits_slots(e0)
    assert sl.tolist() == [[1, 0], [0, 1], [2, 0], [0, 2], [2, 1], [1, 2]]
    assert orc.edgebits_save(sl, e0).tolist() == [0b100111]
    for name in ("counterexample_any_order", "c2_chunglu_279"):
        n, e = graphs[name]
        flipped = e.copy()
        flipped[::3] = flipped[::3, ::-1]                  # some other orientation of the same pairs
        for cur in (e, flipped):
            sl_want, bits_want = _bitoutput_literal(e, cur)
            sl = orc.edgebits_slots(e)
            assert np.array_equal(sl, sl_want), name
            assert np.array_equal(orc.edgebits_save(sl, cur), bits_want), name


def test_save_flag_file_round_trip(tmp_path, graphs):
    n, e = graphs["c2_chunglu_279"]
    f = str(tmp_path / "g.flag")
    ds.save_flag_file(f, n, e)
    n2, e2 = ds.read_flag_file(f)
    n3, e3 = orc.read_flag_file(f)
    assert n2 == n3 == n
    assert np.array_equal(e2, e3)
    assert sorted(map(tuple, e2.tolist())) == sorted(map(tuple, e.tolist()))


def test_cells_oracle_counts_and_order(graphs, expected):
    # for_each_cell restatement: per-dimension row counts == flagser_count, rows in recursion order
    n, e = graphs["counterexample_any_order"]
    fc = orc.count(n, e)
    rows = orc.cells(n, e, 0, len(fc) - 1)
    dims = (rows != 0xFFFFFFFF).sum(axis=1) - 1
    assert [int((dims == d).sum()) for d in range(len(fc))] == fc
    as_tuples = [tuple(int(x) for x in r if x != 0xFFFFFFFF) for r in rows]
    assert as_tuples == sorted(as_tuples)                 # a prefix sorts before its extensions
    e3 = np.array([[0, 1], [1, 2], [0, 2]], dtype=np.uint32)
    assert orc.cells(3, e3, 1, 2).tolist() == [[0, 1, 0xFFFFFFFF], [0, 1, 2], [0, 2, 0xFFFFFFFF], [1, 2, 0xFFFFFFFF]]


def test_edge_neighborhoods_oracle_toy():
    # triangle 0-1-2 plus pendant 3-2: the pair (1,0) has the common neighbour 2, (3,2) has none
    e = np.array([[0, 1], [1, 2], [2, 0], [3, 2]], dtype=np.uint32)
    assert orc.edge_neighborhoods(4, e) == {(1, 0): [2], (2, 0): [1], (2, 1): [0], (3, 2): []}
