#!/usr/bin/env python
"""Generate tests/golden/{fixtures.npz,expected.json}.

Run in the build container (needs /root/reference for the four `.flag` fixtures; the GPU box
does not have it, which is why the outputs are committed).  Everything here is computed WITHOUT
the oracle and WITHOUT the CUDA path, by two independent pure-Python methods:

  A. bitset recursion over Python ints (the flagser recursion)
  B. networkx enumerate_all_cliques on the underlying undirected graph + a subset DP that counts
     the linear orders of each clique that are consistent with the directed edges

plus a numpy restatement of the synthetic-graph recipe.  The script asserts
A == B and that both equal the known answers recorded in SURVEY.md §4 / §8(d).

  python tests/golden/make_golden.py            # fixtures + toy graphs + C2 + C3
"""
import hashlib
import json
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
REF = "/root/reference/flag_file_examples"

# known answers recorded in SURVEY.md §4 (three independent methods in the survey session)
SURVEY_KNOWN = {
    "counterexample_seo_greedy_5_start": ([200, 3937, 7703, 1545, 34], [200, 3937, 10138, 3907, 286, 9, 1]),
    "counterexample_seo_greedy_5_bad": ([200, 3937, 10124, 3851, 254, 2], [200, 3937, 10138, 3907, 286, 9, 1]),
    "bug_calc_relax_de": ([279, 1961, 2637, 1336, 350, 44, 2], [279, 1961, 2858, 1891, 869, 278, 50, 4]),
    "counterexample_any_order": ([8, 18, 15, 5, 1], [8, 18, 16, 6, 1]),
}


def parse_flag(path):
    """src/io.rs:18-35: line 1 skipped, token count of line 2 = n, line 3 skipped, then 'i j [w]'."""
    lines = open(path).read().split("\n")
    n = len([t for t in lines[1].split(" ") if t != ""])
    edges = []
    for ln in lines[3:]:
        t = [x for x in ln.split(" ") if x != ""]
        if len(t) >= 2:
            edges.append((int(t[0]), int(t[1])))
    return n, edges


def count_bitset(n, edges):
    out = [0] * n
    for a, b in edges:
        if a != b:
            out[a] |= 1 << b
    cnt = []

    def bump(d):
        while len(cnt) <= d:
            cnt.append(0)
        cnt[d] += 1

    def rec(P, d):
        Q = P
        while Q:
            low = Q & -Q
            w = low.bit_length() - 1
            Q ^= low
            bump(d)
            R = P & out[w]
            if R:
                rec(R, d + 1)

    for v in range(n):
        bump(0)
        rec(out[v], 1)
    return cnt


def count_networkx(n, edges):
    import networkx as nx

    es = set((a, b) for a, b in edges if a != b)
    G = nx.Graph()
    G.add_nodes_from(range(n))
    G.add_edges_from(es)
    cnt = []
    ncl = []
    for cl in nx.enumerate_all_cliques(G):
        k = len(cl)
        while len(cnt) < k:
            cnt.append(0)
            ncl.append(0)
        ncl[k - 1] += 1
        # f[mask] = number of orders of `mask` where every earlier vertex points to every later one
        f = [0] * (1 << k)
        f[0] = 1
        for mask in range(1, 1 << k):
            tot = 0
            for i in range(k):
                if mask >> i & 1:
                    rest = mask & ~(1 << i)
                    ok = True
                    for j in range(k):
                        if rest >> j & 1 and (cl[i], cl[j]) not in es:
                            ok = False
                            break
                    if ok:
                        tot += f[rest]
            f[mask] = tot
        cnt[k - 1] += f[(1 << k) - 1]
    while cnt and cnt[-1] == 0:
        cnt.pop()
    return cnt, ncl


def normalised(edges):
    """Bounds::calculate's normalised graph (src/lib.rs:125-129): one edge big->small per adjacent pair."""
    s = set((max(a, b), min(a, b)) for a, b in edges if a != b)
    return sorted(s)


# ---- synthetic recipe, numpy u64 arithmetic --------------------------------
M64 = np.uint64(0xFFFFFFFFFFFFFFFF)


def mix(z):
    z = (z ^ (z >> np.uint64(30))) * np.uint64(0xBF58476D1CE4E5B9)
    z = (z ^ (z >> np.uint64(27))) * np.uint64(0x94D049BB133111EB)
    return z ^ (z >> np.uint64(31))


def hij(seed, i, j):
    with np.errstate(over="ignore"):
        a = mix(np.uint64(seed) + np.uint64(0x9E3779B97F4A7C15) * (i.astype(np.uint64) + np.uint64(1)))
        b = np.uint64(0xD1B54A32D192ED03) * (j.astype(np.uint64) + np.uint64(1))
        return mix(a ^ b)


def u01(x):
    return (x >> np.uint64(11)).astype(np.float64) * (1.0 / 9007199254740992.0)


def gen_synth(kind, n, seed, params):
    ii = np.arange(n, dtype=np.uint64)
    edges = []
    for i in range(n):
        j = ii
        u = u01(hij(seed, np.full(n, i, dtype=np.uint64), j))
        if kind == "er":
            p = np.full(n, params[0])
        elif kind == "chunglu":
            w = (np.arange(n, dtype=np.float64) + 1.0) ** (-params[1])
            p = np.minimum(1.0, params[0] * w[i] * w)
        else:
            raise ValueError(kind)
        m = u < p
        m[i] = False
        js = np.nonzero(m)[0]
        edges.extend((i, int(x)) for x in js)
    return edges


def main():
    arrays = {}
    expected = {}
    for name, (known_fc, known_ncl) in SURVEY_KNOWN.items():
        n, edges = parse_flag(os.path.join(REF, name + ".flag"))
        a = count_bitset(n, edges)
        b, ncl = count_networkx(n, edges)
        c = count_bitset(n, normalised(edges))
        assert a == b == known_fc, (name, a, b, known_fc)
        assert c == ncl == known_ncl, (name, c, ncl, known_ncl)
        arrays[name] = np.array(edges, dtype=np.uint32)
        expected[name] = {"n": n, "nedges": len(edges), "flag_count": a, "ncliques": c,
                          "source": "reference fixture flag_file_examples/%s.flag" % name}
        print(name, a, c)

    toys = {
        # flag_file_examples/Testcases.pdf / example_flag_generator.py:42-77
        "ex00_3simplex": (4, [(0, 1), (0, 2), (0, 3), (1, 2), (1, 3), (2, 3)], [4, 6, 4, 1]),
        "ex03_bidirected_k4": (4, [(a, b) for a in range(4) for b in range(4) if a != b], [4, 12, 24, 24]),
        "directed_3cycle": (3, [(0, 1), (1, 2), (2, 0)], [3, 3]),
        "isolated_5": (5, [], [5]),
        "single_double_edge": (2, [(0, 1), (1, 0)], [2, 2]),
    }
    for name, (n, edges, known) in toys.items():
        a = count_bitset(n, edges)
        b, _ = count_networkx(n, edges)
        if not edges:
            b = [n]
        assert a == b == known, (name, a, b, known)
        arrays[name] = np.array(edges, dtype=np.uint32).reshape(-1, 2)
        expected[name] = {"n": n, "nedges": len(edges), "flag_count": a,
                          "ncliques": count_bitset(n, normalised(edges)), "source": "Testcases.pdf toy graph"}

    # C2: Chung-Lu n=279
    e2 = gen_synth("chunglu", 279, 3, (3.1908433806834573, 0.55))
    a = count_bitset(279, e2)
    b, ncl = count_networkx(279, e2)
    assert a == b == [279, 2200, 2743, 2431, 1839, 1054, 379, 64], (a, b)
    arrays["c2_chunglu_279"] = np.array(e2, dtype=np.uint32)
    expected["c2_chunglu_279"] = {"n": 279, "nedges": len(e2), "flag_count": a, "ncliques": ncl,
                                  "source": "synthetic Chung-Lu (SURVEY Appendix A), numpy restatement",
                                  "gen": {"kind": 1, "seed": 3, "params": [3.1908433806834573, 0.55]}}
    print("c2", a)

    # C3: ER n=2000 p=0.05 seed 1 — edges are too many to commit; keep counts + sha256 of the edge array
    e3 = gen_synth("er", 2000, 1, (0.05,))
    a = count_bitset(2000, e3)
    assert a == [2000, 199938, 999447, 250236, 3171, 2], a
    arr3 = np.array(e3, dtype=np.uint32)
    expected["c3_er_2000"] = {"n": 2000, "nedges": len(e3), "flag_count": a,
                              "ncliques": count_bitset(2000, normalised(e3)),
                              "sha256_edges_u32le": hashlib.sha256(arr3.tobytes()).hexdigest(),
                              "source": "synthetic ER (SURVEY Appendix A), numpy restatement",
                              "gen": {"kind": 0, "seed": 1, "params": [0.05]}}
    print("c3", a)

    # C4: too large for pure Python; the numbers are the survey session's (three-method) known answer.
    expected["c4_geo_31346"] = {"n": 31346, "nedges": 8198929,
                                "flag_count": [31346, 8198929, 103022472, 129416796, 24614274, 908696, 7504, 26],
                                "source": "SURVEY.md §8(d) session-computed; regenerated by tools/graphgen.cpp + oracle in tests",
                                "gen": {"kind": 2, "seed": 7, "params": [0.4328, 0.08, 0.4, 0.4, 2.0]}}

    np.savez_compressed(os.path.join(HERE, "fixtures.npz"), **arrays)
    with open(os.path.join(HERE, "expected.json"), "w") as f:
        json.dump(expected, f, indent=1)
    print("wrote", len(arrays), "graphs")


if __name__ == "__main__":
    sys.setrecursionlimit(10000)
    main()
