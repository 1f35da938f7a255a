"""The C++ host side (include/directed_scm.hpp) above the C ABI: host maths on CPU, the reference's
sample flow (src/bin/sample.rs:80-150) on the GPU against the oracle."""
import json
import os
import subprocess

import numpy as np
import pytest

import oracle_binding as orc
from flag_complex_mcmc_b200 import directed_scm as ds

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


@pytest.fixture(scope="module")
def driver(tmp_path_factory):
    out = str(tmp_path_factory.mktemp("cpp") / "host_mirror_check")
    libdir = os.path.join(ROOT, "flag_complex_mcmc_b200")
    cmd = ["g++", "-O2", "-std=c++17", "-Wall", "-Werror", "-I", os.path.join(ROOT, "include"),
           os.path.join(ROOT, "tests", "cpp", "host_mirror_check.cpp"), "-o", out, "-L", libdir, "-lfcm",
           "-Wl,-rpath," + libdir]
    subprocess.check_call(cmd)
    return out


def test_cpp_host_maths_match_python_mirror_and_oracle(driver):
    r = json.loads(subprocess.check_output([driver, "selfcheck"]))
    assert r["abi"] == 2
    assert r["factorial"] == [1, 1, 1, 2, 6]               # src/util.rs:65-71: (x-1)!
    assert r["binomial_5_2"] == ds.binomial(5, 2) == 12
    assert r["oeis"] == list(ds.OEIS_A058298)              # generated as n!/(n-k), equals the reference's table
    sc = [279, 2200, 2743, 2431, 1839, 1054, 379, 64]
    assert r["relax_de"] == ds.calc_relax_de(sc)


def test_cpp_cliques_by_order_matches_oracle(driver, tmp_path, graphs):
    """compute_maximal_cliques + State::new's filing by order (src/lib.rs:41-49) through the C++ mirror: host code, no device"""
    for name in ("counterexample_any_order", "c2_chunglu_279"):
        n, e = graphs[name]
        f = str(tmp_path / (name + ".flag"))
        ds.save_flag_file(f, n, e)
        r = json.loads(subprocess.check_output([driver, "cliques", f]))
        want = orc.maximal_cliques(n, e)
        by = [0] * max(len(c) for c in want)
        h = 1469598103934665603
        for c in want:
            by[len(c) - 1] += 1
            for x in c:
                h = ((h ^ int(x)) * 1099511628211) & 0xFFFFFFFFFFFFFFFF
            h = ((h ^ 0xFF) * 1099511628211) & 0xFFFFFFFFFFFFFFFF
        assert r["by_order"] == by, name
        assert r["fnv"] == h, name


def fnv(edges):
    h = 1469598103934665603
    for x in np.asarray(edges, dtype=np.uint64).reshape(-1):
        h = ((h ^ int(x)) * 1099511628211) & 0xFFFFFFFFFFFFFFFF
    return h


FLAG_TEXTS = {
    # name: (file bytes, parses?)  — the reference splits on ' ' only and unwraps the parse (src/io.rs:25-31)
    "plain": (b"dim 0:\n1 1 1 1\ndim 1:\n0 1 1\n1 2 1\n3 0 1\n", True),
    "crlf": (b"dim 0:\r\n1 1 1 1\r\ndim 1:\r\n0 1 1\r\n1 2 1\r\n3 0 1\r\n", True),        # str::lines() drops the \r
    "blank_lines_and_runs_of_spaces": (b"dim 0:\n1  1 1   1 \ndim 1:\n\n0   1 1\n \n1 2\n3 0 0.5 junk\n", True),
    "no_weights_no_final_newline": (b"dim 0:\n1 1 1 1\ndim 1:\n0 1\n1 2\n3 0", True),
    "tab_separated": (b"dim 0:\n1 1 1 1\ndim 1:\n0\t1\t1\n", None),   # one token "0\t1\t1": < 2 tokens, line skipped
    "tab_inside_token": (b"dim 0:\n1 1 1 1\ndim 1:\n0\t1 2 1\n", False),   # token "0\t1" does not parse: reference panics
}


@pytest.mark.parametrize("name", sorted(FLAG_TEXTS))
def test_flag_reader_tokenises_like_the_reference(name, driver, tmp_path):
    """include/directed_scm.hpp, the Python mirror and the oracle read a .flag file exactly as src/io.rs:18-35 does:
    tokens split on ' ' only, a \r before the newline dropped, lines with fewer than two tokens skipped, a token that
    is not a number is an error (the reference's parse().unwrap() panics)."""
    data, ok = FLAG_TEXTS[name]
    f = str(tmp_path / "g.flag")
    with open(f, "wb") as fh:
        fh.write(data)
    p = subprocess.run([driver, "parse", f], capture_output=True, text=True)
    if ok is False:
        assert p.returncode != 0
        with pytest.raises(ValueError):
            ds.read_flag_file(f)
        with pytest.raises(Exception):
            orc.read_flag_file(f)
        return
    got = json.loads(p.stdout)
    n_py, e_py = ds.read_flag_file(f)
    n_or, e_or = orc.read_flag_file(f)
    want_edges = [] if ok is None else [(0, 1), (1, 2), (3, 0)]
    assert n_py == n_or == got["n"] == 4
    assert e_py.tolist() == [list(x) for x in want_edges] and e_or.tolist() == e_py.tolist()
    assert got["nedges"] == len(want_edges) and got["edges_fnv"] == fnv(e_py)


@pytest.mark.gpu
@pytest.mark.parametrize("name,seed,steps,simple", [("c2_chunglu_279", 3, 4000, 1), ("c2_chunglu_279", 4, 2500, 0),
                                                    ("counterexample_seo_greedy_5_start", 0, 3000, 1)])
def test_cpp_sample_flow_matches_oracle(name, seed, steps, simple, driver, graphs, tmp_path):
    n, e = graphs[name]
    f = str(tmp_path / "g.flag")
    # the reference's save_flag_file sorts the edges; the insertion order decides the trajectory, so the
    # oracle reads the same file
    ds.save_flag_file(f, n, e)
    n2, e2 = orc.read_flag_file(f)
    got = json.loads(subprocess.check_output([driver, "run", f, str(seed), str(steps), str(simple)]))
    o = orc.Sampler(n2, e2, seed, 0.01, move_weights=ds.MOVE_DISTRIBUTION_SIMPLE if simple else ds.MOVE_DISTRIBUTION)
    o.steps(steps, trace=False)
    w = o.read()
    assert got["n"] == n2 and got["nedges"] == e2.shape[0]
    assert got["bounds_min"] == w["bounds_min"] and got["bounds_max"] == w["bounds_max"]
    assert got["sample_distance"] == w["sample_distance"]
    assert got["flag_count"] == w["flag_count"]
    assert (got["sampled"], got["accepted"]) == (w["sampled"], w["accepted"])
    assert got["edges_fnv"] == fnv(w["edges"])
    assert got["bounds_ok"] == 1
