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


def fnv(edges):
    h = 1469598103934665603
    for x in np.asarray(edges, dtype=np.uint64).reshape(-1):
        h = ((h ^ int(x)) * 1099511628211) & 0xFFFFFFFFFFFFFFFF
    return h


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
