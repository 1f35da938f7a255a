"""Probe: one dense fuzz case (24 vertices, 324 edges, simplices up to dimension 11) on a single chain,
timed in chunks, compared with the oracle chunk by chunk."""
import os, sys, time
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests"))
import numpy as np
import oracle_binding as orc
import flag_complex_mcmc_b200 as fcm
from flag_complex_mcmc_b200 import directed_scm as ds
from test_gpu_count import rand_graph

block, want_case = int(sys.argv[1]), int(sys.argv[2])
chunk = int(sys.argv[3]) if len(sys.argv) > 3 else 25
r = np.random.default_rng(1000 + block)
mixes = [ds.MOVE_DISTRIBUTION, (0.0, 0.0, 1.0, 0.0), (0.0, 0.0, 0.0, 1.0), (0.25, 0.25, 0.25, 0.25), (0.0, 0.2, 0.4, 0.4)]
for case in range(12):
    n = int(r.integers(5, 44))
    e = rand_graph(n, float(r.uniform(0.1, 0.65)), int(r.integers(1 << 30)), pdouble=float(r.choice([0.0, 0.2, 0.6])))
    if e.shape[0] == 0:
        continue
    weights = mixes[int(r.integers(len(mixes)))]
    seed = int(r.integers(1 << 40))
    relax = float(r.choice([0.01, 0.05, 0.5]))
    try:
        o = orc.Sampler(n, e, seed, relax, move_weights=weights)
    except RuntimeError:
        continue
    kern = int(r.choice([0, 0, 2]))
    if case != want_case:
        continue
    ctx = fcm.Context(0)
    w = o.read()
    fc0 = fcm.flagser_count(n, e, ctx)
    print("n", n, "m", e.shape[0], "mix", weights, "kernel", kern, "fc0", fc0, flush=True)
    s = fcm.MCMCSampler(ctx, n, e, [seed], ds.Bounds(w["bounds_min"], w["bounds_max"]), fc0, weights, w["sample_distance"])
    s.set_kernel(kern)
    for c in range(500 // chunk):
        t0 = time.time(); want = o.steps(chunk); t1 = time.time()
        got = s.step_trace(chunk, 0); ctx.synchronize(); t2 = time.time()
        same = all(np.array_equal(got[f], want[f]) for f in ("flags", "a", "b", "dm0", "dm1", "K"))
        print("chunk %d: cpu %.3f s gpu %.3f s same %s maxnedit %d fc %s" % (c, t1 - t0, t2 - t1, same, int(want["dm1"].max()), o.read()["flag_count"]), flush=True)
    break
