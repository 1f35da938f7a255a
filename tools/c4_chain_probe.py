"""Probe: chains on the C4-sized graph (31 346 vertices, 8.2 M edges): the large-graph paths
(no clique index, row-gather delta path, lists > 64 -> full pre/post counts), checked against the oracle."""
import sys, os, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "tests"))
import numpy as np
import flag_complex_mcmc_b200 as fcm
import oracle_binding as orc
from tools import synth

nch = int(sys.argv[1]) if len(sys.argv) > 1 else 8
P = int(sys.argv[2]) if len(sys.argv) > 2 else 300
n, e = synth.config("c4_geo_31346")
ctx = fcm.Context(0)
t0 = time.time()
s = fcm.initialize_new_sampler(ctx, n, e, list(range(nch)), target_relaxation=0.01, simple=True)
print("create %.1f s" % (time.time() - t0), s.info(), s.index_info(), flush=True)
t0 = time.time()
tr = s.step_trace(P, 1)
print("GPU: %d chains x %d proposals in %.2f s; max K %d" % (nch, P, time.time() - t0, int(tr["K"].max())), flush=True)
s.status()
ctx.timer_begin(); s.step(P); ms = ctx.timer_end()
print("untraced: %.1f ms -> %.1f k proposals/s" % (ms, nch * P / ms), flush=True)
t0 = time.time()
o = orc.Sampler(n, e, 1, 0.01, nthreads=16)
print("oracle init %.1f s" % (time.time() - t0), flush=True)
w = o.steps(P)
for f in ("flags", "a", "b", "dm0", "dm1", "K"):
    assert np.array_equal(tr[f], w[f]), f
o.steps(P, trace=False)
r, ww = s.read(1), o.read()
assert r["flag_count"] == ww["flag_count"] and r["rng"] == ww["rng"] and np.array_equal(r["edges"], ww["edges"])
print("parity ok: trace of %d proposals + state after %d" % (P, 2 * P), r["flag_count"])
