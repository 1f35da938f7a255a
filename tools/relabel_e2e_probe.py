"""Probe: C4 count with and without the library's locality order of the vertices (FCM_RELABEL=0): kernel time per count
on a built graph, graph build time, and host edge list -> counts through the stateless call."""
import os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import flag_complex_mcmc_b200 as fcm
from tools import synth

n, e = synth.config("c4_geo_31346")
ctx = fcm.Context(0)
for mode in ("0", "default", "0", "default"):
    if mode == "0":
        os.environ["FCM_RELABEL"] = "0"
    else:
        os.environ.pop("FCM_RELABEL", None)
    t0 = time.perf_counter()
    g = fcm.DeviceGraph(ctx, n, e); g.status(); ctx.synchronize()
    t_build = (time.perf_counter() - t0) * 1e3
    g.count_async(0, 1); g.status()
    ctx.timer_begin()
    for _ in range(20):
        g.count_async(0, 1)
    ms = ctx.timer_end() / 20
    g.status()
    c = g.flagser_count()
    g.close()
    ts = []
    for _ in range(4):
        t0 = time.perf_counter(); c2 = fcm.flagser_count(n, e); ts.append((time.perf_counter() - t0) * 1e3)
    print("FCM_RELABEL=%-7s build %.2f ms  count %.3f ms  edge list -> counts %.2f ms (median of %s)  ok=%s" % (
        mode, t_build, ms, float(np.median(ts[1:])), ["%.1f" % t for t in ts], c == c2))
