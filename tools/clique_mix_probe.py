"""Probe: proposals/s with the default move mix (clique moves included) on a config graph."""
import sys, time, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import flag_complex_mcmc_b200 as fcm
from tools import synth

name = sys.argv[1] if len(sys.argv) > 1 else "c3_er_2000"
nch = int(sys.argv[2]) if len(sys.argv) > 2 else 1024
P = int(sys.argv[3]) if len(sys.argv) > 3 else 200
stats = len(sys.argv) > 4 and sys.argv[4] == "stats"      # default: statistic off, the kernel bench.py times
ctx = fcm.Context(0)
n, e = synth.config(name)
t0 = time.time()
s = fcm.initialize_new_sampler(ctx, n, e, list(range(nch)), target_relaxation=0.01, simple=False)
print("create %.2f s" % (time.time() - t0), s.info())
s.set_stats(stats)
s.step(P); ctx.synchronize(); s.status()
ctx.timer_begin()
s.step(P)
ms = ctx.timer_end()
s.status()
r = s.read_all()
print("%s: %d chains x %d proposals: %.2f ms -> %.3f M proposals/s, acceptance %.3f, mean K %.1f" % (
    name, nch, P, ms, nch * P / ms / 1e3, r["accepted"].sum() / r["sampled"].sum(), r["sumK"] / max(r["nonempty"], 1)))
