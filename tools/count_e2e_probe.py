"""Probe: host-to-host latency of the stateless count (fcm_count_unweighted) on a tiny and on the C4 graph."""
import sys, os, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import flag_complex_mcmc_b200 as fcm
from tools import synth
rng = np.random.default_rng(1)
n = 24
e = np.array([(i, j) for i in range(n) for j in range(n) if i != j and rng.random() < 0.3], dtype=np.uint32)
fcm.flagser_count(n, e)
t0 = time.perf_counter()
for _ in range(300):
    c = fcm.flagser_count(n, e)
print("tiny graph (n=24, E=%d): %.1f us per fcm_count_unweighted call" % (len(e), (time.perf_counter() - t0) / 300 * 1e6), c)
n4, e4 = synth.config("c4_geo_31346")
fcm.flagser_count(n4, e4)
t0 = time.perf_counter()
for _ in range(3):
    c = fcm.flagser_count(n4, e4)
print("C4: %.1f ms per fcm_count_unweighted call (host edge list in, counts out)" % ((time.perf_counter() - t0) / 3 * 1e3), c)
