"""Time the C4 recount kernel (count_async x5, CUDA events) with whatever tree this file is run from."""
import sys, os
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import flag_complex_mcmc_b200 as fcm
from tools import synth
ctx = fcm.Context(0)
n, e = synth.config("c4_geo_31346")
g = fcm.DeviceGraph(ctx, n, e)
for _ in range(2):
    g.count_shard(0, 1)
ctx.timer_begin()
for _ in range(5):
    g.count_async(0, 1)
print("%s: %.3f ms per C4 count" % (ROOT, ctx.timer_end() / 5))
