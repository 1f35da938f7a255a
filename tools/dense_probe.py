"""Probe: the dense-row count kernel (fcm_graph_set_kernel(1): one adjacency row of 8*ceil(n/64) bytes read per enumerated
simplex, 128-bit coalesced loads) on the C4 graph - the kernel the dense-row roofline model describes - beside the tile kernel."""
import sys, os, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import flag_complex_mcmc_b200 as fcm
from tools import synth
n, e = synth.config("c4_geo_31346")
ctx = fcm.Context(0)
g = fcm.DeviceGraph(ctx, n, e)
W = (n + 63) // 64
for kind, reps in ((0, 5), (1, 2)):
    g.set_kernel(kind)
    c = g.count_shard(0, 1)
    ctx.timer_begin()
    for _ in range(reps):
        g.count_async(0, 1)
    ms = ctx.timer_end() / reps
    simplices = int(c.sum())
    print("kernel kind %d: %.2f ms per C4 count, %.2f G simplices/s, dense-row model %.1f GB/s (8*%d B x %d simplices)" % (
        kind, ms, simplices / ms / 1e6, 8 * W * simplices / ms / 1e6, W, simplices), flush=True)
g.close()
