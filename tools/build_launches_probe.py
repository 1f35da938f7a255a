"""Probe (run under `ncu --metrics gpu__time_duration.sum`): one C4 graph build + one count, to list the build's kernels."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import flag_complex_mcmc_b200 as fcm
from tools import synth
n, e = synth.config("c4_geo_31346")
ctx = fcm.Context(0)
g = fcm.DeviceGraph(ctx, n, e)
print(g.flagser_count())
g.close()
