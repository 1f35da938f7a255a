"""Probe: phases of fcm_ensemble_create on the C3 graph (FCM_TIMING=1), simple and default move mix."""
import sys, os, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import flag_complex_mcmc_b200 as fcm
from tools import synth
n, e = synth.config("c3_er_2000")
ctx = fcm.Context(0)
for simple in (True, False):
    for rep in range(2):
        t0 = time.perf_counter()
        s = fcm.initialize_new_sampler(ctx, n, e, list(range(4096)), target_relaxation=0.01, simple=simple)
        print("simple=%s rep %d: initialize_new_sampler %.3f s" % (simple, rep, time.perf_counter() - t0), flush=True)
        s.close()
