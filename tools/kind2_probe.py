"""Probe: row-gather kernel (kind 2) on the C2 graph against the oracle, first diverging proposal."""
import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "tests"))
import numpy as np
import flag_complex_mcmc_b200 as fcm
import oracle_binding as orc
z = np.load("tests/golden/fixtures.npz")
e = z["c2_chunglu_279"].reshape(-1, 2).astype(np.uint32)
n = 279
ctx = fcm.Context(0)
for kind in (2,):
    s = fcm.initialize_new_sampler(ctx, n, e, [0], target_relaxation=0.01, simple=True)
    s.set_kernel(kind)
    o = orc.Sampler(n, e, 0, 0.01)
    done = 0
    try:
        for chunk in [1] * 300:
            tr = s.step_trace(chunk, 0)
            w = o.steps(chunk)
            for f in ("flags", "a", "b", "dm0", "dm1", "K"):
                if not np.array_equal(tr[f], w[f]):
                    i = int(np.nonzero(tr[f] != w[f])[0][0])
                    print("kind", kind, "field", f, "differs at", done + i, tr[i], w[i]); raise SystemExit
            r, ww = s.read(0), o.read()
            if r["flag_count"] != ww["flag_count"]:
                print("kind", kind, "flag_count differs after", done + chunk, r["flag_count"], ww["flag_count"]); break
            done += chunk
        print("kind", kind, "ok through", done)
    except fcm.FcmError as ex:
        w = o.steps(1)
        print("kind", kind, "error after", done, ex, "oracle fc before", ww["flag_count"], "after", o.read()["flag_count"], "oracle rec", w[0])
    s.close()
