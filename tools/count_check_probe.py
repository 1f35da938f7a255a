"""Quick GPU check of the recount after a kernel change: golden counts (fixtures, C2, C3, C4 known answer), for_each_cell
on two fixtures against the oracle, C4 time per count with the locality order on and off."""
import json, os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests"))
import numpy as np
import flag_complex_mcmc_b200 as fcm
from flag_complex_mcmc_b200 import directed_scm as ds
from tools import synth
import oracle_binding as orc

exp = json.load(open(os.path.join(ROOT, "tests", "golden", "expected.json")))
z = np.load(os.path.join(ROOT, "tests", "golden", "fixtures.npz"))
graphs = {k: (exp[k]["n"], z[k].reshape(-1, 2).astype(np.uint32)) for k in z.files if k in exp}
graphs["c3_er_2000"] = synth.config("c3_er_2000")
ctx = fcm.Context(0)
bad = 0
for relabel in (None, "1", "0"):
    if relabel is None: os.environ.pop("FCM_RELABEL", None)
    else: os.environ["FCM_RELABEL"] = relabel
    for name, (n, e) in graphs.items():
        g = fcm.DeviceGraph(ctx, n, e)
        ok = g.flagser_count() == exp[name]["flag_count"]
        g.close()
        if not ok: bad += 1; print("MISMATCH", name, relabel)
    for name, dmin, dmax in (("counterexample_any_order", 0, 3), ("c2_chunglu_279", 2, 3)):
        n, e = graphs[name]
        g = ds.DeviceGraph(ctx, n, e)
        ok = np.array_equal(g.cells(dmin, dmax), orc.cells(n, e, dmin, dmax))
        g.close()
        if not ok: bad += 1; print("CELLS MISMATCH", name, relabel)
n, e = synth.config("c4_geo_31346")
for relabel in (None, "0"):
    if relabel is None: os.environ.pop("FCM_RELABEL", None)
    else: os.environ["FCM_RELABEL"] = relabel
    g = fcm.DeviceGraph(ctx, n, e)
    ok = g.flagser_count() == exp["c4_geo_31346"]["flag_count"]
    if not ok: bad += 1
    g.count_async(0, 1); g.status()
    ctx.timer_begin()
    for _ in range(20): g.count_async(0, 1)
    ms = ctx.timer_end() / 20
    g.status(); g.close()
    print("C4 FCM_RELABEL=%s: %.3f ms per count, known answer %s" % (relabel, ms, ok))
print("CHECK", "FAILED %d" % bad if bad else "ok")
