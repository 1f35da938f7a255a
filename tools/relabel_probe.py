"""Probe: what would a locality order of the vertices be worth to the recount?  C4 (geometric graph, ids hash-scattered over
the box) counted as generated and with the vertices renumbered along the long axis of the box (positions recomputed from
SURVEY Appendix A's recipe; a library-side order would be BFS / RCM of the undirected graph - same kind of locality)."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import flag_complex_mcmc_b200 as fcm
from tools import synth

M = np.uint64(0xFFFFFFFFFFFFFFFF)


def mix(z):
    z = (z ^ (z >> np.uint64(30))) * np.uint64(0xBF58476D1CE4E5B9)
    z = (z ^ (z >> np.uint64(27))) * np.uint64(0x94D049BB133111EB)
    return z ^ (z >> np.uint64(31))


def coord(seed, n, k, box):
    with np.errstate(over="ignore"):
        i = np.arange(n, dtype=np.uint64)
        s = np.uint64(seed ^ 0xABCDEF)
        h = mix(mix(s + np.uint64(0x9E3779B97F4A7C15) * (i + np.uint64(1))) ^ (np.uint64(0xD1B54A32D192ED03) * np.uint64(k + 1)))
    return (h >> np.uint64(11)).astype(np.float64) * 2.0 ** -53 * box


def timed(ctx, g, reps=20):
    g.count_async(0, 1); g.status()
    ctx.timer_begin()
    for _ in range(reps):
        g.count_async(0, 1)
    ms = ctx.timer_end() / reps
    g.status()
    return ms


kind, n, seed, params = synth.CONFIGS["c4_geo_31346"]
n, e = synth.config("c4_geo_31346")
ctx = fcm.Context(0)
g = fcm.DeviceGraph(ctx, n, e)
c0 = g.flagser_count()
print("as generated : %.3f ms per count" % timed(ctx, g), c0)
g.close()
x, y, z = (coord(seed, n, k, params[2 + k]) for k in range(3))
for name, keyf in (("z order", z), ("cells of 0.1 along z, then y, then x", np.floor(z / 0.1) * 1e4 + np.floor(y / 0.1) * 1e2 + np.floor(x / 0.1))):
    order = np.argsort(keyf, kind="stable")
    newid = np.empty(n, dtype=np.uint32); newid[order] = np.arange(n, dtype=np.uint32)
    e2 = newid[e]
    g2 = fcm.DeviceGraph(ctx, n, np.ascontiguousarray(e2))
    c2 = g2.flagser_count()
    d = np.abs(e2[:, 0].astype(np.int64) - e2[:, 1].astype(np.int64))
    print("%-40s: %.3f ms per count, counts equal: %s, median |i-j| %d, 95%% %d" % (name, timed(ctx, g2), c2 == c0, np.median(d), np.percentile(d, 95)))
    g2.close()

# orders a library can compute without positions: BFS levels of the undirected graph (ties by id)
import scipy.sparse as sp
from scipy.sparse.csgraph import breadth_first_order, reverse_cuthill_mckee
A = sp.coo_matrix((np.ones(len(e), dtype=np.int8), (e[:, 0].astype(np.int64), e[:, 1].astype(np.int64))), shape=(n, n)).tocsr()
A = ((A + A.T) > 0).astype(np.int8).tocsr()


def levels(root):
    lev = np.full(n, 2 ** 31 - 1, dtype=np.int64); lev[root] = 0
    frontier = np.array([root]); L = 0
    while len(frontier):
        nxt = np.unique(A[frontier].indices)
        nxt = nxt[lev[nxt] > L + 1]
        lev[nxt] = L + 1; frontier = nxt; L += 1
    return lev


l0 = levels(0)
far = int(np.argmax(np.where(l0 < 2 ** 31 - 1, l0, -1)))
l1 = levels(far)
orders = {"BFS levels from vertex 0 (%d levels)" % (l0[l0 < 2 ** 31 - 1].max() + 1): np.argsort(l0, kind="stable"),
          "BFS levels from the farthest vertex (%d levels)" % (l1[l1 < 2 ** 31 - 1].max() + 1): np.argsort(l1, kind="stable"),
          "scipy reverse Cuthill-McKee": np.asarray(reverse_cuthill_mckee(A, symmetric_mode=True))}
for name, order in orders.items():
    newid = np.empty(n, dtype=np.uint32); newid[order] = np.arange(n, dtype=np.uint32)
    e2 = newid[e]
    g2 = fcm.DeviceGraph(ctx, n, np.ascontiguousarray(e2))
    c2 = g2.flagser_count()
    d = np.abs(e2[:, 0].astype(np.int64) - e2[:, 1].astype(np.int64))
    print("%-48s: %.3f ms per count, counts equal: %s, median |i-j| %d, 95%% %d" % (name, timed(ctx, g2), c2 == c0, np.median(d), np.percentile(d, 95)))
    g2.close()
