"""CPU-only estimate: how many 4-byte words / 32-byte sectors / 64-byte DRAM blocks of a chain's orientation bits one
proposal on the C3 graph looks up (own pair, two pairs per common neighbour, one per adjacent pair of common neighbours),
under the library's dense (big, small) pair numbering and under numberings that start every vertex run on a sector / block
boundary.  Result (3000 sampled pairs): 56.7 lookups -> 43 words in 23.3 sectors / 19.8 blocks today; aligned runs would
give 16.6 blocks for a 1.4x larger id space - not built (profiles/README.md, round-2 readings)."""
import sys, numpy as np
import os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from tools import synth
n, e = synth.config("c3_er_2000")
e = np.asarray(e)
big = np.maximum(e[:,0], e[:,1]); small = np.minimum(e[:,0], e[:,1])
key = np.unique(big.astype(np.int64) * n + small)
pb = key // n; ps = key % n
npairs = len(key)
print("pairs", npairs)
adj = np.zeros((n, n), dtype=bool); adj[pb, ps] = True; adj[ps, pb] = True
# numbering schemes: dict name -> array pos[q] (position of pair q in the orientation bit array, in pairs)
def dense(): return np.arange(npairs)
def aligned(gran):          # each big-vertex run starts at a multiple of gran pairs if it would straddle a gran boundary and... greedy
    pos = np.zeros(npairs, dtype=np.int64); cur = 0
    starts = np.searchsorted(pb, np.arange(n + 1))
    for v in range(n):
        L = starts[v + 1] - starts[v]
        if L == 0: continue
        # would straddle more gran-blocks than necessary?
        need = (L + gran - 1) // gran
        span = (cur + L - 1) // gran - cur // gran + 1
        if span > need: cur = (cur + gran - 1) // gran * gran
        pos[starts[v]:starts[v + 1]] = cur + np.arange(L); cur += L
    return pos, cur
schemes = {"dense sorted": (dense(), npairs)}
for g in (128, 256): schemes["greedy aligned %d" % g] = aligned(g)
def always(gran):
    pos = np.zeros(npairs, dtype=np.int64); cur = 0
    starts = np.searchsorted(pb, np.arange(n + 1))
    for v in range(n):
        L = starts[v + 1] - starts[v]
        if L == 0: continue
        cur = (cur + gran - 1) // gran * gran
        pos[starts[v]:starts[v + 1]] = cur + np.arange(L); cur += L
    return pos, cur
schemes["every row at 256"] = always(256)
pid = {}
for q in range(npairs): pid[(int(pb[q]), int(ps[q]))] = q
rng = np.random.default_rng(1)
samples = rng.integers(0, npairs, 3000)
res = {k: [0, 0, 0] for k in schemes}
nl = 0
for q in samples:
    u, v = int(pb[q]), int(ps[q])
    cn = np.nonzero(adj[u] & adj[v])[0]
    look = [q]
    for w in cn:
        look.append(pid[(max(u, w), min(u, w))]); look.append(pid[(max(v, w), min(v, w))])
    for i in range(len(cn)):
        for j in range(i + 1, len(cn)):
            if adj[cn[i], cn[j]]: look.append(pid[(int(cn[j]), int(cn[i]))])
    look = np.array(look); nl += len(look)
    for k, (pos, tot) in schemes.items():
        p = pos[look]
        res[k][0] += len(np.unique(p >> 4)); res[k][1] += len(np.unique(p >> 7)); res[k][2] += len(np.unique(p >> 8))
print("lookups per proposal %.1f" % (nl / len(samples)))
for k, r in res.items():
    print("%-20s words %.1f  32B sectors %.1f  64B blocks %.1f   id space %.2fx" % (k, r[0] / len(samples), r[1] / len(samples), r[2] / len(samples), schemes[k][1] / npairs))
