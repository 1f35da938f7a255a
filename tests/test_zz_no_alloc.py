"""GPU: fcm_graph_count_noalloc (include/fcm.h) - the dense-row count with its scratch in a file-scope device array,
for callers that cannot call cudaMalloc.  No counterpart in the reference.
Called through the C ABI and compared, bit for bit (integer counts: tolerance 0, as everywhere in this suite), with the
existing entry point fcm_graph_count, on two calls in a row - the second call finds the first one's contents in the
shared array.  Collected last on purpose."""
import ctypes as C

import numpy as np
import pytest

import flag_complex_mcmc_b200 as fcm
from flag_complex_mcmc_b200 import _native as N
from tools import synth

pytestmark = pytest.mark.gpu


def count_via(fn, g, part, nparts):
    out = np.zeros(N.FCM_MAXD, dtype=np.uint64)
    ln = C.c_int()
    rc = fn(g.h, part, nparts, N.ptr(out), C.byref(ln))
    return rc, out, ln.value


def test_noalloc_count_equals_the_existing_entry_point_twice_in_a_row():
    L = N.lib()
    ctx = fcm.Context(0)
    import os
    graphs = [synth.config("c3_er_2000")]
    fx = np.load(os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "fixtures.npz"))
    graphs.append((279, fx["c2_chunglu_279"].reshape(-1, 2).astype(np.uint32)))
    graphs.append((5, np.zeros((0, 2), dtype=np.uint32)))              # edgeless: counts = [5]
    for n, e in graphs:
        g = fcm.DeviceGraph(ctx, n, e)
        for part, nparts in ((0, 1), (1, 3)):
            rc, want, wlen = count_via(L.fcm_graph_count, g, part, nparts)
            assert rc == 0
            for call in range(2):                                      # two calls in a row share the static scratch
                rc, got, glen = count_via(L.fcm_graph_count_noalloc, g, part, nparts)
                assert rc == 0, L.fcm_last_error()
                assert np.array_equal(got, want) and glen == wlen, (n, part, nparts, call)
        # a different graph in between must not leak into the next call either
        g.close()
    ctx.close()


def test_noalloc_count_rejects_graphs_beyond_the_static_scratch():
    L = N.lib()
    ctx = fcm.Context(0)
    n = 64 * 64 + 1                                                    # rows of 66 words > FCM_NOALLOC_MAXW = 64
    e = np.array([[0, 1], [1, 2], [0, 2], [n - 1, 0]], dtype=np.uint32)
    g = fcm.DeviceGraph(ctx, n, e)
    rc, _, _ = count_via(L.fcm_graph_count_noalloc, g, 0, 1)
    assert rc == N.FCM_ELIMIT
    assert b"FCM_NOALLOC_MAXW" in L.fcm_last_error()
    rc, want, wlen = count_via(L.fcm_graph_count, g, 0, 1)             # the allocating entry point still serves it
    assert rc == 0 and [int(x) for x in want[:wlen]] == [n, 4, 1]
    g.close()
    g = fcm.DeviceGraph(ctx, 4096, e[:3])                              # the largest graph that fits: rows of 64 words
    rc, got, glen = count_via(L.fcm_graph_count_noalloc, g, 0, 1)
    assert rc == 0 and [int(x) for x in got[:glen]] == [4096, 3, 1]
    g.close()
    ctx.close()
