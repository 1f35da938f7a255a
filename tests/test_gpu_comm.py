"""GPU: the sharded recount's all-reduce inside libfcm.so (include/fcm.h 3b) — State::new's
`graph.flagser_count()` (src/lib.rs:51) split by source vertex.  The summed vector must equal the
single-GPU count (golden vectors) on every rank, for the fused last-CTA epilogue over peer windows
and for the NCCL kind.  Ranks sharing ONE device exercise the whole protocol on a 1-GPU box; the
two-device tests skip below 2 devices."""
import os
import threading

import numpy as np
import pytest

import flag_complex_mcmc_b200 as fcm
from flag_complex_mcmc_b200 import _native as N

pytestmark = pytest.mark.gpu
os.environ.setdefault("FCM_COMM_TIMEOUT_MS", "4000")


def ndev():
    return N.lib().fcm_device_count()


def make_ranks(devices, kind):
    """one context + comm + nothing else per rank, all in this process"""
    nr = len(devices)
    ctxs = [fcm.Context(d) for d in devices]
    comms = [fcm.Comm(ctxs[r], r, nr, kind) for r in range(nr)]
    if kind == fcm.COMM_NCCL:
        hs = [c.handle for c in comms]
        th = [threading.Thread(target=c.connect, args=(hs,)) for c in comms]   # ncclCommInitRank is a collective
        for t in th:
            t.start()
        for t in th:
            t.join()
    else:
        fcm.Comm.connect_local(comms)
    return ctxs, comms


def read_dev(addr, nwords=N.FCM_MAXD + 1):
    """device uint64[nwords] at addr -> numpy (cudaMemcpy D2H through the runtime libfcm.so already loaded)"""
    import ctypes as C
    rt = None
    for nm in ("libcudart.so.12", "libcudart.so"):
        try:
            rt = C.CDLL(nm)
            break
        except OSError:
            pass
    assert rt is not None
    buf = np.zeros(nwords, dtype=np.uint64)
    assert rt.cudaMemcpy(buf.ctypes.data_as(C.c_void_p), C.c_void_p(addr), C.c_size_t(buf.nbytes), C.c_int(2)) == 0
    return buf


def close_all(*groups):
    for g in groups:
        for x in g:
            x.close()


@pytest.mark.parametrize("kind", [fcm.COMM_PEER, fcm.COMM_NCCL])
def test_single_rank_allreduce_is_the_plain_count(kind, expected, graphs):
    ctxs, comms = make_ranks([0], kind)
    for name in ("c2_chunglu_279", "c3_er_2000"):
        n, e = graphs[name]
        g = fcm.DeviceGraph(ctxs[0], n, e)
        for _ in range(3):                                  # slots alternate
            assert fcm.directed_scm.trim(g.count_allreduce(comms[0])) == expected[name]["flag_count"], name
        g.close()
    close_all(comms, ctxs)


@pytest.mark.parametrize("nranks", [2, 3])
def test_peer_allreduce_ranks_on_one_device(nranks, expected, graphs):
    """nranks ranks on device 0, each with its own stream: the kernels overlap, every last CTA pushes into
    every window and waits for the others.  Seven calls in a row reuse both slots several times."""
    ctxs, comms = make_ranks([0] * nranks, fcm.COMM_PEER)
    n, e = graphs["c3_er_2000"]
    gs = [fcm.DeviceGraph(c, n, e) for c in ctxs]
    want = np.zeros(N.FCM_MAXD, dtype=np.uint64)
    want[:len(expected["c3_er_2000"]["flag_count"])] = expected["c3_er_2000"]["flag_count"]
    for it in range(7):
        dev = [gs[r].count_allreduce_async(comms[r]) for r in range(nranks)]
        for r in range(nranks):
            gs[r].status()
        for r in range(nranks):
            buf = read_dev(dev[r])
            assert np.array_equal(buf[:N.FCM_MAXD], want) and buf[N.FCM_MAXD] == 0, (it, r)
    # the synchronous call from one thread per rank
    got = [None] * nranks

    def run(r):
        got[r] = gs[r].count_allreduce(comms[r])
    th = [threading.Thread(target=run, args=(r,)) for r in range(nranks)]
    for t in th:
        t.start()
    for t in th:
        t.join()
    for r in range(nranks):
        assert np.array_equal(got[r], want), r
    close_all(gs, comms, ctxs)


def test_peer_allreduce_device_result_equals_sum(graphs):
    """read the all-reduced device buffer itself (not a recount) on both ranks"""
    ctxs, comms = make_ranks([0, 0], fcm.COMM_PEER)
    n, e = graphs["c2_chunglu_279"]
    gs = [fcm.DeviceGraph(c, n, e) for c in ctxs]
    whole = gs[0].count_shard(0, 1)
    got = [gs[r].count_allreduce_async(comms[r]) for r in range(2)]
    for r in range(2):
        gs[r].status()
    for r in range(2):
        buf = read_dev(got[r])
        assert np.array_equal(buf[:N.FCM_MAXD], whole), r
        assert buf[N.FCM_MAXD] == 0
    close_all(gs, comms, ctxs)


def test_allreduce_over_graphs_built_under_the_locality_order(expected, graphs, monkeypatch):
    """every rank renumbers the vertices on its own (graphs beyond 4096 vertices; forced here): the numbering is integer
    arithmetic on the edge list, so the ranks' shards still partition the sources and the sum is the plain count"""
    monkeypatch.setenv("FCM_RELABEL", "1")
    ctxs, comms = make_ranks([0, 0, 0], fcm.COMM_PEER)
    n, e = graphs["c3_er_2000"]
    gs = [fcm.DeviceGraph(c, n, e) for c in ctxs]
    got = [None] * 3

    def run(r):
        got[r] = gs[r].count_allreduce(comms[r])
    for it in range(2):
        th = [threading.Thread(target=run, args=(r,)) for r in range(3)]
        for t in th:
            t.start()
        for t in th:
            t.join()
        for r in range(3):
            assert fcm.directed_scm.trim(got[r]) == expected["c3_er_2000"]["flag_count"], (it, r)
    close_all(gs, comms, ctxs)


def test_missing_rank_times_out_instead_of_hanging(graphs, monkeypatch):
    monkeypatch.setenv("FCM_COMM_TIMEOUT_MS", "150")
    ctxs, comms = make_ranks([0, 0], fcm.COMM_PEER)
    n, e = graphs["c2_chunglu_279"]
    g = fcm.DeviceGraph(ctxs[0], n, e)
    with pytest.raises(fcm.FcmError) as ei:
        g.count_allreduce(comms[0])                         # rank 1 never calls
    assert ei.value.code == N.FCM_ECUDA
    with pytest.raises(fcm.FcmError):
        g.count_allreduce(comms[0])                         # the comm is broken now
    g.close()
    close_all(comms, ctxs)


def test_mismatched_handles_are_refused():
    ctx = fcm.Context(0)
    a = fcm.Comm(ctx, 0, 2, fcm.COMM_PEER)
    b = fcm.Comm(ctx, 1, 2, fcm.COMM_PEER)
    with pytest.raises(fcm.FcmError):
        a.connect([b.handle, a.handle])                     # wrong order
    close_all([a, b], [ctx])


@pytest.mark.parametrize("kind", [fcm.COMM_PEER, fcm.COMM_NCCL])
def test_two_devices_allreduce(kind, expected, graphs):
    if ndev() < 2:
        pytest.skip("needs 2 CUDA devices")
    nr = min(ndev(), 4)
    ctxs, comms = make_ranks(list(range(nr)), kind)
    n, e = graphs["c3_er_2000"]
    gs = [fcm.DeviceGraph(c, n, e) for c in ctxs]
    want = expected["c3_er_2000"]["flag_count"]
    for it in range(5):
        got = [None] * nr

        def run(r):
            got[r] = gs[r].count_allreduce(comms[r])
        th = [threading.Thread(target=run, args=(r,)) for r in range(nr)]
        for t in th:
            t.start()
        for t in th:
            t.join()
        for r in range(nr):
            assert fcm.directed_scm.trim(got[r]) == want, (it, r)
    close_all(gs, comms, ctxs)
