"""CPU, world_size 2 over gloo: the multi-GPU partition of the hot path — sources of a
full recount dealt round-robin in degree order + allreduce(sum) of the uint64 count vectors; chains
dealt k mod world with no data-path collective.  The per-rank compute here is the oracle (no GPU in
this tier of the suite); bench.py runs the same host logic over NCCL with the CUDA kernels."""
import os
import socket
import sys

import numpy as np
import pytest

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(HERE)


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


def _worker(rank, world, port, n, edges, full, q):
    sys.path.insert(0, ROOT)
    sys.path.insert(0, HERE)
    import torch
    import torch.distributed as dist
    import oracle_binding as orc
    from flag_complex_mcmc_b200 import sharding as sh
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        # ---- recount: this rank's sources, then one allreduce of int64[FCM_MAXD] ----
        order = sh.source_order(n, edges)
        mine = sh.source_shard(order, rank, world)
        part = orc.count_sources(n, edges, mine)
        t = torch.from_numpy(sh.pad_counts(part))
        dist.all_reduce(t)
        tot = [int(x) for x in t.numpy()]
        while tot and tot[-1] == 0:
            tot.pop()
        # every vertex is somebody's source exactly once
        cnt = torch.zeros(n, dtype=torch.int64)
        cnt[torch.from_numpy(mine.astype(np.int64))] = 1
        dist.all_reduce(cnt)
        # ---- chains: k mod world, no collective on the data path; only the summary is reduced ----
        seeds = sh.chain_seeds(100, 5, rank, world)
        acc = 0
        fcs = []
        for sd in seeds:
            o = orc.Sampler(n, edges, int(sd), 0.05)
            o.steps(300, trace=False)
            r = o.read()
            acc += r["accepted"]
            fcs.append((int(sd), r["flag_count"]))
            o.close()
        s = torch.tensor([len(seeds), acc], dtype=torch.int64)
        dist.all_reduce(s)
        q.put((rank, tot == full, bool((cnt == 1).all()), [int(x) for x in s.numpy()], fcs, [int(x) for x in seeds]))
    finally:
        dist.destroy_process_group()


@pytest.mark.timeout(300)
def test_world2_recount_allreduce_and_chain_sharding(fixtures, expected):
    import torch.multiprocessing as mp
    sys.path.insert(0, HERE)
    import oracle_binding as orc
    name = "c2_chunglu_279"
    n, edges = expected[name]["n"], fixtures[name]
    full = expected[name]["flag_count"]
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = _free_port()
    procs = [ctx.Process(target=_worker, args=(r, 2, port, n, edges, full, q)) for r in range(2)]
    for p in procs:
        p.start()
    res = sorted(q.get(timeout=240) for _ in procs)
    for p in procs:
        p.join(60)
        assert p.exitcode == 0
    assert [r[1] for r in res] == [True, True]            # both ranks hold the full flagser_count
    assert [r[2] for r in res] == [True, True]
    assert res[0][5] == [100, 102, 104] and res[1][5] == [101, 103]
    assert res[0][3] == res[1][3] and res[0][3][0] == 5   # 5 chains in total, same summary on both
    # a chain's result does not depend on which rank ran it
    o = orc.Sampler(n, edges, 103, 0.05)
    o.steps(300, trace=False)
    assert dict(res[1][4])[103] == o.read()["flag_count"]


def test_source_order_is_degree_descending_and_stable():
    from flag_complex_mcmc_b200 import sharding as sh
    e = np.array([[0, 1], [0, 2], [3, 1], [3, 2], [3, 0], [2, 1], [2, 1]], dtype=np.uint32)   # duplicate line ignored
    assert sh.source_order(5, e).tolist() == [3, 0, 2, 1, 4]
    assert sh.source_shard(sh.source_order(5, e), 1, 2).tolist() == [0, 1]
    assert sh.chain_seeds(7, 5, 1, 2).tolist() == [8, 10]
