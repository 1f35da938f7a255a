"""Host-side partition of the hot path over the GPUs of one box.  Pure numpy: the
same functions drive bench.py (NCCL) and the world_size-2 gloo tests on CPU.

  chains   independent units (src/bin/all_cxs.rs:33-41 runs them as independent owners): chain k of
           the ensemble lives on rank k mod world; no data-path collective.
  recount  independent per source vertex v0 (every simplex is counted at its first vertex): sources
           in out-degree-descending order are dealt round-robin to the ranks, each rank counts its
           shard (fcm_graph_count(part=rank, nparts=world)), one allreduce(sum) of uint64[FCM_MAXD].
"""
import numpy as np

FCM_MAXD = 32


def chain_seeds(first_seed, nchains_total, rank, world):
    """Seeds of the chains rank `rank` owns: chain k (seed first_seed + k) -> rank k mod world."""
    k = np.arange(rank, nchains_total, world, dtype=np.uint64)
    return k + np.uint64(first_seed)


def source_order(n, edges):
    """Vertices by descending out-degree of the DISTINCT edge set, ties by ascending id — the order
    libfcm builds on the device (fcm_count.cu, `order`)."""
    e = np.asarray(edges, dtype=np.uint64).reshape(-1, 2)
    key = np.unique(e[:, 0] * np.uint64(n) + e[:, 1]) if e.size else np.zeros(0, dtype=np.uint64)
    deg = np.bincount((key // np.uint64(n)).astype(np.int64), minlength=n) if key.size else np.zeros(n, dtype=np.int64)
    return np.lexsort((np.arange(n), -deg)).astype(np.uint32)


def source_shard(order, part, nparts):
    """Sources of shard `part` of `nparts`: order[part], order[part + nparts], ..."""
    return np.asarray(order)[part::nparts]


def pad_counts(v):
    out = np.zeros(FCM_MAXD, dtype=np.int64)
    out[:len(v)] = np.asarray(v, dtype=np.int64)
    return out
