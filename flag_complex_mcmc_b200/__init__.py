"""flag_complex_mcmc_b200 — B200-native (sm_100a) hot path of flag-complex-mcmc:
directed-flag-complex simplex counting and the per-move local delta + bounds test,
behind the C ABI of include/fcm.h (libfcm.so).  No CPU fallback."""
from ._native import FCM_MAXD, FcmError, lib  # noqa: F401
from .directed_scm import (BitOutput, Bounds, COMM_NCCL, COMM_PEER, Comm, Context, DeviceGraph, MCMCSampler, MOVE_DISTRIBUTION,  # noqa: F401
                           MOVE_DISTRIBUTION_SIMPLE, all_le, binomial, calc_relax_de, cliques_by_order, compute_edge_neighborhoods, compute_maximal_cliques, count_cells, factorial,
                           flagser_count, for_each_cell, initialize_new_sampler, load_state, read_flag_file, save_flag_file, save_state, undirected_edges)
