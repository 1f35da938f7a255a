"""ctypes binding of libfcm.so (include/fcm.h).  There is no CPU fallback: if the library is
missing or no CUDA device is usable every compute call raises."""
import ctypes as C
import os

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
# FCM_LIB=<path of an experiment build, csrc/Makefile VARIANT=...> replaces the default library (measurements only)
LIB_PATH = os.environ.get("FCM_LIB") or os.path.join(HERE, "libfcm.so")

FCM_MAXD = 32
FCM_OK, FCM_EINVAL, FCM_ENODEVICE, FCM_ECUDA, FCM_ENOMEM = 0, -1, -2, -3, -4
FCM_EDIM, FCM_ENOTADJ, FCM_EASSERT, FCM_EUNSUPPORTED, FCM_ELIMIT = -5, -6, -7, -8, -9

# every symbol include/fcm.h declares (tests check the library exports all of them)
SYMBOLS = [
    "fcm_last_error", "fcm_abi_version", "fcm_device_count", "fcm_count_unweighted", "fcm_free",
    "fcm_ctx_create", "fcm_ctx_destroy", "fcm_ctx_set_stream", "fcm_ctx_synchronize",
    "fcm_ctx_timer_begin", "fcm_ctx_timer_end", "fcm_ctx_launches",
    "fcm_graph_create", "fcm_graph_create_normalised", "fcm_graph_destroy", "fcm_graph_count",
    "fcm_graph_count_async", "fcm_graph_set_kernel", "fcm_graph_info", "fcm_graph_cells",
    "fcm_edge_neighborhoods", "fcm_ensemble_create", "fcm_ensemble_destroy", "fcm_ensemble_set_bounds", "fcm_ensemble_set_kernel", "fcm_ensemble_step",
    "fcm_ensemble_step_trace", "fcm_ensemble_read", "fcm_ensemble_read_all", "fcm_ensemble_status",
    "fcm_ensemble_info", "fcm_ensemble_set_stats", "fcm_ensemble_index_info", "fcm_apply", "fcm_revert", "fcm_bounds_check",
    "fcm_ensemble_edgebits_info", "fcm_ensemble_edgebits_slots", "fcm_ensemble_edgebits",
    "fcm_ensemble_read_doubles", "fcm_ensemble_restore",
    "fcm_graph_status", "fcm_comm_create", "fcm_comm_connect", "fcm_comm_destroy",
    "fcm_graph_count_allreduce_async", "fcm_graph_count_allreduce", "fcm_graph_count_noalloc", "fcm_ensemble_set_policy", "fcm_maximal_cliques",
]


class FcmError(RuntimeError):
    def __init__(self, code, msg):
        super().__init__("libfcm error %d: %s" % (code, msg))
        self.code = code


class TraceRec(C.Structure):
    _fields_ = [(k, C.c_uint32) for k in ("flags", "a", "b", "dm0", "dm1", "K", "draws", "pad")]


TRACE_DTYPE = np.dtype([(k, np.uint32) for k in ("flags", "a", "b", "dm0", "dm1", "K", "draws", "pad")])

_lib = None


def lib():
    """Load libfcm.so (built in-tree by __graft_entry__.build()); raise loudly when absent."""
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.exists(LIB_PATH):
        raise RuntimeError("libfcm.so is not built (%s): run `python -c 'import __graft_entry__ as g; g.build()'`; "
                           "there is no CPU fallback" % LIB_PATH)
    L = C.CDLL(LIB_PATH)
    vp, u32, u64, i32 = C.c_void_p, C.c_uint32, C.c_uint64, C.c_int
    P = C.POINTER
    L.fcm_last_error.restype = C.c_char_p
    L.fcm_abi_version.restype = i32
    L.fcm_device_count.restype = i32
    L.fcm_count_unweighted.restype = vp
    L.fcm_count_unweighted.argtypes = [C.c_size_t, C.c_size_t, vp, P(C.c_size_t)]
    L.fcm_free.argtypes = [vp]
    L.fcm_free.restype = None
    L.fcm_ctx_create.argtypes = [i32, P(vp)]
    L.fcm_ctx_destroy.argtypes = [vp]
    L.fcm_ctx_set_stream.argtypes = [vp, vp]
    L.fcm_ctx_synchronize.argtypes = [vp]
    L.fcm_ctx_timer_begin.argtypes = [vp]
    L.fcm_ctx_timer_end.argtypes = [vp, P(C.c_float)]
    L.fcm_ctx_launches.argtypes = [vp]
    L.fcm_ctx_launches.restype = u64
    L.fcm_graph_create.argtypes = [vp, u32, u64, vp, P(vp)]
    L.fcm_graph_create_normalised.argtypes = [vp, u32, u64, vp, P(vp)]
    L.fcm_graph_destroy.argtypes = [vp]
    L.fcm_graph_count.argtypes = [vp, u32, u32, vp, P(i32)]
    L.fcm_graph_count_async.argtypes = [vp, u32, u32, P(vp)]
    L.fcm_graph_status.argtypes = [vp]
    L.fcm_comm_create.argtypes = [vp, i32, u32, u32, P(vp), vp]
    L.fcm_comm_connect.argtypes = [vp, vp]
    L.fcm_comm_destroy.argtypes = [vp]
    L.fcm_graph_count_allreduce_async.argtypes = [vp, vp, P(vp)]
    L.fcm_graph_count_allreduce.argtypes = [vp, vp, vp, P(i32)]
    L.fcm_graph_count_noalloc.argtypes = [vp, u32, u32, vp, P(i32)]
    L.fcm_graph_set_kernel.argtypes = [vp, i32]
    L.fcm_graph_info.argtypes = [vp, P(u32), P(u64), P(u32), P(u32)]
    L.fcm_graph_cells.argtypes = [vp, i32, i32, vp, u64, P(u64)]
    L.fcm_edge_neighborhoods.argtypes = [vp, u32, u64, vp, P(u64), P(u64), vp, vp, vp]
    L.fcm_maximal_cliques.argtypes = [u32, u64, vp, P(u64), P(u64), vp, vp]
    L.fcm_ensemble_create.argtypes = [vp, u32, u64, vp, u32, vp, vp, i32, vp, i32, vp, i32, vp, P(vp)]
    L.fcm_ensemble_destroy.argtypes = [vp]
    L.fcm_ensemble_set_bounds.argtypes = [vp, vp, i32, vp, i32]
    L.fcm_ensemble_set_kernel.argtypes = [vp, i32]
    L.fcm_ensemble_set_stats.argtypes = [vp, i32]
    L.fcm_ensemble_set_policy.argtypes = [vp, i32]
    L.fcm_ensemble_index_info.argtypes = [vp, P(i32), P(u64)]
    L.fcm_ensemble_step.argtypes = [vp, u64]
    L.fcm_ensemble_step_trace.argtypes = [vp, u64, u32, vp]
    L.fcm_ensemble_read.argtypes = [vp, u32, vp, P(i32), P(u64), P(u64), vp, P(u64), P(u64), vp]
    L.fcm_ensemble_read_all.argtypes = [vp, vp, vp, vp, vp, vp]
    L.fcm_ensemble_status.argtypes = [vp]
    L.fcm_ensemble_info.argtypes = [vp, P(u32), P(u64), P(u32), P(u32), P(u64)]
    L.fcm_apply.argtypes = [vp, u32, vp, i32, vp, P(i32), vp, P(i32)]
    L.fcm_revert.argtypes = [vp, u32, vp, i32, vp, i32, vp, i32]
    L.fcm_bounds_check.argtypes = [vp, u32, P(i32)]
    L.fcm_ensemble_edgebits_info.argtypes = [vp, P(u64), P(u64)]
    L.fcm_ensemble_edgebits_slots.argtypes = [vp, vp]
    L.fcm_ensemble_edgebits.argtypes = [vp, u32, u32, vp]
    L.fcm_ensemble_read_doubles.argtypes = [vp, u32, vp, P(u64)]
    L.fcm_ensemble_restore.argtypes = [vp, u32, vp, u64, vp, u64, vp, vp, i32, u64, u64]
    for name in SYMBOLS:
        f = getattr(L, name)
        if f.restype is C.c_int and name not in ("fcm_abi_version", "fcm_device_count"):
            f.restype = i32
    _lib = L
    return L


def check(rc):
    if rc != 0:
        raise FcmError(rc, lib().fcm_last_error().decode("utf-8", "replace"))


def edges_array(edges):
    """(E,2) uint32 C-contiguous view/copy of an edge list."""
    a = np.ascontiguousarray(np.asarray(edges, dtype=np.uint32).reshape(-1, 2))
    return a


def ptr(a):
    return a.ctypes.data_as(C.c_void_p) if a is not None and a.size else None
