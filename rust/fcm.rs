// rust/fcm.rs - the reference-side binding of libfcm.so (include/fcm.h), to be dropped into the reference crate as
// `src/fcm.rs` next to its dead precedent `src/flagser.rs` (TheJonny/flag-complex-mcmc, src/flagser.rs:7-21).
// This is the `extern "C"` block of INTEGRATION.md section 2 as a file.  It cannot be compiled in this repository's image
// (no rustc / cargo); it only uses core::ffi types and mirrors the C declarations one to one.
// Entry points added in round 2 (recount all-reduce, status, no-allocation count) are at the end.
#![allow(dead_code)]

use libc::{c_char, c_int, c_void, size_t};
use crate::Node;

#[repr(C)] pub struct FcmCtx { _p: [u8; 0] }
#[repr(C)] pub struct FcmGraph { _p: [u8; 0] }
#[repr(C)] pub struct FcmEnsemble { _p: [u8; 0] }
pub const FCM_MAXD: usize = 32;

#[repr(C)] #[derive(Clone, Copy, Default)]
pub struct FcmTraceRec { pub flags: u32, pub a: u32, pub b: u32, pub dm0: u32, pub dm1: u32,
                         pub k: u32, pub draws: u32, pub pad: u32 }

#[link(name = "fcm")]
extern "C" {
    pub fn fcm_last_error() -> *const c_char;
    pub fn fcm_abi_version() -> c_int;
    pub fn fcm_device_count() -> c_int;
    // src/flagser.rs:9 — identical signature and malloc/free ownership
    pub fn fcm_count_unweighted(nvertices: size_t, nedges: size_t, edges: *const [Node; 2],
                                res_size: *mut size_t) -> *mut size_t;
    pub fn fcm_free(p: *mut c_void);
    pub fn fcm_ctx_create(device: c_int, out: *mut *mut FcmCtx) -> c_int;
    pub fn fcm_ctx_destroy(ctx: *mut FcmCtx) -> c_int;
    pub fn fcm_ctx_synchronize(ctx: *mut FcmCtx) -> c_int;
    pub fn fcm_graph_create(ctx: *mut FcmCtx, n: u32, nedges: u64, edges: *const [Node; 2],
                            out: *mut *mut FcmGraph) -> c_int;
    pub fn fcm_graph_create_normalised(ctx: *mut FcmCtx, n: u32, nedges: u64, edges: *const [Node; 2],
                                       out: *mut *mut FcmGraph) -> c_int;
    pub fn fcm_graph_destroy(g: *mut FcmGraph) -> c_int;
    pub fn fcm_graph_count(g: *mut FcmGraph, part: u32, nparts: u32, counts: *mut u64, len: *mut c_int) -> c_int;
    pub fn fcm_graph_cells(g: *mut FcmGraph, dim_min: c_int, dim_max: c_int, cells_out: *mut u32, cap: u64,
                           ncells: *mut u64) -> c_int;
    pub fn fcm_edge_neighborhoods(ctx: *mut FcmCtx, n: u32, nedges: u64, edges: *const [Node; 2], npairs: *mut u64,
                                  list_len: *mut u64, pairs_out: *mut u32, off_out: *mut u64, list_out: *mut u32) -> c_int;
    pub fn fcm_ensemble_create(ctx: *mut FcmCtx, n: u32, nedges: u64, edges: *const [Node; 2],
                               nchains: u32, seeds: *const u64, flag_count0: *const u64, len0: c_int,
                               bounds_min: *const u64, nmin: c_int, bounds_max: *const u64, nmax: c_int,
                               move_weights: *const f64, out: *mut *mut FcmEnsemble) -> c_int;
    pub fn fcm_ensemble_destroy(e: *mut FcmEnsemble) -> c_int;
    pub fn fcm_ensemble_set_bounds(e: *mut FcmEnsemble, bmin: *const u64, nmin: c_int,
                                   bmax: *const u64, nmax: c_int) -> c_int;
    pub fn fcm_ensemble_step(e: *mut FcmEnsemble, nproposals: u64) -> c_int;
    pub fn fcm_ensemble_step_trace(e: *mut FcmEnsemble, nproposals: u64, chain: u32,
                                   trace_out: *mut FcmTraceRec) -> c_int;
    pub fn fcm_ensemble_read(e: *mut FcmEnsemble, chain: u32, flag_count_out: *mut u64, len_out: *mut c_int,
                             sampled_out: *mut u64, accepted_out: *mut u64, edges_out: *mut u32,
                             nedges_out: *mut u64, ndouble_out: *mut u64, rng_out: *mut u64) -> c_int;
    pub fn fcm_ensemble_read_all(e: *mut FcmEnsemble, flag_counts: *mut u64, lens: *mut c_int,
                                 sampled: *mut u64, accepted: *mut u64, stats: *mut u64) -> c_int;
    pub fn fcm_ensemble_status(e: *mut FcmEnsemble) -> c_int;
    pub fn fcm_apply(e: *mut FcmEnsemble, chain: u32, edits3: *const u32, nedits: c_int,
                     pre: *mut u64, prelen: *mut c_int, post: *mut u64, postlen: *mut c_int) -> c_int;
    pub fn fcm_revert(e: *mut FcmEnsemble, chain: u32, edits3: *const u32, nedits: c_int,
                      pre: *const u64, prelen: c_int, post: *const u64, postlen: c_int) -> c_int;
    pub fn fcm_bounds_check(e: *mut FcmEnsemble, chain: u32, ok: *mut c_int) -> c_int;
    // io::save_state / load_state (src/io.rs:51-62)
    pub fn fcm_ensemble_read_doubles(e: *mut FcmEnsemble, chain: u32, pairs_out: *mut u32, ndouble_out: *mut u64) -> c_int;
    pub fn fcm_ensemble_restore(e: *mut FcmEnsemble, chain: u32, edges: *const u32, nedges: u64, doubles: *const u32,
                                ndouble: u64, rng4: *const u64, flag_count: *const u64, len: c_int,
                                sampled: u64, accepted: u64) -> c_int;
    // io::BitOutput (src/io.rs:128-212)
    pub fn fcm_ensemble_edgebits_info(e: *const FcmEnsemble, nslots: *mut u64, nbytes: *mut u64) -> c_int;
    pub fn fcm_ensemble_edgebits_slots(e: *mut FcmEnsemble, slots_out: *mut u32) -> c_int;
    pub fn fcm_ensemble_edgebits(e: *mut FcmEnsemble, chain_first: u32, nchains: u32, out: *mut u8) -> c_int;
}

/// the reference panics where the ABI returns an error code (src/lib.rs:65,86,104,309)
pub fn check(rc: c_int) {
    if rc != 0 {
        let msg = unsafe { std::ffi::CStr::from_ptr(fcm_last_error()) }.to_string_lossy().into_owned();
        panic!("libfcm: {} ({})", msg, rc);
    }
}

/// drop-in for flagser::count_unweighted (src/flagser.rs:12-21)
pub fn count_unweighted(nvertices: size_t, edges: &[[Node; 2]]) -> Vec<usize> {
    let mut nres: usize = 0;
    let p = unsafe { fcm_count_unweighted(nvertices, edges.len(), edges.as_ptr(), &mut nres) };
    assert!(!p.is_null(), "libfcm: count failed");
    let v = unsafe { std::slice::from_raw_parts(p, nres) }.to_vec();
    unsafe { fcm_free(p as *mut _) };
    v
}

// ---- round 2 additions (include/fcm.h sections 3, 3b) ----
#[repr(C)] pub struct FcmComm { _p: [u8; 0] }
pub const FCM_COMM_HANDLE_BYTES: usize = 256;
pub const FCM_COMM_PEER: c_int = 0;
pub const FCM_COMM_NCCL: c_int = 1;
pub const FCM_NOALLOC_MAXW: u32 = 64;
#[link(name = "fcm")]
extern "C" {
    pub fn fcm_graph_status(g: *mut FcmGraph) -> c_int;
    pub fn fcm_ensemble_set_policy(e: *mut FcmEnsemble, policy: c_int) -> c_int;
    pub fn fcm_comm_create(ctx: *mut FcmCtx, kind: c_int, rank: u32, nranks: u32, out: *mut *mut FcmComm,
                           handle_out: *mut u8) -> c_int;
    pub fn fcm_comm_connect(c: *mut FcmComm, handles: *const u8) -> c_int;
    pub fn fcm_comm_destroy(c: *mut FcmComm) -> c_int;
    pub fn fcm_graph_count_allreduce_async(g: *mut FcmGraph, c: *mut FcmComm, counts_dev: *mut *mut u64) -> c_int;
    pub fn fcm_graph_count_allreduce(g: *mut FcmGraph, c: *mut FcmComm, counts: *mut u64, len: *mut c_int) -> c_int;
    // no counterpart in the reference (no cudaMalloc inside the call)
    pub fn fcm_graph_count_noalloc(g: *mut FcmGraph, part: u32, nparts: u32, counts: *mut u64, len: *mut c_int) -> c_int;
}

// ---- the remaining entry points of include/fcm.h: stream / timer / launch counter of a context, asynchronous count,
//      kernel choice and size queries.  The reference has no counterpart (they are what bench.py and the tests use to
//      measure and to compare kernels); declared so that this file mirrors the header completely
//      (tests/test_abi.py::test_rust_shim_declares_every_entry_point). ----
#[link(name = "fcm")]
extern "C" {
    pub fn fcm_ctx_set_stream(ctx: *mut FcmCtx, cuda_stream: *mut c_void) -> c_int;
    pub fn fcm_ctx_timer_begin(ctx: *mut FcmCtx) -> c_int;
    pub fn fcm_ctx_timer_end(ctx: *mut FcmCtx, ms_out: *mut f32) -> c_int;
    pub fn fcm_ctx_launches(ctx: *const FcmCtx) -> u64;
    pub fn fcm_graph_count_async(g: *mut FcmGraph, part: u32, nparts: u32, counts_dev: *mut *mut u64) -> c_int;
    pub fn fcm_graph_set_kernel(g: *mut FcmGraph, kind: c_int) -> c_int;
    pub fn fcm_graph_info(g: *const FcmGraph, n: *mut u32, nedges: *mut u64, words_per_row: *mut u32,
                          max_out_degree: *mut u32) -> c_int;
    pub fn fcm_ensemble_set_kernel(e: *mut FcmEnsemble, kind: c_int) -> c_int;
    pub fn fcm_ensemble_set_stats(e: *mut FcmEnsemble, on: c_int) -> c_int;
    pub fn fcm_ensemble_index_info(e: *const FcmEnsemble, have_index: *mut c_int, index_bytes: *mut u64) -> c_int;
    pub fn fcm_ensemble_info(e: *const FcmEnsemble, n: *mut u32, npairs: *mut u64, nchains: *mut u32,
                             max_common_nbhd: *mut u32, bytes_per_chain: *mut u64) -> c_int;
    // Graph::compute_maximal_cliques (src/lib.rs:41): host code, no device; two-call sizing like fcm_edge_neighborhoods
    pub fn fcm_maximal_cliques(n: u32, nedges: u64, edges: *const [Node; 2], ncliques: *mut u64, nverts: *mut u64,
                               sizes_out: *mut u32, verts_out: *mut u32) -> c_int;
}
