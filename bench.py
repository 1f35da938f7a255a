#!/usr/bin/env python
"""bench.py — the hot path of flag-complex-mcmc on B200, measured.

Headline metric (BASELINE.json): MCMC proposals/sec.  One *step* = `--proposals` iterations of
the loop body of MCMCSampler::next (src/lib.rs:182-192) on every chain of the ensemble, i.e. one
launch of chain_step_kernel.  Workload at every N: the C3 graph (ER n=2000, p=0.05, seed 1,
E=199 938) with 4096 independent chains per GPU, `--simple` move mix — N=8 is exactly BASELINE
config C5 (8x4096 chains, chains sharded per GPU, no data-path collective => weak scaling).
The second metric of BASELINE.json (full-recount simplices/sec, config C4, sharded by source
vertex + allreduce of the uint64 counts) rides in the same JSON line under "recount".

  python bench.py [--gpus N] [--steps K] [--warmup W]            # this repo's CUDA path
  python bench.py --impl reference ...                           # CPU restatement of the reference

Under torchrun (N>1) one process per GPU; the timed region is bracketed by a barrier and a device
synchronize on both sides, timed with CUDA events on the launching stream, MAX over ranks.
"""
import argparse
import json
import os
import subprocess
import sys
import tempfile
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

METRIC = "mcmc_proposals_per_sec"
UNIT = "proposals/s"
N_C3, W_C3 = 2000, 32                      # vertices, 64-bit words per adjacency row (ceil(n/64))
# ncu --set full summaries of the kernels this file times (tools/ncu_summary.py output, committed); the roofline's
# `traffic` and the issue-slot figures are read from them at run time, nothing is hard-coded here
CHAIN_CAPTURE = os.path.join("profiles", "r02_chain_step.ncu.txt")
COUNT_CAPTURE = os.path.join("profiles", "r02_count_local.ncu.txt")
RANDSECTOR = os.path.join("profiles", "r02_randsector.json")


def read_capture(relpath):
    """-> dict of the raw ncu metrics of a committed capture summary (value in base units) + its '# key=value' header
    fields, or None when the file is absent."""
    path = os.path.join(ROOT, relpath)
    if not os.path.exists(path):
        return None
    scale = {"byte": 1.0, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9, "Tbyte": 1e12, "ns": 1e-9, "us": 1e-6, "ms": 1e-3, "s": 1.0}
    out = {"file": relpath}
    with open(path) as f:
        for ln in f:
            ln = ln.strip()
            if ln.startswith("#"):
                for tok in ln[1:].split():
                    if "=" in tok:
                        k, v = tok.split("=", 1)
                        try:
                            out[k] = float(v)
                        except ValueError:
                            out[k] = v
                continue
            parts = ln.split(",")
            if len(parts) == 3 and parts[0] and not parts[0].startswith("=="):
                try:
                    out[parts[0]] = float(parts[2]) * scale.get(parts[1], 1.0)
                except ValueError:
                    pass
    return out


def random_sector_peak():
    """measured ceiling of independent 32-byte-sector reads out of HBM (tools/randsector.cu), G sectors/s"""
    try:
        with open(os.path.join(ROOT, RANDSECTOR)) as f:
            d = json.load(f)
        return max(r["gsectors_per_s"] for r in d["rows"])
    except Exception:
        return None


def env_int(k, d):
    try:
        return int(os.environ.get(k, d))
    except ValueError:
        return d


def measured_peak():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    try:
        with open(p) as f:
            return float(json.load(f)["hbm_gbs"]), "measured (MEASURED_PEAKS.json hbm_gbs)"
    except Exception:
        return 6650.0, "fallback (B200_PROFILING.md 6.65 TB/s)"


def prop_model_bytes(sumK, nonempty, W):
    """Algorithmic bytes of the proposals: per non-empty proposal the K
    neighbourhood rows of 8*W bytes, the nbhd list 4*(K-2) and <= 2 word updates (16 B)."""
    return 8 * W * sumK + 4 * (sumK - 2 * nonempty) + 16 * nonempty


# ------------------------------------------------------------------------------------------------
# clocks sampler (nvidia-smi during the timed region)
# ------------------------------------------------------------------------------------------------
class Clocks:
    Q = ("clocks.sm,clocks.max.sm,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
         "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, index):
        self.index = index
        self.f = None
        self.p = None

    def start(self):
        # NVML in a thread (a sample every 5 ms: the timed region can be ~100 ms); nvidia-smi -lms as the fallback
        self.rows, self.halt, self.t = [], threading.Event(), None
        try:
            import pynvml as nv
            nv.nvmlInit()
            h = nv.nvmlDeviceGetHandleByIndex(self.index)
            mx = float(nv.nvmlDeviceGetMaxClockInfo(h, nv.NVML_CLOCK_SM))
            bits = [nv.nvmlClocksThrottleReasonHwSlowdown, nv.nvmlClocksThrottleReasonHwThermalSlowdown,
                    nv.nvmlClocksThrottleReasonSwThermalSlowdown, nv.nvmlClocksThrottleReasonSwPowerCap]

            def loop():
                while True:
                    try:
                        r = nv.nvmlDeviceGetCurrentClocksThrottleReasons(h)
                        self.rows.append([str(float(nv.nvmlDeviceGetClockInfo(h, nv.NVML_CLOCK_SM))), str(mx)] +
                                         ["Active" if r & b else "Not Active" for b in bits])
                    except Exception:
                        pass
                    if self.halt.wait(0.005):
                        return
            self.t = threading.Thread(target=loop, daemon=True)
            self.t.start()
            return
        except Exception:
            self.t = None
        try:
            self.f = tempfile.NamedTemporaryFile("w+", suffix=".csv", delete=False)
            self.p = subprocess.Popen(["nvidia-smi", "-i", str(self.index), "--query-gpu=" + self.Q,
                                       "--format=csv,noheader,nounits", "-lms", "100"],
                                      stdout=self.f, stderr=subprocess.DEVNULL)
        except Exception:
            self.p = None

    def stop(self):
        out = {"sm_mhz": None, "sm_max_mhz": None, "reasons": []}
        rows = []
        if getattr(self, "t", None):
            self.halt.set()
            self.t.join(timeout=5)
            rows = list(self.rows)
            out["source"] = "nvml"
        elif self.p:
            try:
                self.p.terminate()
                self.p.wait(timeout=5)
            except Exception:
                pass
            try:
                self.f.flush()
                self.f.seek(0)
                rows = [ln.strip().split(",") for ln in self.f.read().splitlines() if ln.strip()]
                self.f.close()
                os.unlink(self.f.name)
            except Exception:
                pass
            out["source"] = "nvidia-smi"
        rows = [r for r in rows if len(r) >= 6]
        sm = sorted(float(r[0]) for r in rows)
        if sm:
            out["sm_mhz"] = sm[len(sm) // 2]
            out["sm_max_mhz"] = float(rows[0][1])
            names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
            for i, nm in enumerate(names):
                if any(r[2 + i].strip().lower().startswith("active") for r in rows):
                    out["reasons"].append(nm)
            out["samples"] = len(sm)
        return out


# ------------------------------------------------------------------------------------------------
# CPU legs (the oracle = C++ restatement of the reference CPU path; checker/baseline only)
# ------------------------------------------------------------------------------------------------
class CpuChains:
    """`cores` independent reference-shaped chains, one per host thread (the reference runs one
    process per seed); run(budget_s) advances each by the same number of proposals."""

    def __init__(self, n, e, cores, first_seed=0):
        sys.path.insert(0, os.path.join(ROOT, "tests"))
        import oracle_binding as orc
        self.cores = cores
        self.samplers = [orc.Sampler(n, e, first_seed + k, 0.01, nthreads=cores) for k in range(cores)]
        t0 = time.perf_counter()
        self.samplers[0].steps(2000, trace=False)
        self.rate1 = 2000 / max(time.perf_counter() - t0, 1e-6)      # one chain alone, proposals/s

    def run(self, budget_s):
        """-> (proposals/s over all threads, proposals, seconds)"""
        per = max(int(self.rate1 * budget_s), 1000)
        ths = [threading.Thread(target=lambda s=s: s.steps(per, trace=False)) for s in self.samplers]
        t0 = time.perf_counter()
        for t in ths:
            t.start()
        for t in ths:
            t.join()
        dt = time.perf_counter() - t0
        return self.cores * per / dt, self.cores * per, dt

    def close(self):
        for s in self.samplers:
            s.close()


def cpu_recount(n, e, cores, stride):
    """Bounded sample of the full recount: every `stride`-th source vertex of the degree-descending
    order, dynamic over all host threads -> (simplices/s, simplices, seconds)."""
    sys.path.insert(0, os.path.join(ROOT, "tests"))
    import oracle_binding as orc
    deg = np.bincount(e[:, 0], minlength=n)
    order = np.argsort(-deg, kind="stable").astype(np.uint32)
    src = order[::stride]
    t0 = time.perf_counter()
    c = orc.count_sources_mt(n, e, src, cores)
    dt = time.perf_counter() - t0
    return int(c.sum()) / dt, int(c.sum()), dt


def run_reference(args, emit):
    """--impl reference: the reference's CPU algorithm for the path (oracle port; the Rust crate
    cannot be built here: no rustc, un-vendored crates) on all host cores, same config/metric."""
    rank = env_int("RANK", 0)
    if rank != 0:
        return 0
    from tools import synth
    cores = os.cpu_count() or 1
    n, e = synth.config("c3_er_2000")
    cc = CpuChains(n, e, cores)
    for _ in range(args.warmup):
        cc.run(0.5)
    t_all = 0.0
    nprop = 0
    for _ in range(args.steps):
        v, cnt, dt = cc.run(4.0)
        t_all += dt
        nprop += cnt
    cc.close()
    value = nprop / t_all
    sample = "%d chains (one per host thread) x %d proposals per step on the C3 graph, --simple mix" % (cores, nprop // max(args.steps, 1) // cores)
    line = {
        "impl": "reference", "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": args.gpus, "steps": args.steps,
        "warmup": args.warmup, "ms_per_step": 1e3 * t_all / max(args.steps, 1), "higher_is_better": True,
        "scaling": "weak", "vs_baseline": None, "dtype": "u64", "data": "synthetic",
        "config": workload_config(args, args.gpus),
        "cpu_baseline": {"value": value, "unit": UNIT, "cores": cores, "kind": "port", "sample": sample},
        "e2e": {"value": value, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
    }
    emit(line)
    return 0


def workload_config(args, ngpu):
    return {"workload": "C5 shard of BASELINE.json: C3 graph (ER n=2000 p=0.05 seed 1, E=199938), "
                        "%d chains per GPU x %d GPU(s), seeds 0..%d, --simple move mix [0.5,0.5,0,0], t=0.01"
                        % (args.chains, ngpu, args.chains * ngpu - 1),
            "chains_per_gpu": args.chains, "proposals_per_chain_per_step": args.proposals,
            "parallelism": "chains sharded per GPU, no data-path collective",
            "l2": "per-GPU chain state (adjacency rows + edge containers) is ~%.1f GB >> 126 MB L2; "
                  "every step touches fresh random rows" % (args.chains * 5.2e-3)}


# ------------------------------------------------------------------------------------------------
# the CUDA path
# ------------------------------------------------------------------------------------------------
def oracle_chain_after(n, e, seed, nproposals, move_weights=(0.5, 0.5, 0.0, 0.0)):
    """state of the oracle's chain `seed` after nproposals (checker for parity_spot_check)"""
    sys.path.insert(0, os.path.join(ROOT, "tests"))
    import oracle_binding as orc
    o = orc.Sampler(n, e, int(seed), 0.01, move_weights=move_weights)
    o.steps(int(nproposals), trace=False)
    return o.read()


def cpu_single_chain_rate(n, e, move_weights, budget_s=2.0):
    """the oracle port's proposals/s for ONE chain on one host core (the reference runs a chain single-threaded)"""
    sys.path.insert(0, os.path.join(ROOT, "tests"))
    import oracle_binding as orc
    o = orc.Sampler(n, e, 0, 0.01, move_weights=move_weights)
    t0 = time.perf_counter()
    o.steps(500, trace=False)
    r = 500 / max(time.perf_counter() - t0, 1e-6)
    per = max(int(r * budget_s), 500)
    t0 = time.perf_counter()
    o.steps(per, trace=False)
    return per / (time.perf_counter() - t0), per


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=5)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--chains", type=int, default=4096, help="chains per GPU")
    ap.add_argument("--proposals", type=int, default=2000, help="proposals per chain per step")
    ap.add_argument("--no-recount", action="store_true", help="skip the C4 recount leg")
    ap.add_argument("--no-cpu", action="store_true", help="skip the cpu_baseline leg")
    ap.add_argument("--no-clique-mix", action="store_true", help="skip the default-move-mix leg")
    ap.add_argument("--no-configs", action="store_true", help="skip the C1 / C2 / C3-1024 lines")
    ap.add_argument("--no-spot-check", action="store_true", help="skip the oracle comparison of two timed chains")
    ap.add_argument("--recount-graph", default="c4_geo_31346")
    args = ap.parse_args()
    args.warmup = max(args.warmup, 0)
    # stdout carries exactly one JSON line: anything a library prints there (NCCL's version banner ...)
    # goes to stderr instead
    sys.stdout.flush()
    real_stdout = os.dup(1)
    os.dup2(2, 1)

    def emit(obj):
        sys.stdout.flush()
        os.write(real_stdout, (json.dumps(obj) + "\n").encode())
    if args.impl == "reference":
        return run_reference(args, emit)

    import torch
    import flag_complex_mcmc_b200 as fcm
    from flag_complex_mcmc_b200 import directed_scm as ds
    from flag_complex_mcmc_b200 import sharding
    from tools import synth

    world = env_int("WORLD_SIZE", 1)
    rank = env_int("RANK", 0)
    local = env_int("LOCAL_RANK", 0)
    if not torch.cuda.is_available():
        raise SystemExit("bench.py: no CUDA device (the product has no CPU fallback)")
    torch.cuda.set_device(local)
    dist = None
    if world > 1:
        import torch.distributed as dist
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    ngpu = world

    def barrier():
        if dist is not None:
            dist.barrier()
        torch.cuda.synchronize()

    def max_over_ranks(x):
        if dist is None:
            return x
        t = torch.tensor([x], dtype=torch.float64, device="cuda")
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return float(t.item())

    def exchange(blob):
        """host transport of the fcm_comm handles: all_gather over the bench's process group"""
        if dist is None:
            return [blob]
        mine = torch.from_numpy(np.asarray(blob, dtype=np.uint8).copy()).cuda()
        allh = [torch.empty_like(mine) for _ in range(world)]
        dist.all_gather(allh, mine)
        return [t.cpu().numpy() for t in allh]

    ctx = fcm.Context(local)
    peak, peak_src = measured_peak()
    rs_peak = random_sector_peak()

    # ---- chains --------------------------------------------------------------------------------
    n, e = synth.config("c3_er_2000")
    seeds = sharding.chain_seeds(0, args.chains * ngpu, rank, ngpu)   # chain k -> GPU k mod N
    s = fcm.initialize_new_sampler(ctx, n, e, seeds, target_relaxation=0.01, simple=True)
    P = args.proposals
    # warm-up steps run instrumented: they give the mean neighbourhood size K and the non-empty fraction
    # of these chains for the algorithmic-bytes model.  The timed steps run with the statistic off (it is
    # not part of the reference's step; a double-edge move would otherwise intersect two vertex lists
    # only to report K) - "value_instrumented" below is the same measurement with it on.
    st0 = s.read_all()
    for _ in range(args.warmup):
        s.step(P)
    ctx.synchronize()
    st1 = s.read_all()
    sumK_w = st1["sumK"] - st0["sumK"]
    nonempty_w = st1["nonempty"] - st0["nonempty"]
    props_w = float(args.chains) * P * max(args.warmup, 1)
    meanK = sumK_w / max(nonempty_w, 1)
    nonempty_frac = nonempty_w / props_w

    def timed(nsteps):
        barrier()
        ctx.timer_begin()
        for _ in range(nsteps):
            s.step(P)
        t = ctx.timer_end()
        barrier()
        return t

    ms_instr = max_over_ranks(timed(max(1, args.steps // 2)) / max(1, args.steps // 2))
    s.set_stats(False)
    s.step(P)                                                  # the uninstrumented kernel's first launch
    ctx.synchronize()
    st0 = s.read_all()
    l0 = ctx.launches
    clk = Clocks(local)
    barrier()
    clk.start()
    ms = timed(args.steps)
    launches = ctx.launches - l0
    st1 = s.read_all()
    ms_max = max_over_ranks(ms)
    total_props = float(args.chains) * P * args.steps * ngpu
    value = total_props / (ms_max * 1e-3)
    value_instr = float(args.chains) * P * ngpu / (ms_instr * 1e-3)
    nonempty = nonempty_frac * float(args.chains) * P * args.steps
    sumK = meanK * nonempty
    model_bytes = prop_model_bytes(sumK, nonempty, W_C3)       # this rank, whole timed region
    kern_ms = ms / max(args.steps, 1)                          # one chain_step_kernel launch per step
    achieved = model_bytes / max(args.steps, 1) / (kern_ms * 1e-3) / 1e9
    accepted = int((st1["accepted"] - st0["accepted"]).sum())

    # ---- e2e: through the public API with host buffers: step + read_all (D2H) every step ----------
    h2d = 2 * 33 * 8 + 8                                       # bounds vectors + nproposals (kernel parameters)
    d2h = 0
    barrier()
    t0 = time.perf_counter()
    for _ in range(args.steps):
        s.set_bounds(s.bounds)
        s.step(P)
        r = s.read_all()                                       # synchronous D2H of every chain's state
        d2h = r["flag_counts"].nbytes + r["lens"].nbytes + r["sampled"].nbytes + r["accepted"].nbytes + 16
    torch.cuda.synchronize()
    e2e_ms = (time.perf_counter() - t0) * 1e3
    clocks = clk.stop()
    e2e_ms = max_over_ranks(e2e_ms)
    e2e_value = total_props / (e2e_ms * 1e-3)
    info = s.info()
    info.update(s.index_info())

    # ---- parity spot check: two of the TIMED chains (state after every step above) against the oracle ----
    spot = None
    if rank == 0 and not args.no_spot_check and not args.no_cpu:
        try:
            spot = "ok"
            which = [0, min(args.chains - 1, 1777)]
            res = {}
            got = {k: s.read(k) for k in which}               # the handle is read from this thread only

            def chk(k):
                r_ = got[k]
                w_ = oracle_chain_after(n, e, int(seeds[k]), r_["sampled"])
                res[k] = (r_["flag_count"] == w_["flag_count"] and r_["rng"] == w_["rng"] and r_["accepted"] == w_["accepted"]
                          and np.array_equal(r_["edges"], w_["edges"]), r_["sampled"])
            th = [threading.Thread(target=chk, args=(k,)) for k in which]
            for t in th:
                t.start()
            for t in th:
                t.join()
            if not all(res.get(k, (False,))[0] for k in which):
                spot = "MISMATCH on chain(s) %s" % [k for k in which if not res.get(k, (False,))[0]]
            spot_what = "chains %s (seeds %s) after %d proposals each: flag_count, xoshiro state, accepted and the edge container in order equal the oracle's" % (
                which, [int(seeds[k]) for k in which], res[which[0]][1])
        except Exception as ex:                                 # noqa: BLE001
            spot, spot_what = "error: %r" % (ex,), ""
    s.close()

    cap = read_capture(CHAIN_CAPTURE)
    traffic = None
    traffic_src = "no capture of the timed kernel committed yet (%s)" % CHAIN_CAPTURE
    rs = None
    if cap and "dram__bytes_read.sum" in cap and cap.get("chains") and cap.get("proposals"):
        bpp = (cap["dram__bytes_read.sum"] + cap["dram__bytes_write.sum"]) / (cap["chains"] * cap["proposals"])
        traffic = bpp * float(args.chains) * P
        traffic_src = ("%s: ncu --set full of chain_step_kernel, %d chains x %d proposals: dram read %.3f GB + write %.3f GB = %.0f B/proposal, "
                       "scaled to this launch" % (cap["file"], cap["chains"], cap["proposals"], cap["dram__bytes_read.sum"] / 1e9,
                                                  cap["dram__bytes_write.sum"] / 1e9, bpp))
        if rs_peak:
            # HBM moves 64-byte granules per missing sector pair: accesses = traffic / 64
            acc_per_s = bpp / 64.0 * float(args.chains) * P / (kern_ms * 1e-3) / 1e9
            rs = {"bound": "hbm random access", "achieved": acc_per_s, "peak": rs_peak, "unit": "G accesses/s", "frac": acc_per_s / rs_peak,
                  "what": "DRAM accesses of the launch (measured bytes / 64 B) per second against the measured ceiling of independent "
                          "random 32-byte-sector reads (tools/randsector.cu, %s): the chain's state is touched 4 bytes at a time" % RANDSECTOR}
    line = {
        "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": ngpu, "steps": args.steps, "warmup": args.warmup,
        "ms_per_step": ms_max / max(args.steps, 1), "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
        "dtype": "u64", "data": "synthetic", "config": workload_config(args, ngpu),
        "e2e": {"value": e2e_value, "unit": UNIT, "h2d_bytes_per_step": h2d, "d2h_bytes_per_step": d2h,
                "what": "MCMCSampler.set_bounds + step(P) + read_all() per step (flag_count, sampled, accepted of every chain to host)"},
        "gpu_launches": int(launches),
        "clocks": clocks,
        "roofline": {"bound": "hbm", "kernel": "chain_step_kernel", "achieved": achieved, "peak": peak, "unit": "GB/s",
                     "frac": achieved / peak, "traffic": traffic, "traffic_source": traffic_src,
                     "peak_source": peak_src,
                     "model": "8*W*K + 4*(K-2) + 16 bytes per non-empty proposal (SURVEY 8d), W=%d, mean K=%.2f, "
                              "non-empty fraction %.3f (both measured over the instrumented warm-up steps of the same chains)" % (W_C3, meanK, nonempty_frac),
                     "kernel_ms": kern_ms,
                     "issue_active_pct": cap.get("smsp__issue_active.avg.pct_of_peak_sustained_active") if cap else None,
                     "random_access": rs},
        "acceptance_ratio": accepted / (float(args.chains) * P * args.steps),
        "bytes_per_chain": info["bytes_per_chain"],
        "value_instrumented": value_instr,
        "clique_index": {"built": info["have_index"], "bytes": info["index_bytes"]},
        "parity": "bit-exact against oracle/ (C++ restatement); the oracle's counts are pinned by known answers, its trajectory "
                  "semantics are UNPINNED (reference not buildable here: no rustc, un-vendored crates) - DESIGN.md 4",
    }
    if spot is not None:
        line["parity_spot_check"] = spot
        line["parity_spot_check_what"] = spot_what

    # ---- the other BASELINE.json configs: C1 (reference fixture, 1 chain), C2 (1 chain), C3 (1024 chains) -------
    if not args.no_configs and ngpu == 1:
        z = np.load(os.path.join(ROOT, "tests", "golden", "fixtures.npz"))
        cfgs = {}
        cases = [("c1", 200, z["counterexample_seo_greedy_5_start"].reshape(-1, 2).astype(np.uint32), 1,
                  "flag_file_examples/counterexample_seo_greedy_5_start.flag, `sample -s 0` (1 chain, step = min(sample_distance, 20000 / 4000) proposals)"),
                 ("c2", 279, z["c2_chunglu_279"].reshape(-1, 2).astype(np.uint32), 1, "C. elegans-shaped Chung-Lu graph (279 nodes, 2200 edges), 1 chain"),
                 ("c3_1024", n, e, 1024, "ER n=2000 p=0.05 (C3), 1024 batched chains, step = %d proposals per chain" % P)]
        for name, cn, ce, nch, what in cases:
            entry = {"workload": what, "chains": nch}
            for mixname, simple in (("simple", True), ("default_mix", False)):
                sc_ = fcm.initialize_new_sampler(ctx, cn, ce, list(range(nch)), target_relaxation=0.01, simple=simple)
                # a step of a one-chain config = sample_distance proposals (src/bin/sample.rs:102), capped so that the whole
                # default bench stays within minutes: these legs report rates
                Pc = min(sc_.sample_distance, 20000 if simple else 4000) if nch == 1 else (P if simple else max(P // 10, 1))
                sc_.set_stats(False)
                sc_.step(Pc)
                ctx.synchronize()
                ctx.timer_begin()
                for _ in range(2):
                    sc_.step(Pc)
                cms = ctx.timer_end()
                sc_.status()
                v = float(nch) * Pc * 2 / (cms * 1e-3)
                entry[mixname] = {"value": v, "unit": UNIT, "proposals_per_chain_per_step": Pc, "ms_per_step": cms / 2,
                                  "us_per_proposal_per_chain": cms * 1e3 / (2 * Pc)}
                if rank == 0 and not args.no_cpu:
                    mw = ds.MOVE_DISTRIBUTION_SIMPLE if simple else ds.MOVE_DISTRIBUTION
                    cv, cnt = cpu_single_chain_rate(cn, ce, mw, 1.5)
                    entry[mixname]["cpu_baseline"] = {"value": cv, "unit": UNIT, "cores": 1, "kind": "port",
                                                      "sample": "one chain x %d proposals on one host core (oracle/), per-chain rate" % cnt}
                sc_.close()
            cfgs[name] = entry
        line["configs"] = cfgs

    # ---- default move mix [0.1,0.1,0.6,0.2] (clique_permute / clique_swap included, src/bin/sample.rs:17) ----
    if not args.no_clique_mix:
        t0 = time.perf_counter()
        sc = fcm.initialize_new_sampler(ctx, n, e, seeds, target_relaxation=0.01, simple=False)
        create_s = time.perf_counter() - t0
        Pc = max(args.proposals // 10, 1)
        sc.set_stats(False)
        for _ in range(max(args.warmup, 1)):
            sc.step(Pc)
        barrier()
        ctx.timer_begin()
        for _ in range(args.steps):
            sc.step(Pc)
        cms = max_over_ranks(ctx.timer_end())
        barrier()
        sc.status()
        rc_ = sc.read_all()
        line["default_mix"] = {"metric": METRIC, "value": float(args.chains) * Pc * args.steps * ngpu / (cms * 1e-3), "unit": UNIT,
                               "move_distribution": [0.1, 0.1, 0.6, 0.2], "proposals_per_chain_per_step": Pc,
                               "ms_per_step": cms / max(args.steps, 1),
                               "acceptance_ratio": float(rc_["accepted"].sum()) / max(float(rc_["sampled"].sum()), 1.0),
                               "create_s": create_s,
                               "what": "same graph and chains with the CLI's default move mix; clique moves run on the cold path"}
        sc.close()

    # ---- full recount (C4), sharded by source vertex; the all-reduce is part of the library call and of the timed region ----
    if not args.no_recount:
        n4, e4 = synth.config(args.recount_graph)
        g = fcm.DeviceGraph(ctx, n4, e4)
        W4 = (n4 + 63) // 64
        NC = max(args.steps, 50)                               # >= 50 counts: long enough for the clock sampler

        def time_counts(fn):
            for _ in range(max(args.warmup, 1)):
                fn()
            g.status()
            barrier()
            ctx.timer_begin()
            for _ in range(NC):
                fn()
            t = ctx.timer_end()
            g.status()
            barrier()
            return max_over_ranks(t) / NC

        ms_kernel_only = time_counts(lambda: g.count_async(rank, ngpu))      # this rank's shard, no collective
        comm = fcm.Comm(ctx, rank, ngpu, fcm.COMM_PEER, exchange=exchange if ngpu > 1 else None)
        l0 = ctx.launches
        rclk = Clocks(local)
        rclk.start()
        ms_fused = time_counts(lambda: g.count_allreduce_async(comm))       # count + all-reduce in ONE launch (last-CTA epilogue over NVLink)
        rl = (ctx.launches - l0) // (NC + max(args.warmup, 1))
        rclocks = rclk.stop()
        ms_nccl, nccl_err = None, None
        if ngpu > 1:
            try:
                ncomm = fcm.Comm(ctx, rank, ngpu, fcm.COMM_NCCL, exchange=exchange)
                ms_nccl = time_counts(lambda: g.count_allreduce_async(ncomm))   # count kernel, then ncclAllReduce on the same stream
                ncomm.close()
            except Exception as ex:                                 # noqa: BLE001
                nccl_err = repr(ex)
        tot = g.count_allreduce(comm)                          # every rank holds the whole graph's vector
        counts = [int(x) for x in tot]
        while counts and counts[-1] == 0:
            counts.pop()
        simplices = sum(counts)
        rvalue = simplices / (ms_fused * 1e-3)
        rbytes = 8 * W4 * simplices                            # dense-row model, whole graph
        # e2e through the library's sharded call with HOST buffers: every rank uploads the edge list, builds its rows +
        # CSR, counts its shard, all-reduces inside the kernel and reads the vector back
        e2e_runs = []
        for rep in range(4):                                   # one warm-up call (the memory pool grows once), then three timed
            barrier()
            t0 = time.perf_counter()
            g2 = fcm.DeviceGraph(ctx, n4, e4)
            c2 = ds.trim(g2.count_allreduce(comm))
            dt = max_over_ranks(time.perf_counter() - t0)
            g2.close()
            assert c2 == counts, (c2, counts)
            if rep:
                e2e_runs.append(dt)
        re2e = sorted(e2e_runs)[1]                             # median of three
        ccap = read_capture(COUNT_CAPTURE)
        rtraffic, issue, winst = None, None, None
        if ccap and ngpu == 1 and args.recount_graph == ccap.get("graph", "c4_geo_31346"):
            rtraffic = ccap.get("dram__bytes_read.sum", 0.0) + ccap.get("dram__bytes_write.sum", 0.0)
            issue = ccap.get("smsp__issue_active.avg.pct_of_peak_sustained_active")
            winst = ccap.get("smsp__inst_executed.sum")
        line["recount"] = {
            "metric": "full_recount_simplices_per_sec", "value": rvalue, "unit": "simplices/s", "graph": args.recount_graph,
            "n": n4, "nedges": int(e4.shape[0]), "flag_count": counts, "simplices": simplices,
            "ms_per_count": ms_fused, "ms_per_count_kernel_only": ms_kernel_only, "ms_per_count_nccl": ms_nccl, "nccl_error": nccl_err,
            "allreduce": "inside libfcm.so and inside the timed region: fcm_graph_count_allreduce_async - the count kernel's last CTA adds the "
                         "rank's u64[33] into every rank's window with system-scope atomics over NVLink and waits for the others (one launch per "
                         "count); ms_per_count_nccl = same count followed by ncclAllReduce(u64[33]) on the stream; ms_per_count_kernel_only = no collective",
            "gpu_launches": int(rl), "clocks": rclocks,
            "sharding": "source vertices, degree-descending round-robin over %d rank(s)" % ngpu,
            "e2e": {"value": simplices / re2e, "unit": "simplices/s", "h2d_bytes_per_step": int(e4.nbytes),
                    "d2h_bytes_per_step": 8 * 33, "seconds": re2e, "seconds_runs": e2e_runs,
                    "what": "DeviceGraph(host edge list) + count_allreduce on every rank: upload, build bit rows + CSR, count the shard, all-reduce, download; "
                            "median of three calls after one warm-up call"},
            "roofline": {"bound": "issue", "kernel": "count_local_kernel",
                         "dense_row_model": {"achieved": rbytes / (ms_fused * 1e-3) / 1e9 / ngpu, "peak": peak, "unit": "GB/s",
                                             "frac": rbytes / (ms_fused * 1e-3) / 1e9 / ngpu / peak,
                                             "note": "8*ceil(n/64) bytes per enumerated simplex (SURVEY 8d). NOT a bound for this kernel: it gathers each "
                                                     "source's k x k bits once and intersects in shared memory, so this exceeds 1 by design"},
                         "dram": {"traffic": rtraffic, "achieved": (rtraffic / (ms_fused * 1e-3) / 1e9) if rtraffic else None, "peak": peak, "unit": "GB/s",
                                  "frac": (rtraffic / (ms_fused * 1e-3) / 1e9 / peak) if rtraffic else None,
                                  "compulsory_bytes": 8 * (W4 + (W4 & 1)) * n4 + 4 * int(e4.shape[0])},
                         "issue": {"warp_instructions": winst, "issue_active_pct": issue,
                                   "warp_instructions_per_simplex": (winst / simplices) if winst else None},
                         "source": ccap["file"] if ccap else None, "peak_source": peak_src},
        }
        comm.close()
        g.close()

    # ---- CPU baseline beside it (rank 0, N=1 only) ---------------------------------------------
    if rank == 0 and ngpu == 1 and not args.no_cpu:
        cores = os.cpu_count() or 1
        cc = CpuChains(n, e, cores)
        cc.run(0.5)
        v, cnt, dt = cc.run(12.0)
        cc.close()
        line["cpu_baseline"] = {"value": v, "unit": UNIT, "cores": cores, "kind": "port",
                                "sample": "%d chains (one per host thread) x %d proposals on the C3 graph, %.1f s; C++ restatement "
                                          "of the reference CPU path (oracle/), the Rust crate cannot be built here" % (cores, cnt // cores, dt)}
        if not args.no_recount:
            rv, rs_, rdt = cpu_recount(n4, e4, cores, 4)
            line["recount"]["cpu_baseline"] = {"value": rv, "unit": "simplices/s", "cores": cores, "kind": "port",
                                               "sample": "every 4th source vertex (degree-descending order) of %s: %d simplices "
                                                         "in %.2f s incl. bit-matrix build" % (args.recount_graph, rs_, rdt)}
    ctx.close()
    if rank == 0:
        emit(line)
    if dist is not None:
        dist.barrier()
        dist.destroy_process_group()
    return 0


if __name__ == "__main__":
    sys.exit(main())
