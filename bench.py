#!/usr/bin/env python
"""bench.py -- enumeration nodes/s of the lattice search (BASELINE.json metric) on N B200s of one node.

A step = one complete exhaustive search of the workload instance (dense UNSAT BILP, n = m = 60, LLL-reduced,
committed under tests/golden/ with the script that made it).  A node = one layer-distance solve.

  python bench.py --gpus 1 --steps 3 --warmup 3                 (our arm)
  python bench.py --impl reference --gpus 1 --steps 2 --warmup 1 (the reference arm: CPU restatement, see below)
  python -m torch.distributed.run --nproc-per-node N ... bench.py --gpus N ...

`value`  : nodes / device time of the search kernel(s), inputs resident in HBM (CUDA events on the launching stream,
           max over ranks, summed over the K steps).
`e2e`    : the same metric through the plugin call blt_check() (include/blt.h) on a context holding the UN-reduced
           system of the fixture, host data in and out: stretch to the cube, exact LLL (host seam), host->device copy
           of the reduced basis and bounds, search, device->host copy of the result, back-substitution, final exact
           check; wall clock around the call (`lll_ms` / `search_ms` broken out from blt_b200_last_stats).
           At N > 1 the plugin call is one process using N GPUs (BLT_GPUS = N, a host thread per device): rank 0
           makes it while the other ranks wait at a barrier.
N > 1    : every rank runs the deterministic static split and then searches its own share of the subtrees
           (bltcu_opts.rank/world); no collective on the data path, one all-reduce of the counters at the end.
           Total work is fixed => "strong" scaling.

The reference arm: GaloisInc/blt cannot be built on this image (needs NTL, GLPK, Boost; none present, no network),
so `--impl reference` times the CPU restatement in oracle/ (kind "port", single thread like the reference itself)
on a bounded sample of the same workload.
"""
from __future__ import annotations

import argparse
import json
import os
import subprocess
import sys
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

FIXTURE = os.path.join(ROOT, "tests", "golden", os.environ.get("BLT_BENCH_FIXTURE", "bench_dense60_w140.npz"))
METRIC = "enum_nodes_per_s"
UNIT = "nodes/s"


def f_alg(k: int, m: int) -> float:
    """Algorithmic fp64 flops of one node with k unknowns in ambient dimension m."""
    return m * (2 * k + 3) + 4.0 * (k + 1) ** 2


def load_workload():
    z = np.load(FIXTURE)
    return dict(lat=z["lat"].astype(np.int64), con=z["con"].astype(np.int64), A=z["A"].astype(np.int64), C=z["C"].astype(np.int64),
                w=int(z["w"]), seed=int(z["seed"]),
                n=int(z["lat"].shape[0]), m=int(z["lat"].shape[1]), nodes_expected=int(z["nodes"]))


def workload_config(wl):
    return {"workload": f"dense UNSAT BILP n={wl['n']} m={wl['m']} |a|<=100 width={wl['w']} seed={wl['seed']}, "
                        "LLL-reduced, exhaustive enumeration",
            "fixture": "tests/golden/" + os.path.basename(FIXTURE), "nodes_expected": wl["nodes_expected"],
            "l2": "L2 flushed (256 MiB write) between timed steps; inputs are 29 KB, the search's own scratch (per-warp record stacks, "
                  "several GB) exceeds L2"}


class ClockSampler(threading.Thread):
    """Samples SM clock and throttle reasons with nvidia-smi while the timed region runs."""

    def __init__(self, index: int):
        super().__init__(daemon=True)
        self.index = index
        self.samples = []
        self.stop_flag = False
        self.smax = None

    def run(self):
        q = ("clocks.sm,clocks.max.sm,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
             "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")
        while not self.stop_flag:
            try:
                out = subprocess.run(["nvidia-smi", f"--id={self.index}", f"--query-gpu={q}",
                                      "--format=csv,noheader,nounits"], capture_output=True, text=True, timeout=5).stdout
                f = [x.strip() for x in out.strip().split(",")]
                if len(f) >= 6:
                    self.samples.append(f)
            except Exception:
                pass
            time.sleep(0.2)

    def summary(self):
        if not self.samples:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["unsampled"]}
        mhz = sorted(float(s[0]) for s in self.samples if s[0].replace(".", "").isdigit())
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        reasons = [nm for i, nm in enumerate(names) if any(s[2 + i].lower().startswith("active") for s in self.samples)]
        return {"sm_mhz": mhz[len(mhz) // 2] if mhz else None, "sm_max_mhz": float(self.samples[0][1]),
                "reasons": reasons, "samples": len(self.samples)}


def cpu_sample(wl, max_nodes):
    import oracle  # the one place bench.py executes oracle/: the timed CPU baseline

    t0 = time.perf_counter()
    r = oracle.closest(wl["lat"], wl["con"], max_nodes=max_nodes)
    dt = time.perf_counter() - t0
    return r["nodes"] / dt, r["nodes"], dt


def batch_supplement(blt_b200):
    """Supplementary figure, outside the timed region of the headline metric: SURVEY.md section 8d config 5 through
    bltcu_closest_batch -- the 400 committed n = m = 20 instances, tiled to 4000, host buffers in and out."""
    z = np.load(os.path.join(ROOT, "tests", "golden", "batch20.npz"))
    reps = 10
    lats = np.ascontiguousarray(np.tile(z["lat"], (reps, 1, 1)))
    cons = np.ascontiguousarray(np.tile(z["con"], (reps, 1, 1)))
    blt_b200.closest_batch(lats[:400], cons[:400])
    t0 = time.perf_counter()
    r = blt_b200.closest_batch(lats, cons)
    dt = time.perf_counter() - t0
    ok = bool((r["niter"] == np.tile(z["niter"], reps)).all() and (r["nprune"] == np.tile(z["nprune"], reps)).all()
              and (r["status"] == np.tile(z["rc"], reps)).all())
    # the reduction in front of it: 4000 fresh 20 x 20 bases (A^T of config 5) through bltcu_lll_batch
    rng = np.random.default_rng(5)
    Bs = np.stack([rng.integers(-100, 101, size=(20, 20)).T for _ in range(4000)])
    blt_b200.lll_batch(Bs[:64])
    t1 = time.perf_counter()
    _, _, _, linfo = blt_b200.lll_batch(Bs)
    dt_l = time.perf_counter() - t1
    lll = {"workload": "4000 exact LLL reductions of 20x20 integer bases, entries in [-100, 100], one warp each: fp64 decisions + "
                       "exact int64 row operations, deferred problems through the 512-bit kernel",
           "problems_per_s": len(Bs) / dt_l, "kernel_ms": linfo["kernel_ms"], "e2e_ms": 1e3 * dt_l, "deferred": linfo["deferred"],
           "on_host": linfo["on_host"], "hbm_gb_per_s": len(Bs) * 8 * (2 * 20 * 20 + 2 * 20 * 20) / (linfo["kernel_ms"] * 1e-3) / 1e9}
    # time-to-decision on the reference's own SAT regression fixtures (64 x 64, LLL-reduced; SURVEY.md 8d metric ii)
    sat = []
    zj = np.load(os.path.join(ROOT, "tests", "golden", "jpeg_reduced.npz"))
    for name in ("test_95", "test_50", "test_50_8"):
        la, co = zj[f"lat_{name}"], zj[f"con_{name}"]
        blt_b200.closest(la, co, timeout_ms=60000)
        t1 = time.perf_counter()
        g = blt_b200.closest(la, co, timeout_ms=60000)
        sat.append({"instance": name, "found": bool(g["found"]), "gpu_ms": 1e3 * (time.perf_counter() - t1),
                    "gpu_nodes": int(g["nodes"]), "cpu_port_nodes": int(zj[f"nodes_{name}"])})
    # BASELINE.json config 4: dense UNSAT n = m = 80 (10.8 M nodes), exhaustive, counts against the oracle's stored ones
    z80 = np.load(os.path.join(ROOT, "tests", "golden", "bench_dense80_w105.npz"))
    blt_b200.closest(z80["lat"], z80["con"], timeout_ms=600000)
    g80 = blt_b200.closest(z80["lat"], z80["con"], timeout_ms=600000)
    dense80 = {"workload": "config 4: dense UNSAT BILP n=80 m=80 |a|<=100 width=105 seed=0, exhaustive (tests/golden/bench_dense80_w105.npz)",
               "nodes": int(g80["nodes"]), "nodes_per_s": g80["nodes"] / (g80["kernel_ms"] * 1e-3), "kernel_ms": g80["kernel_ms"],
               "count_ok": bool(g80["nodes"] == int(z80["nodes"]) and g80["pivots"] == int(z80["pivots"])), "busy_frac": g80["busy_frac"]}
    # BASELINE.json config 2: 40 variables x 20 rows through blt_check_batch (reduction-bound: time to decision)
    from blt_b200 import api
    rng = np.random.default_rng(2)
    ctxs = []
    for _ in range(30):
        A2 = rng.integers(-100, 101, size=(20, 40))
        c2 = rng.integers(-10**6, 10**6 + 1, size=20)
        cx = api.Context()
        vs = [cx.var(str(j)) for j in range(40)]
        for i in range(20):
            e = cx.const(0)
            for j in range(40):
                e = e + int(A2[i, j]) * vs[j]
            cx.assume(int(c2[i]), e, int(c2[i]) + 8)
        ctxs.append(cx)
    api.check_batch(ctxs[:2])
    t1 = time.perf_counter()
    res2 = api.check_batch(ctxs)
    dt2 = time.perf_counter() - t1
    config2 = {"workload": "config 2: 30 systems of 40 variables x 20 rows, |a|<=100, width 8, through blt_check_batch",
               "ms_per_system": 1e3 * dt2 / len(ctxs), "decisions_sat": int(sum(1 for x in res2 if x == 10)),
               "decisions_unsat": int(sum(1 for x in res2 if x == 11))}
    for cx in ctxs:
        cx.free()
    return {"lll_batch": lll, "sat_time_to_decision": sat, "dense80": dense80, "config2": config2,
            "workload": "config 5: 4000 dense 20x20 instances (tests/golden/batch20.npz x10), a warp per instance plus helpers",
            "instances_per_s": len(lats) / dt,
            "nodes_per_s": (int(r["niter"].sum()) + int(r["nprune"].sum()) + len(lats)) / dt, "kernel_ms": r["kernel_ms"],
            "e2e_ms": 1e3 * dt, "count_ok": ok, "shared_problems": r["donated"], "launches": r["launches"]}


def run_reference(args, rank, world):
    if rank != 0:
        return 0
    wl = load_workload()
    sample_nodes = int(os.environ.get("BLT_BENCH_REF_NODES", "60000"))
    for _ in range(args.warmup):
        cpu_sample(wl, 5000)
    tot_nodes, tot_t = 0, 0.0
    for _ in range(args.steps):
        v, nodes, dt = cpu_sample(wl, sample_nodes)
        tot_nodes += nodes
        tot_t += dt
    value = tot_nodes / tot_t
    # Supplementary: what the host's cores could do if the reference were run once per core on independent work
    # ("all-cores variant").  ctypes releases the GIL, so plain threads run in parallel.
    cores = len(os.sched_getaffinity(0)) if hasattr(os, "sched_getaffinity") else (os.cpu_count() or 1)
    done = []

    def worker():
        done.append(cpu_sample(wl, sample_nodes)[1])

    ths = [threading.Thread(target=worker) for _ in range(cores)]
    t0 = time.perf_counter()
    for t in ths:
        t.start()
    for t in ths:
        t.join()
    all_cores = {"value": sum(done) / (time.perf_counter() - t0), "unit": UNIT, "cores": cores,
                 "sample": f"{cores} concurrent copies of the same {sample_nodes}-node sample, one per core"}
    line = {"impl": "reference", "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": args.gpus,
            "steps": args.steps, "warmup": args.warmup, "ms_per_step": 1e3 * tot_t / args.steps,
            "higher_is_better": True, "scaling": "strong", "vs_baseline": None, "dtype": "f64", "data": "synthetic",
            "config": workload_config(wl),
            "cpu_baseline": {"value": value, "unit": UNIT, "cores": 1, "kind": "port",
                             "sample": f"first {sample_nodes} nodes in DFS order per step; CPU restatement "
                                       "(oracle/blt_oracle.c), single thread like the reference; GaloisInc/blt itself "
                                       "cannot be built here (no NTL/GLPK/Boost)"},
            "e2e": {"value": value, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
            "all_cores": all_cores, "gpu_launches": 0}
    print(json.dumps(line))
    return 0


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=3)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--cpu-nodes", type=int, default=150000, help="size of the CPU baseline sample (nodes)")
    ap.add_argument("--no-cpu", action="store_true")
    args = ap.parse_args()

    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    if args.impl == "reference":
        return run_reference(args, rank, world)

    import torch
    import torch.distributed as dist

    import blt_b200

    if not torch.cuda.is_available():
        raise SystemExit("bench.py: no CUDA device; the search has no CPU fallback")
    torch.cuda.set_device(local_rank)
    host_group = None
    if world > 1:
        dist.init_process_group("nccl", device_id=torch.device("cuda", local_rank))
        # Barriers around the plugin call must not occupy a GPU: rank 0's call runs persistent kernels on EVERY device
        # (one CTA per SM), and an NCCL barrier kernel spinning on another rank's device would keep an SM from them.
        host_group = dist.new_group(backend="gloo")
    wl = load_workload()
    n, m = wl["n"], wl["m"]
    lat_h = torch.from_numpy(wl["lat"]).contiguous().pin_memory()
    con_h = torch.from_numpy(wl["con"]).contiguous().pin_memory()
    lat_d = lat_h.cuda(non_blocking=False)
    con_d = con_h.cuda(non_blocking=False)
    flush = torch.empty(256 << 20, dtype=torch.uint8, device="cuda")
    opts = dict(device=local_rank, rank=rank, world=world, split_depth=(6 if world > 1 else 0),
                timeout_ms=int(os.environ.get("BLT_BENCH_TIMEOUT_MS", "600000")))  # 0 under ncu: replays outlast any deadline

    def barrier():
        torch.cuda.synchronize()
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    shared = None
    if world > 1:  # stop word + on-demand subtree counter shared by the ranks (host memory; no collective)
        from blt_b200.dist import StopFlag
        shared = StopFlag()
        opts["stop_flag"] = shared.address

    def step_dev():
        flush.zero_()
        torch.cuda.synchronize()
        if shared is not None:
            shared.reset()
        r = blt_b200.closest_dev(lat_d.data_ptr(), con_d.data_ptr(), n, m, **opts)
        if r["rc"] != 0:
            raise SystemExit(f"bench.py: search failed with status {r['rc']}")
        return r

    expected_nodes = wl["nodes_expected"]

    # The plugin call: a context with the un-reduced system, as run_blt builds it (libblt/run_blt.cc:151-166); built once,
    # outside the timed region (it is the caller's input), checked K times.
    ctx = None
    if rank == 0:
        from blt_b200 import api
        A, C = wl["A"], wl["C"]
        ctx = api.Context()
        xs = [ctx.var(str(j)) for j in range(A.shape[1])]
        for i in range(A.shape[0]):
            e = ctx.const(0)
            for j in range(A.shape[1]):
                if A[i, j]:
                    e = e + int(A[i, j]) * xs[j]
            if ctx.assume(int(C[i, 0]), e, int(C[i, 1])) != 0:
                raise SystemExit("bench.py: blt_assume failed")

    def step_e2e():
        """One blt_check() of the whole system on rank 0 (all N GPUs from one process at N > 1)."""
        flush.zero_()
        torch.cuda.synchronize()
        if rank != 0:
            return dict(nodes=0, lll_ms=0.0, search_ms=0.0), 0.0
        os.environ["BLT_GPUS"] = str(world)
        os.environ["BLTCU_TIMEOUT_MS"] = str(opts["timeout_ms"])
        t0 = time.perf_counter()
        rc = ctx.check()
        dt = time.perf_counter() - t0
        if rc != 11:  # BLT_CHECK_UNSAT: the workload is an exhaustive (UNSAT) enumeration
            raise SystemExit(f"bench.py: blt_check returned {rc}")
        st = ctx.last_stats()
        return dict(nodes=st["niter"] + st["nprune"] + 2, lll_ms=st["lll_ms"], search_ms=st["search_ms"]), dt

    for _ in range(max(args.warmup, 3)):
        step_dev()
    sampler = ClockSampler(local_rank)
    if rank == 0:
        sampler.start()
    barrier()
    t_wall0 = time.perf_counter()
    kernel_ms, nodes, launches = 0.0, 0, 0
    hist = np.zeros(n + 1, dtype=np.float64)
    last = None
    for _ in range(args.steps):
        r = step_dev()
        kernel_ms += r["kernel_ms"]
        nodes += r["nodes"]
        launches += r["launches"]
        hist += r["hist"].astype(np.float64)
        last = r
    barrier()
    wall_s = time.perf_counter() - t_wall0
    # end to end through the plugin call
    def host_barrier():
        torch.cuda.synchronize()
        if world > 1:
            dist.barrier(group=host_group)

    e2e_s, e2e_nodes, e2e_lll_ms, e2e_search_ms = 0.0, 0, 0.0, 0.0
    host_barrier()
    step_e2e()  # untimed: the other devices' workspaces are allocated on first use
    host_barrier()
    for _ in range(args.steps):
        r, dt = step_e2e()
        host_barrier()
        e2e_s += dt
        e2e_nodes += r["nodes"]
        e2e_lll_ms += r["lll_ms"]
        e2e_search_ms += r["search_ms"]
    sampler.stop_flag = True

    # BASELINE.json config 5 at N > 1: 10,000 instances of 20 x 20, whole problems sharded over the ranks (supplementary)
    batch10k = None
    if world > 1:
        from blt_b200.dist import closest_batch_sharded
        zb = np.load(os.path.join(ROOT, "tests", "golden", "batch20.npz"))
        lats = np.ascontiguousarray(np.tile(zb["lat"], (25, 1, 1)))
        cons = np.ascontiguousarray(np.tile(zb["con"], (25, 1, 1)))
        closest_batch_sharded(lats[:400 * world], cons[:400 * world], device=local_rank)
        barrier()
        tb = time.perf_counter()
        rb = closest_batch_sharded(lats, cons, device=local_rank)
        barrier()
        dtb = torch.tensor([time.perf_counter() - tb], dtype=torch.float64, device="cuda")
        dist.all_reduce(dtb, op=dist.ReduceOp.MAX)
        okb = bool((rb["niter"] == np.tile(zb["niter"], 25)).all() and (rb["nprune"] == np.tile(zb["nprune"], 25)).all()
                   and (rb["status"] == np.tile(zb["rc"], 25)).all())
        batch10k = {"workload": f"config 5: 10,000 dense 20x20 instances (tests/golden/batch20.npz x25), {len(lats) // world} per GPU",
                    "instances_per_s": len(lats) / float(dtb.item()), "e2e_ms": 1e3 * float(dtb.item()),
                    "nodes_per_s": (int(rb["niter"].sum()) + int(rb["nprune"].sum()) + len(lats)) / float(dtb.item()), "count_ok": okb}

    # BASELINE.json config 3 at N > 1: time to the first point of a dimension-60 subset-sum instance, every rank stopped
    # by the shared stop word (supplementary)
    sat_exit = None
    if world > 1:
        from blt_b200.dist import closest_sharded
        zs = np.load(os.path.join(ROOT, "tests", "golden", "subset_sum60_b26.npz"))
        runs = []
        for sd in [int(x) for x in zs["seeds"]]:
            la, co = zs[f"lat{sd}"], zs[f"con{sd}"]
            closest_sharded(la, co, stop=shared, device=local_rank, timeout_ms=120000)
            barrier()
            t1 = time.perf_counter()
            rs = closest_sharded(la, co, stop=shared, device=local_rank, timeout_ms=120000)
            dts = torch.tensor([time.perf_counter() - t1], dtype=torch.float64, device="cuda")
            dist.all_reduce(dts, op=dist.ReduceOp.MAX)
            runs.append({"seed": sd, "found": bool(rs["found"]), "ms_until_every_rank_returned": 1e3 * float(dts.item()),
                         "nodes_all_ranks": int(rs["nodes"])})
        sat_exit = {"workload": "config 3: Lagarias-Odlyzko subset-sum lattice, 60 weights of 26 bits (tests/golden/subset_sum60_b26.npz), "
                                "first point ends the search on every rank (shared stop word)", "runs": runs}

    red = torch.tensor([kernel_ms, e2e_s], dtype=torch.float64, device="cuda")
    tot = torch.tensor([float(nodes), float(e2e_nodes), float(last["niter"]), float(last["nprune"]), float(launches)],
                       dtype=torch.float64, device="cuda")
    hist_t = torch.from_numpy(hist).cuda()
    if world > 1:
        dist.all_reduce(red, op=dist.ReduceOp.MAX)
        dist.all_reduce(tot, op=dist.ReduceOp.SUM)
        dist.all_reduce(hist_t, op=dist.ReduceOp.SUM)
    kernel_ms_max, e2e_s_max = red.tolist()
    nodes_all, e2e_nodes_all, niter_all, nprune_all, launches_all = tot.tolist()
    hist_all = hist_t.cpu().numpy()

    if rank == 0:
        value = nodes_all / (kernel_ms_max * 1e-3)
        flops = float(sum(hist_all[k] * f_alg(k, m) for k in range(n + 1)))
        peak = blt_b200.fp64_peak(local_rank, 5)
        int_peak = blt_b200.int32_peak(local_rank, 3)
        achieved = flops / (kernel_ms_max * 1e-3) / 1e12 / max(world, 1)
        line = {"metric": METRIC, "value": value, "unit": UNIT, "n_gpus": args.gpus, "steps": args.steps,
                "warmup": max(args.warmup, 3), "ms_per_step": kernel_ms_max / args.steps, "higher_is_better": True,
                "scaling": "strong", "vs_baseline": None, "dtype": "f64", "data": "synthetic",
                "config": workload_config(wl),
                "niter": int(niter_all), "nprune": int(nprune_all), "nodes_per_step": nodes_all / args.steps,
                "time_to_decision_ms": 1e3 * e2e_s_max / args.steps, "wall_ms_per_step": 1e3 * wall_s / args.steps,
                "pivots_per_node": last["pivots"] / max(1, last["nodes"]), "donated": last["donated"],
                "warps": last["warps"], "busy_frac": last["busy_frac"],
                "count_ok": bool(int(round(nodes_all / args.steps)) == expected_nodes),
                "e2e": {"value": e2e_nodes_all / e2e_s_max, "unit": UNIT, "h2d_bytes_per_step": (n * m + 2 * m) * 8 * world,
                        "d2h_bytes_per_step": (n * 8 + 1024) * world,
                        "call": "blt_check() on the un-reduced system (stretch, exact LLL on the host, search, back-substitution)"
                                + (f", one process driving {world} GPUs (BLT_GPUS)" if world > 1 else ""),
                        "ms_per_call": 1e3 * e2e_s_max / args.steps, "lll_ms": e2e_lll_ms / args.steps,
                        "search_ms": e2e_search_ms / args.steps, "count_ok": bool(e2e_nodes // args.steps == expected_nodes)},
                "gpu_launches": int(launches_all),
                "roofline": {"bound": "fp64", "achieved": achieved, "peak": peak, "unit": "TFLOP/s",
                             "frac": achieved / peak, "traffic": None, "int32_peak_tiops": int_peak,
                             "note": "achieved = sum_k N_k*F_alg(k,m) per GPU / search-kernel time; peak = FP64 FMA "
                                     "throughput measured on this GPU by bltcu_fp64_peak (MEASURED_PEAKS.json has no "
                                     "FP64 figure)"},
                "clocks": sampler.summary()}
        if batch10k is not None:
            line["batch10k"] = batch10k
        if sat_exit is not None:
            line["sat_early_exit"] = sat_exit
        if world == 1:
            line["batch"] = batch_supplement(blt_b200)
            prof = os.path.join(ROOT, "profiles", "r2_final_summary.json")
            if os.path.exists(prof):  # dram bytes of the search kernel from the committed ncu --set full capture
                pj = json.load(open(prof))
                if pj.get("fixture") == os.path.basename(FIXTURE):
                    line["roofline"]["traffic"] = pj.get("dram_bytes_per_launch")
                    line["roofline"]["traffic_source"] = "profiles/r2_final_summary.json (ncu --set full, one launch)"
                elif pj.get("dram_bytes_per_node"):
                    # the capture was taken on a smaller instance of the same workload (a replay of this one outlasts
                    # the search's watchdog): scaled by the node count, and said so
                    line["roofline"]["traffic"] = pj["dram_bytes_per_node"] * nodes_all / args.steps
                    line["roofline"]["traffic_source"] = (f"profiles/r2_final_summary.json: {pj['dram_bytes_per_node']:.0f} B/node measured "
                                                          f"by ncu --set full on {pj.get('fixture')}, times the nodes of this launch")
        if not args.no_cpu and world == 1:
            v, cn, dt = cpu_sample(wl, args.cpu_nodes)
            line["cpu_baseline"] = {"value": v, "unit": UNIT, "cores": 1, "kind": "port",
                                    "sample": f"first {cn} nodes of the same instance in DFS order, {dt:.1f} s, "
                                              "oracle/blt_oracle.c single thread"}
        print(json.dumps(line))
    if world > 1:
        shared.close()
        dist.destroy_process_group()
    return 0


if __name__ == "__main__":
    sys.exit(main())
