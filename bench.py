#!/usr/bin/env python
"""bench.py -- planner candidates/sec and Plan() latency on a synthetic 10k-node snapshot (BASELINE.json metric).

  python bench.py [--gpus N] [--steps K] [--warmup W] [--impl ours|reference] [--workload NAME]

A "step" is one Plan() over the packed snapshot (nosplan_run: one kernel, k_plan_fused, where the single-launch plan
applies).  `value` is measured with the snapshot already resident in HBM (CUDA events on the engine's stream, L2 flushed
between iterations; median step, slowest rank per step); `e2e` is the same Plan() through nosplan_plan with HOST
buffers, i.e. including the H2D copy of the snapshot and the D2H copy of the plan, by wall clock.  After the timed loops
every rank compares its rows and the job-wide counters with the oracle's plan of the full snapshot: the line carries
"parity" and the exit code is 1 on a mismatch.  One JSON line is printed by rank 0.

--impl reference times the reference's own algorithm on the host cores on the SAME workload: the reference is Go and
there is no Go toolchain on the box, so it is the C++ port (oracle/nos_oracle.cpp) -- the incremental mode at full size
(same plan bit for bit; one core, Plan() is single-goroutine), with the faithful O(N^2 P) mode on a sub-snapshot and 8
independent problems on 8 threads beside it.
"""
import argparse
import json
import os
import subprocess
import sys
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
for p in (ROOT, os.path.join(ROOT, "tests"), os.path.join(ROOT, "oracle")):
    if p not in sys.path:
        sys.path.insert(0, p)

import numpy as np  # noqa: E402

WORKLOADS = {
    # name: (kind, nodes, gpus/node, pods, seed)
    "mig-10k": ("mig", 10000, 8, 10000, 0xC5),      # north star: 10k-node x 8-GPU MIG snapshot, 10k pending pods
    "mig-c2": ("mig", 1000, 8, 5000, 0xC2),         # BASELINE config #2
    "mps-c3": ("mps", 5000, 8, 20000, 0xC3),        # BASELINE config #3
    "mig-100k": ("mig", 100000, 8, 10000, 0xC5),    # scale sweep upper end
}
FILTER_RESOURCES_FIT = 2


def make_workload(name, seed_offset=0, nodes=None, pods=None):
    from nos_b200 import synth
    kind, n, g, p, seed = WORKLOADS[name]
    n = nodes or n
    p = pods or p
    if kind == "mig":
        return synth.mig_cluster(n, g, p, seed=seed + seed_offset, regime="saturating")
    return synth.mps_cluster(n, g, p, seed=seed + seed_offset, regime="saturating")


def algorithmic_bytes(snap, pods):
    """Compulsory traffic of one Plan(): every input element read once, every output written once, with the
    logical row sizes of SURVEY.md section 8(d): 256 B/node (G=8), 32 B/pod, 16 B/GPU of output, tables."""
    n, g, p = snap.n_nodes, snap.gpus_per_node, pods.n_pods
    b_node = g * 16 + 48 + 8 + 64
    return n * b_node + p * 32 + int(snap.geom_counts.size) + n * g * 16


def measured_peak():
    path = os.path.join(ROOT, "MEASURED_PEAKS.json")
    try:
        return float(json.load(open(path))["hbm_gbs"]), "measured (MEASURED_PEAKS.json hbm_gbs)"
    except Exception:
        return 6650.0, "fallback (B200_PROFILING.md)"


def recorded_traffic(kernel):
    """dram__bytes_read + dram__bytes_write per launch of `kernel` from the committed ncu --set full capture of the
    mig-10k workload (profiles/dominant_kernel_traffic.json); None for any other workload or kernel."""
    try:
        return json.load(open(os.path.join(ROOT, "profiles", "dominant_kernel_traffic.json")))[kernel]["dram_bytes_per_launch"]
    except Exception:
        return None


class ClockSampler:
    Q = "clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.active,clocks_event_reasons.hw_slowdown," \
        "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap"

    def __init__(self, index):
        """index: GPU index, or a comma-separated list (one nvidia-smi process polls them all)."""
        self.index, self.proc, self.path = index, None, f"/tmp/nosplan_clocks_{os.getpid()}.csv"

    def start(self):
        try:
            self.f = open(self.path, "w")
            self.proc = subprocess.Popen(["nvidia-smi", "-i", str(self.index), f"--query-gpu={self.Q}",
                                          "--format=csv,noheader,nounits", "-lms", "100"], stdout=self.f, stderr=subprocess.DEVNULL)
        except Exception:
            self.proc = None

    def stop(self):
        if not self.proc:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        time.sleep(0.15)
        self.proc.terminate()
        try:
            self.proc.wait(timeout=5)
        except Exception:
            self.proc.kill()
        self.f.close()
        sm, smax, reasons = [], None, set()
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        for line in open(self.path):
            parts = [x.strip() for x in line.split(",")]
            if len(parts) < 8:
                continue
            try:
                sm.append(float(parts[0])); smax = float(parts[1])
            except ValueError:
                continue
            for nme, val in zip(names, parts[4:8]):
                if val.lower().startswith("active"):
                    reasons.add(nme)
        try:
            os.remove(self.path)
        except OSError:
            pass
        return {"sm_mhz": float(np.median(sm)) if sm else None, "sm_max_mhz": smax, "reasons": sorted(reasons), "samples": len(sm)}


def core_config(args, world):
    """The keys both arms share: they name the workload, nothing else."""
    kind, n_full, gg, p_full, _ = WORKLOADS[args.workload]
    nodes = n_full * world if args.scaling == "weak" else n_full
    return {"workload": args.workload, "kind": kind, "nodes": nodes, "gpus_per_node": gg, "pods": p_full,
            "regime": "saturating", "filter": "resources_fit", "scaling": args.scaling}


def job_workload(args, world):
    """The snapshot and pods of the whole job: weak scaling keeps the workload's node count PER GPU, strong scaling in total."""
    kind, n_full, gg, p_full, _ = WORKLOADS[args.workload]
    return make_workload(args.workload, nodes=n_full * world if args.scaling == "weak" else n_full)


def cpu_reference(snap, pods, steps, budget_s, threads=1):
    """The reference's Plan() on the host: the incremental port (oracle/nos_oracle.cpp: running totals instead of the
    reference's re-summing and cloning -- the same plan, bit for bit, at a cost one core can pay at full size).
    -> (times [s], stats of the last run).  threads > 1: that many INDEPENDENT problems at once (Plan() has no exact parallel
    form: core/planner.go:92-150 is one goroutine), each thread times its own."""
    from support import oracle_lib as ol
    times, st = [], None
    t_end = time.perf_counter() + budget_s
    if threads == 1:
        for _ in range(steps):
            _, _, st, sec = ol.plan(snap, pods, None, FILTER_RESOURCES_FIT, ol.INCREMENTAL)
            times.append(sec)
            if time.perf_counter() > t_end:
                break
        return times, st
    import threading
    out = [None] * threads

    def work(k):
        out[k] = ol.plan(snap, pods, None, FILTER_RESOURCES_FIT, ol.INCREMENTAL)      # ctypes releases the GIL

    t0 = time.perf_counter()
    ths = [threading.Thread(target=work, args=(k,)) for k in range(threads)]
    for t in ths:
        t.start()
    for t in ths:
        t.join()
    return [time.perf_counter() - t0], out[0][2]


def faithful_sample(args, reps):
    """The reference's own asymptotics (re-sum every node per gate, clone every node per fork: O(N^2 P)) cannot run the full
    workload (hours); a 450-node sub-snapshot of the same recipe, and the N^2 P extrapolation, are reported beside the
    full-size figure."""
    from support import oracle_lib as ol
    kind, n_full, gg, p_full, _ = WORKLOADS[args.workload]
    n = min(n_full, 450)
    p = max(10, int(round(n * p_full / n_full)))
    snap, pods = make_workload(args.workload, nodes=n, pods=p)
    secs, st = [], None
    for _ in range(reps):
        _, _, st, sec = ol.plan(snap, pods, None, FILTER_RESOURCES_FIT, ol.FAITHFUL)
        secs.append(sec)
    p50 = float(np.median(secs))
    per = p50 / (float(n) * n * p)
    return {"nodes": n, "pods": p, "repetitions": reps, "seconds_p50": p50, "candidates_per_s": st["candidates"] / p50,
            "fitted_seconds_per_N2P": per, "extrapolated_seconds_full_size": per * float(n_full) ** 2 * p_full,
            "note": "faithful C++ port of core.Planner.Plan (same asymptotics as the Go planner) on a sub-snapshot of the same recipe; "
                    "cost per candidate grows linearly with N"}


def run_reference(args):
    """The reference arm: the reference's algorithm on the host cores, on the SAME workload as our arm (at N GPUs: the
    N-times-larger snapshot of the weak-scaling job).  The reference is Go and there is no Go toolchain on the box, so this is
    the C++ port (oracle/): `value` is the incremental port at full size on one core (Plan() is single-goroutine); the
    faithful O(N^2 P) port on a sub-snapshot and 8 independent problems on 8 threads are extra keys."""
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    import __graft_entry__ as g
    if not os.path.exists(os.path.join(ROOT, "oracle", "libnosoracle.so")):
        g.build()
    world = max(1, args.gpus)
    snap, pods = job_workload(args, world)
    steps = max(1, args.steps)
    for _ in range(min(1, args.warmup)):
        cpu_reference(snap, pods, 1, 1e9)
    times, st = cpu_reference(snap, pods, steps, budget_s=150.0)
    cands = int(st["candidates"])
    p50 = float(np.median(times))
    value = cands / p50
    cores = os.cpu_count() or 1
    nthreads = min(8, cores)
    tt, _ = cpu_reference(snap, pods, 1, 1e9, threads=nthreads)
    sample = (f"incremental C++ port of core.Planner.Plan (same plan as the reference, bit for bit; running totals instead of re-summing "
              f"and cloning) on the FULL workload, {len(times)} of {steps} requested repetitions (wall-time cap 150 s), one thread "
              f"(Plan() is single-goroutine; the Go planner itself cannot run: no Go toolchain on the box)")
    line = {
        "impl": "reference", "metric": "planner_candidates_per_sec", "value": value, "unit": "candidates/s",
        "n_gpus": args.gpus, "steps": len(times), "warmup": min(1, args.warmup), "ms_per_step": 1e3 * p50,
        "higher_is_better": True, "scaling": args.scaling, "vs_baseline": None, "dtype": "int64", "data": "synthetic",
        "config": core_config(args, world),
        "plan_p50_ms": 1e3 * p50, "plan_mean_ms": 1e3 * float(np.mean(times)), "repetitions": len(times),
        "cpu_baseline": {"value": value, "unit": "candidates/s", "cores": 1, "kind": "port", "sample": sample,
                         "host_cores": cores,
                         "independent_problems_on_threads": {"threads": nthreads, "value": nthreads * cands / tt[0], "seconds": tt[0],
                                                             "note": "the same problem planned by every thread at once, labelled replicas: Plan() has no exact parallel form"},
                         "faithful_sub_snapshot": faithful_sample(args, 5)},
        "e2e": {"value": value, "unit": "candidates/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
    }
    print(json.dumps(line), flush=True)


def run_ours(args):
    import torch
    import torch.distributed as dist
    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if world > 1:
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    torch.cuda.set_device(local)
    import __graft_entry__ as g
    from nos_b200 import capi
    if not os.path.exists(g.LIB):
        g.build()

    # N > 1: ONE Plan() over the job's snapshot in slice mode -- every rank holds ONLY its own slice of the name-ordered node list
    # (H2D, the planning kernel and the plan rows are per-rank work); one merged exchange per Plan() (see nosplan_run_sliced).
    kind, n_full, gg, p_full, _ = WORKLOADS[args.workload]
    full_snap, pods = job_workload(args, world)
    snap = full_snap
    lo, hi = 0, full_snap.n_nodes
    if world > 1:
        from nos_b200 import sharded
        chunk = (full_snap.n_nodes + world - 1) // world
        lo, hi = sharded.shard_range(full_snap.n_nodes, rank, world)
        snap = sharded.slice_snapshot(full_snap, lo, hi, pad_to=chunk)
    eng = capi.Engine(local)
    eng.upload(snap, pods, None, FILTER_RESOURCES_FIT)
    dev = torch.device("cuda", local)
    ext = torch.cuda.ExternalStream(eng.stream(), device=dev)
    flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)     # > 126 MB L2
    sharded_run = world > 1
    if sharded_run:                          # the exchange is issued by the library itself (peer-memory mailboxes, or one NCCL all-gather)
        uid = [eng.nccl_unique_id() if rank == 0 else None]
        dist.broadcast_object_list(uid, src=0)
        eng.shard_init(rank, world, uid[0])

    def sync_all():
        torch.cuda.synchronize()
        if world > 1:
            dist.barrier()
            torch.cuda.synchronize()

    tok = torch.zeros(1, dtype=torch.int32, device=dev)

    def one_step(timed):
        if sharded_run:                      # ranks enter the step together: a late peer would otherwise be timed as waiting
            sync_all()
        with torch.cuda.stream(ext):         # L2 flush; the Plan()'s launches queue up behind it exactly as on one GPU
            flush.zero_()
            if sharded_run:
                # the 256 MiB flush does not take the same time on every GPU: a one-element all-reduce on the engine's stream
                # brings the ranks' device timelines back together in front of the timed region (it is not part of it)
                dist.all_reduce(tok)
        s = torch.cuda.Event(enable_timing=True); e = torch.cuda.Event(enable_timing=True)
        s.record(ext)
        if sharded_run:
            eng.run_sliced()
        else:
            eng.run()
        e.record(ext)
        cnt, msk, st = eng.download()        # synchronises; outside the timed pair of events
        return s.elapsed_time(e), st, cnt, msk

    for _ in range(max(3, args.warmup)):
        one_step(False)
    # one poller for the whole job (rank 0, every GPU of the job): NVML queries from 8 processes at once perturb the launches
    sampler = ClockSampler(",".join(str(i) for i in range(world))) if rank == 0 else None
    sync_all()
    if sampler:
        sampler.start()
    step_ms, sweep_ms, par_ms, st = [], [], [], None
    eng.set_kernel_events(False)             # the timed steps run without the library's per-kernel events ...
    for _ in range(args.steps):
        ms, st, cnt, msk = one_step(True)
        step_ms.append(ms)
    sync_all()
    eng.set_kernel_events(True)              # ... which are then taken on the same steps again, for the roofline line
    for _ in range(max(5, args.steps // 2)):
        _, kst, _, _ = one_step(True)
        sweep_ms.append(kst["ms_sweep"]); par_ms.append(kst["ms_parallel"])
    st = dict(st, ms_prepass=kst["ms_prepass"], ms_tracker=kst["ms_tracker"])
    sync_all()
    # ---- end to end: host buffers in, host buffers out, wall clock (H2D + kernels [+ exchange] + D2H inside)
    hsnap = capi.pin_snapshot(snap)      # the caller's snapshot arrays and plan buffers sit in page-locked host memory
    hout = capi.pinned_partitioning(snap)
    # ... and the caller has just packed the snapshot into them, as the shim's packer does before every Plan() (measured on
    # this pool: the copy engine reads page-locked arrays the CPU has written since they were pinned at ~35 GB/s, arrays
    # only ever written by the pinning copy at ~15 GB/s)
    capi.refresh_pinned(hsnap, snap)

    # one GPU: the timed region is the C-ABI call nosplan_plan itself, entered through structs built beforehand (the ctypes
    # marshalling of twenty arrays is this harness's cost, not the caller's)
    call = eng.prepared_plan(hsnap, pods, None, FILTER_RESOURCES_FIT, out=hout) if not sharded_run else None

    def e2e_once():
        if not sharded_run:
            return call()
        eng.upload(hsnap, pods, None, FILTER_RESOURCES_FIT)
        eng.run_sliced()
        return eng.download(out=hout)[2]

    for _ in range(2):
        e2e_once()
    sync_all()
    e2e_s = []
    for _ in range(args.steps):
        if sharded_run:
            sync_all()
        t0 = time.perf_counter()
        est = e2e_once()
        e2e_s.append(time.perf_counter() - t0)
    if not sharded_run:
        call.check(est)
        est = call.result()[2]
    sync_all()
    # the same call with the node side in compact form (nosplan_plan_compact: one record per node, one DMA, expanded on the
    # GPU): what a packer that is free to choose its layout hands over.  It becomes `e2e`; the SoA call stays beside it.
    soa = None
    if not sharded_run and snap.n_profiles <= 8:
        rec_t = torch.empty((snap.n_nodes, capi.compact_record_bytes(snap.gpus_per_node, len(snap.other_resources))), dtype=torch.uint8).pin_memory()
        rec = rec_t.numpy()
        capi.compact_records(snap, out=rec)                      # the caller packs into its page-locked buffer
        ccall = eng.prepared_plan_compact(snap, rec, pods, None, FILTER_RESOURCES_FIT, out=hout)
        for _ in range(3):
            ccall()
        c_s = []
        for _ in range(args.steps):
            t0 = time.perf_counter()
            crc = ccall()
            c_s.append(time.perf_counter() - t0)
        ccall.check(crc)
        ccnt, cmsk, cst = ccall.result()
        if np.array_equal(ccnt, cnt) and np.array_equal(cmsk, msk):
            soa = (e2e_s, est)
            e2e_s, est = c_s, cst
    sync_all()
    # the same call against the snapshot the last e2e call left in HBM (nosplan_plan_resident): only the pods travel in
    res_s = []
    if not sharded_run:
        rcall = eng.prepared_plan(hsnap, pods, None, FILTER_RESOURCES_FIT, out=hout, resident=True)
        for _ in range(2):
            rcall()
        for _ in range(args.steps):
            t0 = time.perf_counter()
            rrc = rcall()
            res_s.append(time.perf_counter() - t0)
        rcall.check(rrc)
        rst = rcall.result()[2]
        sync_all()
    clocks = sampler.stop() if sampler else None

    # ---- parity, outside every timed region: this rank's rows and the job-wide counters against the oracle's plan of the
    #      FULL snapshot (the name-order sweep of core/planner.go:92-150 that the slices must reproduce)
    from support import oracle_lib as ol
    t0 = time.perf_counter()
    ocnt, omsk, ost, osec = ol.plan(full_snap, pods, None, FILTER_RESOURCES_FIT, ol.INCREMENTAL)
    n_rows = (hi - lo) * gg
    same = bool(np.array_equal(cnt[:n_rows], ocnt[lo * gg:hi * gg]) and np.array_equal(msk[:n_rows], omsk[lo * gg:hi * gg]))
    same = same and all(int(st[k]) == int(ost[k]) for k in ("candidates", "visited_nodes", "placements", "geometry_evals", "lacking_at_end"))
    step = np.asarray(step_ms, np.float64)
    e2e = np.asarray(e2e_s, np.float64)
    if world > 1:                          # per step: the slowest rank; parity: every rank
        t = torch.tensor(np.concatenate([step, e2e, [0.0 if same else 1.0]]), dtype=torch.float64, device=f"cuda:{local}")
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        t = t.cpu().numpy()
        step, e2e, same = t[:len(step)], t[len(step):len(step) + len(e2e)], t[-1] == 0.0
    cands = int(st["candidates"])          # visited nodes (all ranks) x pods: already the whole job
    if rank != 0:
        eng.close()
        if world > 1:
            dist.destroy_process_group()
        sys.exit(0 if same else 1)

    p50_ms, mean_ms = float(np.median(step)), float(np.mean(step))
    value = cands / (p50_ms / 1e3)
    e2e_p50 = float(np.median(e2e))
    e2e_value = cands / e2e_p50
    peak, peak_src = measured_peak()
    b_plan = algorithmic_bytes(snap, pods)       # per GPU: its own node rows and the (replicated) pods
    sweep_avg_ms = float(np.mean(sweep_ms))
    # dominant kernel: the single-launch plan, else the node-parallel pass when it produced the plan, else the sequential sweep
    fused = bool(st.get("fused"))
    dom_name = "k_plan_fused" if fused else ("k_plan_parallel" if st["path"] == 1 else "k_plan_sweep")
    dom_ms = float(np.mean(par_ms)) if st["path"] == 1 else sweep_avg_ms
    achieved = b_plan / (dom_ms / 1e3) / 1e9
    if world == 1:
        par = "single GPU"
    else:
        par = (f"node-sliced x{world}: each GPU holds {hi - lo} nodes of the name-ordered list, pods replicated; "
               + ("ONE exchange per Plan(): the last CTA of k_plan_fused stores the rank's block (totals, counters, placed bitmap) into every "
                  "rank's mailbox and k_fused_verify folds them" if fused else "two exchanges per Plan() (cluster totals, then counters + placed-key bitmap)")
               + f"; plan rows stay on their rank; transport: {eng.shard_transport()}")
    line = {
        "metric": "planner_candidates_per_sec", "value": value, "unit": "candidates/s",
        "n_gpus": world, "steps": args.steps, "warmup": max(3, args.warmup), "ms_per_step": p50_ms,
        "higher_is_better": True, "scaling": args.scaling, "vs_baseline": None, "dtype": "int64", "data": "synthetic",
        "config": core_config(args, world),
        "parity": "identical" if same else "MISMATCH",
        "parity_note": "rows of this rank's nodes and the job-wide counters against the oracle's plan of the full snapshot, every rank, after the timed loops",
        "timing": {"value_from": "median step (p50), device time, max over ranks per step", "plan_p50_ms": p50_ms, "plan_mean_ms": mean_ms,
                   "plan_p95_ms": float(np.percentile(step, 95)), "plan_max_ms": float(np.max(step)),
                   "l2": "flushed between iterations (256 MiB device write)"},
        "details": {"parallelism": par, "candidates_per_plan": cands, "visited_nodes": int(st["visited_nodes"]), "placements": int(st["placements"]),
                    "nodes_per_gpu": hi - lo},
        "plan_p50_ms": p50_ms, "plan_p95_ms": float(np.percentile(step, 95)),
        "kernel_ms": {"prepass": float(st["ms_prepass"]), "tracker_init" if world == 1 else "totals_exchange+tracker_init": float(st["ms_tracker"]),
                      "parallel": float(np.mean(par_ms)),
                      "verify_or_sweep" if world == 1 else "exchange+verify": sweep_avg_ms,
                      "path": "single launch (k_plan_fused)" if fused else ("node-parallel" if st["path"] == 1 else "sequential")},
        "roofline": {"bound": "hbm", "kernel": dom_name, "kernel_ms": dom_ms, "achieved": achieved, "peak": peak, "unit": "GB/s",
                     "frac": achieved / peak, "traffic": recorded_traffic(dom_name) if (args.workload == "mig-10k" and world == 1) else None, "algorithmic_bytes": b_plan, "peak_source": peak_src},
        "e2e": {"value": e2e_value, "unit": "candidates/s", "h2d_bytes_per_step": int(est["h2d_bytes"]),
                "d2h_bytes_per_step": int(est["d2h_bytes"]), "ms_per_step": 1e3 * e2e_p50, "ms_mean": 1e3 * float(np.mean(e2e)),
                "ms_h2d": float(est["ms_h2d"]), "ms_kernels": float(est["ms_kernels"]), "ms_d2h": float(est["ms_d2h"]),
                "call": ("nosplan_plan_compact" if soa else "nosplan_plan") if world == 1 else "nosplan_upload + nosplan_run_sliced + nosplan_download",
                "host_buffers": "caller's node records (compact form) and plan buffers page-locked and DMA'd in place; pod arrays (pageable) staged "
                                "for the single-launch kernel in two passes while the node DMA runs"
                                if world == 1 else "per rank: its slice's node arrays and plan rows page-locked and DMA'd in place"},
        "gpu_launches": int(st["kernel_launches"]) * args.steps,
        "clocks": clocks,
    }
    if soa:
        line["e2e_soa"] = {"value": cands / float(np.median(soa[0])), "unit": "candidates/s", "ms_per_step": 1e3 * float(np.median(soa[0])),
                           "h2d_bytes_per_step": int(soa[1]["h2d_bytes"]), "d2h_bytes_per_step": int(soa[1]["d2h_bytes"]),
                           "call": "nosplan_plan (twelve SoA node arrays, page-locked, DMA'd in place)"}
    if res_s:
        line["e2e_resident_snapshot"] = {
            "value": cands / float(np.median(res_s)), "unit": "candidates/s", "ms_per_step": 1e3 * float(np.median(res_s)),
            "h2d_bytes_per_step": int(rst["h2d_bytes"]), "d2h_bytes_per_step": int(rst["d2h_bytes"]),
            "note": "nosplan_plan_resident: node arrays kept in HBM from the previous call (the device-side twin of the reference's "
                    "ClusterState cache), pods in and plan out through host buffers; reported next to e2e, not instead of it"}
    if world == 1 and args.batch > 1:
        line["batched"] = batched_plans(torch, capi, eng, ext, flush, args, local)
    if world == 1 and not args.no_cpu_baseline:
        times, ist = cpu_reference(snap, pods, 5, budget_s=20.0)
        ip50 = float(np.median(times))
        line["cpu_baseline"] = {
            "value": ist["candidates"] / ip50, "unit": "candidates/s", "cores": 1, "kind": "port",
            "sample": f"incremental C++ port of the reference algorithm (same plan, bit for bit) on the full workload, p50 of {len(times)} "
                      f"repetitions, {ip50:.2f} s each, one core (Plan() is single-goroutine)",
            "seconds_p50": ip50, "faithful_sub_snapshot": faithful_sample(args, 5)}
    print(json.dumps(line), flush=True)
    eng.close()
    if world > 1:
        dist.destroy_process_group()
    if not same:
        sys.exit(1)


def quota_algorithmic_bytes(pq, run, pend):
    """Compulsory traffic of one recomputation: every pod row (namespace id, request row, key mask), the
    quota tables (min, max, used) and every output (used, guaranteed, aggregates; one verdict / label byte per pod) once."""
    R = pq.min.shape[1]
    return (run.n_pods + pend.n_pods) * (4 + 8 * R + 4) + pq.min.size * 8 * 3 + run.n_pods + pend.n_pods + pq.min.size * 8 * 3


def run_quota(args):
    """BASELINE config #4: 500 ElasticQuotas, 100 000 running + 10 000 pending pods.  One step = one nosplan_quota_recompute
    (Used, aggregates, guaranteed over-quotas, PreFilter verdicts, operator labels).  The entry point takes host buffers, so
    `value` is pods/s over the kernels alone (the library's CUDA events around them, data already in HBM) and `e2e` the whole
    call by wall clock.  --impl reference: the C++ quota oracle on the same inputs, one core."""
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    import __graft_entry__ as g
    from nos_b200 import synth
    from support import oracle_lib as ol
    pq, run, pend = synth.quota_state(500, 100_000, 10_000, seed=0xC4)
    pods = run.n_pods + pend.n_pods
    config = {"workload": "quota-c4", "quotas": 500, "running_pods": run.n_pods, "pending_pods": pend.n_pods, "resource_columns": int(pq.min.shape[1])}
    if args.impl == "reference":
        if not os.path.exists(os.path.join(ROOT, "oracle", "libnosoracle.so")):
            g.build()
        secs = []
        for _ in range(max(1, args.warmup) + args.steps):
            secs.append(ol.quota(pq, run, pend)["seconds"])
        p50 = float(np.median(secs[max(1, args.warmup):]))
        v = pods / p50
        print(json.dumps({"impl": "reference", "metric": "quota_pods_per_sec", "value": v, "unit": "pods/s", "n_gpus": args.gpus, "steps": args.steps,
                          "warmup": args.warmup, "ms_per_step": 1e3 * p50, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
                          "dtype": "int64", "data": "synthetic", "config": config,
                          "cpu_baseline": {"value": v, "unit": "pods/s", "cores": 1, "kind": "port",
                                           "sample": "C++ quota oracle (oracle/nos_quota_oracle.cpp: reserveResource, aggregates, guaranteed over-quotas, PreFilter, operator labelling) on the full workload, one core"},
                          "e2e": {"value": v, "unit": "pods/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}, "gpu_launches": 0}), flush=True)
        return
    import torch
    from nos_b200 import capi
    if not os.path.exists(g.LIB):
        g.build()
    eng = capi.Engine(0)
    for _ in range(max(3, args.warmup)):
        got, st = eng.quota_recompute(pq, run, pend)
    sampler = ClockSampler("0")
    sampler.start()
    k_ms, call_s = [], []
    for _ in range(args.steps):
        t0 = time.perf_counter()
        got, st = eng.quota_recompute(pq, run, pend)
        call_s.append(time.perf_counter() - t0)
        k_ms.append(st["ms_kernels"])
    clocks = sampler.stop()
    exp = ol.quota(pq, run, pend)
    same = all(np.array_equal(np.asarray(got[k]), np.asarray(exp[k])) for k in exp if k in got and k != "seconds")
    kp50, cp50 = float(np.median(k_ms)), float(np.median(call_s))
    peak, peak_src = measured_peak()
    b_alg = quota_algorithmic_bytes(pq, run, pend)
    achieved = b_alg / (kp50 / 1e3) / 1e9
    line = {"metric": "quota_pods_per_sec", "value": pods / (kp50 / 1e3), "unit": "pods/s", "n_gpus": 1, "steps": args.steps, "warmup": max(3, args.warmup),
            "ms_per_step": kp50, "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "int64 (float64 only in GetGuaranteedOverquotas)",
            "data": "synthetic", "config": config, "parity": "identical" if same else "MISMATCH",
            "timing": {"value_from": "median of the library's CUDA-event time around the kernels of one call; inputs (7.8 MB) larger than nothing cached: every call uploads them afresh",
                       "kernel_launches_per_step": int(st["kernel_launches"])},
            "roofline": {"bound": "hbm", "kernel": "the whole recomputation (all its kernels)", "kernel_ms": kp50, "achieved": achieved, "peak": peak, "unit": "GB/s",
                         "frac": achieved / peak, "traffic": None, "algorithmic_bytes": int(b_alg), "peak_source": peak_src},
            "e2e": {"value": pods / cp50, "unit": "pods/s", "h2d_bytes_per_step": int(st["h2d_bytes"]), "d2h_bytes_per_step": int(st["d2h_bytes"]),
                    "ms_per_step": 1e3 * cp50, "ms_h2d": float(st["ms_h2d"]), "ms_kernels": float(st["ms_kernels"]), "ms_d2h": float(st["ms_d2h"])},
            "cpu_baseline": {"value": pods / exp["seconds"], "unit": "pods/s", "cores": 1, "kind": "port",
                             "sample": f"C++ quota oracle on the full workload, one repetition, {exp['seconds'] * 1e3:.0f} ms, one core"},
            "gpu_launches": int(st["kernel_launches"]) * args.steps, "clocks": clocks}
    print(json.dumps(line), flush=True)
    eng.close()
    if not same:
        sys.exit(1)


def batched_plans(torch, capi, eng0, ext0, flush, args, local):
    """B independent Plan() problems in flight at once on ONE GPU -- what the reference's two controllers
    (MIG and MPS planners, cmd/gpupartitioner/gpupartitioner.go:215-250) or several pending-pod batches give.  Every problem
    is a snapshot of the bench workload's recipe under its own seed on its own handle and stream, resident in HBM; all streams
    are released by one event behind the L2 flush and the step lasts until the last one finishes.  Reported next to the
    single-problem figure, never instead of it."""
    B = args.batch
    engs, exts, cands = [eng0], [ext0], 0
    dev = torch.device("cuda", local)
    for b in range(1, B):
        snap_b, pods_b = make_workload(args.workload, seed_offset=b)
        e = capi.Engine(local)
        e.upload(snap_b, pods_b, None, FILTER_RESOURCES_FIT)
        engs.append(e); exts.append(torch.cuda.ExternalStream(e.stream(), device=dev))
    for e in engs:
        e.set_kernel_events(False)

    def step():
        with torch.cuda.stream(ext0):
            flush.zero_()
        go = torch.cuda.Event(enable_timing=True)
        go.record(ext0)
        for x in exts[1:]:
            x.wait_event(go)
        for e in engs:
            e.run()
        ends = []
        for x in exts:
            ev = torch.cuda.Event(enable_timing=True); ev.record(x); ends.append(ev)
        sts = [e.download()[2] for e in engs]                 # synchronises every stream
        return max(go.elapsed_time(ev) for ev in ends), sum(int(st["candidates"]) for st in sts)

    for _ in range(3):
        step()
    ms = []
    for _ in range(args.steps):
        t, cands = step()
        ms.append(t)
    for e in engs[1:]:
        e.close()
    eng0.set_kernel_events(True)
    return {"problems": B, "value": cands * len(ms) / (sum(ms) / 1e3), "unit": "candidates/s", "ms_per_step": float(np.mean(ms)),
            "ms_p50": float(np.median(ms)), "candidates_per_step": cands,
            "note": f"{B} independent snapshots of the same recipe (different seeds), one handle and stream each, all resident"}


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--workload", default="mig-10k", choices=sorted(WORKLOADS) + ["quota-c4"])
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--scaling", default="weak", choices=["weak", "strong"],
                    help="N > 1: weak = the workload's node count per GPU (default, what the driver's scaling run uses), strong = in total")
    ap.add_argument("--batch", type=int, default=4, help="N=1 only: also time this many independent Plan() problems in flight at once (0/1 = skip)")
    args = ap.parse_args()
    if args.workload == "quota-c4":
        run_quota(args)
        return
    if args.impl == "reference":
        run_reference(args)
    else:
        run_ours(args)


if __name__ == "__main__":
    main()
