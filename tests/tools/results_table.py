"""BASELINE.md sections 3-4 as one results table (run on a GPU box): every BASELINE config, both demand regimes and both
filter modes -- GPU Plan() (kernels with the snapshot resident; end to end with host buffers), the incremental CPU port at
full size, and the faithful CPU port (the reference's own O(N^2 P) asymptotics) at N in {100, 250, 500, 1000} with the
fitted a*N^2*P extrapolated to each config.  Every GPU plan is compared with the oracle before its time is printed.
  python tests/tools/results_table.py > profiles/rNN_results.md"""
import os, sys, time
ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path[:0] = [ROOT, ROOT + "/tests", ROOT + "/oracle"]
import numpy as np
import torch
from nos_b200 import capi, packing as pk, synth
from support import oracle_lib as ol

CONFIGS = [  # name, kind, nodes, gpus, pods, seed
    ("C1", "mig", 1, 1, 10, 0xC1), ("C2", "mig", 1000, 8, 5000, 0xC2), ("C3", "mps", 5000, 8, 20000, 0xC3),
    ("C5@10k", "mig", 10000, 8, 10000, 0xC5), ("C5@100k", "mig", 100000, 8, 10000, 0xC5),
    # demand far below the reconfigurable supply: the lacking vector empties after a few nodes (early exit, planner.go:94-97);
    # the node-parallel speculation is rejected a posteriori and the sequential sweep produces the plan
    ("C2, 300 pods", "mig", 1000, 8, 300, 0xC2), ("C3, 600 pods", "mps", 5000, 8, 600, 0xC3), ("C5@10k, 1000 pods", "mig", 10000, 8, 1000, 0xC5),
    # the middle regime: the lacking vector is non-empty but small, the sweep exits part-way (planner.go:94-97)
    ("C5@10k, 40 pods", "mig", 10000, 8, 40, 0xC5), ("C5@10k, 200 pods", "mig", 10000, 8, 200, 0xC5),
    ("C5@100k, 40 pods", "mig", 100000, 8, 40, 0xC5), ("C5@100k, 200 pods", "mig", 100000, 8, 200, 0xC5),
    # what forces the exact sequential paths: one pod class asking for two profiles; one node whose accounts disagree with its GPUs
    ("C5@10k, 1000 pods, one two-profile class", "mig", 10000, 8, 1000, 0xC5), ("C5@10k, 1000 pods, one inconsistent node", "mig", 10000, 8, 1000, 0xC5),
    ("C5@10k, 40 pods, one inconsistent node", "mig", 10000, 8, 40, 0xC5),
    # C5 as specified: half MIG nodes (C2 recipe), half MPS nodes (C3 recipe), planned per kind as the reference's two controllers do
    ("C5 mixed@10k, MIG half", "mig", 5000, 8, 5000, 0xC5), ("C5 mixed@10k, MPS half", "mps", 5000, 8, 5000, 0xC6),
]


def make(name, kind, n, g, p, seed, regime):
    if name == "C1":
        sys.path.insert(0, ROOT + "/tests")
        import test_plan_gpu
        s, pods, _, _ = test_plan_gpu.c1_snapshot()
        return s, pods
    gen = synth.mig_cluster if kind == "mig" else synth.mps_cluster
    snap, pods = gen(n, g, p, seed=seed, regime=regime)
    if "two-profile" in name:                       # the last tenth of the pods also ask for one device of the next profile
        K = snap.n_profiles
        ptr, col, qty = [0], [], []
        for i in range(pods.n_pods):
            c = int(pods.scalar_col[i])
            col.append(c); qty.append(1)
            if i >= pods.n_pods - pods.n_pods // 10 and c + 1 < 3 + K:
                col.append(c + 1); qty.append(1)
            ptr.append(len(col))
        pods.scalar_ptr = np.array(ptr, np.int32); pods.scalar_col = np.array(col, np.int32); pods.scalar_qty = np.array(qty, np.int64)
    if "inconsistent" in name:
        snap.node_requested[7, 3] += 5
    return snap, pods
eng = capi.Engine(0)
ext = torch.cuda.ExternalStream(eng.stream())
flush = torch.empty(256 << 20, dtype=torch.uint8, device="cuda")


def gpu_time(snap, pods, fm, reps=30):
    """Kernel time of one Plan() with the snapshot resident.  Where the single-launch kernel rejects its own hypothesis the
    plan comes from the classic sequence run afterwards: both are timed (the attempt, then -- with the single-launch path
    switched off -- the classic sequence) and the SUM is reported."""
    def loop(reps):
        ms = []
        for k in range(reps + 3):
            with torch.cuda.stream(ext):
                flush.zero_()
            s = torch.cuda.Event(enable_timing=True); e = torch.cuda.Event(enable_timing=True)
            s.record(ext); eng.run(); e.record(ext)
            cnt, msk, st = eng.download()
            if k >= 3:
                ms.append(s.elapsed_time(e))
        return cnt, msk, st, np.asarray(ms)
    eng.upload(snap, pods, None, fm)
    eng.set_kernel_events(False)
    cnt, msk, st, ms = loop(reps)
    if st["fused_reject"]:
        os.environ["NOSPLAN_NO_FUSED"] = "1"
        eng.upload(snap, pods, None, fm)
        del os.environ["NOSPLAN_NO_FUSED"]
        _, _, _, ms2 = loop(reps if ms.mean() < 5 else 5)
        ms = ms[:len(ms2)] + ms2
    eng.set_kernel_events(True)
    hs, ho = capi.pin_snapshot(snap), capi.pinned_partitioning(snap)
    e2e = []
    for k in range(13):
        t0 = time.perf_counter(); eng.plan(hs, pods, None, fm, out=ho); t1 = time.perf_counter()
        if k >= 3:
            e2e.append(1e3 * (t1 - t0))
    return cnt, msk, st, float(np.median(ms)), float(np.percentile(ms, 95)), float(np.median(e2e))


print("`saturating`: no demanded profile has a free device at t0 and demand exceeds what re-partitioning can supply (every candidate "
      "node is visited).  `satisfiable`: the synthetic recipe with free devices left in place -- for the MIG configs nothing is lacking "
      "at t0 and Plan() returns the unchanged state (planner.go:80-83).  The `N pods` rows are the early-exit regime.")
print()
print("| config | regime | filter | nodes x GPUs, pods | candidates | path | GPU Plan() p50 / p95 (µs) | GPU candidates/s | GPU end-to-end p50 (µs) | CPU incremental port (ms) | CPU candidates/s | parity |")
print("|---|---|---|---|---|---|---|---|---|---|---|---|")
ONLY = [a[7:] for a in sys.argv if a.startswith("--only=")]
for name, kind, n, g, p, seed in CONFIGS:
    if ONLY and not any(o in name for o in ONLY):
        continue
    for regime in (("saturating", "satisfiable") if ("pods" not in name and name != "C1") else ("saturating",)):
        for fm, fname in ((pk.FILTER_NONE, "none"), (pk.FILTER_RESOURCES_FIT, "resources_fit")) if ("pods" not in name or "1000 pods" in name and "," not in name[12:]) else ((pk.FILTER_RESOURCES_FIT, "resources_fit"),):
            snap, pods = make(name, kind, n, g, p, seed, regime)
            ocnt, omsk, ost, osec = ol.plan(snap, pods, None, fm, ol.INCREMENTAL)
            cnt, msk, st, p50, p95, e2e = gpu_time(snap, pods, fm)
            same = np.array_equal(cnt, ocnt) and np.array_equal(msk, omsk) and all(
                st[k] == ost[k] for k in ("candidates", "visited_nodes", "placements", "geometry_evals", "lacking_at_end"))
            c = st["candidates"]
            path = "single launch" if st["fused"] else ("node-parallel" if st["path"] == 1 else (f"sequential, {st['waves']} waves" if st["waves"] else "sequential"))
            if st.get("gate_replayed", -1) >= 0:
                path += f", gate replayed ({st['gate_replayed']} nodes planned again)"
            if st["fused_reject"]:
                path += f" (single launch rejected: {st['fused_reject']})"
            n, g = snap.n_nodes, snap.gpus_per_node
            print(f"| {name} | {regime} | {fname} | {n} x {g}, {pods.n_pods} | {c} | {path} | "
                  f"{1e3 * p50:.1f} / {1e3 * p95:.1f} | {c / (p50 / 1e3):.3e} | {1e3 * e2e:.0f} | {1e3 * osec:.2f} | {c / osec if osec > 0 else 0:.3e} | {'identical' if same else 'MISMATCH'} |",
                  flush=True)

print()
if "--no-faithful" in sys.argv:
    sys.exit(0)
print("Faithful CPU port (clone every node per fork, re-sum every node per gate check: the reference's asymptotics), one core, C2 recipe, saturating, filter resources_fit:")
print()
print("| nodes | pods | seconds | candidates | candidates/s | seconds / (N^2 P) |")
print("|---|---|---|---|---|---|")
fit = []
for n in (100, 250, 500, 1000):
    snap, pods = synth.mig_cluster(n, 8, 5 * n, seed=0xC2, regime="saturating")
    _, _, st, sec = ol.plan(snap, pods, None, pk.FILTER_RESOURCES_FIT, ol.FAITHFUL)
    a = sec / (n * n * 5.0 * n)
    fit.append(a)
    print(f"| {n} | {5 * n} | {sec:.2f} | {st['candidates']} | {st['candidates'] / sec:.3e} | {a:.3e} |", flush=True)
a = float(np.median(fit[1:]))
print()
print(f"Fitted a = {a:.3e} s per N^2 P unit (median of the three larger points).  Extrapolated faithful-port time per Plan(): "
      + "; ".join(f"{name}: {a * n * n * p:.3g} s" for name, kind, n, g, p, seed in CONFIGS[1:]) + ".")
print("The Go reference is not runnable in this image (no toolchain); it is single-goroutine and adds map / allocation overhead on top of these figures.")
