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
]
eng = capi.Engine(0)
ext = torch.cuda.ExternalStream(eng.stream())
flush = torch.empty(256 << 20, dtype=torch.uint8, device="cuda")


def gpu_time(snap, pods, fm, reps=30):
    eng.upload(snap, pods, None, fm)
    eng.set_kernel_events(False)
    ms = []
    for k in range(reps + 3):
        with torch.cuda.stream(ext):
            flush.zero_()
        s = torch.cuda.Event(enable_timing=True); e = torch.cuda.Event(enable_timing=True)
        s.record(ext); eng.run(); e.record(ext)
        cnt, msk, st = eng.download()
        if k >= 3:
            ms.append(s.elapsed_time(e))
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
for name, kind, n, g, p, seed in CONFIGS:
    gen = synth.mig_cluster if kind == "mig" else synth.mps_cluster
    for regime in (("saturating", "satisfiable") if "pods" not in name else ("saturating",)):
        for fm, fname in ((pk.FILTER_NONE, "none"), (pk.FILTER_RESOURCES_FIT, "resources_fit")):
            snap, pods = gen(n, g, p, seed=seed, regime=regime)
            ocnt, omsk, ost, osec = ol.plan(snap, pods, None, fm, ol.INCREMENTAL)
            cnt, msk, st, p50, p95, e2e = gpu_time(snap, pods, fm)
            same = np.array_equal(cnt, ocnt) and np.array_equal(msk, omsk) and all(
                st[k] == ost[k] for k in ("candidates", "visited_nodes", "placements", "geometry_evals", "lacking_at_end"))
            c = st["candidates"]
            print(f"| {name} | {regime} | {fname} | {n} x {g}, {p} | {c} | {'node-parallel' if st['path'] == 1 else 'sequential'} | "
                  f"{1e3 * p50:.1f} / {1e3 * p95:.1f} | {c / (p50 / 1e3):.3e} | {1e3 * e2e:.0f} | {1e3 * osec:.2f} | {c / osec if osec > 0 else 0:.3e} | {'identical' if same else 'MISMATCH'} |",
                  flush=True)

print()
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
