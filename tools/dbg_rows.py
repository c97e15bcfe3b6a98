"""Per-kernel CUDA-event times of the rows of the results table that need the gate replay (classic sequence only)."""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [ROOT, ROOT + "/tests", ROOT + "/oracle"]
import numpy as np
from nos_b200 import capi, packing as pk, synth

def two_profile(s, p):
    K = s.n_profiles
    ptr, col, qty = [0], [], []
    for i in range(p.n_pods):
        c = int(p.scalar_col[i]); col.append(c); qty.append(1)
        if i >= p.n_pods - p.n_pods // 10 and c + 1 < 3 + K:
            col.append(c + 1); qty.append(1)
        ptr.append(len(col))
    p.scalar_ptr = np.array(ptr, np.int32); p.scalar_col = np.array(col, np.int32); p.scalar_qty = np.array(qty, np.int64)

eng = capi.Engine(0)
os.environ["NOSPLAN_NO_FUSED"] = "1"
for name, n, npods in (("two-profile", 10000, 1000), ("inconsistent", 10000, 1000), ("inconsistent", 10000, 40), ("inconsistent", 100000, 1000), ("plain", 10000, 1000)):
    s, p = synth.mig_cluster(n, 8, npods, seed=0xC5)
    if name == "two-profile": two_profile(s, p)
    if name == "inconsistent": s.node_requested[7, 3] += 5
    eng.upload(s, p, None, pk.FILTER_RESOURCES_FIT)
    rows = []
    for k in range(8):
        eng.run(); cnt, msk, st = eng.download()
        rows.append((st["ms_kernels"], st["ms_prepass"], st["ms_tracker"], st["ms_parallel"], st["ms_sweep"]))
    r = np.median(np.array(rows[2:]), axis=0)
    print(f"{name:13s} {n:6d} x {npods:5d}: kernels {1e3*r[0]:7.1f} us = prepass {1e3*r[1]:6.1f} + tracker {1e3*r[2]:6.1f} + parallel {1e3*r[3]:7.1f} + waves/replay {1e3*r[4]:7.1f}"
          f" | path {st['path']} waves {st['waves']} gate_replayed {st['gate_replayed']}")
