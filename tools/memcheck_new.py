"""Small instances of the paths added in round 2, for compute-sanitizer:
  compute-sanitizer --tool memcheck --error-exitcode 9 python tools/memcheck_new.py"""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [ROOT]
import numpy as np
from nos_b200 import capi, packing as pk, synth
eng = capi.Engine(0)
def cols(p): return sorted({int(c) for c in p.scalar_col})
# gate replay behind the node-parallel pass (mode 3), a node planned again
s, p = synth.mig_cluster(300, 8, 1500, seed=21); s.node_requested[7, cols(p)[0]] += 5
print("replay", eng.plan(s, p, None, 2)[2]["gate_replayed"], flush=True)
# waves with the gate
s, p = synth.mig_cluster(600, 8, 30, seed=24); s.node_requested[5, cols(p)[0]] += 3
st = eng.plan(s, p, None, 2)[2]; print("waves", st["waves"], st["gate_replayed"], flush=True)
# a two-profile class
s, p = synth.mig_cluster(200, 8, 800, seed=3)
ptr, col, qty = [0], [], []
for i in range(p.n_pods):
    c = int(p.scalar_col[i]); col.append(c); qty.append(1)
    if i % 7 == 6 and c + 1 < 3 + s.n_profiles: col.append(c + 1); qty.append(1)
    ptr.append(len(col))
p.scalar_ptr = np.array(ptr, np.int32); p.scalar_col = np.array(col, np.int32); p.scalar_qty = np.array(qty, np.int64)
print("two-profile", eng.plan(s, p, None, 2)[2]["gate_replayed"], flush=True)
# MPS in the single-launch kernel, MIG in it through the compact records
s, p = synth.mps_cluster(200, 8, 1500, seed=5); print("mps fused", eng.plan(s, p, None, 2)[2]["fused"], flush=True)
s, p = synth.mig_cluster(200, 8, 1500, seed=5); print("mig compact fused", eng.plan_compact(s, p, None, 2)[2]["fused"], flush=True)
# quota labelling by buckets (one quota cut into several), and the global-sort fallback
pq, run, pend = synth.quota_state(20, 6000, 100, seed=7)
r, st = eng.quota_recompute(pq, run, pend); print("quota", st["kernel_launches"], int(r["over_quota"].sum()), flush=True)
os.environ["NOSPLAN_QUOTA_GLOBAL_SORT"] = "1"
r2, st2 = eng.quota_recompute(pq, run, pend); print("quota global", st2["kernel_launches"], int(r2["over_quota"].sum()), flush=True)
assert np.array_equal(r["over_quota"], r2["over_quota"]) and np.array_equal(r["used"], r2["used"])
eng.close()
print("done")
