"""Operator label-change timing at BASELINE config #4 (run on a GPU box): 500 ElasticQuotas, 100 000 running pods;
nosplan_quota_label_changes (host buffers in and out) against the CPU oracle on the same inputs, for a settled cluster with
1 % of the labels off and for a fresh cluster (every pod of a quota needs its first label)."""
import json, os, sys, time
ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path[:0] = [ROOT, ROOT + "/tests", ROOT + "/oracle"]
import numpy as np
from nos_b200 import capi
from support import oracle_lib as ol
from test_label_changes_gpu import c4_state

eng = capi.Engine(0)
pq, run, settled, in_a_quota = c4_state()
rng = np.random.default_rng(7)
tampered = settled.copy()
hit = rng.random(run.n_pods) < 0.01
tampered[hit] = rng.integers(0, 3, int(hit.sum()))
res = {"config": "C4: 500 quotas, %d running pods, %d resource columns" % (run.n_pods, pq.n_res)}
for name, cur in (("settled_1pct_off", tampered), ("fresh", np.zeros(run.n_pods, np.uint8))):
    for _ in range(3):
        pod, desired, n, _, st = eng.quota_label_changes(pq, run, cur)
    ts, ks = [], []
    for _ in range(20):
        t0 = time.perf_counter(); pod, desired, n, _, st = eng.quota_label_changes(pq, run, cur); ts.append(time.perf_counter() - t0); ks.append(st["ms_kernels"])
    cpu = []
    for _ in range(3):
        t0 = time.perf_counter(); epod, edes = ol.label_changes(pq, run, cur); cpu.append(time.perf_counter() - t0)
    res[name] = {"n_changes": n, "ms_call_p50": 1e3 * float(np.median(ts)), "ms_kernels_p50": float(np.median(ks)), "ms_h2d": st["ms_h2d"],
                 "kernel_launches": st["kernel_launches"], "h2d_bytes": st["h2d_bytes"], "d2h_bytes": st["d2h_bytes"],
                 "cpu_oracle_ms_p50": 1e3 * float(np.median(cpu)), "cpu_cores": 1,
                 "identical_to_oracle": bool(np.array_equal(pod, epod) and np.array_equal(desired, edes))}
print(json.dumps(res))
