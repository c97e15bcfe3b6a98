"""Quota engine timing at BASELINE config #4 (run on a GPU box): 500 ElasticQuotas, 100 000 running pods, 10 000 pending;
nosplan_quota_recompute (host buffers in and out) and nosplan_quota_victims against the CPU oracle on the same inputs."""
import json, os, sys, time
ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path[:0] = [ROOT, ROOT + "/tests", ROOT + "/oracle"]
import numpy as np
from nos_b200 import capi, synth
from support import oracle_lib as ol

pq, run, pend = synth.quota_state(500, 100_000, 10_000, seed=0xC4)
eng = capi.Engine(0)
for _ in range(3):
    got, st = eng.quota_recompute(pq, run, pend)
ts = []
for _ in range(20):
    t0 = time.perf_counter(); got, st = eng.quota_recompute(pq, run, pend); ts.append(time.perf_counter() - t0)
# event-driven: the state stays in HBM, batches of 100 pod events (delete, then add again) with Used / aggregates / guaranteed back
ids = np.arange(run.n_pods, dtype=np.int32)
eng.quota_state_begin(pq, run.n_pods)
eng.quota_state_apply(np.zeros(run.n_pods, np.uint8), ids, run.ns, run.request, run.request_mask, outputs=False)
rng = np.random.default_rng(1)
tev, kev = [], []
for it in range(40):
    b = rng.choice(run.n_pods, 100, replace=False).astype(np.int32)
    op = np.full(100, it & 1 ^ 1, np.uint8)                     # odd iterations add back what the even ones deleted ... roughly
    t0 = time.perf_counter(); sgot, applied, sst = eng.quota_state_apply(op, b, run.ns[b], run.request[b], run.request_mask[b]); tev.append(time.perf_counter() - t0)
    kev.append(sst["ms_kernels"])
eng.quota_state_apply(np.zeros(run.n_pods, np.uint8), ids, run.ns, run.request, run.request_mask, outputs=False)   # everybody accounted again
sfin, _, _ = eng.quota_state_apply([], [], [], np.zeros((0, pq.n_res), np.int64))
t0 = time.perf_counter(); exp = ol.quota(pq, run, pend); t_cpu = time.perf_counter() - t0
same = all(np.array_equal(np.asarray(got[k]), np.asarray(exp[k])) for k in exp if k in got and k != "seconds")
R = pq.min.shape[1]
b_alg = run.n_pods * (4 + 8 * R + 4) + pend.n_pods * (4 + 8 * R + 4) + pq.min.size * 8 * 3 + run.n_pods + pend.n_pods + pq.min.size * 8 * 3
print(json.dumps({"config": "C4: 500 quotas, 100000 running + 10000 pending pods, %d resource columns" % R,
                  "gpu": {"ms_call_p50": 1e3 * float(np.median(ts)), "ms_h2d": st["ms_h2d"], "ms_kernels": st["ms_kernels"], "ms_d2h": st["ms_d2h"],
                          "kernel_launches": st["kernel_launches"], "h2d_bytes": st["h2d_bytes"], "d2h_bytes": st["d2h_bytes"],
                          "pods_per_s_kernels": (run.n_pods + pend.n_pods) / (st["ms_kernels"] / 1e3),
                          "algorithmic_bytes": int(b_alg), "achieved_GBps_kernels": b_alg / (st["ms_kernels"] / 1e3) / 1e9},
                  "resident_state": {"events_per_batch": 100, "ms_call_p50": 1e3 * float(np.median(tev)), "ms_kernels_p50": float(np.median(kev)),
                                     "h2d_bytes": sst["h2d_bytes"], "d2h_bytes": sst["d2h_bytes"],
                                     "identical_to_batch_engine": bool(np.array_equal(sfin["used"], got["used"]) and np.array_equal(sfin["guaranteed"], got["guaranteed"]))},
                  "cpu_oracle_ms": 1e3 * t_cpu, "identical_to_oracle": bool(same)}))
