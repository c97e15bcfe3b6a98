"""Host phase times (NOSPLAN_TRACE=1) of nosplan_plan_compact and nosplan_plan_resident at 10 000 x 8 x 10 000."""
import os, sys, time
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [ROOT]
import numpy as np
from nos_b200 import capi, synth
s, p = synth.mig_cluster(10000, 8, 10000, seed=0xC5)
eng = capi.Engine(0)
out = capi.pinned_partitioning(s)
rec = capi.compact_records(s)
import torch
t = torch.from_numpy(rec.reshape(-1).copy()).pin_memory(); prec = t.numpy().reshape(rec.shape)
for k in range(5): eng.plan_compact(s, p, None, 2, out=out, records=prec)
ts = []
for k in range(20):
    np.copyto(prec, rec)
    t0 = time.perf_counter(); eng.plan_compact(s, p, None, 2, out=out, records=prec); ts.append(time.perf_counter() - t0)
print("compact e2e p50 %.0f us (python call included)" % (1e6 * np.median(ts)), flush=True)
os.environ["NOSPLAN_TRACE"] = "1"
np.copyto(prec, rec)
t0 = time.perf_counter(); st = eng.plan_compact(s, p, None, 2, out=out, records=prec)[2]; t1 = time.perf_counter()
print("traced call %.0f us: h2d %.1f kernels %.1f d2h %.1f" % (1e6 * (t1 - t0), 1e3 * st["ms_h2d"], 1e3 * st["ms_kernels"], 1e3 * st["ms_d2h"]), flush=True)
t0 = time.perf_counter(); st = eng.plan_resident(p, None, 2, out=out)[2]; t1 = time.perf_counter()
print("traced resident %.0f us: h2d %.1f kernels %.1f d2h %.1f" % (1e6 * (t1 - t0), 1e3 * st["ms_h2d"], 1e3 * st["ms_kernels"], 1e3 * st["ms_d2h"]), flush=True)
