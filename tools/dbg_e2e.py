import os, sys, time
sys.path[:0] = ['/root/repo', '/root/repo/tests', '/root/repo/oracle']
import numpy as np
from nos_b200 import capi, synth
s, p = synth.mig_cluster(10000, 8, 10000, seed=0xC5)
eng = capi.Engine(0)
hs = capi.pin_snapshot(s); out = capi.pinned_partitioning(s)
fields = [f for f in capi._PINNED_FIELDS]
for k in range(5): eng.plan(hs, p, None, 2, out=out)
def run(refresh, n=30):
    ts, h2d = [], []
    for k in range(n):
        if refresh:
            for f in fields:
                a = getattr(hs, f); np.copyto(a, getattr(s, f).reshape(a.shape))
        t0 = time.perf_counter(); st = eng.plan(hs, p, None, 2, out=out)[2]; ts.append(time.perf_counter() - t0); h2d.append(st["ms_h2d"])
    return 1e3 * np.median(ts), 1e3 * np.median(h2d)
print("cold buffers   : e2e %.0f us, h2d %.0f us" % run(False))
print("refreshed      : e2e %.0f us, h2d %.0f us" % run(True))
print("cold again     : e2e %.0f us, h2d %.0f us" % run(False))
ts = []
for k in range(30):
    t0 = time.perf_counter(); eng.plan_resident(p, None, 2, out=out); ts.append(time.perf_counter() - t0)
print("resident       : e2e %.0f us" % (1e3 * np.median(ts)))
