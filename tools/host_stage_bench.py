"""Host time of the two-pass pod staging (nosplan_host_stage_pods), no GPU: microseconds per call at P pods."""
import os, sys, time
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [ROOT]
import numpy as np
from nos_b200 import capi, synth
for P in (1000, 10000, 100000):
    s, p = synth.mig_cluster(100, 8, P, seed=0xC5)
    capi.host_stage_pods(s, p, None, repeat=20)
    best = 1e9
    for _ in range(5):
        t0 = time.perf_counter(); capi.host_stage_pods(s, p, None, repeat=200 if P <= 10000 else 20); dt = time.perf_counter() - t0
        best = min(best, dt / ((200 if P <= 10000 else 20) + 1))
    print(f"{P:7d} pods: {1e6 * best:8.1f} us per staging ({1e9 * best / P:5.2f} ns per pod)")
