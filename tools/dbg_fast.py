import os, sys
sys.path[:0] = ['/root/repo', '/root/repo/tests', '/root/repo/oracle']
import numpy as np
from nos_b200 import capi, packing as pk, synth
rng = np.random.default_rng(17)
eng = capi.Engine(0)
for seed, n_pods in ((1, 3000), (2, 64), (3, 9000)):
    s, p = synth.mig_cluster(250, 8, n_pods, seed=100 + seed)
    p.priority[:] = rng.choice([-5, 0, 7, 100, 1000, 2_000_000_000], n_pods)
    pre = (rng.random(n_pods) < 0.8).astype(np.uint8)
    for env in (None, "1"):
        if env: os.environ["NOSPLAN_NO_FAST_PODS"] = env
        else: os.environ.pop("NOSPLAN_NO_FAST_PODS", None)
        cnt, msk, st = eng.plan(s, p, pre, pk.FILTER_RESOURCES_FIT)
        print(seed, n_pods, "nofast" if env else "fast", {k: st[k] for k in ("fused", "fused_reject", "path", "lacking_at_start", "lacking_at_end", "placements", "h2d_bytes")})
import time
s, p = synth.mig_cluster(10000, 8, 10000, seed=0xC5)
os.environ.pop("NOSPLAN_NO_FAST_PODS", None)
hs = capi.pin_snapshot(s); out = capi.pinned_partitioning(s)
for k in range(3): eng.plan(hs, p, None, 2, out=out)
os.environ["NOSPLAN_TRACE"] = "1"
eng.plan_resident(p, None, 2, out=out)
