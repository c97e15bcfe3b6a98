"""Where the end-to-end time of nosplan_plan goes on the host side (run on a GPU box)."""
import sys, time
import os
ROOT=os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path[:0]=[ROOT,ROOT+'/tests',ROOT+'/oracle']
import numpy as np
from nos_b200 import capi, synth
snap,pods=synth.mig_cluster(10000,8,10000,seed=0xC5)
eng=capi.Engine(0)
for _ in range(3): eng.plan(snap,pods,None,2)
def t(f,n=20):
    ts=[]
    for _ in range(n):
        t0=time.perf_counter(); f(); ts.append(time.perf_counter()-t0)
    return 1e3*np.median(ts)
print("plan      ms", t(lambda: eng.plan(snap,pods,None,2)))
print("upload    ms", t(lambda: eng.upload(snap,pods,None,2)))
print("run+dl    ms", t(lambda: (eng.run(), eng.download())))
cs=capi.c_snapshot(snap); cp=capi.c_pods(pods)
print("py structs ms", t(lambda: (capi.c_snapshot(snap), capi.c_pods(pods))))
import ctypes as C
def raw():
    eng.lib.nosplan_upload(eng.h, C.byref(cs[0]), C.byref(cp[0]), None, 2)
print("raw upload ms", t(raw))
os.environ["NOSPLAN_TRACE"] = "1"
sys.stderr.flush()
for pinned in (False, True):
    if pinned:
        snap = capi.pin_snapshot(snap)
        for _ in range(3): eng.plan(snap, pods, None, 2)
        print("plan (node arrays page-locked) ms", t(lambda: eng.plan(snap, pods, None, 2)))
        hout = capi.pinned_partitioning(snap)
        print("plan (+ plan buffers page-locked) ms", t(lambda: eng.plan(snap, pods, None, 2, out=hout)))
    print("---- trace of one nosplan_plan, pinned =", pinned, flush=True)
    os.environ["NOSPLAN_TRACE"] = "1"
    t0 = time.perf_counter(); eng.plan(snap, pods, None, 2); t1 = time.perf_counter()
    del os.environ["NOSPLAN_TRACE"]
    sys.stderr.flush()
    print("that call: %.1f us" % (1e6 * (t1 - t0)), flush=True)
