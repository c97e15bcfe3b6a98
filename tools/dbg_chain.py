"""One Plan() of a snapshot that runs as the single-warp chain (a pod class asking for two profiles), for an ncu capture:
  ncu --set full --import-source on -k regex:k_plan_sweep -c 1 -o gpurun_out/chain python tools/dbg_chain.py [nodes] [pods]"""
import os, sys, time
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [ROOT, ROOT + "/tests", ROOT + "/oracle"]
import numpy as np
from nos_b200 import capi, packing as pk, synth

n = int(sys.argv[1]) if len(sys.argv) > 1 else 2000
npods = int(sys.argv[2]) if len(sys.argv) > 2 else 1000
snap, pods = synth.mig_cluster(n, 8, npods, seed=0xC5)
K = snap.n_profiles
ptr, col, qty = [0], [], []
for i in range(pods.n_pods):
    c = int(pods.scalar_col[i])
    col.append(c); qty.append(1)
    if i >= pods.n_pods - pods.n_pods // 10 and c + 1 < 3 + K:
        col.append(c + 1); qty.append(1)
    ptr.append(len(col))
pods.scalar_ptr = np.array(ptr, np.int32); pods.scalar_col = np.array(col, np.int32); pods.scalar_qty = np.array(qty, np.int64)
eng = capi.Engine(0)
for k in range(2):
    t0 = time.perf_counter()
    cnt, msk, st = eng.plan(snap, pods, None, pk.FILTER_RESOURCES_FIT)
    print("plan", k, f"{1e3 * (time.perf_counter() - t0):.2f} ms", "path", st["path"], "visited", st["visited_nodes"], "placements", st["placements"],
          "ms_kernels", st["ms_kernels"], "ms_sweep", st["ms_sweep"])
