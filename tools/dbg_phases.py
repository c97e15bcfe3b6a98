"""Per-phase cycle counts of k_plan_fused (a library built with -DNOSPLAN_FUSED_PROFILE, NOSPLAN_LIB pointing at it)."""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [ROOT]
import torch
from nos_b200 import capi, packing as pk, synth
n = int(sys.argv[1]) if len(sys.argv) > 1 else 10000
s, p = synth.mig_cluster(n, 8, 10000, seed=0xC5)
eng = capi.Engine(0)
flush = torch.empty(256 << 20, dtype=torch.uint8, device="cuda")
eng.upload(s, p, None, pk.FILTER_RESOURCES_FIT)
for k in range(3):
    flush.zero_(); torch.cuda.synchronize()
    print("run", k, flush=True)
    eng.run(); cnt, msk, st = eng.download()
print(st["fused"], st["ms_kernels"])
