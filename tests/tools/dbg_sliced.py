import sys, os
ROOT=os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path[:0]=[ROOT,ROOT+'/tests',ROOT+'/oracle']
import numpy as np, torch
from nos_b200 import capi, packing as pk, sharded, synth
sys.argv=['x']
import test_sharded_gpu as t
snap, pods = synth.mig_cluster(150, 8, 700, seed=5, regime="satisfiable")
e = capi.Engine(0)
cnt, msk, st = e.plan(snap, pods, None, 2)
print("whole :", {k: st[k] for k in ("path","lacking_at_start","lacking_at_end","visited_nodes","placements","candidate_nodes")})
e.close()
for world in (1, 2):
    out, slices = t.run_sliced_emulated(snap, pods, 2, world)
    for o in out:
        print("sliced", world, None if o is None else {k: o[2][k] for k in ("path","lacking_at_start","lacking_at_end","visited_nodes","placements","candidate_nodes")})
