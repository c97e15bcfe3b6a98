"""Timing of the sequential path (run on a GPU box): problems where the node-parallel pass does not apply -- nothing lacking
(the reference's early return) and unsaturated demand -- at 10 000 and 100 000 nodes, snapshot resident."""
import json, os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path[:0] = [ROOT, ROOT + "/tests", ROOT + "/oracle"]
import numpy as np
from nos_b200 import capi, packing as pk, synth

eng = capi.Engine(0)
out = {}
for n_nodes, n_pods, regime in ((10_000, 10_000, "satisfiable"), (100_000, 10_000, "satisfiable"), (10_000, 3, "saturating"), (100_000, 3, "saturating")):
    snap, pods = synth.mig_cluster(n_nodes, 8, n_pods, seed=0xC5, regime=regime)
    eng.upload(snap, pods, None, pk.FILTER_RESOURCES_FIT)
    ms = []
    for _ in range(12):
        eng.run()
        _, _, st = eng.download()
        ms.append(st["ms_kernels"])
    out[f"{n_nodes} nodes, {n_pods} pods, {regime}"] = {"ms_kernels_p50": float(np.median(ms[2:])), "path": st["path"], "waves": st["waves"],
                                                        "visited_nodes": st["visited_nodes"], "lacking_at_start": st["lacking_at_start"],
                                                        "ms_sweep": st["ms_sweep"], "ms_prepass": st["ms_prepass"]}
print(json.dumps(out))
