"""Random small cases (tests/support/randcases.py) against the oracle, all seeds, listing every mismatch with the path taken."""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [ROOT, ROOT + "/tests", ROOT + "/oracle"]
import numpy as np
import nos_model as m
from nos_b200 import capi, packing as pk
from support import oracle_lib as ol
from support import randcases as rc

lo, hi = int(sys.argv[1]), int(sys.argv[2])
eng = capi.Engine(0)
bad = 0
paths = {}
for seed in range(lo, hi):
    kind_name, nodes, pods, pre, fmode = rc.rand_case(seed)
    kind = pk.KIND_BY_NAME[kind_name]
    if any(q <= 0 and pk.slice_profile_of(kind, r) is not None for p in pods for r, q in pk.compute_pod_request(p).items()):
        continue
    s, p = pk.pack(kind, nodes, pods, m.DEFAULT_KNOWN_GEOMETRIES)
    cnt, msk, st = eng.plan(s, p, pre, fmode)
    ocnt, omsk, ost, _ = ol.plan(s, p, pre, fmode, ol.INCREMENTAL)
    key = (st["path"], st["waves"] > 0, st["gate_replayed"] >= 0, st["fused"])
    paths[key] = paths.get(key, 0) + 1
    same = np.array_equal(cnt, ocnt) and np.array_equal(msk, omsk) and all(st[k] == ost[k] for k in ("candidates", "visited_nodes", "placements", "geometry_evals", "candidate_nodes", "lacking_at_start", "lacking_at_end"))
    if not same:
        bad += 1
        rows = np.nonzero((cnt != ocnt).any(axis=-1).reshape(-1) | (msk != omsk).reshape(-1))[0]
        print("MISMATCH seed", seed, kind_name, "N", s.n_nodes, "G", s.gpus_per_node, "P", p.n_pods, "fm", fmode, "path", st["path"], "waves", st["waves"],
              "gate_replayed", st["gate_replayed"], "fused", st["fused"], st["fused_reject"], "placements", st["placements"], ost["placements"], "rows", rows[:8])
print("mismatches", bad, "paths (path, waves, gate replay, fused):", paths)
