"""Snapshot packer timing (run on a GPU box): status annotations of a 10 000-node x 8-GPU cluster -> per-GPU rows,
on the GPU (nosplan_parse_annotations, host buffers in and out) and with the host packers beside it."""
import json, os, subprocess, sys, time
ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path[:0] = [ROOT, ROOT + "/tests", ROOT + "/oracle"]
import numpy as np
import nos_model as m
from nos_b200 import capi, packing as pk, synth
from support import hostcli as hc

N = int(sys.argv[1]) if len(sys.argv) > 1 else 10000
snap, pods = synth.mig_cluster(N, 8, 10, seed=0xC5)
v1 = hc.logical_to_v1("mig", synth.to_logical(snap, pods)[0], strict=False)
t0 = time.perf_counter(); flat = capi.flatten_annotations(v1, snap.profiles); t_flat = time.perf_counter() - t0
eng = capi.Engine(0)
for _ in range(3):
    out, st = eng.parse_annotations(flat, snap.gpus_per_node)
ts = []
for _ in range(20):
    t0 = time.perf_counter(); out, st = eng.parse_annotations(flat, snap.gpus_per_node); ts.append(time.perf_counter() - t0)
snap, _ = pk.pack(pk.KIND_MIG, pk.nodes_from_v1(pk.KIND_MIG, v1, m.DEFAULT_KNOWN_GEOMETRIES), [], m.DEFAULT_KNOWN_GEOMETRIES)   # what NewNode makes of it
n = snap.n_nodes * snap.gpus_per_node
ok = all(np.array_equal(out[k][:n], getattr(snap, k)[:n]) for k in ("gpu_used", "gpu_free", "gpu_used_mask", "gpu_free_mask", "gpu_index"))
t0 = time.perf_counter(); pk.nodes_from_v1(pk.KIND_MIG, v1, m.DEFAULT_KNOWN_GEOMETRIES); t_py = time.perf_counter() - t0
text = hc.to_text("mig", v1, [], "pack")
t0 = time.perf_counter(); subprocess.run([hc.CLI], input=text, capture_output=True, text=True); t_cli = time.perf_counter() - t0
A = int(flat["node_ptr"][-1])
print(json.dumps({"nodes": N, "annotations": A, "annotation_bytes": int(flat["key_off"][-1] + flat["val_off"][-1]),
                  "gpu": {"ms_call_p50": 1e3 * float(np.median(ts)), "ms_h2d": st["ms_h2d"], "ms_kernels": st["ms_kernels"], "ms_d2h": st["ms_d2h"],
                          "h2d_bytes": st["h2d_bytes"], "d2h_bytes": st["d2h_bytes"], "annotations_per_s": A / float(np.median(ts))},
                  "matches_python_packer": bool(ok),
                  "host": {"python_nodes_from_v1_ms": 1e3 * t_py, "cpp_host_mirror_cli_ms_incl_text_io": 1e3 * t_cli,
                           "python_flatten_for_gpu_ms": 1e3 * t_flat}}))
