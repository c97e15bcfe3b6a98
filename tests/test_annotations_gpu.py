"""Snapshot packer on the GPU (nosplan_parse_annotations): node status annotations -> per-GPU rows, against the Python
packer (nos_b200.packing.nodes_from_v1 + pack), which follows pkg/gpu/annotation.go:29-104 and mig/slicing extractGPUs."""
import json
import os

import numpy as np
import pytest

import nos_model as m
from nos_b200 import capi, packing as pk, synth
from support import hostcli as hc
from support import randcases as rc

pytestmark = pytest.mark.gpu
HERE = os.path.dirname(os.path.abspath(__file__))
PLAN = json.load(open(os.path.join(HERE, "golden", "planner_plan.json")))["cases"]


@pytest.fixture(scope="module")
def eng():
    import __graft_entry__ as g
    g.build()
    e = capi.Engine(0)
    yield e
    e.close()


def check(eng, kind_name, v1_nodes):
    kind = pk.KIND_BY_NAME[kind_name]
    mine = [n for n in v1_nodes if n.get("labels", {}).get(pk.LABEL_PARTITIONING) == kind_name]
    mine.sort(key=lambda n: n["name"].encode())
    snap, _ = pk.pack(kind, pk.nodes_from_v1(kind, mine, m.DEFAULT_KNOWN_GEOMETRIES), [], m.DEFAULT_KNOWN_GEOMETRIES)
    flat = capi.flatten_annotations(mine, snap.profiles)
    out, st = eng.parse_annotations(flat, snap.gpus_per_node)
    n = snap.n_nodes * snap.gpus_per_node
    assert np.array_equal(out["node_gpu_count"], snap.node_gpu_count)
    for k in ("gpu_used", "gpu_free", "gpu_used_mask", "gpu_free_mask", "gpu_index"):
        assert np.array_equal(out[k][:n], getattr(snap, k)[:n]), k
    return st


@pytest.mark.parametrize("c", PLAN, ids=[f"{c['kind']}:{c['name'][:50]}" for c in PLAN])
def test_golden_planner_nodes(eng, c):
    check(eng, c["kind"], c["nodes"])


def test_random_nodes(eng):
    n = 0
    for seed in range(400):
        kind_name, nodes, pods, pre, fmode = rc.rand_case(seed, max_nodes=12)
        v1n = hc.logical_to_v1(kind_name, nodes)
        if v1n is None:
            continue
        check(eng, kind_name, v1n)
        n += 1
    assert n > 100, n


def test_acceptance_rules_of_the_reference(eng):
    # pkg/gpu/annotation.go:29-104: part counts, Atoi of value and index, case-insensitive status, spec keys, foreign keys
    node = {"name": "n1", "labels": {pk.LABEL_PARTITIONING: "mig", pk.LABEL_PRODUCT: "NVIDIA-A30", pk.LABEL_COUNT: "4"},
            "annotations": {"nos.nebuly.com/status-gpu-0-1g.6gb-free": "4",
                            "nos.nebuly.com/status-gpu-1-2g.12gb-USED": "1",
                            "nos.nebuly.com/status-gpu-1-1g.6gb-Free": "0",
                            "nos.nebuly.com/status-gpu-+3-1g.6gb-used": "+2",
                            "nos.nebuly.com/status-gpu-x-1g.6gb-free": "1",
                            "nos.nebuly.com/status-gpu-1-1g.6gb-bogus": "1",
                            "nos.nebuly.com/status-gpu-1-1g.6gb-free-extra": "1",
                            "nos.nebuly.com/status-gpu-2-4g.24gb-unknown": "1",
                            "nos.nebuly.com/status-gpu-0-2g.12gb-free": "one",
                            "nos.nebuly.com/status-gpu-0-2g.12gb-used": "1_0",
                            "nos.nebuly.com/status-gpu-0-2g.12gb-used ": " 1",
                            "nos.nebuly.com/status-gpu-0-4g.24gb-free": "",
                            "nos.nebuly.com/spec-gpu-0-4g.24gb": "1",
                            "nos.nebuly.com/status-gpu": "1",
                            "unrelated": "7"}}
    check(eng, "mig", [node])


def test_large_synthetic_cluster_and_limits(eng):
    snap, pods = synth.mig_cluster(3000, 8, 10, seed=11)
    v1 = hc.logical_to_v1("mig", synth.to_logical(snap, pods)[0], strict=False)
    st = check(eng, "mig", v1)
    assert st["kernel_launches"] == 2 and st["h2d_bytes"] > 0
    # a profile outside the dictionary, a count above 255, more GPUs than the row stride: refused with a clear error
    base = {"name": "n", "labels": {pk.LABEL_PARTITIONING: "mig", pk.LABEL_PRODUCT: "NVIDIA-A30", pk.LABEL_COUNT: "1"}}
    for ann, g, word in (({"nos.nebuly.com/status-gpu-0-9g.99gb-free": "1"}, 1, "dictionary"),
                         ({"nos.nebuly.com/status-gpu-0-1g.6gb-free": "256"}, 1, "[0, 255]"),
                         ({"nos.nebuly.com/status-gpu-0-1g.6gb-free": "1", "nos.nebuly.com/status-gpu-5-1g.6gb-free": "1"}, 1, "more GPUs")):
        flat = capi.flatten_annotations([dict(base, annotations=ann)], ["1g.6gb", "2g.12gb", "4g.24gb"])
        with pytest.raises(capi.NosplanError, match="|".join(map(__import__("re").escape, [word]))):
            eng.parse_annotations(flat, g)
