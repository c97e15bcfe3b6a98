"""The reference's own table tests of node construction and annotation parsing (tests/golden/new_node.json, hand-ported
from pkg/gpu/mig/node_test.go, pkg/gpu/slicing/node_test.go, pkg/gpu/annotation_test.go) against the three packers: the Python
packer (the oracle of the GPU parser), the C++ host mirror, and -- on a GPU -- nosplan_parse_annotations."""
import json
import os

import numpy as np
import pytest

import nos_model as m
from nos_b200 import capi, packing as pk
from support import hostcli as hc

HERE = os.path.dirname(os.path.abspath(__file__))
G = json.load(open(os.path.join(HERE, "golden", "new_node.json")))
NEW_NODE, PARSE, NODE_ANN = G["new_node"], G["parse_status_annotation"], G["node_annotations"]
ids = lambda cs: [f"{c.get('kind', '')}:{c['name'][:48]}" for c in cs]


@pytest.fixture(scope="module", autouse=True)
def built():
    import __graft_entry__ as g
    g.build()


def _gpus_of_logical(n):
    return [{"index": g["index"], "used": g["used"], "free": g["free"], "memory_gb": g.get("memory_gb", 0)} for g in n["gpus"]]


def _same(got, exp):
    assert len(got) == len(exp)
    key = lambda g: g["index"]
    for a, b in zip(sorted(got, key=key), sorted(exp, key=key)):     # assert.ElementsMatch: order-insensitive
        assert a["index"] == b["index"] and a["used"] == b["used"] and a["free"] == b["free"]
        if "memory_gb" in b:
            assert a["memory_gb"] == b["memory_gb"]


@pytest.mark.parametrize("c", NEW_NODE, ids=ids(NEW_NODE))
def test_new_node_python_packer(c):
    kind = pk.KIND_BY_NAME[c["kind"]]
    if c["expect_error"]:
        with pytest.raises(pk.PackError):
            pk.nodes_from_v1(kind, [c["node"]], m.DEFAULT_KNOWN_GEOMETRIES)
        return
    nodes = pk.nodes_from_v1(kind, [c["node"]], m.DEFAULT_KNOWN_GEOMETRIES)
    assert nodes[0]["name"] == c["node"]["name"]
    _same(_gpus_of_logical(nodes[0]), c["expect_gpus"])


def _gpus_of_pack(d):
    profiles = d.get("profiles", [])
    n = int(d["node_gpu_count"][0]) if d.get("node_gpu_count") else 0
    used = np.array([int(x) for x in d["gpu_used"]], np.int64).reshape(-1, 16)
    free = np.array([int(x) for x in d["gpu_free"]], np.int64).reshape(-1, 16)
    out = []
    for j in range(n):
        um, fm = int(d["gpu_used_mask"][j]), int(d["gpu_free_mask"][j])
        out.append({"index": int(d["gpu_index"][j]), "memory_gb": int(d["gpu_memory_gb"][j]),
                    "used": {p: int(used[j, k]) for k, p in enumerate(profiles) if um >> k & 1},
                    "free": {p: int(free[j, k]) for k, p in enumerate(profiles) if fm >> k & 1}})
    return out


@pytest.mark.parametrize("c", NEW_NODE, ids=ids(NEW_NODE))
def test_new_node_cpp_host_mirror(c):
    code, out = hc.run(hc.to_text(c["kind"], [c["node"]], [], "pack"))
    if c["expect_error"]:
        assert code == 1 and out.startswith("error "), out
        return
    assert code == 0, out
    d = hc.parse_pack(out)
    assert d["node_names"] == [c["node"]["name"]]
    _same(_gpus_of_pack(d), c["expect_gpus"])


@pytest.mark.parametrize("c", PARSE, ids=ids(PARSE))
def test_parse_status_annotation(c):
    by_gpu = pk._status_annotations({c["key"]: c["value"]})
    if not c["ok"]:
        assert by_gpu == {}
    else:
        e = c["expect"]
        assert by_gpu == {e["index"]: [(e["profile"], e["status"], e["quantity"])]}
    if c["key"]:                                   # the text driver cannot carry an empty key
        node = {"name": "n", "labels": {pk.LABEL_PARTITIONING: "mig", pk.LABEL_PRODUCT: "NVIDIA-A30", pk.LABEL_COUNT: "0"},
                "annotations": {c["key"]: c["value"]}}
        code, out = hc.run(hc.to_text("mig", [node], [], "pack"))
        assert code == 0, out
        assert hc.parse_pack(out)["node_gpu_count"] == (["1"] if c["ok"] else ["0"])


def test_node_annotations_spec_keys_are_not_status():
    by_gpu = pk._status_annotations(NODE_ANN["annotations"])
    exp = NODE_ANN["expect_status"][0]
    assert by_gpu == {exp["index"]: [(exp["profile"], exp["status"], exp["quantity"])]}


@pytest.mark.gpu
def test_golden_nodes_on_the_gpu_parser():
    eng = capi.Engine(0)
    try:
        for c in NEW_NODE:
            if c["expect_error"]:
                continue                           # label errors are the host's (labels never reach the device)
            kind = pk.KIND_BY_NAME[c["kind"]]
            snap, _ = pk.pack(kind, pk.nodes_from_v1(kind, [c["node"]], m.DEFAULT_KNOWN_GEOMETRIES), [], m.DEFAULT_KNOWN_GEOMETRIES)
            out, _ = eng.parse_annotations(capi.flatten_annotations([c["node"]], snap.profiles), snap.gpus_per_node)
            n = snap.gpus_per_node
            d = {"profiles": snap.profiles, "node_gpu_count": [int(out["node_gpu_count"][0])], "gpu_used": out["gpu_used"][:n].reshape(-1),
                 "gpu_free": out["gpu_free"][:n].reshape(-1), "gpu_used_mask": out["gpu_used_mask"][:n], "gpu_free_mask": out["gpu_free_mask"][:n],
                 "gpu_index": out["gpu_index"][:n], "gpu_memory_gb": snap.gpu_memory_gb[:n]}
            _same(_gpus_of_pack(d), c["expect_gpus"])
        for c in PARSE + [dict(NODE_ANN, key=None)]:
            ann = c["annotations"] if c.get("key") is None else {c["key"]: c["value"]}
            node = {"name": "n", "labels": {pk.LABEL_COUNT: "0"}, "annotations": ann}
            out, _ = eng.parse_annotations(capi.flatten_annotations([node], ["1g.10gb", "2g.10gb"]), 4)
            valid = c.get("ok", True)
            assert int(out["node_gpu_count"][0]) == (1 if valid else 0), c["name"]
            if valid and c.get("key") is None:
                assert out["gpu_free"][0, 0] == 3 and out["gpu_index"][0] == 0 and out["gpu_used_mask"][0] == 0
    finally:
        eng.close()


def test_allowed_geometries_group_by_model_golden():
    """pkg/gpu/mig/allowed_geometries_test.go:133-203 (TestAllowedMigGeometriesList__GroupByModel)."""
    from nos_b200 import packing as pk
    assert pk.group_by_model([]) == {}
    got = pk.group_by_model([{"models": ["m1", "m2"], "allowedGeometries": [{"p1": 1, "p2": 2}, {"p1": 3}]},
                             {"models": ["m3"], "allowedGeometries": [{"p3": 2, "p2": 2}]}])
    assert got == {"m1": [{"p1": 1, "p2": 2}, {"p1": 3}], "m2": [{"p1": 1, "p2": 2}, {"p1": 3}], "m3": [{"p3": 2, "p2": 2}]}
    got["m1"][0]["p1"] = 9
    assert got["m2"][0]["p1"] == 1                                    # the models of one entry do not share their lists


def test_parse_spec_annotation_golden():
    """pkg/gpu/annotation_test.go:391-450 (TestParseSpecAnnotation) through the Python model and nos::gpu::ParseSpecAnnotation."""
    import nos_model as m
    from support import hostcli as hc
    prefix = "nos.nebuly.com/spec-gpu"
    cases = [("nos.nebuly.com/foo", "1", None), (prefix + "foo", "1", None), (f"{prefix}-0-1g.10gb", "foo", None),
             (f"{prefix}-1-1g.10gb", "1", (1, "1g.10gb", 1))]
    assert m.parse_spec_annotation("", "") is None                       # "Empty key and value" (the text driver cannot carry an empty key)
    for key, value, exp in cases:
        assert m.parse_spec_annotation(key, value) == exp, key
        code, out = hc.run(f"kind mig\nnode n\nannotation {key} {value}\nrun parsespec\n")
        assert code == 0, out
        want = f"spec {exp[0]} {exp[1]} {exp[2]}" if exp else f"notspec {key}"
        assert out.strip() == want, (key, out)
