"""The oracle's resource and profile arithmetic against the reference's own table tests (tests/golden/arithmetic.json), and the
same facts through the packers where they surface there (profile memory / Gi columns, gpu-memory coefficients)."""
import json
import os

import pytest

import nos_model as m
import nos_quota_model as qm
from nos_b200 import packing as pk

G = json.load(open(os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "arithmetic.json")))


def R(d):
    return m.Resource(d["milli_cpu"], d["memory"], d["ephemeral"], d["allowed_pods"], dict(d["scalars"]))


@pytest.mark.parametrize("op,fn", [("sum", m.res_sum), ("subtract", m.res_subtract), ("subtract_non_negative", m.res_subtract_non_negative)])
def test_resource_arithmetic(op, fn):
    for c in G[op]:
        assert fn(R(c["r1"]), R(c["r2"])) == R(c["expected"]), (op, c["name"])


def test_profile_comparisons():
    for c in G["mig_smaller_than"]:
        assert m.mig_smaller_than(c["a"], c["b"]) == c["expected"], c["name"]
    for c in G["mig_gi_slices"]:
        assert m.mig_gi_slices(c["profile"]) == c["expected"] and pk.mig_gi(c["profile"]) == c["expected"]
    for c in G["mps_memory_size_gb"]:
        assert m.mps_memory_gb(c["profile"]) == c["expected"] and pk.mps_size(c["profile"]) == c["expected"], c["name"]
    for c in G["mps_smaller_than"]:
        assert m.mps_smaller_than(c["a"], c["b"]) == c["expected"], c["name"]


def test_required_gpu_memory():
    for c in G["required_gpu_memory_gb"]:
        pod = m.Pod("ns", "p", containers=[dict(c["request"])])
        assert qm.quota_pod_request(pod, c["nvidia_gpu_memory_gb"])[m.RESOURCE_GPU_MEMORY] == c["expected"], c["name"]


def test_mig_resource_name_functions():
    for c in G["is_nvidia_mig_device"]:
        assert m.is_mig_resource(c["resource"]) == c["expected"], c["name"]
        assert (pk.slice_profile_of(pk.KIND_MIG, c["resource"]) is not None) == c["expected"], c["name"]
    for c in G["extract_memory_gb_from_mig_format"]:
        # the model folds the error into 0, which is how every caller on the path uses it (pkg/gpu/util/resource.go:76-82)
        assert m.mig_memory_gb_from_resource(c["resource"]) == (0 if c["error"] else c["expected"]), c["name"]


def test_validate_configs():
    from support import hostcli as hc
    import __graft_entry__ as g
    g.build()
    node = {"name": "n", "labels": {pk.LABEL_PARTITIONING: "mig", pk.LABEL_PRODUCT: "NVIDIA-A30", pk.LABEL_COUNT: "1"}}
    for c in G["validate_configs"]:
        configs = m.DEFAULT_KNOWN_GEOMETRIES if c["configs"] == "default" else c["configs"]
        if c["error"]:
            with pytest.raises(pk.PackError):
                pk.validate_known_geometries(configs)
        else:
            pk.validate_known_geometries(configs)
        if configs:                                   # the text driver cannot state an empty table (it falls back to the default)
            code, out = hc.run(hc.to_text("mig", [node], [], "state", geometries=configs))
            assert (code == 1 and out.startswith("error ")) == c["error"], (c["name"], out)
