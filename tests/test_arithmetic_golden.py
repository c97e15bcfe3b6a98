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
