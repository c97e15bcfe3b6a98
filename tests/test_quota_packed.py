"""The packed C++ quota oracle (oracle/nos_quota_oracle.cpp) through the product's quota packer, against the golden vectors
and against the faithful Python model on random quota states."""
import json
import os

import numpy as np
import pytest

from nos_b200 import quota_packing as qp
from support import oracle_lib as ol
from support import quotacases as qc

GOLD = json.load(open(os.path.join(os.path.dirname(__file__), "golden", "quota.json")))


def dense(pq, d):
    return np.array([d.get(c, 0) for c in pq.columns], np.int64)


@pytest.mark.parametrize("c", GOLD["guaranteed_overquotas"][1:], ids=[c["name"][:40] for c in GOLD["guaranteed_overquotas"][1:]])
def test_golden_guaranteed(c):
    quotas = {k: {"namespaces": [k], **v} for k, v in c["quotas"].items()}
    pq, _ = qp.pack_quotas(quotas, [], 16)
    r = ol.quota(pq)
    e = pq.names.index(c["quota"])
    assert np.array_equal(r["guaranteed"][e], dense(pq, c["expected"]))


@pytest.mark.parametrize("c", GOLD["aggregated_overquotas"][1:], ids=[c["name"][:40] for c in GOLD["aggregated_overquotas"][1:]])
def test_golden_aggregated_overquotas(c):
    quotas = {k: {"namespaces": [k], **v} for k, v in c["quotas"].items()}
    pq, _ = qp.pack_quotas(quotas, [], 16)
    assert np.array_equal(ol.quota(pq)["agg_overquota"], dense(pq, c["expected"]))


@pytest.mark.parametrize("c", GOLD["reserve"], ids=[c["name"][:40] for c in GOLD["reserve"]])
def test_golden_reserve(c):
    quotas = {"eq": {"namespaces": ["ns"], "min": {}, "used": {k: v for k, v in c["used"].items()}}}
    pods = [{"namespace": "ns", "name": f"p{i}", "request": dict(p)} for i, p in enumerate(c["pods"])]
    pq, (run,) = qp.pack_quotas(quotas, [pods], c["gpu_memory_gb"])
    assert np.array_equal(ol.quota(pq, run)["used"][0], dense(pq, c["expected"]))


@pytest.mark.parametrize("c", GOLD["prefilter"], ids=[c["name"][:40] for c in GOLD["prefilter"]])
def test_golden_prefilter(c):
    pods = [{"namespace": p["namespace"], "name": f"p{i}", "request": {"memory": p["memory"], "cpu": p["cpu"], "nvidia.com/gpu": p["gpu"]}}
            for i, p in enumerate(c["pods"])]
    pq, (pend,) = qp.pack_quotas(c["quotas"], [pods], c["gpu_memory_gb"])
    r = ol.quota(pq, None, pend)
    assert [int(x != 0) for x in r["prefilter"]] == c["expected"]


@pytest.mark.parametrize("block", range(4))
def test_random_vs_python_model(block):
    for seed in range(block * 150, (block + 1) * 150):
        quotas, running, pending, gb = qc.rand_quota_case(seed)
        pq, (run, pend) = qp.pack_quotas(quotas, [running, pending], gb)
        got = ol.quota(pq, run, pend)
        exp = qc.model_eval(quotas, running, pending, gb, pq)
        for k, v in exp.items():
            assert np.array_equal(got[k], v), f"seed {seed}: {k}\n{got[k]}\n{v}"
        assert got["ambiguous_order"] == 0
