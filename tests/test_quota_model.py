"""The faithful Python quota model (oracle/nos_quota_model.py) against the vectors ported from the reference's
capacity-scheduling and elasticquota-operator tests (tests/golden/quota.json)."""
import json
import os

import pytest

import nos_model as m
import nos_quota_model as q

GOLD = json.load(open(os.path.join(os.path.dirname(__file__), "golden", "quota.json")))


def ids(cs):
    return [c["name"][:50] for c in cs]


def res(d):
    return m.resource_from_list(d) if d is not None else None


def infos_from(quotas):
    out = q.ElasticQuotaInfos()
    for name, e in quotas.items():
        info = q.ElasticQuotaInfo(name=name, namespaces=e.get("namespaces", [name]), min=res(e.get("min")), max=res(e.get("max")),
                                  used=res(e.get("used")), max_enforced=e.get("max_enforced", False))
        for ns in info.namespaces:
            out[ns] = info
    return out


def make_pod(d, i=0):
    c = {"memory": d["memory"], "cpu": d["cpu"], "nvidia.com/gpu": d.get("nvidia.com/gpu", d.get("gpu", 0))}
    return m.Pod(d.get("namespace", "ns"), d.get("name", f"p{i}"), d.get("priority", 0), [c])


def as_list(r):
    return {k: v for k, v in m.resource_to_list(r).items()}


@pytest.mark.parametrize("c", GOLD["reserve"], ids=ids(GOLD["reserve"]))
def test_reserve(c):
    e = q.ElasticQuotaInfo(used=res(c["used"]))
    for i, p in enumerate(c["pods"]):
        e.reserve(m.resource_from_list(q.quota_pod_request(make_pod(p, i), c["gpu_memory_gb"])))
    assert e.used == res(c["expected"])


@pytest.mark.parametrize("c", GOLD["unreserve"], ids=ids(GOLD["unreserve"]))
def test_unreserve(c):
    e = q.ElasticQuotaInfo(used=res(c["used"]))
    for i, p in enumerate(c["pods"]):
        e.unreserve(m.resource_from_list(q.quota_pod_request(make_pod(p, i), c["gpu_memory_gb"])))
    assert e.used == res(c["expected"])


@pytest.mark.parametrize("c", GOLD["used_over_max_with"], ids=ids(GOLD["used_over_max_with"]))
def test_used_over_max_with(c):
    e = q.ElasticQuotaInfo(used=res(c["used"]), max=res(c["max"]), max_enforced=c["max_enforced"])
    assert e.used_over_max_with(res(c["request"])) == c["expected"]


@pytest.mark.parametrize("c", GOLD["guaranteed_overquotas"], ids=ids(GOLD["guaranteed_overquotas"]))
def test_guaranteed_overquotas(c):
    got = infos_from(c["quotas"]).guaranteed_overquotas(c["quota"])
    if c["expected"] is None:
        assert got is None
    else:
        assert got == res(c["expected"])


@pytest.mark.parametrize("c", GOLD["percentages"], ids=ids(GOLD["percentages"]))
def test_percentages(c):
    infos = infos_from(c["quotas"])
    got = infos.guaranteed_overquota_percentages(infos[c["quota"]])
    assert got == {k: float(v) for k, v in c["expected"].items()}       # exact float equality, as assert.Equal in Go


@pytest.mark.parametrize("c", GOLD["aggregated_overquotas"], ids=ids(GOLD["aggregated_overquotas"]))
def test_aggregated_overquotas(c):
    infos = infos_from(c["quotas"])
    got = infos.aggregated_overquotas()
    assert got == res(c["expected"])
    mn = infos.aggregated_min()
    assert got.milli_cpu <= mn.milli_cpu and got.memory <= mn.memory


@pytest.mark.parametrize("c", GOLD["used_lte_with"], ids=ids(GOLD["used_lte_with"]))
def test_used_lte_with(c):
    e = q.ElasticQuotaInfo(used=res(c["used"]))
    assert e.used_lte_with(res(c["resource"]), res(c["request"])) == c["expected"]


@pytest.mark.parametrize("c", GOLD["aggregated_used_over_min_with"], ids=ids(GOLD["aggregated_used_over_min_with"]))
def test_aggregated_used_over_min_with(c):
    assert infos_from(c["quotas"]).aggregated_used_over_min_with(res(c["request"])) == c["expected"]


@pytest.mark.parametrize("c", GOLD["prefilter"], ids=ids(GOLD["prefilter"]))
def test_prefilter(c):
    infos = infos_from(c["quotas"])
    got = [int(q.prefilter(infos, make_pod(p, i), c["gpu_memory_gb"]) != 0) for i, p in enumerate(c["pods"])]
    assert got == c["expected"]


def test_operator_sort_golden():
    """internal/controllers/elasticquota/elasticquota_test.go:30-133"""
    def pod(name, ts, prio, gpu=None):
        return m.Pod("ns-1", name, prio, [{"nvidia.com/gpu": gpu}] if gpu is not None else [], creation_ts=ts)
    y95, y96, y00 = 788918400, 820454400, 946684800
    cases = [
        ([], []),
        ([pod("pd-1", y95, 100), pod("pd-2", y00, 10), pod("pd-3", y96, 50)], ["pd-1", "pd-3", "pd-2"]),
        ([pod("pd-1", y95, 100), pod("pd-2", y95, 10), pod("pd-3", y95, 50)], ["pd-2", "pd-3", "pd-1"]),
        ([pod("pd-1", y95, 10, 10), pod("pd-2", y95, 10, 1), pod("pd-3", y95, 10, 5)], ["pd-2", "pd-3", "pd-1"]),
        ([pod("pd-1", y95, 10), pod("pd-3", y95, 10), pod("pd-2", y95, 10)], ["pd-1", "pd-2", "pd-3"]),
    ]
    for pods, exp in cases:
        assert [p.name for p in q.operator_sort(pods, 16)] == exp


def test_patch_pods_and_compute_used():
    pods = [m.Pod("ns", f"p{i}", 0, [{"cpu": 1000, "nvidia.com/gpu": 1}], creation_ts=i) for i in range(4)]
    used, labels = q.patch_pods_and_compute_used(pods, {"cpu": 2500, "nos.nebuly.com/gpu-memory": 48}, {}, 16)
    assert [o for _, o in labels] == [False, False, True, True]
    assert used == {"cpu": 4000, "nos.nebuly.com/gpu-memory": 64}
