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


def _info(name, ns_name, namespaces):
    e = q.ElasticQuotaInfo(name=name, namespaces=list(namespaces))
    e.resource_namespace = ns_name
    return e


def test_elastic_quota_infos_add_update_golden():
    """capacity_scheduling_test.go:562-700 (TestElasticQuotaInfos_Add / _Update): which namespace keys point at which quota after
    Add and Update -- the map the packer's ns_quota row encodes."""
    from nos_b200 import quota_packing as qp
    # Add
    infos = q.ElasticQuotaInfos(); infos.add(q.ElasticQuotaInfo())
    assert len(infos) == 0                                                          # "Empty ElasticQuotaInfos, empty ElasticQuotaInfo"
    infos = q.ElasticQuotaInfos({"ns-1": q.ElasticQuotaInfo()}); infos.add(q.ElasticQuotaInfo())
    assert list(infos) == ["ns-1"]                                                  # "Empty ElasticQuotaInfo"
    infos = q.ElasticQuotaInfos({"ns-1": _info("test", "test", [])})
    infos.add(_info("updated", "updated", ["ns-2", "ns-3", "ns-4"]))               # "ElasticQuotaInfo with some namespaces not present"
    assert sorted(infos) == ["ns-1", "ns-2", "ns-3", "ns-4"] and infos["ns-1"].name == "test"
    assert all(infos[k].name == "updated" and sorted(infos[k].namespaces) == ["ns-2", "ns-3", "ns-4"] for k in ("ns-2", "ns-3", "ns-4"))
    assert infos["ns-2"] is infos["ns-3"] is infos["ns-4"]                          # one object under several keys
    # Update
    infos = q.ElasticQuotaInfos()
    infos.update(q.ElasticQuotaInfo(), _info("new-name", "new-ns", ["ns-1", "ns-2"]))   # "Empty ElasticQuotaInfos"
    assert sorted(infos) == ["ns-1", "ns-2"] and all(e.name == "new-name" and e.resource_namespace == "new-ns" for e in infos.values())
    infos = q.ElasticQuotaInfos()
    infos.update(_info("old-name", "old-ns", ["ns-1", "ns-2"]), _info("new-name", "new-ns", ["ns-2", "ns-3"]))
    assert sorted(infos) == ["ns-2", "ns-3"] and all(sorted(e.namespaces) == ["ns-2", "ns-3"] for e in infos.values())
    # Update carries Used and the accounted pods over (elasticquotainfo.go:54-57), Delete drops every key of the quota (:45-49)
    old = _info("eq", "ns", ["ns-1"]); old.used = m.resource_from_list({"cpu": 500}); old.pods = {"uid-1"}
    infos = q.ElasticQuotaInfos(); infos.add(old)
    new = _info("eq", "ns", ["ns-1", "ns-9"]); new.used = m.Resource()
    infos.update(old, new)
    assert infos["ns-9"].used.milli_cpu == 500 and infos["ns-9"].pods == {"uid-1"}
    infos.delete(new)
    assert len(infos) == 0
    # the packer's map: every namespace of a composite quota points at the quota's one row
    pq, _ = qp.pack_quotas({"test": {"namespaces": ["ns-1"], "min": {"cpu": 1}}, "updated": {"namespaces": ["ns-2", "ns-3", "ns-4"], "min": {"cpu": 1}}}, [], 8)
    row = {ns: int(pq.ns_quota[i]) for i, ns in enumerate(pq.namespaces)}
    assert row["ns-1"] == pq.names.index("test") and row["ns-2"] == row["ns-3"] == row["ns-4"] == pq.names.index("updated")
