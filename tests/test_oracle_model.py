"""The faithful Python model (oracle/nos_model.py) against the golden vectors ported from the
reference's own tests (tests/golden/*.json)."""
import json
import os

import pytest

import nos_model as m

GOLD = os.path.join(os.path.dirname(__file__), "golden")


def _load(name):
    with open(os.path.join(GOLD, name)) as f:
        return json.load(f)


PLAN = _load("planner_plan.json")["cases"]
COMP = _load("components.json")
CORE = _load("core.json")


def _ids(cases):
    return [f"{c.get('kind', '')}:{c['name'][:60]}" for c in cases]


def overall(kind, state):
    got = []
    for _node, gl in state.items():
        for idx, res in gl:
            got.append((0 if kind == "mig" else idx, tuple(sorted(res.items()))))
    return sorted(got)


def expected_overall(kind, exp):
    return sorted((0 if kind == "mig" else e["gpu_index"], tuple(sorted(e["resources"].items()))) for e in exp)


@pytest.mark.parametrize("c", PLAN, ids=_ids(PLAN))
def test_planner_plan_golden(c):
    snap = m.take_snapshot(c["kind"], c["nodes"])
    pods = [m.pod_from(p) for p in c["pods"]]
    state, _ = m.plan(snap, pods, [c["prefilter_ok"]] * len(pods), m.FILTER_NONE, c["filter_ok"])
    assert overall(c["kind"], state) == expected_overall(c["kind"], c["expected"])


def _mig_gpu(c):
    return m.MigGPU(c["model"], c.get("index", 0), dict(c["used"]), dict(c["free"]), m.DEFAULT_KNOWN_GEOMETRIES[c["model"]])


@pytest.mark.parametrize("c", COMP["mig_gpu_add_pod"], ids=_ids(COMP["mig_gpu_add_pod"]))
def test_mig_gpu_add_pod(c):
    g = _mig_gpu(c)
    ok = g.add_pod(m.pod_from(c["pod"]))
    assert ok == (not c["expect_err"])
    assert g.used == c["expect_used"] and g.free == c["expect_free"]


@pytest.mark.parametrize("c", COMP["mig_gpu_apply_geometry"], ids=_ids(COMP["mig_gpu_apply_geometry"]))
def test_mig_gpu_apply_geometry(c):
    g = _mig_gpu(c)
    ok = g.apply_geometry(c["geometry"])
    assert ok == (not c["expect_err"])
    assert g.used == c["expect_used"] and g.free == c["expect_free"]


@pytest.mark.parametrize("c", COMP["mig_gpu_update_geometry"], ids=_ids(COMP["mig_gpu_update_geometry"]))
def test_mig_gpu_update_geometry(c):
    g = _mig_gpu(c)
    upd, _ = g.update_geometry_for(dict(c["required"]))
    assert upd == c["expect_updated"]
    assert g.geometry() == c["expect_geometry"]


@pytest.mark.parametrize("c", COMP["mig_node_update_geometry"], ids=_ids(COMP["mig_node_update_geometry"]))
def test_mig_node_update_geometry(c):
    node = m.MigNode("test", [_mig_gpu(g) for g in c["gpus"]], m.NodeInfo())
    upd, _ = node.update_geometry_for(dict(c["required"]))
    assert upd == c["expect_updated"]
    assert node.geometry() == c["expect_geometry"]


@pytest.mark.parametrize("c", COMP["mig_node_has_free_capacity"], ids=_ids(COMP["mig_node_has_free_capacity"]))
def test_mig_node_has_free_capacity(c):
    node = m.MigNode("test", [_mig_gpu(g) for g in c["gpus"]], m.NodeInfo())
    assert node.has_free_capacity() == c["expect"]


@pytest.mark.parametrize("c", COMP["node_add_pod"], ids=_ids(COMP["node_add_pod"]))
def test_node_add_pod(c):
    snap = m.take_snapshot(c["kind"], [c["node"]])
    node = snap.get_node(c["node"]["name"])
    ok = node.add_pod(m.pod_from(c["pod"]))
    assert ok == (not c["expect_err"])
    used, free = {}, {}
    for g in node.gpus:
        for p, q in g.used.items():
            used[p] = used.get(p, 0) + q
        for p, q in g.free.items():
            free[p] = free.get(p, 0) + q
    assert used == c["expect_used"] and free == c["expect_free"]
    assert m.resource_to_list(node.node_info.requested) == {
        **{"cpu": 0, "memory": 0, "pods": 0, "ephemeral-storage": 0}, **c["expect_requested"]}
    assert node.node_info.num_pods == 0  # the Pods append is lost on the value copy (mig/node.go:204-205)


@pytest.mark.parametrize("c", COMP["mps_gpu_new"], ids=_ids(COMP["mps_gpu_new"]))
def test_mps_gpu_new(c):
    g = m.MpsGPU("NVIDIA-A100-80GB-PCIe", 0, c["memory_gb"], dict(c["used"]), dict(c["free"]))
    assert (g.validate() is not None) == c["expect_err"]


@pytest.mark.parametrize("c", COMP["mps_gpu_update_geometry"], ids=_ids(COMP["mps_gpu_update_geometry"]))
def test_mps_gpu_update_geometry(c):
    g = m.MpsGPU("NVIDIA-A100-80GB-PCIe", 0, c["memory_gb"], dict(c["used"]), dict(c["free"]))
    upd = g.update_geometry_for(dict(c["required"]))
    assert upd == c["expect_updated"]
    assert g.geometry() == c["expect_geometry"]


@pytest.mark.parametrize("c", COMP["mps_node_has_free_capacity"], ids=_ids(COMP["mps_node_has_free_capacity"]))
def test_mps_node_has_free_capacity(c):
    gpus = [m.MpsGPU("NVIDIA-A30", i, g["memory_gb"], dict(g["used"]), dict(g["free"])) for i, g in enumerate(c["gpus"])]
    assert m.MpsNode("test", gpus, m.NodeInfo()).has_free_capacity() == c["expect"]


@pytest.mark.parametrize("c", COMP["mps_node_update_geometry"], ids=_ids(COMP["mps_node_update_geometry"]))
def test_mps_node_update_geometry(c):
    node = m.new_mps_node({"name": "node-1", "labels": c["labels"], "annotations": c["annotations"],
                           "allocatable": {"nebuly.com/foo": 1}})
    upd, _ = node.update_geometry_for(dict(c["required"]))
    assert upd == c["expect_updated"]
    if node.gpus and c["required"]:
        assert node.node_info.allocatable.scalars.get("nebuly.com/foo") == 1   # non-slice scalars survive (node.go:157-161)


@pytest.mark.parametrize("c", CORE["lacking_slices"], ids=_ids(CORE["lacking_slices"]))
def test_lacking_slices(c):
    nodes = {n["name"]: m.MigNode(n["name"], [], m.node_info_from(n)) for n in c["node_infos"]}
    snap = m.ClusterSnapshot(c["kind"], nodes)
    assert snap.lacking_slices(m.pod_from(c["pod"])) == c["expected"]


@pytest.mark.parametrize("c", CORE["tracker_remove"], ids=_ids(CORE["tracker_remove"]))
def test_tracker_remove(c):
    snap = m.ClusterSnapshot("mig", {})
    tr = m.SliceTracker(snap, [m.pod_from(p) for p in c["pods"]])
    tr.remove(m.pod_from(c["remove"]))
    assert tr.requested == c["expect_requested"] and tr.lacking == c["expect_lacking"]


@pytest.mark.parametrize("c", CORE["pod_sorter"], ids=_ids(CORE["pod_sorter"]))
def test_pod_sorter(c):
    out = m.sort_pods("mig", [m.pod_from(p) for p in c["pods"]])
    assert [p.key() for p in out] == c["expected"]


def test_go_stable_sort_matches_python_sorted_on_strict_weak_orders():
    import random
    rng = random.Random(7)
    for n in (0, 1, 2, 19, 20, 21, 39, 40, 41, 100, 257, 1000):
        xs = [(rng.randrange(8), i) for i in range(n)]
        got = m.go_stable_sort(xs, lambda a, b: a[0] < b[0])
        assert got == sorted(xs, key=lambda t: t[0])


def test_snapshot_fork_commit_revert():
    """core/snapshot_test.go:152-231: fork twice fails; revert drops the fork; commit keeps it."""
    node = {"name": "node-1", "labels": {"nvidia.com/gpu.product": "NVIDIA-A30", "nvidia.com/gpu.count": "1",
                                         "nos.nebuly.com/gpu-partitioning": "mig"},
            "annotations": {"nos.nebuly.com/status-gpu-0-1g.6gb-free": "4"}}
    snap = m.take_snapshot("mig", [node])
    snap.fork()
    with pytest.raises(RuntimeError):
        snap.fork()
    pod = m.pod_from({"namespace": "a", "name": "b", "containers": [{"nvidia.com/mig-1g.6gb": 1}]})
    assert snap.add_pod("node-1", pod)
    assert snap.get_node("node-1").gpus[0].free == {"1g.6gb": 3}
    snap.revert()
    assert snap.get_node("node-1").gpus[0].free == {"1g.6gb": 4}
    snap.fork()
    assert snap.add_pod("node-1", pod)
    snap.commit()
    assert snap.get_node("node-1").gpus[0].free == {"1g.6gb": 3}
    assert snap.get_node("node-1").gpus[0].used == {"1g.6gb": 1}
