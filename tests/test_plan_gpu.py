"""Parity of the CUDA planner (libnosplan.so through the C-ABI) against the oracle -- bit-exact plans and counters."""
import json
import os

import numpy as np
import pytest

import nos_model as m
from nos_b200 import capi, packing as pk, synth
from nos_b200.geometries import DEFAULT_KNOWN_GEOMETRIES, HELM_KNOWN_GEOMETRIES
from support import oracle_lib as ol
from support import randcases as rc

pytestmark = pytest.mark.gpu

GOLD = os.path.join(os.path.dirname(__file__), "golden")
PLAN = json.load(open(os.path.join(GOLD, "planner_plan.json")))["cases"]
COUNTERS = ("candidates", "visited_nodes", "placements", "geometry_evals", "candidate_nodes", "lacking_at_start", "lacking_at_end")


@pytest.fixture(scope="module")
def eng():
    import __graft_entry__ as g
    g.build()
    e = capi.Engine(0)
    yield e
    e.close()


def check(eng, snap, pods, pre, fm, what=""):
    cnt, msk, st = eng.plan(snap, pods, pre, fm)
    ocnt, omsk, ost, _ = ol.plan(snap, pods, pre, fm, ol.INCREMENTAL)
    assert np.array_equal(msk, omsk), f"{what}: key masks differ"
    assert np.array_equal(cnt, ocnt), f"{what}: counts differ"
    for k in COUNTERS:
        assert st[k] == ost[k], f"{what}: {k} {st[k]} != {ost[k]}"
    return st


@pytest.mark.parametrize("c", PLAN, ids=[f"{c['kind']}:{c['name'][:50]}" for c in PLAN])
def test_golden_planner(eng, c):
    kind = pk.KIND_BY_NAME[c["kind"]]
    nodes = pk.nodes_from_v1(kind, c["nodes"], m.DEFAULT_KNOWN_GEOMETRIES)
    s, p = pk.pack(kind, nodes, c["pods"], m.DEFAULT_KNOWN_GEOMETRIES)
    fm = pk.FILTER_NONE if c["filter_ok"] else pk.FILTER_REJECT_ALL
    cnt, msk, _ = eng.plan(s, p, [1 if c["prefilter_ok"] else 0] * p.n_pods, fm)
    got = sorted((0 if c["kind"] == "mig" else i, tuple(sorted(r.items()))) for gl in pk.unpack_plan(s, cnt, msk).values() for i, r in gl)
    exp = sorted((0 if c["kind"] == "mig" else e["gpu_index"], tuple(sorted(e["resources"].items()))) for e in c["expected"])
    assert got == exp


def _has_zero_qty_slice(kind, pods):
    for p in pods:
        for r, q in pk.compute_pod_request(p).items():
            if q <= 0 and pk.slice_profile_of(kind, r) is not None:
                return True
    return False


@pytest.mark.parametrize("block", range(8))
def test_random_small_vs_oracle(eng, block):
    placed = 0
    for seed in range(block * 400, (block + 1) * 400):
        kind_name, nodes, pods, pre, fmode = rc.rand_case(seed)
        kind = pk.KIND_BY_NAME[kind_name]
        if _has_zero_qty_slice(kind, pods):
            continue
        s, p = pk.pack(kind, nodes, pods, m.DEFAULT_KNOWN_GEOMETRIES)
        st = check(eng, s, p, pre, fmode, f"seed {seed}")
        placed += st["placements"] > 0
    assert placed > 30


@pytest.mark.parametrize("fm", [pk.FILTER_NONE, pk.FILTER_RESOURCES_FIT], ids=["none", "fit"])
@pytest.mark.parametrize("shape", [(1, 1, 10), (8, 8, 40), (64, 8, 256), (200, 8, 1000), (333, 5, 777)])
def test_synth_mig(eng, shape, fm):
    for regime in ("saturating", "satisfiable"):
        s, p = synth.mig_cluster(*shape, seed=0xC2 + shape[0], regime=regime)
        check(eng, s, p, None, fm, f"mig {shape} {regime}")


@pytest.mark.parametrize("fm", [pk.FILTER_NONE, pk.FILTER_RESOURCES_FIT], ids=["none", "fit"])
@pytest.mark.parametrize("shape", [(1, 1, 10), (8, 8, 40), (64, 8, 256), (200, 8, 1000)])
def test_synth_mps(eng, shape, fm):
    for regime in ("saturating", "satisfiable"):
        s, p = synth.mps_cluster(*shape, seed=0xC3 + shape[0], regime=regime)
        check(eng, s, p, None, fm, f"mps {shape} {regime}")


def test_helm_table_order(eng):
    """Same model, production table (different row order -> different canonical tie-breaks)."""
    s, p = synth.mig_cluster(100, 8, 500, seed=7, known=HELM_KNOWN_GEOMETRIES, model="NVIDIA-A100-80GB-PCIe")
    check(eng, s, p, None, pk.FILTER_RESOURCES_FIT, "helm table")


def c1_snapshot():
    """BASELINE config #1 exactly as SURVEY.md section 8d specifies it: 1 node, 1 x NVIDIA-A100-40GB-SXM4 (the 12 allowed
    geometries of pkg/gpu/mig/known_configs.go:41-90), free {7g.40gb: 1}, allocatable cpu 64 000 m / memory 512 GiB / pods 110;
    10 pods, priority 0, alternating 1g.5gb:1 / 2g.10gb:1, cpu 100 m.  No RNG."""
    from nos_b200.geometries import A100_40
    s, p = synth.mig_cluster(1, 1, 10, seed=0, model=A100_40, regime="satisfiable", p_full_free=0.0, p_used=0.0)
    prof = {n: i for i, n in enumerate(s.profiles)}
    K = s.n_profiles
    s.gpu_used[:] = 0; s.gpu_free[:] = 0; s.gpu_used_mask[:] = 0
    s.gpu_free[0, prof["7g.40gb"]] = 1; s.gpu_free_mask[0] = 1 << prof["7g.40gb"]
    s.node_alloc[:] = 0; s.node_requested[:] = 0
    s.node_alloc[0, 0] = 64000; s.node_alloc[0, 1] = 512 << 30; s.node_alloc[0, 3 + prof["7g.40gb"]] = 1
    s.node_allowed_pods[:] = 110; s.node_num_pods[:] = 0
    p.priority[:] = 0
    p.base[:] = 0; p.base[:, 0] = 100
    p.scalar_col[:] = [3 + prof["1g.5gb" if i % 2 == 0 else "2g.10gb"] for i in range(10)]
    p.scalar_qty[:] = 1
    return s, p, prof, K


def test_config_c1_exact(eng):
    s, p, prof, K = c1_snapshot()
    assert s.geom_counts.shape[0] == 12
    for fm in (pk.FILTER_NONE, pk.FILTER_RESOURCES_FIT):
        cnt, msk, st = eng.plan(s, p, None, fm)
        ocnt, omsk, ost, _ = ol.plan(s, p, None, fm, ol.INCREMENTAL)
        assert np.array_equal(cnt, ocnt) and np.array_equal(msk, omsk)
        for k in COUNTERS:
            assert st[k] == ost[k], k
        # the one GPU is re-partitioned away from {7g.40gb: 1}; both requested profiles appear and pods are placed on it
        assert cnt[0, prof["7g.40gb"]] == 0 and cnt[0, prof["1g.5gb"]] > 0 and cnt[0, prof["2g.10gb"]] > 0
        assert st["visited_nodes"] == 1 and st["candidates"] == 10 and st["placements"] > 0
    # the faithful mode of the oracle (the reference's own asymptotics) agrees as well
    fcnt, fmsk, fst, _ = ol.plan(s, p, None, pk.FILTER_RESOURCES_FIT, ol.FAITHFUL)
    assert np.array_equal(cnt, fcnt) and np.array_equal(msk, fmsk)


def test_config_c5_mixed_mig_and_mps(eng):
    """BASELINE config #5 at its lower end in miniature and in full: a snapshot that is half MIG nodes (C2 recipe) and half
    MPS nodes (C3 recipe), 50/50 requests, planned per kind as the reference's two planners do
    (cmd/gpupartitioner/gpupartitioner.go:215-250: one controller per partitioning kind over the nodes of that kind)."""
    for n, pods in ((200, 400), (10000, 10000)):
        smig, pmig = synth.mig_cluster(n // 2, 8, pods // 2, seed=0xC5)
        smps, pmps = synth.mps_cluster(n // 2, 8, pods // 2, seed=0xC5 + 1)
        a = check(eng, smig, pmig, None, pk.FILTER_RESOURCES_FIT, f"C5 mig half {n}")
        b = check(eng, smps, pmps, None, pk.FILTER_RESOURCES_FIT, f"C5 mps half {n}")
        assert a["candidates"] == a["visited_nodes"] * (pods // 2) and b["candidates"] == b["visited_nodes"] * (pods // 2)


def test_config_c2_full_size(eng):
    """BASELINE config #2: 1,000 nodes x 8 A100-80GB, 5,000 pending pods."""
    s, p = synth.mig_cluster(1000, 8, 5000, seed=0xC2)
    for fm in (pk.FILTER_NONE, pk.FILTER_RESOURCES_FIT):
        st = check(eng, s, p, None, fm, "C2")
        assert st["candidates"] == 5_000_000


def test_config_c3_full_size(eng):
    """BASELINE config #3: 5,000 nodes x 8 GPUs MPS, 20,000 pending pods."""
    s, p = synth.mps_cluster(5000, 8, 20000, seed=0xC3)
    st = check(eng, s, p, None, pk.FILTER_RESOURCES_FIT, "C3")
    assert st["candidates"] == st["visited_nodes"] * 20000


def test_north_star_size_properties(eng):
    """10k nodes x 8 GPUs, 10k pods: compared with the incremental oracle and through size-independent
    properties: used devices are never deleted, geometry of every touched GPU is an allowed one, idempotence of
    a second call on the same handle."""
    s, p = synth.mig_cluster(10000, 8, 10000, seed=0xC5)
    st = check(eng, s, p, None, pk.FILTER_RESOURCES_FIT, "10k x 10k")
    assert st["candidates"] == st["visited_nodes"] * 10000
    cnt, msk, st1 = eng.plan(s, p, None, pk.FILTER_RESOURCES_FIT)
    cnt2, msk2, st2 = eng.plan(s, p, None, pk.FILTER_RESOURCES_FIT)
    assert np.array_equal(cnt, cnt2) and np.array_equal(msk, msk2)
    assert (cnt.astype(int) >= s.gpu_used.astype(int)).all()
    init = s.gpu_used + s.gpu_free
    changed = (cnt != init).any(1)
    rows = {tuple(r) for r in s.geom_counts}
    assert changed.any()
    assert all(tuple(r) in rows for r in cnt[changed])


def test_prefilter_vector_and_duplicates(eng):
    s, p = synth.mig_cluster(50, 4, 300, seed=11)
    rng = np.random.default_rng(3)
    pre = (rng.random(300) < 0.7).astype(np.uint8)
    p.tracker_key[:] = rng.integers(0, 40, 300)          # many pods share a namespace/name
    check(eng, s, p, pre, pk.FILTER_RESOURCES_FIT, "prefilter+dups")
    check(eng, s, p, pre, pk.FILTER_REJECT_ALL, "reject all")


def test_errors_are_reported_not_crashes(eng):
    s, p = synth.mig_cluster(4, 2, 8, seed=1)
    p.scalar_qty[0] = 0
    with pytest.raises(capi.NosplanError):
        eng.plan(s, p, None, 0)
    p.scalar_qty[0] = 1 << 31                              # the lacking quantities would not fit the kernels' int32
    with pytest.raises(capi.NosplanError, match="2\\^31"):
        eng.plan(s, p, None, 0)
    p.scalar_qty[0] = 1
    with pytest.raises(capi.NosplanError):
        eng.plan(s, p, None, 7)
    eng.plan(s, p, None, 0)                               # the handle stays usable


def test_errors_while_the_node_dma_is_in_flight(eng):
    """On a snapshot large enough for its node arrays to be in flight while the host validates: an error found in a node, an
    error found in a pod, and both at once come back as codes, and the handle keeps working -- with results identical to
    the oracle's."""
    s, p = synth.mig_cluster(2500, 8, 3000, seed=21)
    good_model, good_qty = int(s.gpu_model[8 * 1700]), int(p.scalar_qty[5])
    s.gpu_model[8 * 1700] = 99
    with pytest.raises(capi.NosplanError, match="node 1700: GPU model id out of range"):
        eng.plan(s, p, None, pk.FILTER_RESOURCES_FIT)
    p.scalar_qty[5] = 0                                   # a pod error as well: whichever is reported, it is an error
    with pytest.raises(capi.NosplanError):
        eng.plan(s, p, None, pk.FILTER_RESOURCES_FIT)
    s.gpu_model[8 * 1700] = good_model
    with pytest.raises(capi.NosplanError, match="pod 5"):
        eng.plan(s, p, None, pk.FILTER_RESOURCES_FIT)
    p.scalar_qty[5] = good_qty
    for _ in range(3):
        check(eng, s, p, None, pk.FILTER_RESOURCES_FIT, "after errors")


def test_paths_taken(eng):
    """The node-parallel pass must be the one that produced the saturating plans, the sequential sweep the others."""
    s, p = synth.mig_cluster(200, 8, 1000, seed=5)
    st = check(eng, s, p, None, pk.FILTER_RESOURCES_FIT, "saturating")
    assert st["path"] == 1
    s, p = synth.mig_cluster(200, 8, 1000, seed=5, regime="satisfiable")
    st = check(eng, s, p, None, pk.FILTER_RESOURCES_FIT, "satisfiable")
    assert st["path"] == 0
    # inconsistent node accounts -> the cluster-wide gate matters: still the node-parallel pass, with the gate replayed over
    # its records and a few nodes planned again (tests/test_gate_replay_gpu.py)
    s, p = synth.mig_cluster(200, 8, 1000, seed=5)
    s.node_requested[7, 3] += 5
    st = check(eng, s, p, None, pk.FILTER_RESOURCES_FIT, "inconsistent")
    assert st["path"] == 1 and st["gate_replayed"] >= 0


def test_saturation_hypothesis_fails_a_posteriori(eng):
    """lack starts just above the saturation bound (G+2)*M = 70 and is driven below it by the placements: the
    node-parallel result must be discarded and the sequential sweep must produce the plan."""
    for n_pods in (71, 75, 90, 150):
        s, p = synth.mig_cluster(12, 8, n_pods, seed=9, demand=["1g.10gb"], demand_probs=[1.0])
        st = check(eng, s, p, None, pk.FILTER_NONE, f"posteriori {n_pods}")
        assert st["lacking_at_start"] == 1
    assert st["path"] in (0, 1)


@pytest.mark.parametrize("kind", ["mig", "mps"])
def test_sweep_as_parallel_waves(eng, kind):
    """Few lacking slices, many nodes: the exact sweep runs as parallel waves (k_plan_waves) -- same plan and counters as
    the oracle's sequential sweep, including the early exit (planner.go:94-97) and demand that can never be met."""
    gen = synth.mig_cluster if kind == "mig" else synth.mps_cluster
    waves_seen = 0
    for n_nodes, n_pods, seed in ((300, 20, 1), (1000, 60, 2), (2500, 300, 3), (5000, 600, 0xC3), (700, 5, 4), (64, 3, 5)):
        for fm in (pk.FILTER_NONE, pk.FILTER_RESOURCES_FIT):
            snap, pods = gen(n_nodes, 8, n_pods, seed=seed, regime="saturating")
            st = check(eng, snap, pods, None, fm, f"{kind} {n_nodes}x{n_pods} fm={fm}")
            waves_seen += st["waves"] > 0
            assert st["path"] == 1 or st["waves"] > 0 or st["visited_nodes"] * 45 >= 0
    assert waves_seen >= 4


def test_compiled_in_limits(eng):
    """At the limits of include/nosplan.h: 16 profiles, 16 GPUs per node, 32 allowed geometries of one model, 12 non-slice
    scalar resources -- the wide code paths (4 words of byte-SIMD scoring, 32 geometry lanes per GPU) in every regime."""
    import random
    rng = random.Random(7)
    profiles = [f"{i}g.{2 * i}gb" for i in range(1, 17)]
    geoms = [{p: rng.randint(1, 7) for p in rng.sample(profiles, rng.randint(1, 4))} for _ in range(32)]
    geoms[0] = {profiles[-1]: 1}
    known = {"TEST-GPU": geoms}
    paths = set()
    for n_pods, seed in ((3000, 5), (40, 6), (300, 7)):
        snap, pods = synth.mig_cluster(200, 16, n_pods, seed=seed, model="TEST-GPU", known=known, regime="saturating",
                                       demand=profiles, demand_probs=[1 / 16] * 16)
        assert snap.n_profiles == 16 and snap.gpus_per_node == 16 and snap.geom_counts.shape == (32, 16)
        for fm in (pk.FILTER_NONE, pk.FILTER_RESOURCES_FIT):
            st = check(eng, snap, pods, None, fm, f"limits {n_pods} fm={fm}")
            paths.add((st["path"], st["waves"] > 0))
        # the same with 12 other scalar resources on nodes and on a third of the pods (exact per-pod path)
        other = [f"example.com/res-{k:02d}" for k in range(12)]
        snap.other_resources = other
        N = snap.n_nodes
        extra_a = np.full((N, 12), 50, np.int64); extra_r = np.zeros((N, 12), np.int64)
        snap.node_alloc = np.concatenate([snap.node_alloc, extra_a], 1); snap.node_requested = np.concatenate([snap.node_requested, extra_r], 1)
        K = snap.n_profiles
        ptr, col, qty = [0], [], []
        for i in range(pods.n_pods):
            col.append(int(pods.scalar_col[i])); qty.append(1)
            if i % 3 == 0:
                col.append(3 + K + (i % 12)); qty.append(1 + i % 4)
            ptr.append(len(col))
        pods.scalar_ptr = np.array(ptr, np.int32); pods.scalar_col = np.array(col, np.int32); pods.scalar_qty = np.array(qty, np.int64)
        check(eng, snap, pods, None, pk.FILTER_RESOURCES_FIT, f"limits+other {n_pods}")
    assert (1, False) in paths
