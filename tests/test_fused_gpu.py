"""The single-launch Plan() (k_plan_fused, nos_b200/csrc/fused.cu): one kernel that speculates on the lacking set and
verifies itself.  Where it stands its plan and counters must be the oracle's bit for bit (and the classic kernel sequence's);
where its hypothesis fails it must say so and the classic sequence must produce the plan."""
import os

import numpy as np
import pytest

from nos_b200 import capi, packing as pk, synth
from support import oracle_lib as ol

pytestmark = pytest.mark.gpu

COUNTERS = ("candidates", "visited_nodes", "placements", "geometry_evals", "candidate_nodes", "lacking_at_start", "lacking_at_end")
REJ_COUNTS, REJ_REM, REJ_HYPOTHESIS, REJ_ACCOUNTS = 1, 2, 4, 8


@pytest.fixture(scope="module")
def eng():
    import __graft_entry__ as g
    g.build()
    e = capi.Engine(0)
    yield e
    e.close()


def classic(eng, snap, pods, pre, fm):
    os.environ["NOSPLAN_NO_FUSED"] = "1"
    try:
        return eng.plan(snap, pods, pre, fm)
    finally:
        del os.environ["NOSPLAN_NO_FUSED"]


def check(eng, snap, pods, pre, fm, what=""):
    cnt, msk, st = eng.plan(snap, pods, pre, fm)
    ocnt, omsk, ost, _ = ol.plan(snap, pods, pre, fm, ol.INCREMENTAL)
    assert np.array_equal(msk, omsk), f"{what}: key masks differ"
    assert np.array_equal(cnt, ocnt), f"{what}: counts differ"
    for k in COUNTERS:
        assert st[k] == ost[k], f"{what}: {k} {st[k]} != {ost[k]}"
    return cnt, msk, st


@pytest.mark.parametrize("fm", [pk.FILTER_NONE, pk.FILTER_REJECT_ALL, pk.FILTER_RESOURCES_FIT], ids=["none", "reject", "fit"])
@pytest.mark.parametrize("shape", [(1, 1, 200), (16, 8, 700), (17, 3, 900), (200, 8, 1000), (1000, 8, 5000), (333, 5, 7777)])
def test_fused_stands_on_saturating_demand(eng, shape, fm):
    s, p = synth.mig_cluster(*shape, seed=0xF0 + shape[0])
    cnt, msk, st = check(eng, s, p, None, fm, f"{shape}")
    assert st["fused"] == 1 and st["fused_reject"] == 0 and st["path"] == 1 and st["kernel_launches"] == 1
    ccnt, cmsk, cst = classic(eng, s, p, None, fm)
    assert cst["fused"] == 0 and np.array_equal(cnt, ccnt) and np.array_equal(msk, cmsk)
    for k in COUNTERS:
        assert st[k] == cst[k]


def test_fused_with_prefilter_vector_quantities_and_slice_less_pods(eng):
    rng = np.random.default_rng(11)
    s, p = synth.mig_cluster(300, 8, 4000, seed=31)
    pre = (rng.random(4000) < 0.6).astype(np.uint8)
    _, _, st = check(eng, s, p, pre, pk.FILTER_RESOURCES_FIT, "prefilter vector")
    assert st["fused"] == 1
    # the pods of one profile ask for two devices each, the pods of one priority level for no slice at all
    p.scalar_qty[p.scalar_col == p.scalar_col[0]] = 2
    ptr, col, qty = [0], [], []
    for i in range(p.n_pods):
        if p.priority[i] != 100:
            col.append(int(p.scalar_col[i])); qty.append(int(p.scalar_qty[i]))
        ptr.append(len(col))
    p.scalar_ptr = np.array(ptr, np.int32); p.scalar_col = np.array(col, np.int32); p.scalar_qty = np.array(qty, np.int64)
    for fm in (pk.FILTER_NONE, pk.FILTER_RESOURCES_FIT):
        _, _, st = check(eng, s, p, pre, fm, f"quantities + slice-less fm={fm}")
        assert st["fused"] == 1 or st["fused_reject"] == 0
    # tight nodes: little cpu and memory left, pod counts at the limit on some
    p.scalar_qty[:] = 1
    s.node_requested[:, 0] = rng.integers(0, 64001, s.n_nodes)
    s.node_requested[:, 1] = rng.integers(0, 513, s.n_nodes) << 30
    s.node_num_pods[:] = rng.integers(100, 112, s.n_nodes)
    _, _, st = check(eng, s, p, pre, pk.FILTER_RESOURCES_FIT, "tight nodes")
    assert st["fused"] == 1


def test_fused_units_follow_the_requests(eng):
    """cpu / memory requests that share no power of two (unit 1), and byte-granular memory on small nodes."""
    rng = np.random.default_rng(5)
    s, p = synth.mig_cluster(150, 8, 2000, seed=41)
    p.base[:, 0] = rng.integers(1, 4001, p.n_pods)
    p.base[:, 1] = rng.integers(1, 1 << 27, p.n_pods)                # odd byte counts, below the int32 cost limit
    s.node_alloc[:, 1] = 1 << 30
    _, _, st = check(eng, s, p, None, pk.FILTER_RESOURCES_FIT, "unit 1")
    assert st["fused"] == 1
    s.node_alloc[:, 1] = 512 << 30                                   # now the nodes do not fit the unit: rejected, classic plans
    _, _, st = check(eng, s, p, None, pk.FILTER_RESOURCES_FIT, "unit 1, large nodes")
    assert st["fused"] == 0 and st["fused_reject"] & REJ_REM


def test_fused_rejects_itself(eng):
    # nothing (or not everything) lacking: the hypothesis on the lacking set fails
    s, p = synth.mig_cluster(200, 8, 1000, seed=5, regime="satisfiable")
    _, _, st = check(eng, s, p, None, pk.FILTER_RESOURCES_FIT, "satisfiable")
    assert st["fused"] == 0 and st["fused_reject"] & REJ_HYPOTHESIS
    # saturation lost on the way (test_saturation_hypothesis_fails_a_posteriori of test_plan_gpu.py)
    rejected = 0
    for n_pods in (71, 75, 90, 150, 400):
        s, p = synth.mig_cluster(12, 8, n_pods, seed=9, demand=["1g.10gb"], demand_probs=[1.0])
        _, _, st = check(eng, s, p, None, pk.FILTER_NONE, f"posteriori {n_pods}")
        rejected += st["fused_reject"] != 0
    assert rejected >= 3
    # a node whose accounts disagree with its GPUs
    s, p = synth.mig_cluster(200, 8, 1000, seed=5)
    s.node_requested[7, 3] += 5
    _, _, st = check(eng, s, p, None, pk.FILTER_RESOURCES_FIT, "inconsistent")
    assert st["fused"] == 0 and st["fused_reject"] & REJ_ACCOUNTS and st["path"] == 1 and st["gate_replayed"] >= 0
    # counts above 31 on a GPU
    s, p = synth.mig_cluster(50, 8, 1000, seed=6)
    s.gpu_used[3, 6] = 40; s.gpu_used_mask[3] |= 1 << 6
    s.node_alloc[0, 3 + 6] += 40; s.node_requested[0, 3 + 6] += 40
    _, _, st = check(eng, s, p, None, pk.FILTER_RESOURCES_FIT, "large count")
    assert st["fused"] == 0 and st["fused_reject"] & REJ_COUNTS


def test_not_eligible_goes_straight_to_the_classic_sequence(eng):
    s, p = synth.mig_cluster(100, 8, 600, seed=12)
    p.tracker_key[5] = p.tracker_key[6]                               # two pods share a tracker entry
    _, _, st = check(eng, s, p, None, pk.FILTER_RESOURCES_FIT, "duplicate keys")
    assert st["fused"] == 0 and st["fused_reject"] == 0
    s, p = synth.mps_cluster(100, 8, 600, seed=12)
    p.scalar_qty[:] = 2                                               # two devices per pod: the first-fit GPU choice shows in the plan
    _, _, st = check(eng, s, p, None, pk.FILTER_RESOURCES_FIT, "mps, quantity 2")
    assert st["fused"] == 0 and st["fused_reject"] == 0
    s, p = synth.mig_cluster(500, 8, 40, seed=12)                     # 40 pods cannot lack more than (G+2)*M_p devices of anything
    _, _, st = check(eng, s, p, None, pk.FILTER_RESOURCES_FIT, "too few pods to saturate")
    assert st["fused"] == 0 and st["fused_reject"] == 0
    s, p = synth.mig_cluster(100, 8, 600, seed=12)
    p.base[3, 2] = 1 << 20                                            # an ephemeral-storage request
    _, _, st = check(eng, s, p, None, pk.FILTER_RESOURCES_FIT, "ephemeral")
    assert st["fused"] == 0 and st["fused_reject"] == 0


def test_fused_and_classic_alternate_on_one_handle(eng):
    """The position-indexed bitmap of the fused kernel and the key-indexed one of the classic kernels share memory: runs
    of both kinds on one uploaded problem, in any order, give the same plan."""
    s, p = synth.mig_cluster(400, 8, 3000, seed=77)
    ocnt, omsk, ost, _ = ol.plan(s, p, None, pk.FILTER_RESOURCES_FIT, ol.INCREMENTAL)
    eng.upload(s, p, None, pk.FILTER_RESOURCES_FIT)
    for k in range(6):
        eng.run()
        cnt, msk, st = eng.download()
        assert st["fused"] == 1 and np.array_equal(cnt, ocnt) and np.array_equal(msk, omsk) and st["placements"] == ost["placements"]
        assert st["lacking_at_end"] == ost["lacking_at_end"]
    pods2 = synth.mig_cluster(400, 8, 2500, seed=78)[1]
    cnt, msk, st = eng.plan(s, p, None, pk.FILTER_RESOURCES_FIT)
    assert st["fused"] == 1
    ccnt, cmsk, cst = classic(eng, s, p, None, pk.FILTER_RESOURCES_FIT)
    rcnt, rmsk, rst = eng.plan_resident(pods2, None, pk.FILTER_RESOURCES_FIT)          # fused again, after a classic run
    o2 = ol.plan(s, pods2, None, pk.FILTER_RESOURCES_FIT, ol.INCREMENTAL)
    assert rst["fused"] == 1 and np.array_equal(rcnt, o2[0]) and np.array_equal(rmsk, o2[1]) and rst["lacking_at_end"] == o2[2]["lacking_at_end"]


def test_fused_at_the_compiled_in_profile_limit(eng):
    """16 profiles (the wide code path), 8 GPUs, 32 geometries of one model."""
    import random
    rng = random.Random(7)
    profiles = [f"{i}g.{2 * i}gb" for i in range(1, 17)]
    geoms = [{p: rng.randint(1, 7) for p in rng.sample(profiles, rng.randint(1, 4))} for _ in range(32)]
    geoms[0] = {profiles[-1]: 1}
    snap, pods = synth.mig_cluster(300, 8, 6000, seed=5, model="TEST-GPU", known={"TEST-GPU": geoms}, regime="saturating",
                                   demand=profiles, demand_probs=[1 / 16] * 16)
    assert snap.n_profiles == 16
    for fm in (pk.FILTER_NONE, pk.FILTER_RESOURCES_FIT):
        _, _, st = check(eng, snap, pods, None, fm, f"wide fm={fm}")
        assert st["fused"] == 1


def test_fused_with_a_pod_list_larger_than_shared_memory(eng):
    """40 000 pods: the pod pairs (320 KB) stay in global memory, only the minima are staged in shared memory; 150 000 pods:
    the per-8 minima stay in global memory too."""
    for n_pods in (40000, 150000):
        s, p = synth.mig_cluster(300, 8, n_pods, seed=91)
        _, _, st = check(eng, s, p, None, pk.FILTER_RESOURCES_FIT, f"{n_pods} pods")
        assert st["fused"] == 1


def test_fast_and_general_pod_paths_agree(eng):
    """nosplan_plan stages the pods for the single-launch kernel in two passes when it can; NOSPLAN_NO_FAST_PODS sends them
    through the general path (NewPodSorter as a sort, runs, blob).  Same plan, same counters, whatever the pods look like;
    and a plan rejected after the fast path is re-planned through the general path inside the same call."""
    rng = np.random.default_rng(17)
    for seed, n_pods in ((1, 3000), (2, 1200), (3, 9000)):
        s, p = synth.mig_cluster(250, 8, n_pods, seed=100 + seed)
        p.priority[:] = rng.choice([-5, 0, 7, 100, 1000, 2_000_000_000], n_pods)
        pre = (rng.random(n_pods) < 0.8).astype(np.uint8)
        cnt, msk, st = check(eng, s, p, pre, pk.FILTER_RESOURCES_FIT, f"fast path seed {seed}")
        assert st["fused"] == 1
        os.environ["NOSPLAN_NO_FAST_PODS"] = "1"
        try:
            gcnt, gmsk, gst = eng.plan(s, p, pre, pk.FILTER_RESOURCES_FIT)
        finally:
            del os.environ["NOSPLAN_NO_FAST_PODS"]
        assert gst["fused"] == 1 and np.array_equal(cnt, gcnt) and np.array_equal(msk, gmsk)
        for k in COUNTERS:
            assert st[k] == gst[k]
        assert st["h2d_bytes"] <= gst["h2d_bytes"]
    # rejected after the fast path: satisfiable demand, and resident calls on top of it
    s, p = synth.mig_cluster(200, 8, 1000, seed=5, regime="satisfiable")
    _, _, st = check(eng, s, p, None, pk.FILTER_RESOURCES_FIT, "satisfiable after fast path")
    assert st["fused"] == 0 and st["fused_reject"] & REJ_HYPOTHESIS
    p2 = synth.mig_cluster(200, 8, 700, seed=6, regime="satisfiable")[1]
    rcnt, rmsk, rst = eng.plan_resident(p2, None, pk.FILTER_RESOURCES_FIT)
    o = ol.plan(s, p2, None, pk.FILTER_RESOURCES_FIT, ol.INCREMENTAL)
    assert np.array_equal(rcnt, o[0]) and np.array_equal(rmsk, o[1]) and rst["placements"] == o[2]["placements"]


def test_fused_north_star_size(eng):
    s, p = synth.mig_cluster(10000, 8, 10000, seed=0xC5)
    _, _, st = check(eng, s, p, None, pk.FILTER_RESOURCES_FIT, "10k x 10k")
    assert st["fused"] == 1 and st["kernel_launches"] == 1


@pytest.mark.parametrize("fm", [pk.FILTER_NONE, pk.FILTER_REJECT_ALL, pk.FILTER_RESOURCES_FIT], ids=["none", "reject", "fit"])
@pytest.mark.parametrize("shape", [(1, 1, 200), (16, 8, 700), (17, 3, 900), (200, 8, 1000), (1000, 8, 5000), (5000, 8, 20000)])
def test_fused_mps(eng, shape, fm):
    """MPS: slicing.GPU.UpdateGeometryFor (pkg/gpu/slicing/gpu.go:162-232) per GPU lane inside the single-launch kernel --
    spare capacity, the delete-all-free-keys quirk, all-or-nothing restore -- against the oracle and the classic kernels."""
    s, p = synth.mps_cluster(*shape, seed=0x30 + shape[0])
    cnt, msk, st = check(eng, s, p, None, fm, f"mps {shape}")
    assert st["fused"] == 1 and st["kernel_launches"] == 1, st
    ccnt, cmsk, cst = classic(eng, s, p, None, fm)
    assert cst["fused"] == 0 and np.array_equal(cnt, ccnt) and np.array_equal(msk, cmsk)
    for k in COUNTERS:
        assert st[k] == cst[k]


def test_fused_mps_odd_shapes(eng):
    """GPUs of different sizes, free slices of several sizes in place, a PreFilter vector, tight nodes, satisfiable demand."""
    rng = np.random.default_rng(3)
    for seed in range(6):
        n = int(rng.integers(30, 400))
        s, p = synth.mps_cluster(n, int(rng.integers(1, 9)), int(rng.integers(200, 4000)), seed=0x60 + seed,
                                 gpu_memory_mb=int(rng.choice([24576, 40960, 81920])))
        s.gpu_memory_gb[rng.random(s.gpu_memory_gb.shape) < 0.3] = 48
        pre = (rng.random(p.n_pods) < 0.7).astype(np.uint8)
        s.node_requested[:, 0] = rng.integers(0, 64001, s.n_nodes)
        for fm in (pk.FILTER_NONE, pk.FILTER_RESOURCES_FIT):
            check(eng, s, p, pre, fm, f"mps odd {seed} fm {fm}")
    s, p = synth.mps_cluster(300, 8, 1500, seed=5, regime="satisfiable")
    _, _, st = check(eng, s, p, None, pk.FILTER_RESOURCES_FIT, "mps satisfiable")
    s, p = synth.mps_cluster(300, 8, 1500, seed=5)
    p2 = synth.mps_cluster(300, 8, 900, seed=6)[1]
    check(eng, s, p, None, pk.FILTER_RESOURCES_FIT, "mps, then resident")
    rcnt, rmsk, rst = eng.plan_resident(p2, None, pk.FILTER_RESOURCES_FIT)
    o = ol.plan(s, p2, None, pk.FILTER_RESOURCES_FIT, ol.INCREMENTAL)
    assert np.array_equal(rcnt, o[0]) and np.array_equal(rmsk, o[1]) and rst["placements"] == o[2]["placements"]
