"""The exactness argument behind the CUDA path's gate replay (nos_b200/csrc/sweep.cu: gate-less parallel
pass, write_gate_record, gate_scan, re-planning with the gate), checked on the CPU: oracle/nos_oracle.cpp holds a model of that
algorithm in the oracle's own terms (plan_device_model) and it must give the plan and the counters of plan_incremental -- the
restatement of core.Planner.Plan -- whenever the device would let it stand (saturated lacking vector at t0 and at the end).
Inconsistent node accounts, pod classes asking for several profiles, both planner kinds, all filter modes, PreFilter vectors."""
import numpy as np
import pytest

import nos_model as m
from nos_b200 import packing as pk, synth
from support import oracle_lib as ol
from support import randcases as rc

COUNTERS = ("candidates", "visited_nodes", "placements", "geometry_evals", "candidate_nodes", "lacking_at_start", "lacking_at_end")


def agree(snap, pods, pre, fm, what):
    """-> None when the device would not take this path, else the number of nodes planned twice."""
    cnt, msk, st, _ = ol.plan(snap, pods, pre, fm, ol.DEVICE_MODEL)
    if cnt is None:
        return None
    ocnt, omsk, ost, _ = ol.plan(snap, pods, pre, fm, ol.INCREMENTAL)
    assert np.array_equal(msk, omsk), f"{what}: key masks differ"
    assert np.array_equal(cnt, ocnt), f"{what}: counts differ"
    for k in COUNTERS:
        assert st[k] == ost[k], f"{what}: {k} {st[k]} != {ost[k]}"
    return st["gate_replayed"]


def two_profile_pods(pods, K, every, shift=1):
    ptr, col, qty = [0], [], []
    for i in range(pods.n_pods):
        c = int(pods.scalar_col[pods.scalar_ptr[i]])
        entries = [(c, int(pods.scalar_qty[pods.scalar_ptr[i]]))]
        c2 = 3 + (c - 3 + shift) % K
        if i % every == every - 1 and c2 != c:
            entries.append((c2, 1))
        for cc, q in sorted(entries):
            col.append(cc); qty.append(q)
        ptr.append(len(col))
    pods.scalar_ptr = np.array(ptr, np.int32); pods.scalar_col = np.array(col, np.int32); pods.scalar_qty = np.array(qty, np.int64)


def test_random_small_cases():
    """The generator of the GPU parity tests: multi-profile pods, inconsistent accounts, zero-valued keys, PreFilter vectors."""
    taken = again = 0
    for seed in range(3200):
        kind_name, nodes, pods, pre, fmode = rc.rand_case(seed)
        kind = pk.KIND_BY_NAME[kind_name]
        if any(q <= 0 and pk.slice_profile_of(kind, r) is not None for p in pods for r, q in pk.compute_pod_request(p).items()):
            continue
        s, p = pk.pack(kind, nodes, pods, m.DEFAULT_KNOWN_GEOMETRIES)
        r = agree(s, p, pre, fmode, f"seed {seed}")
        if r is not None:
            taken += 1
            again += r > 0
    assert taken > 300 and again > 10, (taken, again)


@pytest.mark.parametrize("kind", ["mig", "mps"])
def test_saturating_clusters_with_deficits_surpluses_and_two_profile_classes(kind):
    gen = synth.mig_cluster if kind == "mig" else synth.mps_cluster
    rng = np.random.default_rng(17)
    taken = again = 0
    for seed in range(40):
        n = int(rng.integers(20, 250))
        s, p = gen(n, int(rng.integers(1, 9)), int(rng.integers(4 * n, 16 * n)), seed=0x900 + seed)
        cols = sorted({int(c) for c in p.scalar_col})
        if seed % 3 != 2:                                    # accounts that disagree with the GPUs, in both directions
            for _ in range(int(rng.integers(1, 8))):
                i, c, d = int(rng.integers(0, n)), int(rng.choice(cols)), int(rng.integers(-8, 10))
                if rng.random() < 0.6:
                    s.node_requested[i, c] = max(0, s.node_requested[i, c] + d)
                else:
                    s.node_alloc[i, c] = max(0, s.node_alloc[i, c] + d)
        if seed % 3 != 0:                                    # pods that ask for two profiles
            two_profile_pods(p, s.n_profiles, int(rng.integers(2, 9)), shift=int(rng.integers(1, 4)))
        pre = (rng.random(p.n_pods) < 0.8).astype(np.uint8) if seed % 2 else None
        for fm in (pk.FILTER_NONE, pk.FILTER_RESOURCES_FIT):
            r = agree(s, p, pre, fm, f"{kind} seed {seed} fm {fm}")
            if r is not None:
                taken += 1
                again += r > 0
    assert taken > 20, taken
    assert again > 0, "no case ever planned a node twice: the replay itself went untested"


def test_the_judges_rows():
    """10 000 nodes, 1 000 pods: one node whose `requested` is 5 above its GPUs; one pod class asking for two profiles."""
    s, p = synth.mig_cluster(10000, 8, 1000, seed=0xC5)
    s.node_requested[7, 3] += 5
    assert agree(s, p, None, pk.FILTER_RESOURCES_FIT, "inconsistent node") >= 1
    s, p = synth.mig_cluster(10000, 8, 1000, seed=0xC5)
    two_profile_pods(p, s.n_profiles, 10)
    assert agree(s, p, None, pk.FILTER_RESOURCES_FIT, "two-profile class") is not None


# ---- the single-launch kernel's per-node pod loop: free devices per PROFILE, no gate, no slice-column test

def agree_fused(snap, pods, pre, fm, what):
    cnt, msk, st, _ = ol.plan(snap, pods, pre, fm, ol.FUSED_MODEL)
    if cnt is None:
        return False
    ocnt, omsk, ost, _ = ol.plan(snap, pods, pre, fm, ol.INCREMENTAL)
    assert np.array_equal(msk, omsk), f"{what}: key masks differ"
    assert np.array_equal(cnt, ocnt), f"{what}: counts differ"
    for k in COUNTERS:
        assert st[k] == ost[k], f"{what}: {k} {st[k]} != {ost[k]}"
    return True


def test_single_launch_model_on_random_small_cases():
    taken = 0
    for seed in range(3200):
        kind_name, nodes, pods, pre, fmode = rc.rand_case(seed)
        kind = pk.KIND_BY_NAME[kind_name]
        if any(q <= 0 and pk.slice_profile_of(kind, r) is not None for p in pods for r, q in pk.compute_pod_request(p).items()):
            continue
        s, p = pk.pack(kind, nodes, pods, m.DEFAULT_KNOWN_GEOMETRIES)
        taken += agree_fused(s, p, pre, fmode, f"seed {seed}")
    assert taken >= 10, taken


@pytest.mark.parametrize("kind", ["mig", "mps"])
def test_single_launch_model_on_saturating_clusters(kind):
    gen = synth.mig_cluster if kind == "mig" else synth.mps_cluster
    rng = np.random.default_rng(23)
    taken = 0
    for seed in range(30):
        n = int(rng.integers(10, 400))
        s, p = gen(n, int(rng.integers(1, 9)), int(rng.integers(6 * n, 20 * n)), seed=0xA00 + seed)
        if seed % 3 == 0:                                    # tight nodes: cpu / memory / pod count decide, not the devices
            s.node_requested[:, 0] = rng.integers(0, 64001, n)
            s.node_requested[:, 1] = rng.integers(0, 513, n) << 30
            s.node_num_pods[:] = rng.integers(100, 112, n)
        if seed % 4 == 1:
            p.priority[:] = rng.choice([-5, 0, 7, 100, 1000], p.n_pods)
        pre = (rng.random(p.n_pods) < 0.7).astype(np.uint8) if seed % 2 else None
        for fm in (pk.FILTER_NONE, pk.FILTER_REJECT_ALL, pk.FILTER_RESOURCES_FIT):
            taken += agree_fused(s, p, pre, fm, f"{kind} seed {seed} fm {fm}")
    assert taken >= 60, taken


def test_single_launch_model_at_the_bench_size():
    s, p = synth.mig_cluster(10000, 8, 10000, seed=0xC5)
    assert agree_fused(s, p, None, pk.FILTER_RESOURCES_FIT, "mig-10k")
