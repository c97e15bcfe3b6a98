"""A snapshot that stays in HBM between Plan() calls (nosplan_plan_resident / nosplan_update_nodes): every resident call
must give what a full nosplan_plan on the same (updated) snapshot gives -- and so what the oracle gives -- bit for bit."""
import copy

import numpy as np
import pytest

from nos_b200 import capi, packing as pk, sharded, synth
from support import oracle_lib as ol

pytestmark = pytest.mark.gpu
KEYS = ("candidates", "visited_nodes", "placements", "geometry_evals", "lacking_at_start", "lacking_at_end", "candidate_nodes")


@pytest.fixture(scope="module")
def eng():
    import __graft_entry__ as g
    g.build()
    e = capi.Engine(0)
    yield e
    e.close()


def same_as_oracle(got, snap, pods, fm, pre=None):
    cnt, msk, st = got
    ocnt, omsk, ost, _ = ol.plan(snap, pods, pre, fm, ol.INCREMENTAL)
    assert np.array_equal(cnt, ocnt) and np.array_equal(msk, omsk)
    for k in KEYS:
        assert st[k] == ost[k], (k, st[k], ost[k])
    return st


@pytest.mark.parametrize("kind", ["mig", "mps"])
def test_new_pods_against_the_resident_snapshot(eng, kind):
    gen = synth.mig_cluster if kind == "mig" else synth.mps_cluster
    snap, pods = gen(600, 8, 900, seed=41, regime="saturating")
    st0 = same_as_oracle(eng.plan(snap, pods, None, pk.FILTER_RESOURCES_FIT), snap, pods, pk.FILTER_RESOURCES_FIT)
    for seed, n_pods, regime in ((42, 900, "saturating"), (43, 40, "saturating"), (44, 5000, "saturating"), (45, 700, "satisfiable")):
        _, pods2 = gen(600, 8, n_pods, seed=seed, regime=regime)         # only the pods of the new problem are used
        st = same_as_oracle(eng.plan_resident(pods2, None, pk.FILTER_RESOURCES_FIT), snap, pods2, pk.FILTER_RESOURCES_FIT)
        assert st["h2d_bytes"] < 80 * n_pods + 16384                       # the pods only: the node arrays did not travel again
    pre = (np.random.default_rng(1).random(pods.n_pods) < 0.6).astype(np.uint8)
    same_as_oracle(eng.plan_resident(pods, pre, pk.FILTER_NONE), snap, pods, pk.FILTER_NONE, pre)


def test_a_growing_pod_list_moves_the_resident_snapshot_with_the_arena():
    e = capi.Engine(0)                                                     # a fresh handle: its arena is sized by the first call
    snap, pods = synth.mig_cluster(300, 8, 50, seed=5, regime="saturating")
    e.plan(snap, pods, None, pk.FILTER_RESOURCES_FIT)
    _, big = synth.mig_cluster(300, 8, 60_000, seed=6, regime="saturating")
    same_as_oracle(e.plan_resident(big, None, pk.FILTER_RESOURCES_FIT), snap, big, pk.FILTER_RESOURCES_FIT)
    same_as_oracle(e.plan_resident(pods, None, pk.FILTER_RESOURCES_FIT), snap, pods, pk.FILTER_RESOURCES_FIT)
    e.close()


def test_update_nodes(eng):
    snap, pods = synth.mig_cluster(500, 8, 800, seed=51, regime="saturating")
    other, _ = synth.mig_cluster(500, 8, 800, seed=52, regime="saturating")   # rows to take the updates from
    eng.plan(snap, pods, None, pk.FILTER_RESOURCES_FIT)
    cur = copy.copy(snap)
    for f in sharded._PER_NODE + sharded._PER_GPU:
        setattr(cur, f, np.array(getattr(snap, f), copy=True))
    G = snap.gpus_per_node
    for lo, hi in ((0, 1), (17, 49), (499, 500), (100, 400)):
        rows = sharded.slice_snapshot(other, lo, hi)
        eng.update_nodes(lo, rows)
        for f in sharded._PER_NODE:
            getattr(cur, f)[lo:hi] = getattr(other, f)[lo:hi]
        for f in sharded._PER_GPU:
            getattr(cur, f)[lo * G:hi * G] = getattr(other, f)[lo * G:hi * G]
        same_as_oracle(eng.plan_resident(pods, None, pk.FILTER_RESOURCES_FIT), cur, pods, pk.FILTER_RESOURCES_FIT)


def test_sliced_ranks_keep_their_slices_resident():
    """Slice mode on two emulated ranks: the slices are uploaded once, then two more pod lists are planned with
    nosplan_upload_pods + the sliced protocol; rows and counters are the oracle's each time."""
    import torch
    dev = torch.device("cuda", 0)
    world = 2
    snap, pods = synth.mig_cluster(400, 8, 900, seed=71, regime="saturating")
    G = snap.gpus_per_node
    engines, planners, slices = [], [], []
    for r in range(world):
        lo, hi = sharded.shard_range(snap.n_nodes, r, world)
        e = capi.Engine(0)
        e.upload(sharded.slice_snapshot(snap, lo, hi, pad_to=(snap.n_nodes + world - 1) // world), pods, None, pk.FILTER_RESOURCES_FIT)
        engines.append(e); slices.append((lo, hi))
    for step, seed in enumerate((71, 72, 73)):
        _, pp = synth.mig_cluster(400, 8, 900 if step < 2 else 3000, seed=seed, regime="saturating")
        ocnt, omsk, ost, _ = ol.plan(snap, pp, None, pk.FILTER_RESOURCES_FIT, ol.INCREMENTAL)
        planners = []
        for r, e in enumerate(engines):
            if step:
                e.upload_pods(pp, None, pk.FILTER_RESOURCES_FIT)
            p = sharded.SlicedPlanner(e, r, world, dev, all_gather=lambda dst, src: None)
            p.prepare()
            planners.append(p)
        for e, p in zip(engines, planners):
            e.slice_begin(p.tot_local.data_ptr())
        torch.cuda.synchronize()
        tot_all = torch.cat([p.tot_local for p in planners])
        for r, (e, p) in enumerate(zip(engines, planners)):
            e.slice_mid(r, world, tot_all.data_ptr(), p.local.data_ptr(), p.nbytes)
        torch.cuda.synchronize()
        gathered = torch.cat([p.local for p in planners])
        for e, (lo, hi) in zip(engines, slices):
            e.shard_end(gathered.data_ptr(), world)
            cnt, msk, st = e.download()
            n = (hi - lo) * G
            assert np.array_equal(cnt[:n], ocnt[lo * G:hi * G]) and np.array_equal(msk[:n], omsk[lo * G:hi * G]), step
            for k in ("candidates", "visited_nodes", "placements", "geometry_evals", "lacking_at_end"):
                assert st[k] == ost[k], (step, k)
    for e in engines:
        e.close()


def test_errors(eng):
    fresh = capi.Engine(0)
    snap, pods = synth.mig_cluster(20, 4, 30, seed=3)
    with pytest.raises(capi.NosplanError, match="no snapshot resident"):
        fresh._snap = snap
        fresh.plan_resident(pods, None, 0)
    fresh.plan(snap, pods, None, 0)
    with pytest.raises(capi.NosplanError, match="outside the resident snapshot"):
        fresh.update_nodes(15, sharded.slice_snapshot(snap, 0, 10))
    bad = sharded.slice_snapshot(snap, 0, 2)
    bad.gpu_model = np.array(bad.gpu_model, copy=True)
    bad.gpu_model[0] = 77
    with pytest.raises(capi.NosplanError, match="node 3: GPU model id out of range"):
        fresh.update_nodes(3, bad)
    mps, _ = synth.mps_cluster(20, 4, 30, seed=3)
    with pytest.raises(capi.NosplanError, match="another shape"):
        fresh.update_nodes(0, sharded.slice_snapshot(mps, 0, 2))
    same_as_oracle(fresh.plan_resident(pods, None, 0), snap, pods, 0)      # the handle and its snapshot are intact
    fresh.close()
