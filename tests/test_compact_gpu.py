"""nosplan_plan_compact: the node side as one array of records (128 + 24*G bytes per node), expanded on the GPU.  Same plans
and counters as nosplan_plan and the oracle; the snapshot stays resident afterwards."""
import numpy as np
import pytest

from nos_b200 import capi, packing as pk, synth
from support import oracle_lib as ol

pytestmark = pytest.mark.gpu
COUNTERS = ("candidates", "visited_nodes", "placements", "geometry_evals", "candidate_nodes", "lacking_at_start", "lacking_at_end")


@pytest.fixture(scope="module")
def eng():
    import __graft_entry__ as g
    g.build()
    e = capi.Engine(0)
    yield e
    e.close()


def check(eng, snap, pods, pre, fm, what):
    cnt, msk, st = eng.plan_compact(snap, pods, pre, fm)
    ocnt, omsk, ost, _ = ol.plan(snap, pods, pre, fm, ol.INCREMENTAL)
    assert np.array_equal(msk, omsk) and np.array_equal(cnt, ocnt), what
    for k in COUNTERS:
        assert st[k] == ost[k], f"{what}: {k}"
    return st


@pytest.mark.parametrize("kind", ["mig", "mps"])
def test_compact_matches_the_oracle(eng, kind):
    gen = synth.mig_cluster if kind == "mig" else synth.mps_cluster
    for shape, regime in (((1, 1, 10), "saturating"), ((300, 8, 2000), "saturating"), ((77, 5, 900), "satisfiable"), ((2000, 8, 40), "saturating")):
        snap, pods = gen(*shape, seed=0xCA + shape[0], regime=regime)
        for fm in (pk.FILTER_NONE, pk.FILTER_RESOURCES_FIT):
            st = check(eng, snap, pods, None, fm, f"{kind} {shape} {regime} fm={fm}")
    assert st["h2d_bytes"] > 0


def test_compact_with_other_scalars_short_nodes_and_resident_calls(eng):
    rng = np.random.default_rng(3)
    snap, pods = synth.mig_cluster(120, 8, 800, seed=7)
    snap.node_gpu_count[::7] = 5                                          # nodes with fewer GPUs than the row stride
    for i in range(0, 120, 7):
        snap.gpu_used[i * 8 + 5:i * 8 + 8] = 0; snap.gpu_free[i * 8 + 5:i * 8 + 8] = 0
        snap.gpu_used_mask[i * 8 + 5:i * 8 + 8] = 0; snap.gpu_free_mask[i * 8 + 5:i * 8 + 8] = 0
    K = snap.n_profiles
    tot = (snap.gpu_used.astype(np.int64) + snap.gpu_free).reshape(120, 8, 16).sum(1)
    snap.node_alloc[:, 3:3 + K] = tot[:, :K]
    snap.node_requested[:, 3:3 + K] = snap.gpu_used.astype(np.int64).reshape(120, 8, 16).sum(1)[:, :K]
    snap.other_resources = ["example.com/a", "example.com/b"]
    snap.node_alloc = np.concatenate([snap.node_alloc, rng.integers(0, 9, (120, 2))], 1).astype(np.int64)
    snap.node_requested = np.concatenate([snap.node_requested, rng.integers(0, 3, (120, 2))], 1).astype(np.int64)
    check(eng, snap, pods, (rng.random(800) < 0.7).astype(np.uint8), pk.FILTER_RESOURCES_FIT, "other scalars, short nodes")
    pods2 = synth.mig_cluster(120, 8, 500, seed=8)[1]
    cnt, msk, st = eng.plan_resident(pods2, None, pk.FILTER_RESOURCES_FIT)     # the expanded snapshot is resident
    o = ol.plan(snap, pods2, None, pk.FILTER_RESOURCES_FIT, ol.INCREMENTAL)
    assert np.array_equal(cnt, o[0]) and np.array_equal(msk, o[1]) and st["placements"] == o[2]["placements"]


def test_compact_limits_and_errors(eng):
    import random
    rng = random.Random(7)
    profiles = [f"{i}g.{2 * i}gb" for i in range(1, 17)]
    geoms = [{p: rng.randint(1, 7) for p in rng.sample(profiles, rng.randint(1, 4))} for _ in range(32)]
    geoms[0] = {profiles[-1]: 1}
    snap, pods = synth.mig_cluster(10, 8, 50, seed=5, model="TEST-GPU", known={"TEST-GPU": geoms}, demand=profiles, demand_probs=[1 / 16] * 16)
    assert snap.n_profiles == 16
    with pytest.raises(ValueError):
        capi.compact_records(snap)                                        # 16 profiles
    snap, pods = synth.mig_cluster(30, 8, 100, seed=5)
    rec = capi.compact_records(snap)
    rec[3, 128 + 24 * 2 + 18] = 200                                       # model id out of range on node 3, GPU 2
    with pytest.raises(capi.NosplanError, match="node 3: GPU model id out of range"):
        eng.plan_compact(snap, pods, None, 0, records=rec)
    check(eng, snap, pods, None, pk.FILTER_RESOURCES_FIT, "after an error")
