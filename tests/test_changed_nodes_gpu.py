"""nosplan_changed_nodes (CUDA, through the C-ABI): the nodes whose desired partitioning differs from the snapshot's current
one, as NodePartitioning.Equal compares them (internal/partitioning/state/partitioning.go:33-38; zero-count keys included).
Reference: the same comparison in numpy on the plan the oracle produces."""
import numpy as np
import pytest

from nos_b200 import capi, packing as pk, synth
from support import oracle_lib as ol

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def eng():
    import __graft_entry__ as g
    g.build()
    e = capi.Engine(0)
    yield e
    e.close()


def expected_changed(snap, cnt, msk):
    G, N = snap.gpus_per_node, snap.n_nodes
    cur_cnt = (snap.gpu_used.astype(np.int32) + snap.gpu_free.astype(np.int32)).astype(np.uint8)
    cur_msk = snap.gpu_used_mask | snap.gpu_free_mask
    differs = ((cur_cnt != cnt).any(1) | (cur_msk != msk)).reshape(N, G)
    live = np.arange(G)[None, :] < snap.node_gpu_count[:, None]
    return np.flatnonzero((differs & live).any(1)).astype(np.int32)


@pytest.mark.parametrize("kind,shape,regime", [("mig", (700, 8, 900), "saturating"), ("mig", (300, 8, 40), "saturating"),
                                               ("mig", (400, 8, 600), "satisfiable"), ("mps", (500, 8, 2000), "saturating"),
                                               ("mig", (333, 5, 777), "saturating")])
def test_changed_nodes_vs_numpy_on_the_oracle_plan(eng, kind, shape, regime):
    gen = synth.mig_cluster if kind == "mig" else synth.mps_cluster
    snap, pods = gen(*shape, seed=61, regime=regime)
    ocnt, omsk, ost, _ = ol.plan(snap, pods, None, pk.FILTER_RESOURCES_FIT, ol.INCREMENTAL)
    exp = expected_changed(snap, ocnt, omsk)
    cnt, msk, st = eng.plan(snap, pods, None, pk.FILTER_RESOURCES_FIT)
    assert np.array_equal(cnt, ocnt) and np.array_equal(msk, omsk)
    got = eng.changed_nodes()
    assert np.array_equal(got, exp), (len(got), len(exp))
    if regime == "saturating":
        assert 0 < len(exp)
    # the resident path leaves the same rows behind
    eng.plan_resident(pods, None, pk.FILTER_RESOURCES_FIT)
    assert np.array_equal(eng.changed_nodes(), exp)


def test_zero_count_keys_and_gpu_less_rows(eng):
    """A GPU whose count bytes stay the same but whose key set grows (a zero-valued key appears) has changed; rows beyond a
    node's gpu_count never count."""
    snap, pods = synth.mig_cluster(64, 8, 200, seed=7, regime="saturating")
    snap.node_gpu_count = np.array(snap.node_gpu_count, copy=True)
    snap.node_gpu_count[::3] = 5                           # rows 5..7 of these nodes are padding
    ocnt, omsk, _, _ = ol.plan(snap, pods, None, pk.FILTER_NONE, ol.INCREMENTAL)
    eng.plan(snap, pods, None, pk.FILTER_NONE)
    assert np.array_equal(eng.changed_nodes(), expected_changed(snap, ocnt, omsk))


def test_capacity(eng):
    snap, pods = synth.mig_cluster(200, 8, 500, seed=8, regime="saturating")
    eng.plan(snap, pods, None, pk.FILTER_RESOURCES_FIT)
    full = eng.changed_nodes()
    assert len(full) > 3
    with pytest.raises(capi.NosplanError, match="capacity") as ei:
        eng.changed_nodes(capacity=3)
    part, n = ei.value.partial
    assert n == len(full) and np.array_equal(part, full[:3])
    assert np.array_equal(eng.changed_nodes(capacity=len(full)), full)
