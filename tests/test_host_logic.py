"""Host-side logic of libnosplan that needs no GPU: the pod order (NewPodSorter) on both of its paths -- counting sort
where the comparator is a strict weak order, Go's stable sort restated otherwise -- against the oracle's sorter."""
import numpy as np
import pytest

import nos_model as m
from nos_b200 import capi, packing as pk, synth
from support import oracle_lib as ol
from support import randcases as rc


@pytest.fixture(scope="module", autouse=True)
def built():
    import __graft_entry__ as g
    g.build()


def test_pod_order_random_small():
    n_nontrivial = 0
    for seed in range(600):
        kind_name, nodes, pods, _, _ = rc.rand_case(seed, max_pods=80)
        s, p = pk.pack(pk.KIND_BY_NAME[kind_name], nodes, pods, m.DEFAULT_KNOWN_GEOMETRIES)
        got = capi.host_pod_order(s, p)
        exp = ol.sort_pods(s, p)
        assert np.array_equal(got, exp), f"seed {seed}"
        n_nontrivial += int(not np.array_equal(exp, np.arange(p.n_pods)))
    assert n_nontrivial > 100


def test_pod_order_mixed_mig_families_is_not_a_strict_weak_order():
    """4g.24gb vs 3g.40gb: SmallerThan holds both ways (pkg/gpu/mig/profile.go:84-96) -> the general path must run and must
    still agree with the restated Go algorithm."""
    rng = np.random.default_rng(1)
    profs = ["1g.6gb", "4g.24gb", "3g.40gb", "2g.20gb", "7g.79gb"]
    pods = [{"namespace": "ns", "name": f"p{i}", "priority": int(rng.choice([0, 0, 5])),
             "containers": [{pk.MIG_PREFIX + str(rng.choice(profs)): 1}] if rng.random() < 0.9 else []} for i in range(500)]
    s, p = pk.pack(pk.KIND_MIG, [], pods, m.DEFAULT_KNOWN_GEOMETRIES)
    assert np.array_equal(capi.host_pod_order(s, p), ol.sort_pods(s, p))


def test_pod_order_full_size_fast_path():
    s, p = synth.mig_cluster(10, 8, 20000, seed=3)
    got = capi.host_pod_order(s, p)
    assert np.array_equal(got, ol.sort_pods(s, p))
    assert sorted(got.tolist()) == list(range(20000))


def test_slice_snapshot_partitions_and_pads():
    from nos_b200 import sharded, synth
    snap, _ = synth.mig_cluster(23, 4, 10, seed=3)
    G = snap.gpus_per_node
    chunk = (23 + 3) // 4
    seen = 0
    for r in range(4):
        lo, hi = sharded.shard_range(23, r, 4)
        sl = sharded.slice_snapshot(snap, lo, hi, pad_to=chunk)
        assert sl.n_nodes == chunk and sl.gpu_used.shape == (chunk * G, 16) and sl.node_alloc.shape == (chunk, snap.n_res)
        assert np.array_equal(sl.gpu_free[:(hi - lo) * G], snap.gpu_free[lo * G:hi * G])
        assert np.array_equal(sl.node_requested[:hi - lo], snap.node_requested[lo:hi])
        assert not sl.node_gpu_count[hi - lo:].any() and not sl.node_alloc[hi - lo:].any()      # padding: GPU-less, resource-less
        assert sl.profiles == snap.profiles and sl.geom_counts is snap.geom_counts
        seen += hi - lo
    assert seen == 23
