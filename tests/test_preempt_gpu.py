"""nosplan_preempt_dryrun (CUDA, through the C-ABI) against the packed oracle of the whole SelectVictimsOnNode
(capacity_scheduling.go:468-675): victims, per-node status and numViolatingVictim, bit-exact."""
import collections

import numpy as np
import pytest

from nos_b200 import capi, preempt_packing as pp, synth, victim_packing as vp
from support import oracle_lib as ol
from support import preemptcases as pc

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def eng():
    import __graft_entry__ as g
    g.build()
    e = capi.Engine(0)
    yield e
    e.close()


def engine_used_fn(eng):
    def used(pq, run):
        r, _ = eng.quota_recompute(pq, run)
        return r["used"], r["used_mask"]
    return used


@pytest.mark.parametrize("block", range(4))
def test_random_vs_oracle(eng, block):
    seen = collections.Counter()
    for seed in range(block * 250, (block + 1) * 250):
        case = pc.rand_preempt_case(seed)
        pq, query = pc.packed_dryrun_query(*case, engine_used_fn(eng))
        gv, gs, gn, st = eng.preempt_dryrun(pq, query)
        ev, es, en = ol.preempt_dryrun(pq, query)
        assert np.array_equal(gs, es), f"seed {seed} status\n{gs}\n{es}"
        assert np.array_equal(gv, ev), f"seed {seed} victims\n{gv}\n{ev}"
        assert np.array_equal(gn, en), f"seed {seed} numViolating\n{gn}\n{en}"
        assert st["kernel_launches"] == 2
        seen.update(int(x) for x in es)
        seen["victims"] += int(ev.sum())
    assert seen[0] > 20 and seen[1] > 20 and seen[2] > 5 and seen["victims"] > 10, seen


def c4_query(eng, n_nodes=10_000, n_pdbs=16, seed=4, fill=None):
    """100,000 running pods of 500 quotas over 10,000 nodes with node accounts and PodDisruptionBudget matches.  `fill`: scale
    every Min column so that the cluster-wide Used is that fraction of the aggregated Min (popular namespaces borrow, the
    others lend) instead of the recipe's heavily over-subscribed state; Max = 2 * Min as in the recipe."""
    pq, run, _ = synth.quota_state(500, 100_000, 0, seed=0xC4)
    r, _ = eng.quota_recompute(pq, run)
    if fill is not None:
        live = np.flatnonzero(pq.min.sum(0) > 0)
        rows = pq.ns_quota[pq.ns_quota >= 0]               # the aggregates count a composite quota once per namespace
        used_tot, min_tot = r["used"][rows].sum(0)[live], pq.min[rows].sum(0)[live]
        pq.min[:, live] = pq.min[:, live] * (used_tot // (fill * min_tot).astype(np.int64) + 1)
        pq.max[:, live] = 2 * pq.min[:, live]
        r, _ = eng.quota_recompute(pq, run)
    pq.used_init, pq.used_mask = r["used"].copy(), r["used_mask"].copy()
    rng = np.random.default_rng(seed)
    node = rng.integers(0, n_nodes, run.n_pods)
    order = np.lexsort((run.priority, node))              # by node, least important (lowest priority) first
    ptr = np.zeros(n_nodes + 1, np.int32)
    ptr[1:] = np.cumsum(np.bincount(node, minlength=n_nodes))
    req = run.request[order]
    R = pq.n_res
    node_req = np.zeros((n_nodes, R), np.int64)
    np.add.at(node_req, node[order], req)
    node_req[:, pq.gpu_memory_col] = 0
    node_req[:, 3] = np.diff(ptr)
    # a third of the nodes exactly full, a third with room for a few more pods, a third roomy
    kind = rng.integers(0, 3, n_nodes)
    alloc = node_req.copy()
    alloc[kind == 1] += np.array([2000, 8 << 30, 0, 2, 0, 1], np.int64)[:R]
    alloc[kind == 2] = alloc[kind == 2] * 4 + np.array([64000, 256 << 30, 0, 110, 0, 8], np.int64)[:R]
    matches = rng.random((run.n_pods, n_pdbs)) < 0.08
    pdb_ptr = np.zeros(run.n_pods + 1, np.int32)
    pdb_ptr[1:] = np.cumsum(matches.sum(1))
    pod_pdb = np.nonzero(matches)[1].astype(np.int32)
    pdb_allowed = rng.integers(0, 2, n_pdbs).astype(np.int32)   # per node: the budgets are counted afresh for every node
    return pq, run, r, order, ptr, req, node_req, alloc, pdb_allowed, pdb_ptr, pod_pdb


@pytest.mark.parametrize("fill", [None, 0.98])
def test_config_c4_size(eng, fill):
    pq, run, r, order, ptr, req, node_req, alloc, pdb_allowed, pdb_ptr, pod_pdb = c4_query(eng, fill=fill)
    seen = collections.Counter()
    popular = int(np.bincount(run.ns).argmax())
    for pre_ns, pre_prio in ((int(run.ns[0]), 1000), (int(run.ns[17]), 0), (len(pq.namespaces) - 1, 100), (popular, 1000), (popular, 100)):
        nom = np.zeros(pq.n_res, np.int64)
        nom[0], nom[1], nom[5] = 4000, 8 << 30, 1
        nom[pq.gpu_memory_col] = 32
        fit = nom.copy()
        fit[pq.gpu_memory_col] = 0
        nominated = nom.copy()
        nominated[0] += 12000
        v = vp.PackedVictimQuery(pre_ns, pre_prio, nom, ptr, run.ns[order], run.priority[order], r["over_quota"][order],
                                 req, run.request_mask[order])
        q = pp.PackedPreemptQuery(v, nom.copy(), nominated, fit, 1 << 5, alloc, node_req, pdb_allowed, pdb_ptr, pod_pdb)
        gv, gs, gn, st = eng.preempt_dryrun(pq, q)
        ev, es, en = ol.preempt_dryrun(pq, q)
        assert np.array_equal(gs, es) and np.array_equal(gv, ev) and np.array_equal(gn, en)
        # properties that hold whatever the inputs: victims only on nodes that returned Success, violating <= victims
        per_node = np.add.reduceat(np.append(gv, 0).astype(np.int64), ptr[:-1])[: len(ptr) - 1] * (np.diff(ptr) > 0)
        assert np.all(per_node[gs != pp.VICTIMS_OK] == 0) and np.all(gn <= per_node)
        seen.update(int(x) for x in gs)
        seen["victims"] += int(gv.sum())
    assert gv.shape == (100_000,) and seen[0] > 0 and seen["victims"] > 0, seen


def test_limits_and_errors(eng):
    case = pc.rand_preempt_case(3)
    pq, query = pc.packed_dryrun_query(*case, engine_used_fn(eng))
    query.pdb_allowed = np.zeros(65, np.int32)
    with pytest.raises(capi.NosplanError, match="NOSPLAN_PREEMPT_MAX_PDBS"):
        eng.preempt_dryrun(pq, query)
    pq, query = pc.packed_dryrun_query(*case, engine_used_fn(eng))
    if len(query.pod_pdb):
        query.pod_pdb = query.pod_pdb.copy()
        query.pod_pdb[0] = len(query.pdb_allowed)
        with pytest.raises(capi.NosplanError, match="outside"):
            eng.preempt_dryrun(pq, query)


def test_empty_query(eng):
    pq, run, _ = synth.quota_state(4, 10, 0, seed=1)
    z = np.zeros(pq.n_res, np.int64)
    v = vp.PackedVictimQuery(0, 0, z, np.zeros(1, np.int32), np.zeros(0, np.int32), np.zeros(0, np.int32), np.zeros(0, np.uint8),
                             np.zeros((0, pq.n_res), np.int64), np.zeros(0, np.uint32))
    q = pp.PackedPreemptQuery(v, z, z, z, 0, np.zeros((0, pq.n_res), np.int64), np.zeros((0, pq.n_res), np.int64),
                              np.zeros(0, np.int32), np.zeros(1, np.int32), np.zeros(0, np.int32))
    gv, gs, gn, _ = eng.preempt_dryrun(pq, q)
    assert gv.size == 0 and gs.size == 0 and gn.size == 0


def test_next_preemptors_against_the_resident_cluster_state(eng):
    """nosplan_preempt_dryrun_again: several preemptors in a row against the quotas / nodes / pods one full call uploaded, other
    entry points of the handle used in between; every answer is the oracle's for that preemptor on the full query."""
    pq, run, r, order, ptr, req, node_req, alloc, pdb_allowed, pdb_ptr, pod_pdb = c4_query(eng, n_nodes=2000, fill=0.98)
    popular = int(np.bincount(run.ns).argmax())

    def query(pre_ns, pre_prio, cpu, extra):
        nom = np.zeros(pq.n_res, np.int64)
        nom[0], nom[1], nom[5] = cpu, 8 << 30, 1
        nom[pq.gpu_memory_col] = 32
        fit = nom.copy()
        fit[pq.gpu_memory_col] = 0
        nominated = nom.copy()
        nominated[0] += extra
        v = vp.PackedVictimQuery(pre_ns, pre_prio, nom, ptr, run.ns[order], run.priority[order], r["over_quota"][order],
                                 req, run.request_mask[order])
        return pp.PackedPreemptQuery(v, nom.copy(), nominated, fit, 1 << 5, alloc, node_req, pdb_allowed, pdb_ptr, pod_pdb)

    first = query(popular, 1000, 4000, 12000)
    gv, gs, gn, st0 = eng.preempt_dryrun(pq, first)
    ev, es, en = ol.preempt_dryrun(pq, first)
    assert np.array_equal(gv, ev) and np.array_equal(gs, es) and np.array_equal(gn, en)
    seen = collections.Counter()
    for k, (pre_ns, pre_prio, cpu, extra) in enumerate(((popular, 100, 500, 0), (int(run.ns[3]), 1000, 64000, 0),
                                                         (len(pq.namespaces) - 1, 100, 4000, 10 ** 9), (popular, 1000, 4000, 12000))):
        if k == 2:
            eng.quota_recompute(pq, run)                     # another entry point of the handle must not disturb the resident state
        q = query(pre_ns, pre_prio, cpu, extra)
        gv, gs, gn, st = eng.preempt_dryrun_again(q)
        ev, es, en = ol.preempt_dryrun(pq, q)
        assert np.array_equal(gs, es) and np.array_equal(gv, ev) and np.array_equal(gn, en), k
        assert st["h2d_bytes"] == 0 and st["kernel_launches"] == 2
        seen.update(int(x) for x in gs)
    assert seen[0] > 0 and seen[1] > 0


def test_again_without_a_resident_state():
    fresh = capi.Engine(0)
    case = pc.rand_preempt_case(3)
    pq, query = pc.packed_dryrun_query(*case, lambda pq_, run_: (ol.quota(pq_, run_)["used"], ol.quota(pq_, run_)["used_mask"]))
    with pytest.raises(capi.NosplanError, match="no cluster state resident"):
        fresh.preempt_dryrun_again(query)
    fresh.close()


def test_reference_dryrun_preemption_fixtures(eng):
    """capacity_scheduling_test.go:249-535 (TestDryRunPreemption) through the C-ABI: the candidate count the reference asserts and
    the victims its `want` names where they do not depend on the order of equal-priority pods."""
    from support import dryrun_golden as dg
    for c, quotas, running, nodes, preemptor, gb, pdb_allowed, extra in dg.cases():
        pq, query = pc.packed_dryrun_query(quotas, running, nodes, preemptor, gb, pdb_allowed, extra, engine_used_fn(eng))
        gv, gs, gn, _ = eng.preempt_dryrun(pq, query)
        ev, es, en = ol.preempt_dryrun(pq, query)
        assert np.array_equal(gv, ev) and np.array_equal(gs, es) and np.array_equal(gn, en), c["name"]
        assert int((gs == pp.VICTIMS_OK).sum()) == c["expected_candidates"], (c["name"], gs)
        if c["expected_victims"] is not None:
            assert [p["name"] for p, v in zip(nodes[0]["pods"], gv) if v] == c["expected_victims"], c["name"]
        assert int(gn.sum()) == 0
