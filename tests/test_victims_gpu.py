"""nosplan_quota_victims (CUDA, through the C-ABI) against the packed oracle: potential-victim flags, bit-exact."""
import ctypes as C

import numpy as np
import pytest

from nos_b200 import capi, quota_packing as qp, synth, victim_packing as vp
from support import oracle_lib as ol
from support import victimcases as vc
from test_victims_packed import oracle_victims

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def eng():
    import __graft_entry__ as g
    g.build()
    e = capi.Engine(0)
    yield e
    e.close()


@pytest.mark.parametrize("block", range(4))
def test_random_vs_oracle(eng, block):
    n_victims = 0
    for seed in range(block * 200, (block + 1) * 200):
        quotas, running, node_pods, preemptor, gb = vc.rand_victim_case(seed)

        def engine_used(pq, run):
            r, _ = eng.quota_recompute(pq, run)
            return r["used"], r["used_mask"]

        pq, query = vc.packed_query(quotas, running, node_pods, preemptor, gb, engine_used)
        got, _ = eng.quota_victims(pq, query)
        exp = oracle_victims(pq, query)
        assert np.array_equal(got, exp), f"seed {seed}\n{got}\n{exp}"
        n_victims += int(exp.sum())
    assert n_victims > 50


def test_config_c4_size(eng):
    """100,000 running pods of 500 quotas spread over 10,000 nodes, one preemptor per query."""
    pq, run, _ = synth.quota_state(500, 100_000, 0, seed=0xC4)
    r, _ = eng.quota_recompute(pq, run)
    pq.used_init, pq.used_mask = r["used"].copy(), r["used_mask"].copy()
    rng = np.random.default_rng(4)
    node = rng.integers(0, 10_000, run.n_pods)
    order = np.lexsort((run.priority, node))              # by node, least important (lowest priority) first
    ptr = np.zeros(10_001, np.int32)
    ptr[1:] = np.cumsum(np.bincount(node, minlength=10_000))
    for pre_ns, pre_prio in ((int(run.ns[0]), 1000), (int(run.ns[17]), 0), (len(pq.namespaces) - 1, 100)):
        nom = np.zeros(pq.n_res, np.int64)
        nom[0], nom[1], nom[5] = 4000, 8 << 30, 1
        nom[pq.gpu_memory_col] = 32
        q = vp.PackedVictimQuery(pre_ns, pre_prio, nom, ptr, run.ns[order], run.priority[order], r["over_quota"][order],
                                 run.request[order], run.request_mask[order])
        got, st = eng.quota_victims(pq, q)
        exp = oracle_victims(pq, q)
        assert np.array_equal(got, exp)
    assert got.shape == (100_000,)
