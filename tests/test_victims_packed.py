"""First pass of SelectVictimsOnNode: packed C++ oracle vs the faithful Python model on random preemption queries."""
import ctypes as C

import numpy as np
import pytest

from nos_b200 import quota_packing as qp, victim_packing as vp
from support import oracle_lib as ol
from support import victimcases as vc


def oracle_victims(pq, query):
    L = ol.lib()
    L.nos_oracle_quota_victims.argtypes = [C.POINTER(qp.CQuotas), C.POINTER(vp.CVictimQuery), C.c_void_p]
    L.nos_oracle_quota_victims.restype = C.c_int
    cq, k1 = qp.c_quotas(pq)
    cv, k2 = query.c_struct()
    out = np.zeros(max(1, query.n_pods), np.uint8)
    assert L.nos_oracle_quota_victims(C.byref(cq), C.byref(cv), out.ctypes.data) == 0
    return out[:query.n_pods]


def oracle_used(pq, run):
    r = ol.quota(pq, run)
    return r["used"], r["used_mask"]


@pytest.mark.parametrize("block", range(4))
def test_random_vs_python_model(block):
    n_victims = 0
    for seed in range(block * 200, (block + 1) * 200):
        quotas, running, node_pods, preemptor, gb = vc.rand_victim_case(seed)
        exp, _ = vc.model_victims(quotas, running, node_pods, preemptor, gb)
        pq, query = vc.packed_query(quotas, running, node_pods, preemptor, gb, oracle_used)
        got = oracle_victims(pq, query)
        assert np.array_equal(got, exp), f"seed {seed}\n{got}\n{exp}"
        n_victims += int(exp.sum())
    assert n_victims > 50
