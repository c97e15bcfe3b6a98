"""The whole of SelectVictimsOnNode (first pass, Filter, quota checks, PDB split, reprieve loops): packed C++ oracle vs the
faithful Python model on random cases -- every status the reference can return must occur."""
import collections

import numpy as np
import pytest

from support import oracle_lib as ol
from support import preemptcases as pc
from test_victims_packed import oracle_used


TOTAL = collections.Counter()


@pytest.mark.parametrize("block", range(4))
def test_random_vs_python_model(block):
    seen = collections.Counter()
    for seed in range(block * 250, (block + 1) * 250):
        case = pc.rand_preempt_case(seed)
        ev, es, en = pc.model_dryrun(*case)
        pq, query = pc.packed_dryrun_query(*case, oracle_used)
        gv, gs, gn = ol.preempt_dryrun(pq, query)
        assert np.array_equal(gs, es), f"seed {seed} status\n{gs}\n{es}"
        assert np.array_equal(gv, ev), f"seed {seed} victims\n{gv}\n{ev}"
        assert np.array_equal(gn, en), f"seed {seed} numViolating\n{gn}\n{en}"
        seen.update(int(x) for x in es)
        seen["victims"] += int(ev.sum()); seen["violating"] += int(en.sum())
    assert seen[0] > 20 and seen[1] > 20 and seen[2] > 5 and seen["victims"] > 10, seen
    TOTAL.update(seen); TOTAL["blocks"] += 1


def test_every_status_occurs():
    if TOTAL["blocks"] < 4:
        pytest.skip("runs after the four blocks above, in one process")
    assert all(TOTAL[s] > 0 for s in range(6)) and TOTAL["violating"] > 0, TOTAL


def test_reference_dryrun_preemption_fixtures():
    """capacity_scheduling_test.go:249-535 (TestDryRunPreemption): the candidate count the reference asserts, and the victims its
    `want` names where they do not depend on the order of equal-priority pods -- Python model and packed oracle."""
    from support import dryrun_golden as dg
    for c, quotas, running, nodes, preemptor, gb, pdb_allowed, extra in dg.cases():
        ev, es, en = pc.model_dryrun(quotas, running, nodes, preemptor, gb, pdb_allowed, extra)
        pq, query = pc.packed_dryrun_query(quotas, running, nodes, preemptor, gb, pdb_allowed, extra, oracle_used)
        gv, gs, gn = ol.preempt_dryrun(pq, query)
        assert np.array_equal(gv, ev) and np.array_equal(gs, es) and np.array_equal(gn, en), c["name"]
        assert int((es == 0).sum()) == c["expected_candidates"], (c["name"], es)
        if c["expected_victims"] is not None:
            assert [p["name"] for p, v in zip(nodes[0]["pods"], ev) if v] == c["expected_victims"], c["name"]
        assert int(en.sum()) == 0                                           # NumPDBViolations: 0 in every `want`
