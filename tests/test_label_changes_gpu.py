"""nosplan_quota_label_changes (CUDA, through the C-ABI) against the packed oracle: the pods whose nos.nebuly.com/capacity
label has to be patched and the value to set, in the reference's patch order -- bit-exact."""
import numpy as np
import pytest

from nos_b200 import capi, quota_packing as qp, synth
from support import oracle_lib as ol
from support import quotacases as qc
from test_label_changes_packed import rand_current

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
    n_changes = 0
    for seed in range(block * 150, (block + 1) * 150):
        quotas, running, _, gb = qc.rand_quota_case(seed)
        pq, (run, _) = qp.pack_quotas(quotas, [running, []], gb)
        current = rand_current(seed, len(running))
        pod, desired, n, outs, _ = eng.quota_label_changes(pq, run, current, with_outputs=True)
        epod, edes = ol.label_changes(pq, run, current)
        assert n == len(epod) and np.array_equal(pod, epod) and np.array_equal(desired, edes), f"seed {seed}"
        if run.n_pods:
            exp = ol.quota(pq, run)
            assert np.array_equal(outs["over_quota"], exp["over_quota"]) and np.array_equal(outs["operator_used"], exp["operator_used"])
        n_changes += n
    assert n_changes > 200


def c4_state():
    pq, run, _ = synth.quota_state(500, 100_000, 0, seed=0xC4)
    desired = ol.quota(pq, run)["over_quota"]
    in_a_quota = pq.ns_quota[run.ns] >= 0
    settled = np.where(in_a_quota, 1 + desired, 0).astype(np.uint8)
    return pq, run, settled, in_a_quota


def test_config_c4_size(eng):
    """100 000 running pods of 500 quotas: a fresh cluster (every pod of a quota is patched), a settled one (nothing to patch),
    and a settled one where 1 % of the labels were tampered with."""
    pq, run, settled, in_a_quota = c4_state()
    fresh = np.zeros(run.n_pods, np.uint8)
    pod, desired, n, _, _ = eng.quota_label_changes(pq, run, fresh)
    epod, edes = ol.label_changes(pq, run, fresh)
    assert n == int(in_a_quota.sum()) and np.array_equal(pod, epod) and np.array_equal(desired, edes)
    assert np.array_equal(np.sort(pod), np.flatnonzero(in_a_quota))           # every pod of a quota exactly once
    assert np.all(np.diff(pq.ns_quota[run.ns[pod]]) >= 0)                      # by quota row
    pod, desired, n, _, _ = eng.quota_label_changes(pq, run, settled)
    assert n == 0 and pod.size == 0                                            # idempotence: applying the patches leaves nothing to do
    rng = np.random.default_rng(7)
    tampered = settled.copy()
    hit = rng.random(run.n_pods) < 0.01
    tampered[hit] = rng.integers(0, 3, int(hit.sum()))
    pod, desired, n, _, st = eng.quota_label_changes(pq, run, tampered)
    epod, edes = ol.label_changes(pq, run, tampered)
    assert np.array_equal(pod, epod) and np.array_equal(desired, edes) and 0 < n < 1200
    assert np.array_equal(desired, settled[pod])
    assert st["d2h_bytes"] < 64 * 1024                                         # the list, not a label per pod, comes back


def test_capacity_and_errors(eng):
    quotas, running, _, gb = qc.rand_quota_case(5, n_quotas=3, n_running=40, n_pending=0)
    pq, (run, _) = qp.pack_quotas(quotas, [running, []], gb)
    fresh = np.zeros(run.n_pods, np.uint8)
    epod, edes = ol.label_changes(pq, run, fresh)
    assert len(epod) > 4
    with pytest.raises(capi.NosplanError, match="capacity") as ei:
        eng.quota_label_changes(pq, run, fresh, capacity=4)
    pod, desired, n = ei.value.partial
    assert n == len(epod) and np.array_equal(pod, epod[:4]) and np.array_equal(desired, edes[:4])
    pod, desired, n, _, _ = eng.quota_label_changes(pq, run, fresh, capacity=len(epod))      # exactly enough
    assert np.array_equal(pod, epod)


def test_no_running_pods(eng):
    quotas, running, _, gb = qc.rand_quota_case(5, n_quotas=3, n_running=0, n_pending=0)
    pq, (run, _) = qp.pack_quotas(quotas, [running, []], gb)
    pod, desired, n, _, _ = eng.quota_label_changes(pq, run, np.zeros(0, np.uint8))
    assert n == 0 and pod.size == 0
