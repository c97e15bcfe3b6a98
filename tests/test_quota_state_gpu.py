"""Event-driven quota accounting on a resident state (nosplan_quota_state_begin / apply) against the faithful Python model's
addPodIfNotPresent / deletePodIfPresent (elasticquotainfo.go:276-310) applied event by event, and against the batch engine."""
import random

import numpy as np
import pytest

import nos_model as m
import nos_quota_model as q
from nos_b200 import capi, quota_packing as qp, synth, victim_packing as vp
from support import oracle_lib as ol
from support import quotacases as qc
from support import victimcases as vc

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def eng():
    import __graft_entry__ as g
    g.build()
    e = capi.Engine(0)
    yield e
    e.close()


def model_rows(infos, objs, pq):
    cols = pq.columns

    def row(r):
        d = m.resource_to_list(r)
        return np.array([d.get(c, 0) for c in cols], np.int64)

    out = {"used": np.stack([row(o.used) for o in objs]), "agg_min": row(infos.aggregated_min()), "agg_used": row(infos.aggregated_used()),
           "agg_overquota": row(infos.aggregated_overquotas()),
           "guaranteed": np.stack([row(infos.guaranteed_overquotas(o.namespaces[0])) for o in objs])}
    return out


@pytest.mark.parametrize("block", range(3))
def test_random_event_streams_vs_python_model(eng, block):
    applied_total = ignored_total = 0
    for seed in range(block * 40, (block + 1) * 40):
        rng = random.Random(seed)
        quotas, running, _, gb = qc.rand_quota_case(seed, n_quotas=rng.randint(1, 6), n_running=rng.randint(5, 60), n_pending=0)
        for e in quotas.values():
            if e.get("min") is None:
                e["min"] = {"cpu": 1000}
        pq, (run, _) = qp.pack_quotas(quotas, [running, []], gb)
        infos = vc.build_infos(quotas, [], gb)
        objs, seen_ids = [], set()
        for name in quotas:
            objs.append(next(o for o in infos.values() if o.name == name))
        eng.quota_state_begin(pq, len(running))
        for _ in range(rng.randint(2, 8)):
            batch = rng.sample(range(len(running)), rng.randint(0, len(running)))
            ops = [rng.choice([capi.EVENT_ADD, capi.EVENT_ADD, capi.EVENT_DELETE]) for _ in batch]
            req = run.request[batch].copy() if batch else np.zeros((0, pq.n_res), np.int64)
            expect_applied = 0
            for k, (i, op) in enumerate(zip(batch, ops)):
                p = running[i]
                r = dict(p["request"])
                if op == capi.EVENT_DELETE and rng.random() < 0.3:         # the pod changed since it was accounted: the reference
                    r["cpu"] = r.get("cpu", 0) + 250                      # unreserves what the pod requests NOW
                    req[k, 0] += 250
                info = infos.get(p["namespace"])
                if info is None:
                    continue
                pod = m.Pod(p["namespace"], p["name"], p["priority"], [r], creation_ts=p["creation_ts"])
                rr = m.resource_from_list(q.quota_pod_request(pod, gb))
                before = len(info.pods)
                if op == capi.EVENT_ADD:
                    info.add_pod_if_not_present(f"uid-{i}", rr)
                else:
                    info.delete_pod_if_present(f"uid-{i}", rr)
                expect_applied += len(info.pods) != before
            got, applied, st = eng.quota_state_apply(ops, batch, run.ns[batch] if batch else [], req, run.request_mask[batch] if batch else None)
            exp = model_rows(infos, objs, pq)
            for k, v in exp.items():
                assert np.array_equal(got[k], v), f"seed {seed}: {k}\n{got[k]}\n{v}"
            assert applied == expect_applied, (seed, applied, expect_applied)
            applied_total += applied; ignored_total += len(batch) - applied
    assert applied_total > 300 and ignored_total > 100


def test_c4_size_against_the_batch_engine(eng):
    """100 000 pods added in one batch = nosplan_quota_recompute of those pods; a third deleted = recompute of the rest; adding
    everybody again changes only the deleted ones."""
    pq, run, _ = synth.quota_state(500, 100_000, 0, seed=0xC4)
    n = run.n_pods
    ids = np.arange(n, dtype=np.int32)
    eng.quota_state_begin(pq, n)
    got, applied, st = eng.quota_state_apply(np.zeros(n, np.uint8), ids, run.ns, run.request, run.request_mask)
    ref, _ = eng.quota_recompute(pq, run)
    in_quota = int((pq.ns_quota[run.ns] >= 0).sum())
    assert applied == in_quota
    for k in ("used", "agg_min", "agg_used", "agg_overquota", "guaranteed"):
        assert np.array_equal(got[k], ref[k]), k
    assert np.array_equal(got["used_mask"] & 0x7ffffff0, ref["used_mask"] & 0x7ffffff0)
    gone = np.flatnonzero(np.arange(n) % 3 == 0).astype(np.int32)
    got, applied, _ = eng.quota_state_apply(np.ones(len(gone), np.uint8), gone, run.ns[gone], run.request[gone], run.request_mask[gone])
    from nos_b200 import sharded
    rest = np.flatnonzero(np.arange(n) % 3 != 0)
    ref, _ = eng.quota_recompute(pq, sharded.take_pods(run, rest))
    for k in ("used", "agg_used", "agg_overquota", "guaranteed"):
        assert np.array_equal(got[k], ref[k]), k
    _, applied, st = eng.quota_state_apply(np.zeros(n, np.uint8), ids, run.ns, run.request, run.request_mask, outputs=False)
    assert applied == int((pq.ns_quota[run.ns[gone]] >= 0).sum())
    got, applied, st = eng.quota_state_apply([], [], [], np.zeros((0, pq.n_res), np.int64))        # no events: just read the state
    full, _ = eng.quota_recompute(pq, run)
    assert applied == 0 and np.array_equal(got["used"], full["used"]) and np.array_equal(got["guaranteed"], full["guaranteed"])


def test_prefilter_against_the_resident_state(eng):
    """PreFilter verdicts read from the resident state = the batch engine's verdicts for the same accounted pods."""
    pq, run, pend = synth.quota_state(300, 30_000, 5_000, seed=12)
    eng.quota_state_begin(pq, run.n_pods)
    ids = np.arange(run.n_pods, dtype=np.int32)
    eng.quota_state_apply(np.zeros(run.n_pods, np.uint8), ids, run.ns, run.request, run.request_mask, outputs=False)
    verdict, _ = eng.quota_state_prefilter(pend)
    exp = ol.quota(pq, run, pend)["prefilter"]
    assert np.array_equal(verdict, exp) and len(set(exp.tolist())) > 1
    from nos_b200 import sharded
    gone = np.flatnonzero(ids % 2 == 0).astype(np.int32)
    eng.quota_state_apply(np.ones(len(gone), np.uint8), gone, run.ns[gone], run.request[gone], run.request_mask[gone], outputs=False)
    verdict, _ = eng.quota_state_prefilter(pend)
    exp2 = ol.quota(pq, sharded.take_pods(run, np.flatnonzero(ids % 2 == 1)), pend)["prefilter"]
    assert np.array_equal(verdict, exp2) and not np.array_equal(exp, exp2)


def test_errors(eng):
    fresh = capi.Engine(0)
    pq, run, _ = synth.quota_state(5, 20, 0, seed=2)
    fresh._state_pq = pq
    with pytest.raises(capi.NosplanError, match="no quota state resident"):
        fresh.quota_state_apply([0], [0], run.ns[:1], run.request[:1])
    fresh.quota_state_begin(pq, 20)
    with pytest.raises(capi.NosplanError, match="twice in one batch"):
        fresh.quota_state_apply([0, 1], [3, 3], run.ns[[3, 3]], run.request[[3, 3]])
    with pytest.raises(capi.NosplanError, match="outside"):
        fresh.quota_state_apply([0], [20], run.ns[:1], run.request[:1])
    got, applied, _ = fresh.quota_state_apply([0, 0], [3, 4], run.ns[[3, 4]], run.request[[3, 4]], run.request_mask[[3, 4]])   # still usable
    assert applied == int((pq.ns_quota[run.ns[[3, 4]]] >= 0).sum())
    fresh.close()
