"""Parity of the CUDA quota engine (nosplan_quota_recompute through the C-ABI) with the oracle: int64 fields bit-exact,
the float64 guaranteed-overquota path bit-exact too (same IEEE operations in the same order)."""
import json
import os

import numpy as np
import pytest

from nos_b200 import capi, quota_packing as qp, synth
from support import oracle_lib as ol
from support import quotacases as qc

pytestmark = pytest.mark.gpu
GOLD = json.load(open(os.path.join(os.path.dirname(__file__), "golden", "quota.json")))
FIELDS = ("used", "used_mask", "agg_min", "agg_used", "agg_overquota", "agg_masks", "guaranteed", "prefilter", "over_quota",
          "operator_used", "ambiguous_order")


@pytest.fixture(scope="module")
def eng():
    import __graft_entry__ as g
    g.build()
    e = capi.Engine(0)
    yield e
    e.close()


def compare(got, exp, what):
    for k in FIELDS:
        assert np.array_equal(got[k], exp[k]), f"{what}: {k}\n{got[k]}\n{exp[k]}"


def dense(pq, d):
    return np.array([d.get(c, 0) for c in pq.columns], np.int64)


def test_golden_guaranteed_overquotas(eng):
    for c in GOLD["guaranteed_overquotas"][1:]:
        quotas = {k: {"namespaces": [k], **v} for k, v in c["quotas"].items()}
        pq, _ = qp.pack_quotas(quotas, [], 16)
        r, _ = eng.quota_recompute(pq)
        assert np.array_equal(r["guaranteed"][pq.names.index(c["quota"])], dense(pq, c["expected"])), c["name"]


def test_golden_aggregated_and_reserve(eng):
    for c in GOLD["aggregated_overquotas"][1:]:
        quotas = {k: {"namespaces": [k], **v} for k, v in c["quotas"].items()}
        pq, _ = qp.pack_quotas(quotas, [], 16)
        assert np.array_equal(eng.quota_recompute(pq)[0]["agg_overquota"], dense(pq, c["expected"])), c["name"]
    for c in GOLD["reserve"]:
        quotas = {"eq": {"namespaces": ["ns"], "min": {}, "used": dict(c["used"])}}
        pods = [{"namespace": "ns", "name": f"p{i}", "request": dict(p)} for i, p in enumerate(c["pods"])]
        pq, (run,) = qp.pack_quotas(quotas, [pods], c["gpu_memory_gb"])
        assert np.array_equal(eng.quota_recompute(pq, run)[0]["used"][0], dense(pq, c["expected"]))


def test_golden_prefilter(eng):
    for c in GOLD["prefilter"]:
        pods = [{"namespace": p["namespace"], "name": f"p{i}", "request": {"memory": p["memory"], "cpu": p["cpu"], "nvidia.com/gpu": p["gpu"]}}
                for i, p in enumerate(c["pods"])]
        pq, (pend,) = qp.pack_quotas(c["quotas"], [pods], c["gpu_memory_gb"])
        r, _ = eng.quota_recompute(pq, None, pend)
        assert [int(x != 0) for x in r["prefilter"]] == c["expected"], c["name"]


@pytest.mark.parametrize("block", range(4))
def test_random_vs_oracle(eng, block):
    for seed in range(block * 200, (block + 1) * 200):
        quotas, running, pending, gb = qc.rand_quota_case(seed, unique_ts=(seed % 3 != 0))
        pq, (run, pend) = qp.pack_quotas(quotas, [running, pending], gb)
        got, _ = eng.quota_recompute(pq, run, pend)
        exp = ol.quota(pq, run, pend)
        compare(got, exp, f"seed {seed}")


def test_config_c4_full_size(eng):
    """BASELINE config #4: 500 ElasticQuotas / namespaces, 100,000 pods."""
    pq, run, pend = synth.quota_state(500, 100_000, 10_000, seed=0xC4)
    got, st = eng.quota_recompute(pq, run, pend)
    exp = ol.quota(pq, run, pend)
    compare(got, exp, "C4")
    assert got["ambiguous_order"] == 0
    # size-independent properties: used is a checksum of the requests; aggregates are sums of rows weighted by namespaces
    e_of = pq.ns_quota[run.ns]
    assert got["used"][:, 0].sum() == run.request[e_of >= 0, 0].sum()
    w = np.bincount(pq.ns_quota[pq.ns_quota >= 0], minlength=500)
    assert np.array_equal(got["agg_min"], (pq.min * w[:, None]).sum(0))
    assert (got["guaranteed"] >= 0).all() and (got["guaranteed"].sum(0) <= got["agg_overquota"] * w.max()).all()
    assert 0 < got["over_quota"].sum() < 100_000 and (got["prefilter"] <= 2).all()


def test_zero_sizes(eng):
    pq, run, pend = synth.quota_state(3, 0, 0, seed=1)
    got, _ = eng.quota_recompute(pq, run, pend)
    compare(got, ol.quota(pq, run, pend), "no pods")
    pq, (run,) = qp.pack_quotas({}, [[{"namespace": "a", "name": "p", "request": {"cpu": 1}}]], 16)
    got, _ = eng.quota_recompute(pq, run)
    compare(got, ol.quota(pq, run), "no quotas")


def test_labelling_with_key_ranges_beyond_one_word(eng):
    """creationTimestamps and priorities spanning (almost) their whole types: quota | timestamp | priority no longer fit one
    64-bit sort key, so the three stable radix passes run instead of the fused one -- same labels either way."""
    for seed in (3, 11, 42):
        pq, run, pend = synth.quota_state(40, 3000, 50, seed=seed)
        i = np.arange(run.n_pods)
        run.creation_ts = np.where(i % 2 == 0, run.creation_ts - (1 << 61), run.creation_ts + (1 << 61)).astype(np.int64)
        run.priority = np.where(i % 3 == 0, np.int64(-2**31), np.where(i % 3 == 1, np.int64(2**31 - 1), run.priority)).astype(np.int32)
        got, _ = eng.quota_recompute(pq, run, pend)
        compare(got, ol.quota(pq, run, pend), f"wide keys seed {seed}")


def test_labelling_by_segments_and_by_global_sort_agree(eng):
    """The labelling sorts inside buckets (counting sort by quota and leading timestamp bits, one CTA per bucket) and falls back
    to a device-wide radix sort when a bucket overflows (or under NOSPLAN_QUOTA_GLOBAL_SORT): same labels, same status.used
    either way; including tie groups and one quota holding most of the pods."""
    import os
    rng = np.random.default_rng(77)
    for n_q, n_run, seed, squeeze in ((500, 100_000, 0xC4, False), (60, 30_000, 5, False), (7, 9000, 6, True), (300, 2500, 7, False), (1, 2048, 8, False)):
        pq, run, pend = synth.quota_state(n_q, n_run, 100, seed=seed)
        if squeeze:                                       # most pods in one namespace: beyond the segment limit -> global sort
            run.ns[: n_run * 3 // 4] = run.ns[0]
        # tie groups; in the squeezed case thousands of pods of one quota share a timestamp, more than one bucket holds:
        # the bucket path gives up on the device and the call is repeated with the global sort
        tie = rng.random(n_run) < (0.3 if squeeze else 0.002)
        run.creation_ts[tie] = run.creation_ts[rng.integers(0, n_run, int(tie.sum())) // 50 * 50]
        if squeeze:
            run.creation_ts[: n_run // 3] = run.creation_ts[0]
        run.priority[rng.random(n_run) < 0.5] = 0
        exp = ol.quota(pq, run, pend)
        got, st1 = eng.quota_recompute(pq, run, pend)
        compare(got, exp, f"segments {n_q} x {n_run}")
        os.environ["NOSPLAN_QUOTA_GLOBAL_SORT"] = "1"
        try:
            got2, st2 = eng.quota_recompute(pq, run, pend)
        finally:
            del os.environ["NOSPLAN_QUOTA_GLOBAL_SORT"]
        compare(got2, exp, f"global sort {n_q} x {n_run}")
        if not squeeze and n_run / n_q < 1000:
            assert st1["kernel_launches"] < st2["kernel_launches"]


def test_sharded_entry_point_with_one_rank():
    """nosplan_quota_recompute_sharded with a one-rank communicator: the NCCL all-reduce / all-gather + fold of a single rank
    must leave the rows as nosplan_quota_recompute computes them (the two-rank form runs in tests/run_sharded_nccl.py)."""
    import __graft_entry__ as g
    g.build()
    from nos_b200 import capi, synth
    from support import oracle_lib as ol
    eng = capi.Engine(0)
    try:
        eng.shard_init(0, 1, eng.nccl_unique_id())
    except capi.NosplanError as e:
        eng.close()
        pytest.skip(f"NCCL not available: {e}")
    pq, run, pend = synth.quota_state(120, 20_000, 2_000, seed=0xC4)
    exp = ol.quota(pq, run, pend)
    got, st = eng.quota_recompute(pq, run, pend, sharded=True)
    for k in ("used", "agg_min", "agg_used", "agg_overquota", "guaranteed", "prefilter", "over_quota", "operator_used"):
        assert np.array_equal(np.asarray(got[k]), np.asarray(exp[k])), k
    eng.close()

