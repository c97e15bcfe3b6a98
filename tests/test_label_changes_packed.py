"""The operator's label patches (patchCapacityInfoIfDifferent, internal/controllers/elasticquota/elasticquota.go:44-60,
:120-149): packed C++ oracle against the faithful Python model on random quota states and random current labels."""
import random

import numpy as np
import pytest

import nos_model as m
import nos_quota_model as q
from nos_b200 import quota_packing as qp
from support import oracle_lib as ol
from support import quotacases as qc

LABEL_TEXT = {1: "in-quota", 2: "over-quota"}


def rand_current(seed, n):
    rng = random.Random(seed * 31 + 7)
    mode = rng.random()
    if mode < 0.15:
        return np.zeros(n, np.uint8)                        # a fresh cluster: no pod labelled yet
    return np.array([rng.choice([0, 1, 1, 2, 2]) for _ in range(n)], np.uint8)


def model_changes(quotas, running, current, gb):
    """One PatchPodsAndComputeUsedQuota per quota, in quota row order -> [(running index, desired code)]"""
    out = []
    for name, e in quotas.items():
        idx = [i for i, p in enumerate(running) if p["namespace"] in e["namespaces"]]
        pods = []
        for i in idx:
            p = running[i]
            labels = {"nos.nebuly.com/capacity": LABEL_TEXT[int(current[i])]} if current[i] else ({} if i % 2 else {"nos.nebuly.com/capacity": "foo"})
            pods.append(m.Pod(p["namespace"], p["name"], p["priority"], [dict(p["request"])], labels=labels, creation_ts=p["creation_ts"]))
        key_to_i = {f"{running[i]['namespace']}/{running[i]['name']}": i for i in idx}
        for key, want in q.capacity_info_patches(pods, e.get("min") or {}, e.get("max") or {}, gb):
            out.append((key_to_i[key], 1 if want == "in-quota" else 2))
    return out


@pytest.mark.parametrize("block", range(4))
def test_random_vs_python_model(block):
    n_changes = 0
    for seed in range(block * 150, (block + 1) * 150):
        quotas, running, _, gb = qc.rand_quota_case(seed)
        pq, (run, _) = qp.pack_quotas(quotas, [running, []], gb)
        current = rand_current(seed, len(running))
        exp = model_changes(quotas, running, current, gb)
        pod, desired = ol.label_changes(pq, run, current)
        assert list(zip(pod.tolist(), desired.tolist())) == exp, f"seed {seed}"
        n_changes += len(exp)
    assert n_changes > 200


def test_nothing_to_patch_when_labels_are_current():
    quotas, running, _, gb = qc.rand_quota_case(11, n_quotas=4, n_running=50, n_pending=0)
    pq, (run, _) = qp.pack_quotas(quotas, [running, []], gb)
    r = ol.quota(pq, run)
    in_a_quota = pq.ns_quota[run.ns] >= 0
    current = np.where(in_a_quota, 1 + r["over_quota"], 0).astype(np.uint8)
    pod, desired = ol.label_changes(pq, run, current)
    assert pod.size == 0 and desired.size == 0
