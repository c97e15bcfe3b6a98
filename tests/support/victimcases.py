"""Random preemption queries (test plumbing): quota state + a preemptor + pods grouped by node, and their evaluation with
the faithful Python model."""
import random

import numpy as np

import nos_model as m
import nos_quota_model as q
from nos_b200 import quota_packing as qp, victim_packing as vp
from support import quotacases as qc


def rand_victim_case(seed, n_nodes=None, pods_per_node=8):
    rng = random.Random(seed)
    quotas, running, _, gb = qc.rand_quota_case(seed, n_quotas=rng.randint(1, 6), n_running=rng.randint(5, 60), n_pending=0)
    for e in quotas.values():                      # SelectVictimsOnNode dereferences Min
        if e.get("min") is None:
            e["min"] = {"cpu": 1000}
        if rng.random() < 0.5:                     # tighter quotas so that over-min states occur
            e["min"] = {k: v // rng.choice([1, 4, 64]) for k, v in e["min"].items()}
    for p in running:
        p["over_quota"] = rng.random() < 0.5
    n_nodes = n_nodes if n_nodes is not None else rng.randint(1, 6)
    node_pods = [[] for _ in range(n_nodes)]
    for p in running:
        node_pods[rng.randrange(n_nodes)].append(p)
    for pods in node_pods:                          # reference order: least important first (priority asc)
        pods.sort(key=lambda p: p["priority"])
    all_ns = sorted({ns for e in quotas.values() for ns in e["namespaces"]}) + ["free-ns"]
    preemptor = {"namespace": rng.choice(all_ns), "name": "preemptor", "priority": rng.choice([0, 100, 1000, 5000]), "creation_ts": 0,
                 "request": {"cpu": rng.choice([100, 4000, 64000]), "memory": rng.choice([1, 64]) << 30, "nvidia.com/gpu": rng.randint(0, 2)}}
    return quotas, running, node_pods, preemptor, gb


def build_infos(quotas, running, gb):
    infos = q.ElasticQuotaInfos()
    for name, e in quotas.items():
        info = q.ElasticQuotaInfo(name=name, namespaces=list(e["namespaces"]), min=qc._res(e.get("min")), max=qc._res(e.get("max")),
                                  used=qc._res(e.get("used")) or m.Resource(), max_enforced=bool(e.get("max_enforced")))
        for ns in info.namespaces:
            infos[ns] = info
    for i, p in enumerate(running):
        info = infos.get(p["namespace"])
        if info is not None:
            info.add_pod_if_not_present(f"uid-{i}", m.resource_from_list(q.quota_pod_request(mk_pod(p), gb)))
    return infos


def mk_pod(p):
    labels = {"nos.nebuly.com/capacity": "over-quota" if p.get("over_quota") else "in-quota"}
    return m.Pod(p["namespace"], p["name"], p["priority"], [dict(p["request"])], labels=labels, creation_ts=p.get("creation_ts", 0))


def model_victims(quotas, running, node_pods, preemptor, gb):
    infos = build_infos(quotas, running, gb)
    pre = mk_pod(preemptor)
    nom = m.Resource()
    nom.add_list(m.resource_to_list(m.resource_from_list(q.quota_pod_request(pre, gb))))
    out = []
    for pods in node_pods:
        out += q.select_potential_victims(infos, pre, nom, [mk_pod(p) for p in pods], gb)
    return np.array(out, np.uint8), infos


def packed_query(quotas, running, node_pods, preemptor, gb, engine_or_oracle_used):
    """-> (PackedQuotas with used_init = current Used, PackedVictimQuery).  `engine_or_oracle_used(pq, run)` returns the
    recomputed (used, used_mask)."""
    pq, (run, _) = qp.pack_quotas(quotas, [running, [preemptor]], gb)        # the preemptor's resources need columns too
    used, umask = engine_or_oracle_used(pq, run)
    pq.used_init = used.copy()
    pq.used_mask = umask.copy()
    nom_row = vp.quota_request_row(pq, preemptor["request"])
    nom = {c: int(v) for c, v in zip(pq.columns, nom_row)}
    return pq, vp.pack_victim_query(pq, preemptor, nom, node_pods)
