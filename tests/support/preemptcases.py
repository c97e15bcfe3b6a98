"""Random preemption dry-run cases (test plumbing): the victim cases of victimcases plus node accounts, PodDisruptionBudgets
and nominated requests, and their evaluation with the faithful Python model."""
import random

import numpy as np

import nos_model as m
import nos_quota_model as q
from nos_b200 import preempt_packing as pp, quota_packing as qp, victim_packing as vp
from support import victimcases as vc


def rand_preempt_case(seed):
    rng = random.Random(seed * 7919 + 13)
    quotas, running, node_pods, preemptor, gb = vc.rand_victim_case(seed)
    if rng.random() < 0.5:                                  # a small preemptor: the Filter passes once a victim or two are gone
        preemptor["request"] = {"cpu": rng.choice([100, 500]), "memory": rng.choice([1 << 20, 1 << 28]), "nvidia.com/gpu": rng.randint(0, 1)}
    if rng.random() < 0.4:                                  # enforce Max on some quotas so that "max quota exceeded" occurs
        for e in quotas.values():
            if e.get("max") is not None and rng.random() < 0.7:
                e["max_enforced"] = True
    if rng.random() < 0.5:                                  # one roomy quota: the aggregated Min stops being the binding check
        e = rng.choice(list(quotas.values()))
        e["min"] = {k: v * 1000 + 10 ** 6 for k, v in e["min"].items()}
    n_pdbs = rng.randint(0, 3)
    pdb_allowed = [rng.randint(0, 2) for _ in range(n_pdbs)]
    nodes = []
    for pods in node_pods:
        tot = {}
        for p in pods:
            for r, v in p["request"].items():
                tot[r] = tot.get(r, 0) + v
        alloc = {}
        mode = rng.random()
        for r in set(tot) | set(preemptor["request"]):
            if mode < 0.4:                                  # tight node: the preemptor fits only while enough victims stay away
                alloc[r] = tot.get(r, 0) + rng.choice([0, 0, 1]) * preemptor["request"].get(r, 0) // 2
            else:
                want = tot.get(r, 0) + (preemptor["request"].get(r, 0) if rng.random() < 0.5 else 0)
                alloc[r] = max(0, want + rng.choice([-3, -1, 0, 0, 1, 5]) * max(1, want // 8))
        alloc["pods"] = len(pods) + rng.choice([-1, 0, 1, 1, 5])
        if 0.4 <= mode < 0.6:                               # roomy node: every victim can be reprieved on node resources alone
            alloc = {r: 10 * (v + 1000) for r, v in alloc.items()}
            alloc["pods"] = len(pods) + 10
        nodes.append({"allocatable": alloc, "pods": pods,
                      "pdbs_of_pod": [[b for b in range(n_pdbs) if rng.random() < 0.35] for _ in pods]})
    extra = {"cpu": rng.choice([0, 0, 500, 20000, 10 ** 9]), "memory": rng.choice([0, 1 << 30])}
    return quotas, running, nodes, preemptor, gb, pdb_allowed, extra


def model_dryrun(quotas, running, nodes, preemptor, gb, pdb_allowed, extra):
    infos = vc.build_infos(quotas, running, gb)
    pre = vc.mk_pod(preemptor)
    pod_req = m.resource_from_list(q.quota_pod_request(pre, gb))
    nom_in_eq = m.Resource(); nom_in_eq.add_list(m.resource_to_list(pod_req))
    nominated = nom_in_eq.clone(); nominated.add_list(extra)
    victim, status, nviol = [], [], []
    for n in nodes:
        node = q.PreemptionNode(m.resource_from_list(n["allocatable"]), m.Resource(), [vc.mk_pod(p) for p in n["pods"]], n["pdbs_of_pod"])
        ni = m.NodeInfo(node.allocatable, node.requested, 0)
        for p in node.pods:
            ni.add_pod(p)
        st, vs, nv = q.select_victims_on_node(infos, pre, pod_req, nom_in_eq, nominated, node, pdb_allowed, gb)
        flags = [0] * len(node.pods)
        for i in vs:
            flags[i] = 1
        victim += flags; status.append(st); nviol.append(nv)
    return np.array(victim, np.uint8), np.array(status, np.int32), np.array(nviol, np.int32)


def packed_dryrun_query(quotas, running, nodes, preemptor, gb, pdb_allowed, extra, used_fn):
    pq, sets = qp.pack_quotas(quotas, [running, [preemptor]], gb)
    run = sets[0]
    used, umask = used_fn(pq, run)
    pq.used_init = used.copy(); pq.used_mask = umask.copy()
    nom_row = vp.quota_request_row(pq, preemptor["request"])
    nom_in_eq = {c: int(v) for c, v in zip(pq.columns, nom_row)}
    nominated = dict(nom_in_eq)
    for r, v in extra.items():
        nominated[r] = nominated.get(r, 0) + v
    return pq, pp.pack_preempt_query(pq, preemptor, nom_in_eq, nominated, nodes, pdb_allowed)
