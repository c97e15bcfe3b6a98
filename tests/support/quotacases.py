"""Random logical quota states (test plumbing) + the faithful Python evaluation of everything nosplan_quota_recompute
returns, so packed implementations can be compared field by field."""
import random

import numpy as np

import nos_model as m
import nos_quota_model as q
from nos_b200 import quota_packing as qp

SCALARS = ["nvidia.com/gpu", "nos.nebuly.com/gpu-memory", "nvidia.com/mig-1g.10gb", "nvidia.com/mig-3g.40gb", "example.com/foo"]


def rand_quota_case(seed, n_quotas=None, n_running=None, n_pending=None, unique_ts=True):
    rng = random.Random(seed)
    E = n_quotas if n_quotas is not None else rng.randint(0, 6)
    quotas, all_ns = {}, []
    for i in range(E):
        k = 1 if rng.random() < 0.8 else rng.randint(2, 3)          # composite quota over k namespaces
        nss = [f"ns-{i}-{j}" for j in range(k)]
        all_ns += nss
        mn = {"cpu": rng.choice([0, 8000, 64000]), "memory": rng.choice([0, 16, 256]) << 30}
        if rng.random() < 0.3:
            mn["pods"] = rng.randint(0, 50)
        if rng.random() < 0.2:
            mn["ephemeral-storage"] = rng.randint(0, 100)
        for s in SCALARS:
            if rng.random() < 0.5:
                mn[s] = rng.choice([0, 2, 40, 640])
        e = {"namespaces": nss, "min": mn, "used": {}}
        if rng.random() < 0.6:
            e["max"] = {k_: v * 2 for k_, v in mn.items() if rng.random() < 0.8}
            e["max_enforced"] = rng.random() < 0.7
        if rng.random() < 0.3:
            e["used"] = {"cpu": rng.randint(0, 4000), "pods": rng.randint(0, 3), **({"example.com/foo": 1} if rng.random() < 0.3 else {})}
        if rng.random() < 0.05:
            e["min"] = None
        quotas[f"eq-{i}"] = e
    all_ns += ["free-ns"]

    def pods(n, tag):
        out = []
        for i in range(n):
            r = {}
            if rng.random() < 0.8:
                r["cpu"] = rng.choice([100, 1000, 4000])
            if rng.random() < 0.7:
                r["memory"] = rng.choice([1, 8, 32]) << 30
            if rng.random() < 0.5:
                r["nvidia.com/gpu"] = rng.randint(0, 2)
            if rng.random() < 0.3:
                r[rng.choice(SCALARS[2:4])] = rng.randint(1, 2)
            if rng.random() < 0.1:
                r["example.com/foo"] = rng.randint(0, 2)
            out.append({"namespace": rng.choice(all_ns), "name": f"{tag}-{rng.randint(0, 10**6)}-{i}", "priority": rng.choice([0, 0, 100, 1000]),
                        "creation_ts": (1_600_000_000 + i) if unique_ts else 1_600_000_000 + rng.randint(0, max(1, n // 3)),
                        "request": r})
        return out

    running = pods(n_running if n_running is not None else rng.randint(0, 40), "run")
    pending = pods(n_pending if n_pending is not None else rng.randint(0, 15), "pend")
    return quotas, running, pending, rng.choice([8, 16, 32])


def _res(d):
    return m.resource_from_list(d) if d is not None else None


def model_eval(quotas, running, pending, gpu_gb, pq):
    """Faithful evaluation with oracle/nos_quota_model.py -> same dict layout as QuotaResult.trimmed() (dense rows in pq.columns order)."""
    infos = q.ElasticQuotaInfos()
    objs = []
    for name, e in quotas.items():
        info = q.ElasticQuotaInfo(name=name, namespaces=list(e["namespaces"]), min=_res(e.get("min")), max=_res(e.get("max")),
                                  used=_res(e.get("used")) or m.Resource(), max_enforced=bool(e.get("max_enforced")))
        objs.append(info)
        for ns in info.namespaces:
            infos[ns] = info

    def mk(p):
        return m.Pod(p["namespace"], p["name"], p["priority"], [dict(p["request"])], creation_ts=p["creation_ts"])

    for i, p in enumerate(running):
        info = infos.get(p["namespace"])
        if info is not None:
            info.add_pod_if_not_present(f"uid-{i}", m.resource_from_list(q.quota_pod_request(mk(p), gpu_gb)))
    cols = pq.columns

    def row(r):
        d = m.resource_to_list(r)
        return np.array([d.get(c, 0) for c in cols], np.int64)

    out = {"used": np.stack([row(o.used) for o in objs]) if objs else np.zeros((0, len(cols)), np.int64)}
    out["agg_min"] = row(infos.aggregated_min()); out["agg_used"] = row(infos.aggregated_used())
    has_nil_min = any(o.min is None for o in infos.values())
    if not has_nil_min:
        out["agg_overquota"] = row(infos.aggregated_overquotas())
        g = []
        for o in objs:
            r = infos.guaranteed_overquotas(o.namespaces[0])
            g.append(row(r))
        out["guaranteed"] = np.stack(g) if g else np.zeros((0, len(cols)), np.int64)
    out["prefilter"] = np.array([q.prefilter(infos, mk(p), gpu_gb) for p in pending], np.uint8)
    over = np.zeros(len(running), np.uint8)
    op_used = np.zeros((len(objs), len(cols)), np.int64)
    for ei, (name, e) in enumerate(quotas.items()):
        idx = [i for i, p in enumerate(running) if p["namespace"] in e["namespaces"]]
        if e.get("min") is None:
            continue
        used, labels = q.patch_pods_and_compute_used([mk(running[i]) for i in idx], e["min"], e.get("max") or {}, gpu_gb)
        key_to_i = {f"{running[i]['namespace']}/{running[i]['name']}": i for i in idx}
        for key, ov in labels:
            over[key_to_i[key]] = 1 if ov else 0
        for r_, v in used.items():
            op_used[ei, cols.index(r_)] = v
    out["over_quota"] = over
    out["operator_used"] = op_used
    return out
