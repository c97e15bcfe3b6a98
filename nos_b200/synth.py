"""Synthetic cluster snapshots in packed (SoA) form, generated with a counter-based SplitMix64 so the same
seed gives the same bytes everywhere.  All data is synthetic; shapes follow BASELINE.json.

Demand regimes
  saturating   devices of the demanded profiles are never free at t0 (used, or carved out), so every pod lacks
               its slice cluster-wide and the sweep visits every candidate node: N_cand * P candidates.
  satisfiable  only `n_pods` ~ a small fraction of the reconfigurable supply: the sweep exits early
               (core/planner.go:94-97).
"""
from __future__ import annotations

from typing import Dict, List, Optional, Tuple

import numpy as np

from . import packing as pk
from .geometries import A100_80, DEFAULT_KNOWN_GEOMETRIES

_GAMMA = np.uint64(0x9E3779B97F4A7C15)


def splitmix64(seed: int, n: int, stream: int = 0) -> np.ndarray:
    """n outputs of SplitMix64 started at `seed` (+ a stream offset so independent fields do not overlap)."""
    with np.errstate(over="ignore"):
        s = np.uint64(seed) + np.uint64(stream) * np.uint64(0xD1342543DE82EF95)
        z = s + _GAMMA * (np.arange(1, n + 1, dtype=np.uint64))
        z = (z ^ (z >> np.uint64(30))) * np.uint64(0xBF58476D1CE4E5B9)
        z = (z ^ (z >> np.uint64(27))) * np.uint64(0x94D049BB133111EB)
        return z ^ (z >> np.uint64(31))


def _uniform(seed, n, stream):
    return (splitmix64(seed, n, stream) >> np.uint64(11)).astype(np.float64) * (1.0 / (1 << 53))


def _randint(seed, n, stream, lo, hi):
    """uniform integers in [lo, hi]"""
    return lo + (splitmix64(seed, n, stream) % np.uint64(hi - lo + 1)).astype(np.int64)


def _choice(seed, n, stream, probs):
    u = _uniform(seed, n, stream)
    return np.searchsorted(np.cumsum(probs), u, side="right").clip(0, len(probs) - 1)


def mig_cluster(n_nodes: int, gpus_per_node: int, n_pods: int, seed: int, model: str = A100_80,
                known: Optional[dict] = None, regime: str = "saturating", p_full_free: float = 0.3,
                p_used: float = 0.6, demand=None, demand_probs=None,
                cpu_milli: int = 64000, mem_gib: int = 512, allowed_pods: int = 110,
                unique_pod_requests: bool = True) -> Tuple[pk.PackedSnapshot, pk.PackedPods]:
    """N nodes x G GPUs of one MIG model (the C2 / C5 recipe).  Node names are node-%06d (byte order == numeric)."""
    known = known or DEFAULT_KNOWN_GEOMETRIES
    geoms = known[model]
    profiles = sorted({p for g in geoms for p in g})
    K = len(profiles)
    pid = {p: i for i, p in enumerate(profiles)}
    if demand is None:
        demand = ["1g.10gb", "2g.20gb", "3g.40gb", "4g.40gb", "7g.79gb"] if model == A100_80 else profiles[:]
        demand_probs = [.4, .25, .15, .1, .1] if model == A100_80 else [1.0 / len(demand)] * len(demand)
    if regime == "saturating" and model == A100_80 and len(demand) == 5:
        demand, demand_probs = demand[:4], [.45, .25, .15, .15]
    N, G = n_nodes, gpus_per_node
    NG = N * G
    rows = np.zeros((len(geoms), 16), np.uint8)
    for r, g in enumerate(geoms):
        for p, q in g.items():
            rows[r, pid[p]] = q
    demanded = np.zeros(16, bool)
    for p in demand:
        demanded[pid[p]] = True
    biggest = int(np.argmax(rows.sum(1) == 1)) if (rows.sum(1) == 1).any() else 0   # the single-device geometry

    gsel = _randint(seed, NG, 1, 0, len(geoms) - 1)
    full_free = _uniform(seed, NG, 2) < p_full_free
    gsel = np.where(full_free, biggest, gsel)
    geom = rows[gsel].astype(np.int64)                                  # [NG,16] device counts
    # each device used w.p. p_used: binomial via per-device uniforms (<= 7 devices per profile)
    used = np.zeros((NG, 16), np.int64)
    for d in range(7):
        u = _uniform(seed, NG * 16, 10 + d).reshape(NG, 16)
        used += ((u < p_used) & (geom > d))
    used = np.where(full_free[:, None], 0, used)
    free = geom - used
    if regime == "saturating":
        free = np.where(demanded[None, :], 0, free)                     # demanded profiles are never free at t0
    umask = ((used > 0) << np.arange(16)).sum(1).astype(np.uint16)
    fmask = ((free > 0) << np.arange(16)).sum(1).astype(np.uint16)

    R = 3 + K
    alloc = np.zeros((N, R), np.int64)
    req = np.zeros((N, R), np.int64)
    alloc[:, 0] = cpu_milli
    alloc[:, 1] = mem_gib << 30
    tot = (used + free).reshape(N, G, 16).sum(1)
    alloc[:, 3:3 + K] = tot[:, :K]                                      # Status.Allocatable consistent with annotations
    req[:, 3:3 + K] = used.reshape(N, G, 16).sum(1)[:, :K]
    snap = pk.PackedSnapshot(
        kind=pk.KIND_MIG, profiles=profiles,
        profile_mem_gb=np.array([pk.mig_mem(p) for p in profiles], np.int32),
        profile_gi=np.array([pk.mig_gi(p) for p in profiles], np.int32),
        models=[model], model_geom_offset=np.array([0, len(geoms)], np.int32), geom_counts=rows,
        node_names=[f"node-{i:06d}" for i in range(N)], gpus_per_node=G,
        node_gpu_count=np.full(N, G, np.int32), gpu_model=np.zeros(NG, np.int32),
        gpu_index=np.tile(np.arange(G, dtype=np.int32), N), gpu_memory_gb=np.zeros(NG, np.int32),
        gpu_used=used.astype(np.uint8), gpu_free=free.astype(np.uint8), gpu_used_mask=umask, gpu_free_mask=fmask,
        other_resources=[], node_alloc=alloc, node_requested=req,
        node_allowed_pods=np.full(N, allowed_pods, np.int32), node_num_pods=np.zeros(N, np.int32))

    P = n_pods
    prof = _choice(seed, P, 20, demand_probs)
    prof_col = np.array([3 + pid[p] for p in demand], np.int32)[prof]
    prio = np.array([0, 100, 1000], np.int32)[_choice(seed, P, 21, [.7, .2, .1])]
    base = np.zeros((P, 3), np.int64)
    if unique_pod_requests:
        base[:, 0] = _randint(seed, P, 22, 100, 4000)
        base[:, 1] = _randint(seed, P, 23, 1, 32) << 30
    pods = pk.PackedPods(keys=[f"ns/pod-{i}" for i in range(P)], priority=prio,
                         tracker_key=np.arange(P, dtype=np.int32), base=base,
                         scalar_ptr=np.arange(P + 1, dtype=np.int32), scalar_col=prof_col.astype(np.int32),
                         scalar_qty=np.ones(P, np.int64))
    return snap, pods


def mps_cluster(n_nodes: int, gpus_per_node: int, n_pods: int, seed: int, gpu_memory_mb: int = 81920,
                sizes=(10, 20, 40), demand_sizes=(10, 20, 30, 40), regime: str = "saturating",
                cpu_milli: int = 64000, mem_gib: int = 512, allowed_pods: int = 110):
    """N nodes x G GPUs sliced by memory (the C3 recipe): gpu.memory label in MB -> MemoryGB = ceil(MB/1000)."""
    mem_gb = -(-gpu_memory_mb // 1000)
    all_sizes = sorted(set(sizes) | set(demand_sizes))
    profiles = [f"{s}gb" for s in all_sizes]
    K = len(profiles)
    pid = {s: i for i, s in enumerate(all_sizes)}
    N, G = n_nodes, gpus_per_node
    NG = N * G
    fill = _randint(seed, NG, 1, 0, mem_gb - 2)                         # GB already carved into slices
    used = np.zeros((NG, 16), np.int64)
    free = np.zeros((NG, 16), np.int64)
    left = fill.copy()
    for it in range(8):                                                 # carve slices of random sizes until `fill` is spent
        sidx = _randint(seed, NG, 30 + it, 0, len(sizes) - 1)
        sz = np.array(sizes, np.int64)[sidx]
        ok = sz <= left
        is_used = _uniform(seed, NG, 40 + it) < (1.0 if regime == "saturating" else 0.5)
        col = np.array([pid[s] for s in sizes], np.int64)[sidx]
        np.add.at(used, (np.nonzero(ok & is_used)[0], col[ok & is_used]), 1)
        np.add.at(free, (np.nonzero(ok & ~is_used)[0], col[ok & ~is_used]), 1)
        left = np.where(ok, left - sz, left)
    umask = ((used > 0) << np.arange(16)).sum(1).astype(np.uint16)
    fmask = ((free > 0) << np.arange(16)).sum(1).astype(np.uint16)
    R = 3 + K
    alloc = np.zeros((N, R), np.int64)
    req = np.zeros((N, R), np.int64)
    alloc[:, 0] = cpu_milli
    alloc[:, 1] = mem_gib << 30
    alloc[:, 3:3 + K] = (used + free).reshape(N, G, 16).sum(1)[:, :K]
    req[:, 3:3 + K] = used.reshape(N, G, 16).sum(1)[:, :K]
    snap = pk.PackedSnapshot(
        kind=pk.KIND_MPS, profiles=profiles, profile_mem_gb=np.array(all_sizes, np.int32), profile_gi=np.zeros(K, np.int32),
        models=["NVIDIA-A100-80GB-PCIe"], model_geom_offset=np.zeros(2, np.int32), geom_counts=np.zeros((0, 16), np.uint8),
        node_names=[f"node-{i:06d}" for i in range(N)], gpus_per_node=G,
        node_gpu_count=np.full(N, G, np.int32), gpu_model=np.zeros(NG, np.int32),
        gpu_index=np.tile(np.arange(G, dtype=np.int32), N), gpu_memory_gb=np.full(NG, mem_gb, np.int32),
        gpu_used=used.astype(np.uint8), gpu_free=free.astype(np.uint8), gpu_used_mask=umask, gpu_free_mask=fmask,
        other_resources=[], node_alloc=alloc, node_requested=req,
        node_allowed_pods=np.full(N, allowed_pods, np.int32), node_num_pods=np.zeros(N, np.int32))
    P = n_pods
    d = _randint(seed, P, 20, 0, len(demand_sizes) - 1)
    col = np.array([3 + pid[s] for s in demand_sizes], np.int32)[d]
    prio = np.array([0, 100, 1000], np.int32)[_choice(seed, P, 21, [.7, .2, .1])]
    base = np.zeros((P, 3), np.int64)
    base[:, 0] = _randint(seed, P, 22, 100, 4000)
    base[:, 1] = _randint(seed, P, 23, 1, 32) << 30
    pods = pk.PackedPods(keys=[f"ns/pod-{i}" for i in range(P)], priority=prio, tracker_key=np.arange(P, dtype=np.int32),
                         base=base, scalar_ptr=np.arange(P + 1, dtype=np.int32), scalar_col=col.astype(np.int32),
                         scalar_qty=np.ones(P, np.int64))
    return snap, pods


def to_logical(snap: pk.PackedSnapshot, pods: pk.PackedPods):
    """packed -> logical dicts (small cases only; used to feed the faithful Python model in tests)."""
    G, K = snap.gpus_per_node, snap.n_profiles
    cols = ["cpu", "memory", "ephemeral-storage"] + [pk.resource_of_profile(snap.kind, p) for p in snap.profiles] + snap.other_resources
    nodes = []
    for i, name in enumerate(snap.node_names):
        gpus = []
        for j in range(int(snap.node_gpu_count[i])):
            x = i * G + j
            used = {snap.profiles[k]: int(snap.gpu_used[x, k]) for k in range(K) if snap.gpu_used_mask[x] >> k & 1}
            free = {snap.profiles[k]: int(snap.gpu_free[x, k]) for k in range(K) if snap.gpu_free_mask[x] >> k & 1}
            gpus.append({"model": snap.models[int(snap.gpu_model[x])], "index": int(snap.gpu_index[x]),
                         "memory_gb": int(snap.gpu_memory_gb[x]), "used": used, "free": free})
        alloc = {c: int(v) for c, v in zip(cols, snap.node_alloc[i]) if v}
        alloc["pods"] = int(snap.node_allowed_pods[i])
        req = {c: int(v) for c, v in zip(cols, snap.node_requested[i]) if v}
        nodes.append({"name": name, "gpus": gpus, "allocatable": alloc, "requested": req, "num_pods": int(snap.node_num_pods[i])})
    lpods = []
    for i in range(pods.n_pods):
        ns, nm = pods.keys[i].split("/", 1)
        c = {}
        for k, v in zip(("cpu", "memory", "ephemeral-storage"), pods.base[i]):
            if v:
                c[k] = int(v)
        for e in range(int(pods.scalar_ptr[i]), int(pods.scalar_ptr[i + 1])):
            c[cols[int(pods.scalar_col[e])]] = int(pods.scalar_qty[e])
        lpods.append({"namespace": ns, "name": nm, "priority": int(pods.priority[i]), "containers": [c]})
    return nodes, lpods


def quota_state(n_quotas: int, n_running: int, n_pending: int, seed: int, nvidia_gpu_memory_gb: int = 32, zipf_s: float = 1.1):
    """BASELINE config #4 recipe in packed form: n_quotas ElasticQuotas (10 % composite over 2-4 namespaces), Min of
    cpu U{8..256} cores / memory / gpu-memory U{40..640}, Max = 2*Min enforced on half of them; running and pending pods
    with namespaces ~ Zipf(s), unique creation timestamps, C2 priorities and requests plus nvidia.com/gpu U{0..2}.
    -> (PackedQuotas, PackedQuotaPods running, PackedQuotaPods pending)"""
    from . import quota_packing as qp
    cols = ["cpu", "memory", "ephemeral-storage", "pods", "nos.nebuly.com/gpu-memory", "nvidia.com/gpu"]
    R, gm, gc = len(cols), 4, 5
    E = n_quotas
    composite = _uniform(seed, E, 1) < 0.10
    n_ns_of = np.where(composite, _randint(seed, E, 2, 2, 4), 1)
    ns_first = np.concatenate([[0], np.cumsum(n_ns_of)])[:-1]
    NS = int(n_ns_of.sum()) + 1                                   # + one namespace without quota
    ns_quota = np.full(NS, -1, np.int32)
    for e in range(E):
        ns_quota[ns_first[e]:ns_first[e] + n_ns_of[e]] = e
    mn = np.zeros((E, R), np.int64)
    mn[:, 0] = _randint(seed, E, 3, 8, 256) * 1000
    mn[:, 1] = _randint(seed, E, 4, 16, 2048) << 30
    mn[:, gm] = _randint(seed, E, 5, 40, 640)
    mx = 2 * mn
    enforced = (_uniform(seed, E, 6) < 0.5).astype(np.uint8)
    bits = np.uint32(qp.NONNIL | 0b11 | (1 << gm))
    pq = qp.PackedQuotas([f"eq-{e}" for e in range(E)], [f"ns-{i}" for i in range(NS)], cols, gm, ns_quota, mn, mx,
                         np.zeros((E, R), np.int64), np.full(E, bits, np.uint32), np.full(E, bits, np.uint32),
                         np.full(E, qp.NONNIL, np.uint32), enforced, np.array([0, 0, 0, 0, 0, nvidia_gpu_memory_gb], np.int64))

    w = 1.0 / np.arange(1, NS + 1, dtype=np.float64) ** zipf_s
    cdf = np.cumsum(w / w.sum())
    perm = np.argsort(splitmix64(seed, NS, 7))                    # which namespace is the popular one is random

    def pods(P, stream, t0):
        ns = perm[np.searchsorted(cdf, _uniform(seed, P, stream), side="right").clip(0, NS - 1)].astype(np.int32)
        req = np.zeros((P, R), np.int64)
        req[:, 0] = _randint(seed, P, stream + 1, 100, 4000)
        req[:, 1] = _randint(seed, P, stream + 2, 1, 32) << 30
        req[:, gc] = _randint(seed, P, stream + 3, 0, 2)
        prio = np.array([0, 100, 1000], np.int32)[_choice(seed, P, stream + 4, [.7, .2, .1])]
        ts = t0 + np.argsort(splitmix64(seed, P, stream + 5)).astype(np.int64)          # unique timestamps
        return qp.PackedQuotaPods([f"ns/pod-{i}" for i in range(P)], ns, req, np.full(P, 1 << gc, np.uint32), prio, ts,
                                  np.arange(P, dtype=np.int32))

    return pq, pods(n_running, 20, 1_700_000_000), pods(n_pending, 40, 1_800_000_000)
