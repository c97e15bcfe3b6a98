"""Logical ElasticQuota state -> packed quota tensors of include/nosplan.h (nosplan_quotas / nosplan_quota_pods).

Logical schema:
  quotas = {quota name: {"namespaces": [ns, ...], "min": {res: n} | None, "max": {res: n} | None, "used": {res: n} | None,
                         "max_enforced": bool}}
  pod    = {"namespace", "name", "priority", "creation_ts", "request": {res: n}}   (request WITHOUT gpu-memory)
cpu in milli-units, "pods" = AllowedPodNumber, other names are scalar resources.
"""
from __future__ import annotations

import ctypes as C
import re
from dataclasses import dataclass
from typing import Dict, List, Optional, Sequence

import numpy as np

CPU, MEMORY, EPHEMERAL, PODS = "cpu", "memory", "ephemeral-storage", "pods"
GPU_MEMORY = "nos.nebuly.com/gpu-memory"
NVIDIA_GPU = "nvidia.com/gpu"
MAX_RES = 16
NONNIL = 0x80000000
_RE_MIG = re.compile(r"nvidia\.com\/mig-\d+g\.\d+gb")
_RE_GB = re.compile(r"\d+gb")
_FIXED = {CPU: 0, MEMORY: 1, EPHEMERAL: 2, PODS: 3}


def mig_memory_gb(resource: str) -> int:
    """mig.ExtractMemoryGBFromMigFormat (pkg/gpu/mig/util.go:68-86); 0 where the reference returns an error."""
    if _RE_MIG.search(resource) is None:
        return 0
    m = _RE_GB.findall(resource)
    return int(re.search(r"\d+", m[0]).group(0)) if len(m) == 1 else 0


@dataclass
class PackedQuotas:
    names: List[str]
    namespaces: List[str]
    columns: List[str]
    gpu_memory_col: int
    ns_quota: np.ndarray
    min: np.ndarray
    max: np.ndarray
    used_init: np.ndarray
    min_mask: np.ndarray
    max_mask: np.ndarray
    used_mask: np.ndarray
    max_enforced: np.ndarray
    gpu_memory_coeff: np.ndarray

    @property
    def n_res(self):
        return len(self.columns)


@dataclass
class PackedQuotaPods:
    keys: List[str]
    ns: np.ndarray
    request: np.ndarray
    request_mask: np.ndarray
    priority: np.ndarray
    creation_ts: np.ndarray
    name_rank: np.ndarray

    @property
    def n_pods(self):
        return len(self.ns)


def pack_quotas(quotas: Dict[str, dict], pod_sets: Sequence[Sequence[dict]], nvidia_gpu_memory_gb: int,
                extra_namespaces: Sequence[str] = ()):
    """-> (PackedQuotas, [PackedQuotaPods per pod set])"""
    scalars: List[str] = []

    def note(rl):
        for r in (rl or {}):
            if r not in _FIXED and r not in scalars:
                scalars.append(r)

    for e in quotas.values():
        note(e.get("min")); note(e.get("max")); note(e.get("used"))
    for pods in pod_sets:
        for p in pods:
            note(p.get("request"))
    if GPU_MEMORY not in scalars:
        scalars.append(GPU_MEMORY)
    scalars.sort()
    columns = [CPU, MEMORY, EPHEMERAL, PODS] + scalars
    if len(columns) > MAX_RES:
        raise ValueError(f"{len(columns)} quota resource columns > {MAX_RES}")
    col = {r: i for i, r in enumerate(columns)}
    R = len(columns)
    gm = col[GPU_MEMORY]
    coeff = np.zeros(R, np.int64)
    for r, c in col.items():
        if r == NVIDIA_GPU:
            coeff[c] = nvidia_gpu_memory_gb
        elif _RE_MIG.search(r):
            coeff[c] = mig_memory_gb(r)

    names = list(quotas)
    namespaces: List[str] = []
    for e in quotas.values():
        for ns in e.get("namespaces", []):
            if ns not in namespaces:
                namespaces.append(ns)
    for pods in pod_sets:
        for p in pods:
            if p["namespace"] not in namespaces:
                namespaces.append(p["namespace"])
    for ns in extra_namespaces:
        if ns not in namespaces:
            namespaces.append(ns)
    nsid = {n: i for i, n in enumerate(namespaces)}
    E = len(names)
    ns_quota = np.full(len(namespaces), -1, np.int32)
    mn = np.zeros((E, R), np.int64); mx = np.zeros((E, R), np.int64); us = np.zeros((E, R), np.int64)
    mmask = np.zeros(E, np.uint32); xmask = np.zeros(E, np.uint32); umask = np.zeros(E, np.uint32)
    enforced = np.zeros(E, np.uint8)

    def fill(row, rl):
        m = 0
        if rl is None:
            return 0
        for r, v in rl.items():
            row[col[r]] += v
            m |= 1 << col[r]
        return m | NONNIL

    for i, (name, e) in enumerate(quotas.items()):
        for ns in e.get("namespaces", []):
            ns_quota[nsid[ns]] = i                      # later quotas overwrite, like ElasticQuotaInfos.Add
        mmask[i] = fill(mn[i], e.get("min"))
        xmask[i] = fill(mx[i], e.get("max"))
        umask[i] = fill(us[i], e.get("used"))
        enforced[i] = 1 if e.get("max_enforced") else 0
    pq = PackedQuotas(names, namespaces, columns, gm, ns_quota, mn, mx, us, mmask, xmask, umask, enforced, coeff)

    packed_pods = []
    for pods in pod_sets:
        P = len(pods)
        req = np.zeros((P, R), np.int64)
        rmask = np.zeros(P, np.uint32)
        for i, p in enumerate(pods):
            for r, v in (p.get("request") or {}).items():
                if r == GPU_MEMORY:
                    continue
                req[i, col[r]] += v
                if col[r] >= 4:
                    rmask[i] |= np.uint32(1 << col[r])
        order = sorted(range(P), key=lambda i: pods[i]["name"].encode())
        rank = np.zeros(P, np.int32)
        prev, r = None, -1
        for i in order:
            if pods[i]["name"] != prev:
                r += 1
                prev = pods[i]["name"]
            rank[i] = r
        packed_pods.append(PackedQuotaPods(
            [f"{p['namespace']}/{p['name']}" for p in pods],
            np.array([nsid[p["namespace"]] for p in pods], np.int32).reshape(P), req, rmask,
            np.array([p.get("priority") or 0 for p in pods], np.int32).reshape(P),
            np.array([p.get("creation_ts", 0) for p in pods], np.int64).reshape(P), rank))
    return pq, packed_pods


# ------------------------------------------------------------------------------------------------------------
# ctypes mirrors
# ------------------------------------------------------------------------------------------------------------
class CQuotas(C.Structure):
    _fields_ = [("n_quotas", C.c_int32), ("n_res", C.c_int32), ("n_namespaces", C.c_int32), ("gpu_memory_col", C.c_int32),
                ("ns_quota", C.c_void_p), ("min", C.c_void_p), ("max", C.c_void_p), ("used_init", C.c_void_p),
                ("min_mask", C.c_void_p), ("max_mask", C.c_void_p), ("used_mask", C.c_void_p), ("max_enforced", C.c_void_p),
                ("gpu_memory_coeff", C.c_void_p)]


class CQuotaPods(C.Structure):
    _fields_ = [("n_pods", C.c_int32), ("ns", C.c_void_p), ("request", C.c_void_p), ("request_mask", C.c_void_p),
                ("priority", C.c_void_p), ("creation_ts", C.c_void_p), ("name_rank", C.c_void_p)]


class CQuotaOut(C.Structure):
    _fields_ = [("used", C.c_void_p), ("used_mask", C.c_void_p), ("agg_min", C.c_void_p), ("agg_used", C.c_void_p),
                ("agg_overquota", C.c_void_p), ("agg_masks", C.c_void_p), ("guaranteed", C.c_void_p), ("prefilter", C.c_void_p),
                ("over_quota", C.c_void_p), ("operator_used", C.c_void_p), ("ambiguous_order", C.c_void_p)]


def c_quotas(pq: PackedQuotas):
    keep = []

    def p(a, dt):
        a = np.ascontiguousarray(a, dtype=dt)
        keep.append(a)
        return a.ctypes.data

    cq = CQuotas(len(pq.names), pq.n_res, len(pq.namespaces), pq.gpu_memory_col,
                 p(pq.ns_quota, np.int32), p(pq.min, np.int64), p(pq.max, np.int64), p(pq.used_init, np.int64),
                 p(pq.min_mask, np.uint32), p(pq.max_mask, np.uint32), p(pq.used_mask, np.uint32), p(pq.max_enforced, np.uint8),
                 p(pq.gpu_memory_coeff, np.int64))
    return cq, keep


def c_quota_pods(pp: Optional[PackedQuotaPods]):
    if pp is None:
        return CQuotaPods(0, None, None, None, None, None, None), []
    keep = []

    def p(a, dt):
        a = np.ascontiguousarray(a, dtype=dt)
        keep.append(a)
        return a.ctypes.data

    return CQuotaPods(pp.n_pods, p(pp.ns, np.int32), p(pp.request, np.int64), p(pp.request_mask, np.uint32),
                      p(pp.priority, np.int32), p(pp.creation_ts, np.int64), p(pp.name_rank, np.int32)), keep


class QuotaResult:
    def __init__(self, pq: PackedQuotas, n_running: int, n_pending: int):
        E, R = len(pq.names), pq.n_res
        self.used = np.zeros((max(E, 1), R), np.int64)
        self.used_mask = np.zeros(max(E, 1), np.uint32)
        self.agg_min = np.zeros(R, np.int64); self.agg_used = np.zeros(R, np.int64); self.agg_overquota = np.zeros(R, np.int64)
        self.agg_masks = np.zeros(3, np.uint32)
        self.guaranteed = np.zeros((max(E, 1), R), np.int64)
        self.prefilter = np.zeros(max(n_pending, 1), np.uint8)
        self.over_quota = np.zeros(max(n_running, 1), np.uint8)
        self.operator_used = np.zeros((max(E, 1), R), np.int64)
        self.ambiguous_order = np.zeros(1, np.int32)
        self.E, self.n_running, self.n_pending = E, n_running, n_pending

    def c_struct(self):
        return CQuotaOut(*[getattr(self, n).ctypes.data for n, _ in CQuotaOut._fields_])

    def trimmed(self):
        return {"used": self.used[:self.E], "used_mask": self.used_mask[:self.E], "agg_min": self.agg_min, "agg_used": self.agg_used,
                "agg_overquota": self.agg_overquota, "agg_masks": self.agg_masks, "guaranteed": self.guaranteed[:self.E],
                "prefilter": self.prefilter[:self.n_pending], "over_quota": self.over_quota[:self.n_running],
                "operator_used": self.operator_used[:self.E], "ambiguous_order": int(self.ambiguous_order[0])}
