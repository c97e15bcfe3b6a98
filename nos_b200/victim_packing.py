"""nosplan_victim_query for nosplan_quota_victims: ctypes mirror + packer from logical node pod lists.

  node_pods = [[pod, ...] per node], pod = {"namespace", "priority", "over_quota": bool, "request": {res: n}}
Pods of a node must already be in the reference's processing order (least important first,
capacity_scheduling.go:516)."""
from __future__ import annotations

import ctypes as C
from dataclasses import dataclass
from typing import Dict, List, Sequence

import numpy as np

from . import quota_packing as qp


class CVictimQuery(C.Structure):
    _fields_ = [("preemptor_ns", C.c_int32), ("preemptor_priority", C.c_int32), ("nominated_in_eq", C.c_void_p),
                ("n_nodes", C.c_int32), ("node_ptr", C.c_void_p), ("n_pods", C.c_int32), ("ns", C.c_void_p), ("priority", C.c_void_p),
                ("over_quota", C.c_void_p), ("request", C.c_void_p), ("request_mask", C.c_void_p)]


@dataclass
class PackedVictimQuery:
    preemptor_ns: int
    preemptor_priority: int
    nominated_in_eq: np.ndarray
    node_ptr: np.ndarray
    ns: np.ndarray
    priority: np.ndarray
    over_quota: np.ndarray
    request: np.ndarray
    request_mask: np.ndarray

    @property
    def n_pods(self):
        return len(self.ns)

    def c_struct(self):
        keep = []

        def p(a, dt):
            a = np.ascontiguousarray(a, dtype=dt)
            keep.append(a)
            return a.ctypes.data

        cs = CVictimQuery(self.preemptor_ns, self.preemptor_priority, p(self.nominated_in_eq, np.int64), len(self.node_ptr) - 1,
                          p(self.node_ptr, np.int32), self.n_pods, p(self.ns, np.int32), p(self.priority, np.int32),
                          p(self.over_quota, np.uint8), p(self.request, np.int64), p(self.request_mask, np.uint32))
        return cs, keep


def quota_request_row(pq: qp.PackedQuotas, request: Dict[str, int]) -> np.ndarray:
    """Dense quota request with the gpu-memory column filled (pkg/gpu/util/resource.go:60-86)."""
    row = np.zeros(pq.n_res, np.int64)
    for r, v in request.items():
        if r != qp.GPU_MEMORY:
            row[pq.columns.index(r)] += v
    row[pq.gpu_memory_col] = int((pq.gpu_memory_coeff * row).sum())
    return row


def pack_victim_query(pq: qp.PackedQuotas, preemptor: dict, nominated_in_eq: Dict[str, int],
                      node_pods: Sequence[Sequence[dict]]) -> PackedVictimQuery:
    R = pq.n_res
    nsid = {n: i for i, n in enumerate(pq.namespaces)}
    flat = [p for pods in node_pods for p in pods]
    ptr = np.zeros(len(node_pods) + 1, np.int32)
    ptr[1:] = np.cumsum([len(p) for p in node_pods])
    P = len(flat)
    req = np.zeros((P, R), np.int64)
    mask = np.zeros(P, np.uint32)
    for i, p in enumerate(flat):
        for r, v in p["request"].items():
            if r == qp.GPU_MEMORY:
                continue
            c = pq.columns.index(r)
            req[i, c] += v
            if c >= 4:
                mask[i] |= np.uint32(1 << c)
    nom = np.zeros(R, np.int64)
    for r, v in nominated_in_eq.items():
        nom[pq.columns.index(r)] += v
    return PackedVictimQuery(nsid.get(preemptor["namespace"], -1), int(preemptor.get("priority") or 0), nom, ptr,
                             np.array([nsid.get(p["namespace"], -1) for p in flat], np.int32).reshape(P),
                             np.array([p.get("priority") or 0 for p in flat], np.int32).reshape(P),
                             np.array([1 if p.get("over_quota") else 0 for p in flat], np.uint8).reshape(P), req, mask)
