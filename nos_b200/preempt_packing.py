"""nosplan_preempt_query for nosplan_preempt_dryrun: ctypes mirror + packer (the victim query of victim_packing plus the node
accounts, the preemptor's requests and the PodDisruptionBudget matches)."""
from __future__ import annotations

import ctypes as C
from dataclasses import dataclass
from typing import Dict, List, Sequence

import numpy as np

from . import quota_packing as qp, victim_packing as vp

VICTIMS_OK, VICTIMS_NONE, VICTIMS_FILTER, VICTIMS_MAX_QUOTA, VICTIMS_MIN_QUOTA, VICTIMS_ERROR = 0, 1, 2, 3, 4, 5


class CPreemptQuery(C.Structure):
    _fields_ = [("v", vp.CVictimQuery), ("pod_req", C.c_void_p), ("nominated", C.c_void_p), ("fit_request", C.c_void_p),
                ("fit_mask", C.c_uint32), ("n_pdbs", C.c_int32), ("node_alloc", C.c_void_p), ("node_requested", C.c_void_p),
                ("pdb_allowed", C.c_void_p), ("pod_pdb_ptr", C.c_void_p), ("pod_pdb", C.c_void_p)]


@dataclass
class PackedPreemptQuery:
    v: vp.PackedVictimQuery
    pod_req: np.ndarray
    nominated: np.ndarray
    fit_request: np.ndarray
    fit_mask: int
    node_alloc: np.ndarray
    node_requested: np.ndarray
    pdb_allowed: np.ndarray
    pod_pdb_ptr: np.ndarray
    pod_pdb: np.ndarray

    def c_struct(self):
        cv, keep = self.v.c_struct()

        def p(a, dt):
            a = np.ascontiguousarray(a, dtype=dt)
            if a.size == 0:
                a = np.zeros(1, dt)
            keep.append(a)
            return a.ctypes.data

        cs = CPreemptQuery(cv, p(self.pod_req, np.int64), p(self.nominated, np.int64), p(self.fit_request, np.int64), int(self.fit_mask),
                           len(self.pdb_allowed), p(self.node_alloc, np.int64), p(self.node_requested, np.int64),
                           p(self.pdb_allowed, np.int32), p(self.pod_pdb_ptr, np.int32), p(self.pod_pdb, np.int32))
        return cs, keep


def _row(pq: qp.PackedQuotas, rl: Dict[str, int], pods_key: str = "pods"):
    """Dense row over the quota columns of a plain resource list (+ the mask of its scalar keys)."""
    row = np.zeros(pq.n_res, np.int64)
    mask = 0
    for r, v in rl.items():
        if r == qp.GPU_MEMORY:
            continue
        c = pq.columns.index(r)
        row[c] += v
        if c >= 4:
            mask |= 1 << c
    return row, mask


def pack_preempt_query(pq: qp.PackedQuotas, preemptor: dict, nominated_in_eq: Dict[str, int], nominated: Dict[str, int],
                       nodes: Sequence[dict], pdb_allowed: Sequence[int]) -> PackedPreemptQuery:
    """nodes = [{"allocatable": {res: n, "pods": n}, "pods": [pod...], "pdbs_of_pod": [[pdb id...] per pod]}]; Requested and
    len(Pods) are derived from the pods (NodeInfo keeps them consistent)."""
    v = vp.pack_victim_query(pq, preemptor, nominated_in_eq, [n["pods"] for n in nodes])
    R = pq.n_res
    alloc = np.zeros((len(nodes), R), np.int64)
    req = np.zeros((len(nodes), R), np.int64)
    ptr, ids = [0], []
    for k, n in enumerate(nodes):
        alloc[k], _ = _row(pq, n["allocatable"])
        for p in n["pods"]:
            r, _ = _row(pq, {a: b for a, b in p["request"].items() if a != "pods"})
            req[k] += r
        req[k, 3] = len(n["pods"])
        for matches in n["pdbs_of_pod"]:
            ids += list(matches)
            ptr.append(len(ids))
    fit, fit_mask = _row(pq, {a: b for a, b in preemptor["request"].items() if a != "pods"})
    nom = np.zeros(R, np.int64)
    for r, val in nominated.items():
        nom[pq.columns.index(r)] += val
    return PackedPreemptQuery(v, vp.quota_request_row(pq, preemptor["request"]), nom, fit, fit_mask, alloc, req,
                              np.array(list(pdb_allowed), np.int32), np.array(ptr, np.int32), np.array(ids, np.int32))
