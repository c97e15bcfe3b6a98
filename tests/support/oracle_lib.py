"""ctypes wrapper of the CPU oracle (oracle/libnosoracle.so) -- test infrastructure."""
import ctypes as C
import os
import subprocess

import numpy as np

from nos_b200 import capi

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
ORACLE_DIR = os.path.join(ROOT, "oracle")
FAITHFUL, INCREMENTAL, DEVICE_MODEL, FUSED_MODEL = 0, 1, 2, 3
_lib = None


def lib():
    global _lib
    if _lib is None:
        path = os.path.join(ORACLE_DIR, "libnosoracle.so")
        srcs = [os.path.join(ORACLE_DIR, f) for f in ("nos_oracle.cpp", "nos_quota_oracle.cpp", "nos_victims_oracle.cpp", "nos_preempt_oracle.cpp")] + \
               [os.path.join(ROOT, "include", "nosplan.h")]
        if not os.path.exists(path) or any(os.path.getmtime(s) > os.path.getmtime(path) for s in srcs if os.path.exists(s)):
            subprocess.check_call(["make", "-C", ORACLE_DIR, "-s"])
        _lib = C.CDLL(path)
        _lib.nos_oracle_plan.argtypes = [C.POINTER(capi.CSnapshot), C.POINTER(capi.CPods), C.c_void_p, C.c_int32, C.c_int32,
                                         C.POINTER(capi.CPartitioning), C.POINTER(capi.CStats), C.POINTER(C.c_double)]
        _lib.nos_oracle_plan.restype = C.c_int
        _lib.nos_oracle_sort_pods.argtypes = [C.POINTER(capi.CSnapshot), C.POINTER(capi.CPods), C.c_void_p]
        _lib.nos_oracle_sort_pods.restype = C.c_int
    return _lib


def plan(snap, pods, prefilter_ok=None, filter_mode=0, mode=INCREMENTAL):
    """-> (gpu_count, gpu_mask, stats dict, seconds)"""
    cs, k1 = capi.c_snapshot(snap)
    cp, k2 = capi.c_pods(pods)
    out, cnt, msk = capi.alloc_partitioning(snap)
    st = capi.CStats()
    sec = C.c_double()
    pre = None
    if prefilter_ok is not None:
        pa = np.ascontiguousarray(np.asarray(prefilter_ok, dtype=np.uint8))
        pre = pa.ctypes.data
    rc = lib().nos_oracle_plan(C.byref(cs), C.byref(cp), pre, filter_mode, mode, C.byref(out), C.byref(st), C.byref(sec))
    if rc == 1 and mode in (DEVICE_MODEL, FUSED_MODEL):
        return None, None, None, sec.value          # the device would take the exact sweep: nothing to model
    if rc != 0:
        raise RuntimeError(f"nos_oracle_plan -> {rc}")
    return cnt, msk, st.as_dict(), sec.value


def sort_pods(snap, pods):
    cs, k1 = capi.c_snapshot(snap)
    cp, k2 = capi.c_pods(pods)
    order = np.zeros(max(1, pods.n_pods), np.int32)
    rc = lib().nos_oracle_sort_pods(C.byref(cs), C.byref(cp), order.ctypes.data)
    if rc != 0:
        raise RuntimeError(f"nos_oracle_sort_pods -> {rc}")
    return order[:pods.n_pods]


def quota(pq, running=None, pending=None):
    """CPU quota oracle on packed quota structs -> dict of arrays (see nos_b200.quota_packing.QuotaResult.trimmed)."""
    from nos_b200 import quota_packing as qp
    L = lib()
    L.nos_oracle_quota.argtypes = [C.POINTER(qp.CQuotas), C.POINTER(qp.CQuotaPods), C.POINTER(qp.CQuotaPods),
                                   C.POINTER(qp.CQuotaOut), C.POINTER(C.c_double)]
    L.nos_oracle_quota.restype = C.c_int
    cq, k1 = qp.c_quotas(pq)
    cr, k2 = qp.c_quota_pods(running)
    cp, k3 = qp.c_quota_pods(pending)
    res = qp.QuotaResult(pq, running.n_pods if running else 0, pending.n_pods if pending else 0)
    out = res.c_struct()
    sec = C.c_double()
    rc = L.nos_oracle_quota(C.byref(cq), C.byref(cr), C.byref(cp), C.byref(out), C.byref(sec))
    if rc != 0:
        raise RuntimeError(f"nos_oracle_quota -> {rc}")
    d = res.trimmed()
    d["seconds"] = sec.value
    return d


def preempt_dryrun(pq, query):
    """CPU oracle of the whole SelectVictimsOnNode on packed structs -> (victim [P] u8, node_status [N] i32, num_violating [N] i32)."""
    from nos_b200 import preempt_packing as pp, quota_packing as qp
    L = lib()
    L.nos_oracle_preempt_dryrun.argtypes = [C.POINTER(qp.CQuotas), C.POINTER(pp.CPreemptQuery), C.c_void_p, C.c_void_p, C.c_void_p]
    L.nos_oracle_preempt_dryrun.restype = C.c_int
    cq, k1 = qp.c_quotas(pq)
    cv, k2 = query.c_struct()
    n_nodes = len(query.v.node_ptr) - 1
    victim = np.zeros(max(1, query.v.n_pods), np.uint8)
    status = np.zeros(max(1, n_nodes), np.int32)
    nviol = np.zeros(max(1, n_nodes), np.int32)
    rc = L.nos_oracle_preempt_dryrun(C.byref(cq), C.byref(cv), victim.ctypes.data, status.ctypes.data, nviol.ctypes.data)
    if rc != 0:
        raise RuntimeError(f"nos_oracle_preempt_dryrun -> {rc}")
    return victim[:query.v.n_pods], status[:n_nodes], nviol[:n_nodes]


def label_changes(pq, running, current):
    """CPU oracle of patchCapacityInfoIfDifferent over every quota -> (pod indices int32, desired labels uint8), in patch order."""
    from nos_b200 import quota_packing as qp
    L = lib()
    L.nos_oracle_label_changes.argtypes = [C.POINTER(qp.CQuotas), C.POINTER(qp.CQuotaPods), C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p]
    L.nos_oracle_label_changes.restype = C.c_int
    cq, k1 = qp.c_quotas(pq)
    cr, k2 = qp.c_quota_pods(running)
    cur = np.ascontiguousarray(current, dtype=np.uint8)
    if cur.size == 0:
        cur = np.zeros(1, np.uint8)
    pod = np.zeros(max(1, running.n_pods), np.int32)
    desired = np.zeros(max(1, running.n_pods), np.uint8)
    n = np.zeros(1, np.int32)
    rc = L.nos_oracle_label_changes(C.byref(cq), C.byref(cr), cur.ctypes.data, pod.ctypes.data, desired.ctypes.data, n.ctypes.data)
    if rc != 0:
        raise RuntimeError(f"nos_oracle_label_changes -> {rc}")
    return pod[:n[0]].copy(), desired[:n[0]].copy()
