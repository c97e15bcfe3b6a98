"""ctypes binding of include/nosplan.h.

`Engine` wraps libnosplan.so (the CUDA product).  It fails loudly if the library is missing or no CUDA
device is present: there is no CPU fallback.  The struct builders are also used by the test-suite to
call the CPU oracle (oracle/libnosoracle.so) with the very same packed buffers.
"""
from __future__ import annotations

import ctypes as C
import os
from typing import Tuple

import numpy as np

from .packing import PackedPods, PackedSnapshot

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.environ.get("NOSPLAN_LIB") or os.path.join(_HERE, "libnosplan.so")

EXPORTED_SYMBOLS = [
    "nosplan_create", "nosplan_destroy", "nosplan_last_error", "nosplan_plan",
    "nosplan_upload", "nosplan_run", "nosplan_download", "nosplan_stream",
    "nosplan_lacking_static", "nosplan_quota_recompute", "nosplan_quota_victims",
    "nosplan_shard_exchange_bytes", "nosplan_shard_begin", "nosplan_shard_end", "nosplan_device_outputs",
    "nosplan_nccl_unique_id", "nosplan_shard_init", "nosplan_run_sharded", "nosplan_download_sharded",
    "nosplan_host_pod_order", "nosplan_host_stage_pods", "nosplan_set_kernel_events",
    "nosplan_slice_begin", "nosplan_slice_mid", "nosplan_run_sliced", "nosplan_parse_annotations", "nosplan_shard_transport",
    "nosplan_preempt_dryrun", "nosplan_quota_label_changes", "nosplan_slice_chain_bytes", "nosplan_slice_chain",
    "nosplan_plan_resident", "nosplan_update_nodes", "nosplan_upload_pods", "nosplan_changed_nodes", "nosplan_preempt_dryrun_again", "nosplan_quota_state_begin", "nosplan_quota_state_apply", "nosplan_quota_state_prefilter",
    "nosplan_quota_recompute_sharded", "nosplan_plan_compact", "nosplan_compact_record_bytes",
]
SLICE_TOTALS_BYTES = 256
SLICED_NOT_APPLICABLE = 1


def host_pod_order(snap: PackedSnapshot, pods: PackedPods, lib_path: str = None) -> np.ndarray:
    """nosplan_host_pod_order: NewPodSorter order as the engine computes it (host only, no GPU needed)."""
    lib = C.CDLL(lib_path or LIB_PATH)
    lib.nosplan_host_pod_order.argtypes = [C.POINTER(CSnapshot), C.POINTER(CPods), C.c_void_p]
    lib.nosplan_host_pod_order.restype = C.c_int
    cs, k1 = c_snapshot(snap)
    cp, k2 = c_pods(pods)
    order = np.zeros(max(1, pods.n_pods), np.int32)
    rc = lib.nosplan_host_pod_order(C.byref(cs), C.byref(cp), order.ctypes.data)
    if rc != 0:
        raise NosplanError(f"nosplan_host_pod_order failed ({rc})")
    return order[:pods.n_pods]


STAGE_META = ("key_masks", "model_ranges", "runs", "runs_per_profile", "l3", "l2", "l1", "sfx", "cum", "pods", "end")


def host_stage_pods(snap: PackedSnapshot, pods: PackedPods, prefilter_ok=None, lib_path: str = None, repeat: int = 1):
    """nosplan_host_stage_pods: the pod structure of the single-launch kernel as nosplan_plan stages it (host only, no GPU).
    -> (blob uint8 array, meta dict with the section offsets, n_runs, pad_pods, cpu_shift, mem_shift, rset), or (None, None)
    when these pods take the general path.  repeat > 1 stages that many times (for timing)."""
    lib = C.CDLL(lib_path or LIB_PATH)
    lib.nosplan_host_stage_pods.argtypes = [C.POINTER(CSnapshot), C.POINTER(CPods), C.c_void_p, C.c_void_p, C.c_int64, C.c_void_p]
    lib.nosplan_host_stage_pods.restype = C.c_int64
    cs, k1 = c_snapshot(snap)
    cp, k2 = c_pods(pods)
    pre = None if prefilter_ok is None else np.ascontiguousarray(prefilter_ok, np.uint8)
    meta = np.zeros(16, np.int32)
    need = lib.nosplan_host_stage_pods(C.byref(cs), C.byref(cp), None if pre is None else pre.ctypes.data, None, 0, meta.ctypes.data)
    if need < 0:
        raise NosplanError(f"nosplan_host_stage_pods failed ({need})")
    if need == 0:
        return None, None
    blob = np.zeros(need, np.uint8)
    for _ in range(max(1, repeat)):
        got = lib.nosplan_host_stage_pods(C.byref(cs), C.byref(cp), None if pre is None else pre.ctypes.data, blob.ctypes.data, need, meta.ctypes.data)
    assert got == need
    m = {name: int(meta[k]) for k, name in enumerate(STAGE_META)}
    m.update(n_runs=int(meta[11]), pad_pods=int(meta[12]), cpu_shift=int(meta[13]), mem_shift=int(meta[14]), rset=int(meta[15]) & 0xffffffff)
    return blob, m


STATIC_MAX_KEYS = 16384


class CSnapshot(C.Structure):
    _fields_ = [
        ("kind", C.c_int32), ("n_profiles", C.c_int32),
        ("profile_mem_gb", C.c_void_p), ("profile_gi", C.c_void_p),
        ("n_models", C.c_int32), ("model_geom_offset", C.c_void_p), ("geom_counts", C.c_void_p),
        ("n_nodes", C.c_int32), ("gpus_per_node", C.c_int32),
        ("node_gpu_count", C.c_void_p), ("gpu_model", C.c_void_p), ("gpu_index", C.c_void_p),
        ("gpu_memory_gb", C.c_void_p), ("gpu_used", C.c_void_p), ("gpu_free", C.c_void_p),
        ("gpu_used_mask", C.c_void_p), ("gpu_free_mask", C.c_void_p),
        ("n_other", C.c_int32), ("node_alloc", C.c_void_p), ("node_requested", C.c_void_p),
        ("node_allowed_pods", C.c_void_p), ("node_num_pods", C.c_void_p),
    ]


class CSnapshotCompact(C.Structure):
    _fields_ = [
        ("kind", C.c_int32), ("n_profiles", C.c_int32), ("profile_mem_gb", C.c_void_p), ("profile_gi", C.c_void_p),
        ("n_models", C.c_int32), ("model_geom_offset", C.c_void_p), ("geom_counts", C.c_void_p),
        ("n_nodes", C.c_int32), ("gpus_per_node", C.c_int32), ("n_other", C.c_int32), ("record_bytes", C.c_int32),
        ("records", C.c_void_p),
    ]


def compact_record_bytes(gpus_per_node: int, n_other: int) -> int:
    return (128 + 16 * n_other + 24 * gpus_per_node + 15) & ~15


def compact_records(s: PackedSnapshot, out: np.ndarray = None) -> np.ndarray:
    """The node side of `s` as [N][record_bytes] uint8 (layout in include/nosplan.h, nosplan_snapshot_compact): what a packer
    writes instead of the twelve SoA arrays.  K <= 8 profiles, model ids < 256."""
    N, G, K, n_other = s.n_nodes, s.gpus_per_node, s.n_profiles, len(s.other_resources)
    if K > 8:
        raise ValueError("the compact form holds 8 profiles")
    rb = compact_record_bytes(G, n_other)
    rec = out if out is not None else np.zeros((N, rb), np.uint8)
    rec[:] = 0
    head = rec[:, 0:16].view(np.int32)
    head[:, 0], head[:, 1], head[:, 2] = s.node_gpu_count, s.node_allowed_pods, s.node_num_pods
    base = rec[:, 16:64].view(np.int64)
    base[:, 0:3], base[:, 3:6] = s.node_alloc[:, 0:3], s.node_requested[:, 0:3]
    sl = rec[:, 64:128].view(np.int32)
    sl[:, 0:K], sl[:, 8:8 + K] = s.node_alloc[:, 3:3 + K], s.node_requested[:, 3:3 + K]
    if n_other:
        ot = rec[:, 128:128 + 16 * n_other].view(np.int64)
        ot[:, 0:n_other], ot[:, n_other:2 * n_other] = s.node_alloc[:, 3 + K:], s.node_requested[:, 3 + K:]
    g0 = 128 + 16 * n_other
    gp = rec[:, g0:g0 + 24 * G].reshape(N, G, 24)
    gp[:, :, 0:8] = np.asarray(s.gpu_used).reshape(N, G, 16)[:, :, 0:8]
    gp[:, :, 8:16] = np.asarray(s.gpu_free).reshape(N, G, 16)[:, :, 0:8]
    gp[:, :, 16] = np.asarray(s.gpu_used_mask).reshape(N, G) & 0xff
    gp[:, :, 17] = np.asarray(s.gpu_free_mask).reshape(N, G) & 0xff
    gp[:, :, 18] = np.asarray(s.gpu_model).reshape(N, G) & 0xff
    mem = np.asarray(s.gpu_memory_gb).reshape(N, G).astype(np.uint16)
    gp[:, :, 20] = mem & 0xff
    gp[:, :, 21] = mem >> 8
    return rec


class CPods(C.Structure):
    _fields_ = [
        ("n_pods", C.c_int32), ("priority", C.c_void_p), ("tracker_key", C.c_void_p), ("base", C.c_void_p),
        ("scalar_ptr", C.c_void_p), ("scalar_col", C.c_void_p), ("scalar_qty", C.c_void_p),
    ]


class CPartitioning(C.Structure):
    _fields_ = [("gpu_count", C.c_void_p), ("gpu_mask", C.c_void_p)]


class CStats(C.Structure):
    _fields_ = [
        ("candidates", C.c_int64), ("visited_nodes", C.c_int64), ("geometry_evals", C.c_int64),
        ("placements", C.c_int64), ("candidate_nodes", C.c_int64),
        ("lacking_at_start", C.c_int32), ("lacking_at_end", C.c_int32),
        ("ms_h2d", C.c_float), ("ms_kernels", C.c_float), ("ms_d2h", C.c_float),
        ("ms_prepass", C.c_float), ("ms_tracker", C.c_float), ("ms_parallel", C.c_float), ("ms_sweep", C.c_float),
        ("waves", C.c_int32),
        ("kernel_launches", C.c_int32), ("path", C.c_int32),
        ("h2d_bytes", C.c_int64), ("d2h_bytes", C.c_int64),
        ("fused", C.c_int32), ("fused_reject", C.c_int32), ("gate_replayed", C.c_int32), ("reserved0", C.c_int32),
    ]

    def as_dict(self):
        return {n: getattr(self, n) for n, _ in self._fields_}


def _ptr(a: np.ndarray, dtype) -> Tuple[int, np.ndarray]:
    a = np.ascontiguousarray(a, dtype=dtype)
    return a.ctypes.data, a


def c_snapshot(s: PackedSnapshot):
    """-> (CSnapshot, keepalive list).  The arrays must stay alive while the struct is in use."""
    keep = []

    def p(a, dt):
        addr, arr = _ptr(a, dt)
        keep.append(arr)
        return addr

    cs = CSnapshot()
    cs.kind, cs.n_profiles = s.kind, s.n_profiles
    cs.profile_mem_gb, cs.profile_gi = p(s.profile_mem_gb, np.int32), p(s.profile_gi, np.int32)
    cs.n_models = len(s.models)
    cs.model_geom_offset, cs.geom_counts = p(s.model_geom_offset, np.int32), p(s.geom_counts, np.uint8)
    cs.n_nodes, cs.gpus_per_node = s.n_nodes, s.gpus_per_node
    cs.node_gpu_count = p(s.node_gpu_count, np.int32)
    cs.gpu_model, cs.gpu_index = p(s.gpu_model, np.int32), p(s.gpu_index, np.int32)
    cs.gpu_memory_gb = p(s.gpu_memory_gb, np.int32)
    cs.gpu_used, cs.gpu_free = p(s.gpu_used, np.uint8), p(s.gpu_free, np.uint8)
    cs.gpu_used_mask, cs.gpu_free_mask = p(s.gpu_used_mask, np.uint16), p(s.gpu_free_mask, np.uint16)
    cs.n_other = len(s.other_resources)
    cs.node_alloc, cs.node_requested = p(s.node_alloc, np.int64), p(s.node_requested, np.int64)
    cs.node_allowed_pods, cs.node_num_pods = p(s.node_allowed_pods, np.int32), p(s.node_num_pods, np.int32)
    return cs, keep


def c_pods(pp: PackedPods):
    keep = []

    def p(a, dt):
        addr, arr = _ptr(a, dt)
        keep.append(arr)
        return addr

    cp = CPods()
    cp.n_pods = pp.n_pods
    cp.priority, cp.tracker_key = p(pp.priority, np.int32), p(pp.tracker_key, np.int32)
    cp.base = p(pp.base, np.int64)
    cp.scalar_ptr, cp.scalar_col = p(pp.scalar_ptr, np.int32), p(pp.scalar_col, np.int32)
    cp.scalar_qty = p(pp.scalar_qty, np.int64)
    return cp, keep


_PINNED_FIELDS = ("node_gpu_count", "gpu_model", "gpu_index", "gpu_memory_gb", "gpu_used", "gpu_free", "gpu_used_mask",
                  "gpu_free_mask", "node_alloc", "node_requested", "node_allowed_pods", "node_num_pods")


def pin_snapshot(s: PackedSnapshot) -> PackedSnapshot:
    """Copy of `s` whose per-node arrays live in page-locked host memory (allocated through torch), so nosplan_upload can
    DMA them without staging.  The torch tensors are kept alive on the returned object."""
    import copy

    import torch
    out = copy.copy(s)
    keep = []
    for name in _PINNED_FIELDS:
        a = np.ascontiguousarray(getattr(s, name))
        t = torch.from_numpy(a.view(np.uint8).reshape(-1).copy() if a.size else np.zeros(0, np.uint8)).pin_memory()
        keep.append(t)
        setattr(out, name, t.numpy().view(a.dtype).reshape(a.shape))
    out._pinned_keep = keep
    return out


def refresh_pinned(pinned: PackedSnapshot, src: PackedSnapshot) -> None:
    """Write the per-node arrays of `src` into the page-locked arrays of `pinned` again -- what a caller does when it packs a
    fresh snapshot into the buffers it keeps (the Go shim's packer writes into its cudaHostAlloc'ed slices on every Plan())."""
    for name in _PINNED_FIELDS:
        a = getattr(pinned, name)
        if a.size:
            np.copyto(a, np.asarray(getattr(src, name)).reshape(a.shape))


def alloc_partitioning(s: PackedSnapshot, out=None):
    n = max(1, s.n_nodes * s.gpus_per_node)
    if out is None:
        cnt = np.zeros((n, 16), np.uint8)
        msk = np.zeros(n, np.uint16)
    else:
        cnt, msk = out
        assert cnt.dtype == np.uint8 and cnt.shape == (n, 16) and msk.dtype == np.uint16 and msk.shape == (n,)
        assert cnt.flags.c_contiguous and msk.flags.c_contiguous
    cp = CPartitioning(cnt.ctypes.data, msk.ctypes.data)
    return cp, cnt, msk


def pinned_partitioning(s: PackedSnapshot):
    """Plan output buffers in page-locked host memory (through torch): nosplan_download DMAs into them directly.
    -> (gpu_count [N*G,16] u8, gpu_mask [N*G] u16); pass as out= to Engine.plan / download."""
    import torch
    n = max(1, s.n_nodes * s.gpus_per_node)
    tc = torch.zeros(n * 16, dtype=torch.uint8).pin_memory()
    tm = torch.zeros(n * 2, dtype=torch.uint8).pin_memory()
    cnt = tc.numpy().reshape(n, 16)
    msk = tm.numpy().view(np.uint16)
    _KEEP_PINNED.append((tc, tm))
    return cnt, msk


_KEEP_PINNED = []


class CAnnotations(C.Structure):
    _fields_ = [("n_nodes", C.c_int32), ("node_ptr", C.c_void_p), ("key_off", C.c_void_p), ("key_bytes", C.c_void_p),
                ("val_off", C.c_void_p), ("val_bytes", C.c_void_p), ("gpu_count", C.c_void_p), ("n_profiles", C.c_int32),
                ("profile_off", C.c_void_p), ("profile_bytes", C.c_void_p), ("gpus_per_node", C.c_int32)]


class CParsedGpus(C.Structure):
    _fields_ = [("gpu_used", C.c_void_p), ("gpu_free", C.c_void_p), ("gpu_used_mask", C.c_void_p), ("gpu_free_mask", C.c_void_p),
                ("gpu_index", C.c_void_p), ("node_gpu_count", C.c_void_p), ("errors", C.c_void_p)]


def flatten_annotations(v1_nodes, profiles):
    """[{annotations: {k: v}, labels: {...}}] -> the flat buffers of nosplan_annotations (numpy arrays kept by the caller).
    gpu_count comes from the nvidia.com/gpu.count label (0 when absent or not an integer: NewNode would have failed)."""
    import re
    keys, vals, ptr, cnt = [], [], [0], []
    for n in v1_nodes:
        for k, v in n.get("annotations", {}).items():
            keys.append(k.encode()); vals.append(str(v).encode())
        ptr.append(len(keys))
        c = n.get("labels", {}).get("nvidia.com/gpu.count", "0")
        cnt.append(int(c) if re.fullmatch(r"[+-]?\d+", str(c)) else 0)

    def pack(strs):
        off = np.zeros(len(strs) + 1, np.int32)
        if strs:
            off[1:] = np.cumsum([len(x) for x in strs])
        return off, np.frombuffer(b"".join(strs) or b"\0", np.uint8).copy()

    ko, kb = pack(keys); vo, vb = pack(vals); po, pb = pack([p.encode() for p in profiles])
    return {"node_ptr": np.array(ptr, np.int32), "key_off": ko, "key_bytes": kb, "val_off": vo, "val_bytes": vb,
            "gpu_count": np.array(cnt, np.int32).reshape(len(v1_nodes)), "profile_off": po, "profile_bytes": pb,
            "n_nodes": len(v1_nodes), "n_profiles": len(profiles)}


class CLabelChanges(C.Structure):
    _fields_ = [("current", C.c_void_p), ("capacity", C.c_int32), ("pod", C.c_void_p), ("desired", C.c_void_p), ("n_changes", C.c_void_p)]


LABEL_OTHER, LABEL_IN_QUOTA, LABEL_OVER_QUOTA = 0, 1, 2


class CQuotaEvents(C.Structure):
    _fields_ = [("n_events", C.c_int32), ("op", C.c_void_p), ("pod_id", C.c_void_p), ("ns", C.c_void_p), ("request", C.c_void_p),
                ("request_mask", C.c_void_p)]


EVENT_ADD, EVENT_DELETE = 0, 1


class NosplanError(RuntimeError):
    pass


class SlicedNotApplicable(NosplanError):
    """nosplan_download returned NOSPLAN_SLICED_NOT_APPLICABLE (same verdict on every rank)."""


def load_library(path: str = LIB_PATH) -> C.CDLL:
    if not os.path.exists(path):
        raise NosplanError(f"{path} not found: build it with `python -c 'import __graft_entry__ as g; g.build()'` "
                           "(there is no CPU fallback)")
    lib = C.CDLL(path)
    lib.nosplan_create.argtypes = [C.c_int, C.POINTER(C.c_void_p)]
    lib.nosplan_create.restype = C.c_int
    lib.nosplan_destroy.argtypes = [C.c_void_p]
    lib.nosplan_destroy.restype = None
    lib.nosplan_last_error.argtypes = [C.c_void_p]
    lib.nosplan_last_error.restype = C.c_char_p
    lib.nosplan_plan.argtypes = [C.c_void_p, C.POINTER(CSnapshot), C.POINTER(CPods), C.c_void_p, C.c_int32,
                                 C.POINTER(CPartitioning), C.POINTER(CStats)]
    lib.nosplan_plan.restype = C.c_int
    lib.nosplan_upload.argtypes = [C.c_void_p, C.POINTER(CSnapshot), C.POINTER(CPods), C.c_void_p, C.c_int32]
    lib.nosplan_upload.restype = C.c_int
    lib.nosplan_run.argtypes = [C.c_void_p]
    lib.nosplan_run.restype = C.c_int
    lib.nosplan_download.argtypes = [C.c_void_p, C.POINTER(CPartitioning), C.POINTER(CStats)]
    lib.nosplan_download.restype = C.c_int
    lib.nosplan_stream.argtypes = [C.c_void_p]
    lib.nosplan_stream.restype = C.c_void_p
    lib.nosplan_lacking_static.argtypes = [C.c_int, C.c_void_p, C.c_int32, C.c_int32, C.c_int32, C.c_void_p, C.c_void_p,
                                           C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_char_p, C.c_int32]
    lib.nosplan_lacking_static.restype = C.c_int
    from . import quota_packing as qp
    lib.nosplan_quota_recompute.argtypes = [C.c_void_p, C.POINTER(qp.CQuotas), C.POINTER(qp.CQuotaPods), C.POINTER(qp.CQuotaPods),
                                            C.POINTER(qp.CQuotaOut), C.POINTER(CStats)]
    lib.nosplan_quota_recompute.restype = C.c_int
    lib.nosplan_quota_recompute_sharded.argtypes = lib.nosplan_quota_recompute.argtypes
    lib.nosplan_quota_recompute_sharded.restype = C.c_int
    lib.nosplan_shard_exchange_bytes.argtypes = [C.c_void_p]
    lib.nosplan_shard_exchange_bytes.restype = C.c_int64
    lib.nosplan_shard_begin.argtypes = [C.c_void_p, C.c_int32, C.c_int32, C.c_void_p, C.c_int64]
    lib.nosplan_shard_begin.restype = C.c_int
    lib.nosplan_shard_end.argtypes = [C.c_void_p, C.c_void_p, C.c_int32]
    lib.nosplan_shard_end.restype = C.c_int
    lib.nosplan_device_outputs.argtypes = [C.c_void_p, C.POINTER(C.c_void_p), C.POINTER(C.c_void_p)]
    lib.nosplan_device_outputs.restype = C.c_int
    lib.nosplan_nccl_unique_id.argtypes = [C.c_void_p]
    lib.nosplan_nccl_unique_id.restype = C.c_int
    lib.nosplan_shard_init.argtypes = [C.c_void_p, C.c_int32, C.c_int32, C.c_void_p]
    lib.nosplan_shard_init.restype = C.c_int
    lib.nosplan_run_sharded.argtypes = [C.c_void_p, C.c_int32]
    lib.nosplan_run_sharded.restype = C.c_int
    lib.nosplan_download_sharded.argtypes = [C.c_void_p, C.POINTER(CPartitioning), C.POINTER(CStats)]
    lib.nosplan_download_sharded.restype = C.c_int
    lib.nosplan_slice_begin.argtypes = [C.c_void_p, C.c_void_p]
    lib.nosplan_slice_begin.restype = C.c_int
    lib.nosplan_slice_mid.argtypes = [C.c_void_p, C.c_int32, C.c_int32, C.c_void_p, C.c_void_p, C.c_int64]
    lib.nosplan_slice_mid.restype = C.c_int
    lib.nosplan_run_sliced.argtypes = [C.c_void_p]
    lib.nosplan_run_sliced.restype = C.c_int
    lib.nosplan_shard_transport.argtypes = [C.c_void_p]
    lib.nosplan_shard_transport.restype = C.c_int
    lib.nosplan_parse_annotations.argtypes = [C.c_void_p, C.POINTER(CAnnotations), C.POINTER(CParsedGpus), C.POINTER(CStats)]
    lib.nosplan_parse_annotations.restype = C.c_int
    from . import victim_packing as vp
    lib.nosplan_quota_victims.argtypes = [C.c_void_p, C.POINTER(qp.CQuotas), C.POINTER(vp.CVictimQuery), C.c_void_p, C.POINTER(CStats)]
    lib.nosplan_quota_victims.restype = C.c_int
    from . import preempt_packing as pp
    lib.nosplan_preempt_dryrun.argtypes = [C.c_void_p, C.POINTER(qp.CQuotas), C.POINTER(pp.CPreemptQuery), C.c_void_p, C.c_void_p, C.c_void_p,
                                           C.POINTER(CStats)]
    lib.nosplan_preempt_dryrun.restype = C.c_int
    lib.nosplan_quota_label_changes.argtypes = [C.c_void_p, C.POINTER(qp.CQuotas), C.POINTER(qp.CQuotaPods), C.POINTER(CLabelChanges),
                                                C.POINTER(qp.CQuotaOut), C.POINTER(CStats)]
    lib.nosplan_quota_label_changes.restype = C.c_int
    return lib


def lacking_static(lib, device, stream, n_pods, n_keys, n_profiles, d_key, d_ptr, d_col, d_qty, d_totals, d_lack, d_entries):
    """nosplan_lacking_static on raw device addresses (ints).  Allocates nothing on the device; asynchronous on `stream`.
    Returns (rc, message)."""
    buf = C.create_string_buffer(256)
    rc = lib.nosplan_lacking_static(device, stream, n_pods, n_keys, n_profiles, d_key, d_ptr, d_col, d_qty, d_totals,
                                    d_lack, d_entries, buf, 256)
    return rc, buf.value.decode()


class Engine:
    """One nosplan handle on one CUDA device (nosplan_create .. nosplan_destroy)."""

    def __init__(self, device: int = 0, lib_path: str = LIB_PATH):
        self.h = None
        self.lib = load_library(lib_path)
        h = C.c_void_p()
        rc = self.lib.nosplan_create(device, C.byref(h))
        if rc != 0 or not h.value:
            raise NosplanError(f"nosplan_create(device={device}) failed with {rc}: no usable CUDA device "
                               "(libnosplan has no CPU fallback)")
        self.h = h
        self._snap = None

    def close(self):
        if getattr(self, "h", None) is not None and self.h.value:
            self.lib.nosplan_destroy(self.h)
            self.h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def _check(self, rc: int, what: str):
        if rc != 0:
            msg = self.lib.nosplan_last_error(self.h)
            raise NosplanError(f"{what} failed ({rc}): {msg.decode() if msg else ''}")

    @staticmethod
    def _pre(prefilter_ok, n):
        if prefilter_ok is None:
            return None, None
        a = np.ascontiguousarray(np.asarray(prefilter_ok, dtype=np.uint8).reshape(n))
        return a.ctypes.data, a

    def plan(self, snap: PackedSnapshot, pods: PackedPods, prefilter_ok=None, filter_mode: int = 0, out=None):
        """nosplan_plan: host buffers in, host buffers out.  -> (gpu_count [N*G,16], gpu_mask [N*G], stats dict)"""
        cs, k1 = c_snapshot(snap)
        cp, k2 = c_pods(pods)
        out, cnt, msk = alloc_partitioning(snap, out)
        st = CStats()
        pre, k3 = self._pre(prefilter_ok, pods.n_pods)
        rc = self.lib.nosplan_plan(self.h, C.byref(cs), C.byref(cp), pre, filter_mode, C.byref(out), C.byref(st))
        self._check(rc, "nosplan_plan")
        self._snap = snap
        del k1, k2, k3
        return cnt, msk, st.as_dict()

    def prepared_plan(self, snap: PackedSnapshot, pods: PackedPods, prefilter_ok=None, filter_mode: int = 0, out=None, resident=False):
        """The C structs of one nosplan_plan (or nosplan_plan_resident) call built once: -> call() that does nothing but enter
        the C-ABI (returns the raw return code; raise through .check), and result() -> (gpu_count, gpu_mask, stats dict).
        For timing the call itself, not this Python wrapper."""
        cs, k1 = c_snapshot(snap)
        cp, k2 = c_pods(pods)
        cout, cnt, msk = alloc_partitioning(snap, out)
        st = CStats()
        pre, k3 = self._pre(prefilter_ok, pods.n_pods)
        lib, h = self.lib, self.h
        lib.nosplan_plan_resident.argtypes = [C.c_void_p, C.POINTER(CPods), C.c_void_p, C.c_int32, C.POINTER(CPartitioning), C.POINTER(CStats)]
        lib.nosplan_plan_resident.restype = C.c_int
        a_cs, a_cp, a_out, a_st = C.byref(cs), C.byref(cp), C.byref(cout), C.byref(st)
        if resident:
            def call():
                return lib.nosplan_plan_resident(h, a_cp, pre, filter_mode, a_out, a_st)
        else:
            def call():
                return lib.nosplan_plan(h, a_cs, a_cp, pre, filter_mode, a_out, a_st)
        call.keep = (cs, cp, cout, st, k1, k2, k3, snap, pods)
        call.check = lambda rc: self._check(rc, "nosplan_plan")
        call.result = lambda: (cnt, msk, st.as_dict())
        self._snap = snap
        return call

    def prepared_plan_compact(self, snap: PackedSnapshot, records: np.ndarray, pods: PackedPods, prefilter_ok=None, filter_mode: int = 0, out=None):
        """nosplan_plan_compact with its C structs built once (see prepared_plan): `records` = compact_records(snap), possibly
        in page-locked memory."""
        keep = []

        def p(a, dt):
            addr, arr = _ptr(a, dt)
            keep.append(arr)
            return addr
        cc = CSnapshotCompact()
        cc.kind, cc.n_profiles = snap.kind, snap.n_profiles
        cc.profile_mem_gb, cc.profile_gi = p(snap.profile_mem_gb, np.int32), p(snap.profile_gi, np.int32)
        cc.n_models = len(snap.models)
        cc.model_geom_offset, cc.geom_counts = p(snap.model_geom_offset, np.int32), p(snap.geom_counts, np.uint8)
        cc.n_nodes, cc.gpus_per_node, cc.n_other = snap.n_nodes, snap.gpus_per_node, len(snap.other_resources)
        cc.record_bytes = records.shape[1] if records.ndim == 2 and records.shape[0] else compact_record_bytes(snap.gpus_per_node, len(snap.other_resources))
        cc.records = records.ctypes.data
        cp, k2 = c_pods(pods)
        cout, cnt, msk = alloc_partitioning(snap, out)
        st = CStats()
        pre, k3 = self._pre(prefilter_ok, pods.n_pods)
        lib, h = self.lib, self.h
        lib.nosplan_plan_compact.argtypes = [C.c_void_p, C.POINTER(CSnapshotCompact), C.POINTER(CPods), C.c_void_p, C.c_int32,
                                             C.POINTER(CPartitioning), C.POINTER(CStats)]
        lib.nosplan_plan_compact.restype = C.c_int
        a_cc, a_cp, a_out, a_st = C.byref(cc), C.byref(cp), C.byref(cout), C.byref(st)

        def call():
            return lib.nosplan_plan_compact(h, a_cc, a_cp, pre, filter_mode, a_out, a_st)
        call.keep = (cc, cp, cout, st, keep, k2, k3, snap, pods, records)
        call.check = lambda rc: self._check(rc, "nosplan_plan_compact")
        call.result = lambda: (cnt, msk, st.as_dict())
        self._snap = snap
        return call

    def plan_compact(self, snap: PackedSnapshot, pods: PackedPods, prefilter_ok=None, filter_mode: int = 0, out=None, records=None):
        """nosplan_plan_compact: the node side as one array of records.  -> (gpu_count, gpu_mask, stats dict)"""
        call = self.prepared_plan_compact(snap, records if records is not None else compact_records(snap), pods, prefilter_ok, filter_mode, out)
        call.check(call())
        return call.result()

    def plan_resident(self, pods: PackedPods, prefilter_ok=None, filter_mode: int = 0, out=None):
        """nosplan_plan_resident: new pods against the snapshot the last plan() / upload() left in HBM."""
        snap = self._snap
        cp, k2 = c_pods(pods)
        out, cnt, msk = alloc_partitioning(snap, out)
        st = CStats()
        pre, k3 = self._pre(prefilter_ok, pods.n_pods)
        self.lib.nosplan_plan_resident.argtypes = [C.c_void_p, C.POINTER(CPods), C.c_void_p, C.c_int32, C.POINTER(CPartitioning), C.POINTER(CStats)]
        self.lib.nosplan_plan_resident.restype = C.c_int
        rc = self.lib.nosplan_plan_resident(self.h, C.byref(cp), pre, filter_mode, C.byref(out), C.byref(st))
        self._check(rc, "nosplan_plan_resident")
        del k2, k3
        return cnt, msk, st.as_dict()

    def changed_nodes(self, capacity=None):
        """nosplan_changed_nodes -> indices (int32, ascending) of the nodes whose desired partitioning differs from the current one."""
        cap = self._snap.n_nodes if capacity is None else int(capacity)
        idx = np.zeros(max(1, cap), np.int32)
        n = np.zeros(1, np.int32)
        self.lib.nosplan_changed_nodes.argtypes = [C.c_void_p, C.c_void_p, C.c_int32, C.c_void_p]
        self.lib.nosplan_changed_nodes.restype = C.c_int
        rc = self.lib.nosplan_changed_nodes(self.h, idx.ctypes.data, cap, n.ctypes.data)
        if rc != 0:
            msg = self.lib.nosplan_last_error(self.h)
            e = NosplanError(f"nosplan_changed_nodes failed ({rc}): {msg.decode() if msg else ''}")
            e.partial = (idx[:min(int(n[0]), cap)].copy(), int(n[0]))
            raise e
        return idx[:int(n[0])]

    def upload_pods(self, pods: PackedPods, prefilter_ok=None, filter_mode: int = 0):
        """nosplan_upload_pods: new pods against the resident snapshot, kernels not run yet."""
        cp, k2 = c_pods(pods)
        pre, k3 = self._pre(prefilter_ok, pods.n_pods)
        self.lib.nosplan_upload_pods.argtypes = [C.c_void_p, C.POINTER(CPods), C.c_void_p, C.c_int32]
        self.lib.nosplan_upload_pods.restype = C.c_int
        self._check(self.lib.nosplan_upload_pods(self.h, C.byref(cp), pre, filter_mode), "nosplan_upload_pods")
        del k2, k3

    def update_nodes(self, first: int, rows: PackedSnapshot):
        """nosplan_update_nodes: overwrite nodes [first, first + rows.n_nodes) of the resident snapshot."""
        cs, k1 = c_snapshot(rows)
        self.lib.nosplan_update_nodes.argtypes = [C.c_void_p, C.c_int32, C.POINTER(CSnapshot)]
        self.lib.nosplan_update_nodes.restype = C.c_int
        self._check(self.lib.nosplan_update_nodes(self.h, first, C.byref(cs)), "nosplan_update_nodes")
        del k1

    def upload(self, snap: PackedSnapshot, pods: PackedPods, prefilter_ok=None, filter_mode: int = 0):
        cs, k1 = c_snapshot(snap)
        cp, k2 = c_pods(pods)
        pre, k3 = self._pre(prefilter_ok, pods.n_pods)
        self._check(self.lib.nosplan_upload(self.h, C.byref(cs), C.byref(cp), pre, filter_mode), "nosplan_upload")
        self._snap = snap
        del k1, k2, k3

    def run(self):
        self._check(self.lib.nosplan_run(self.h), "nosplan_run")

    def download(self, out=None):
        out, cnt, msk = alloc_partitioning(self._snap, out)
        st = CStats()
        rc = self.lib.nosplan_download(self.h, C.byref(out), C.byref(st))
        if rc == SLICED_NOT_APPLICABLE:      # slice mode only: this problem needs the whole snapshot on one GPU
            raise SlicedNotApplicable("the node-parallel pass does not apply: plan this problem unsharded")
        self._check(rc, "nosplan_download")
        return cnt, msk, st.as_dict()

    def parse_annotations(self, flat: dict, gpus_per_node: int):
        """nosplan_parse_annotations on the buffers of flatten_annotations -> (dict of per-GPU arrays, stats dict)."""
        N, G = flat["n_nodes"], gpus_per_node
        n = max(1, N * G)
        out = {"gpu_used": np.zeros((n, 16), np.uint8), "gpu_free": np.zeros((n, 16), np.uint8), "gpu_used_mask": np.zeros(n, np.uint16),
               "gpu_free_mask": np.zeros(n, np.uint16), "gpu_index": np.zeros(n, np.int32), "node_gpu_count": np.zeros(max(1, N), np.int32),
               "errors": np.zeros(4, np.int32)}
        ptr = lambda a: a.ctypes.data
        ca = CAnnotations(N, ptr(flat["node_ptr"]), ptr(flat["key_off"]), ptr(flat["key_bytes"]), ptr(flat["val_off"]), ptr(flat["val_bytes"]),
                          ptr(flat["gpu_count"]) if N else None, flat["n_profiles"], ptr(flat["profile_off"]), ptr(flat["profile_bytes"]), G)
        co = CParsedGpus(*[ptr(out[k]) for k in ("gpu_used", "gpu_free", "gpu_used_mask", "gpu_free_mask", "gpu_index", "node_gpu_count", "errors")])
        st = CStats()
        rc = self.lib.nosplan_parse_annotations(self.h, C.byref(ca), C.byref(co), C.byref(st))
        out["node_gpu_count"] = out["node_gpu_count"][:N]
        if rc != 0:
            msg = self.lib.nosplan_last_error(self.h)
            e = NosplanError(f"nosplan_parse_annotations failed ({rc}): {msg.decode() if msg else ''}")
            e.errors = out["errors"].copy()
            raise e
        return out, st.as_dict()

    def slice_begin(self, d_totals: int):
        self._check(self.lib.nosplan_slice_begin(self.h, d_totals), "nosplan_slice_begin")

    def slice_mid(self, rank: int, world: int, d_totals_gathered: int, d_exchange: int, nbytes: int):
        self._check(self.lib.nosplan_slice_mid(self.h, rank, world, d_totals_gathered, d_exchange, nbytes), "nosplan_slice_mid")

    def slice_chain_bytes(self) -> int:
        self.lib.nosplan_slice_chain_bytes.argtypes = [C.c_void_p]
        self.lib.nosplan_slice_chain_bytes.restype = C.c_int64
        return int(self.lib.nosplan_slice_chain_bytes(self.h))

    def slice_chain(self, d_state_in: int, d_state_out: int):
        self.lib.nosplan_slice_chain.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p]
        self.lib.nosplan_slice_chain.restype = C.c_int
        self._check(self.lib.nosplan_slice_chain(self.h, d_state_in or None, d_state_out), "nosplan_slice_chain")

    def shard_transport(self) -> str:
        return "peer-memory" if self.lib.nosplan_shard_transport(self.h) else "nccl"

    def run_sliced(self):
        self._check(self.lib.nosplan_run_sliced(self.h), "nosplan_run_sliced")

    def set_kernel_events(self, on: bool):
        self.lib.nosplan_set_kernel_events.argtypes = [C.c_void_p, C.c_int32]
        self._check(self.lib.nosplan_set_kernel_events(self.h, 1 if on else 0), "nosplan_set_kernel_events")

    def stream(self) -> int:
        return int(self.lib.nosplan_stream(self.h) or 0)

    # ---- node-sharded Plan() (one Engine per rank; see nos_b200/sharded.py for the collective around it)
    def shard_exchange_bytes(self) -> int:
        n = int(self.lib.nosplan_shard_exchange_bytes(self.h))
        if n < 0:
            raise NosplanError("nosplan_shard_exchange_bytes before nosplan_upload")
        return n

    def shard_begin(self, rank: int, world: int, d_exchange: int, nbytes: int):
        self._check(self.lib.nosplan_shard_begin(self.h, rank, world, d_exchange, nbytes), "nosplan_shard_begin")

    def shard_end(self, d_gathered: int, world: int):
        self._check(self.lib.nosplan_shard_end(self.h, d_gathered, world), "nosplan_shard_end")

    def nccl_unique_id(self) -> bytes:
        buf = (C.c_uint8 * 128)()
        rc = self.lib.nosplan_nccl_unique_id(buf)
        if rc != 0:
            raise NosplanError(f"nosplan_nccl_unique_id failed ({rc}): libnccl.so.2 not loadable")
        return bytes(buf)

    def shard_init(self, rank: int, world: int, unique_id: bytes):
        buf = (C.c_uint8 * 128).from_buffer_copy(unique_id)
        self._check(self.lib.nosplan_shard_init(self.h, rank, world, buf), "nosplan_shard_init")

    def run_sharded(self, gather_plan: bool = True):
        self._check(self.lib.nosplan_run_sharded(self.h, 1 if gather_plan else 0), "nosplan_run_sharded")

    def download_sharded(self):
        out, cnt, msk = alloc_partitioning(self._snap)
        st = CStats()
        self._check(self.lib.nosplan_download_sharded(self.h, C.byref(out), C.byref(st)), "nosplan_download_sharded")
        return cnt, msk, st.as_dict()

    def device_outputs(self):
        a, b = C.c_void_p(), C.c_void_p()
        self._check(self.lib.nosplan_device_outputs(self.h, C.byref(a), C.byref(b)), "nosplan_device_outputs")
        return int(a.value or 0), int(b.value or 0)

    def quota_victims(self, pq, query):
        """nosplan_quota_victims -> (victim flags [n_pods] uint8, stats dict).  pq.used_init / used_mask = current Used."""
        from . import quota_packing as qp
        cq, k1 = qp.c_quotas(pq)
        cv, k2 = query.c_struct()
        victim = np.zeros(max(1, query.n_pods), np.uint8)
        st = CStats()
        rc = self.lib.nosplan_quota_victims(self.h, C.byref(cq), C.byref(cv), victim.ctypes.data, C.byref(st))
        self._check(rc, "nosplan_quota_victims")
        del k1, k2
        return victim[:query.n_pods], st.as_dict()

    def quota_label_changes(self, pq, running, current, capacity=None, with_outputs=False):
        """nosplan_quota_label_changes -> (pod indices to patch int32, desired labels uint8, n_changes, outputs dict or None, stats).
        A capacity smaller than the number of changes raises NosplanError after filling what fits (read `.partial`)."""
        from . import quota_packing as qp
        cq, k1 = qp.c_quotas(pq)
        cr, k2 = qp.c_quota_pods(running)
        cur = np.ascontiguousarray(current, dtype=np.uint8)
        cap = running.n_pods if capacity is None else int(capacity)
        pod = np.zeros(max(1, cap), np.int32)
        desired = np.zeros(max(1, cap), np.uint8)
        n = np.zeros(1, np.int32)
        lc = CLabelChanges(cur.ctypes.data if cur.size else None, cap, pod.ctypes.data, desired.ctypes.data, n.ctypes.data)
        res = qp.QuotaResult(pq, running.n_pods, 0) if with_outputs else None
        out = res.c_struct() if with_outputs else None
        st = CStats()
        rc = self.lib.nosplan_quota_label_changes(self.h, C.byref(cq), C.byref(cr), C.byref(lc), C.byref(out) if with_outputs else None,
                                                  C.byref(st))
        take = min(int(n[0]), cap)
        if rc != 0:
            msg = self.lib.nosplan_last_error(self.h)
            e = NosplanError(f"nosplan_quota_label_changes failed ({rc}): {msg.decode() if msg else ''}")
            e.partial = (pod[:take].copy(), desired[:take].copy(), int(n[0]))
            raise e
        del k1, k2
        return pod[:take], desired[:take], int(n[0]), (res.trimmed() if with_outputs else None), st.as_dict()

    def preempt_dryrun(self, pq, query):
        """nosplan_preempt_dryrun -> (victim [n_pods] uint8, node_status [n_nodes] int32, num_violating [n_nodes] int32, stats dict)."""
        from . import quota_packing as qp
        cq, k1 = qp.c_quotas(pq)
        cv, k2 = query.c_struct()
        n_pods, n_nodes = query.v.n_pods, len(query.v.node_ptr) - 1
        victim = np.zeros(max(1, n_pods), np.uint8)
        status = np.zeros(max(1, n_nodes), np.int32)
        nviol = np.zeros(max(1, n_nodes), np.int32)
        st = CStats()
        rc = self.lib.nosplan_preempt_dryrun(self.h, C.byref(cq), C.byref(cv), victim.ctypes.data, status.ctypes.data, nviol.ctypes.data,
                                             C.byref(st))
        self._check(rc, "nosplan_preempt_dryrun")
        del k1, k2
        return victim[:n_pods], status[:n_nodes], nviol[:n_nodes], st.as_dict()

    def quota_state_begin(self, pq, pod_capacity: int):
        """nosplan_quota_state_begin: quota tables into HBM, Used = used_init, no pod accounted."""
        from . import quota_packing as qp
        cq, k1 = qp.c_quotas(pq)
        self.lib.nosplan_quota_state_begin.argtypes = [C.c_void_p, C.POINTER(qp.CQuotas), C.c_int32]
        self.lib.nosplan_quota_state_begin.restype = C.c_int
        self._check(self.lib.nosplan_quota_state_begin(self.h, C.byref(cq), int(pod_capacity)), "nosplan_quota_state_begin")
        self._state_pq = pq
        del k1

    def quota_state_apply(self, op, pod_id, ns, request, request_mask=None, outputs=True):
        """nosplan_quota_state_apply: one batch of AddPodIfNotPresent / DeletePodIfPresent events.
        -> (dict with used, used_mask, agg_*, guaranteed -- or None --, number of events that changed the state, stats)"""
        from . import quota_packing as qp
        pq = self._state_pq
        n = len(op)
        keep = [np.ascontiguousarray(op, np.uint8), np.ascontiguousarray(pod_id, np.int32), np.ascontiguousarray(ns, np.int32),
                np.ascontiguousarray(request, np.int64).reshape(n, pq.n_res) if n else np.zeros((1, pq.n_res), np.int64)]
        mask = np.ascontiguousarray(request_mask, np.uint32) if request_mask is not None else None
        p = lambda a: a.ctypes.data if a.size else None
        ev = CQuotaEvents(n, p(keep[0]), p(keep[1]), p(keep[2]), keep[3].ctypes.data, mask.ctypes.data if mask is not None and n else None)
        res = qp.QuotaResult(pq, 0, 0) if outputs else None
        out = res.c_struct() if outputs else None
        if outputs:
            out.prefilter = None; out.over_quota = None; out.operator_used = None; out.ambiguous_order = None
        applied = np.zeros(1, np.int32)
        st = CStats()
        self.lib.nosplan_quota_state_apply.argtypes = [C.c_void_p, C.POINTER(CQuotaEvents), C.POINTER(qp.CQuotaOut), C.c_void_p, C.POINTER(CStats)]
        self.lib.nosplan_quota_state_apply.restype = C.c_int
        rc = self.lib.nosplan_quota_state_apply(self.h, C.byref(ev), C.byref(out) if outputs else None, applied.ctypes.data, C.byref(st))
        self._check(rc, "nosplan_quota_state_apply")
        d = None
        if outputs:
            t = res.trimmed()
            d = {k: t[k] for k in ("used", "used_mask", "agg_min", "agg_used", "agg_overquota", "agg_masks", "guaranteed")}
        return d, int(applied[0]), st.as_dict()

    def quota_state_prefilter(self, pending):
        """nosplan_quota_state_prefilter -> verdict per pending pod (uint8) against the resident quota state."""
        from . import quota_packing as qp
        cp, k = qp.c_quota_pods(pending)
        verdict = np.zeros(max(1, pending.n_pods), np.uint8)
        st = CStats()
        self.lib.nosplan_quota_state_prefilter.argtypes = [C.c_void_p, C.POINTER(qp.CQuotaPods), C.c_void_p, C.POINTER(CStats)]
        self.lib.nosplan_quota_state_prefilter.restype = C.c_int
        self._check(self.lib.nosplan_quota_state_prefilter(self.h, C.byref(cp), verdict.ctypes.data, C.byref(st)), "nosplan_quota_state_prefilter")
        del k
        return verdict[:pending.n_pods], st.as_dict()

    def preempt_dryrun_again(self, query):
        """nosplan_preempt_dryrun_again: `query`'s preemptor against the cluster state the last preempt_dryrun() uploaded."""
        cv, k2 = query.c_struct()
        n_pods, n_nodes = query.v.n_pods, len(query.v.node_ptr) - 1
        victim = np.zeros(max(1, n_pods), np.uint8)
        status = np.zeros(max(1, n_nodes), np.int32)
        nviol = np.zeros(max(1, n_nodes), np.int32)
        st = CStats()
        self.lib.nosplan_preempt_dryrun_again.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.POINTER(CStats)]
        self.lib.nosplan_preempt_dryrun_again.restype = C.c_int
        rc = self.lib.nosplan_preempt_dryrun_again(self.h, C.byref(cv), victim.ctypes.data, status.ctypes.data, nviol.ctypes.data, C.byref(st))
        self._check(rc, "nosplan_preempt_dryrun_again")
        del k2
        return victim[:n_pods], status[:n_nodes], nviol[:n_nodes], st.as_dict()

    def quota_recompute(self, pq, running=None, pending=None, sharded=False):
        """nosplan_quota_recompute (sharded: nosplan_quota_recompute_sharded, this rank's share of the pods, after shard_init)
        -> (dict of arrays as QuotaResult.trimmed(), stats dict)"""
        from . import quota_packing as qp
        cq, k1 = qp.c_quotas(pq)
        cr, k2 = qp.c_quota_pods(running)
        cp, k3 = qp.c_quota_pods(pending)
        res = qp.QuotaResult(pq, running.n_pods if running else 0, pending.n_pods if pending else 0)
        out = res.c_struct()
        st = CStats()
        fn = self.lib.nosplan_quota_recompute_sharded if sharded else self.lib.nosplan_quota_recompute
        rc = fn(self.h, C.byref(cq), C.byref(cr), C.byref(cp), C.byref(out), C.byref(st))
        self._check(rc, "nosplan_quota_recompute_sharded" if sharded else "nosplan_quota_recompute")
        del k1, k2, k3
        return res.trimmed(), st.as_dict()
