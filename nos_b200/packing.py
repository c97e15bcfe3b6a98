"""Logical snapshot -> packed SoA tensors of include/nosplan.h (and back).

Host-side mirror of what the Go shim does before calling nosplan_plan: intern profile names, GPU
models and resource names into small ints, classify resource names with the reference's regexes,
compute pod requests, and lay everything out as flat numpy arrays.  Citations are relative to the
reference root (WindowsXp-Beta/nos v0.1.2).

Logical schema (plain dicts, JSON-friendly):
  node = {"name", "gpus": [{"model", "index", "memory_gb", "used": {profile: n}, "free": {profile: n}}],
          "allocatable": {resource: n}, "requested": {resource: n}, "num_pods": n}
  pod  = {"namespace", "name", "priority" (optional), "containers": [{resource: n}], "init_containers": [...]}
cpu quantities are milli-units, everything else integer Value()s.
"""
from __future__ import annotations

import math
import re
from dataclasses import dataclass, field
from typing import Dict, List, Optional, Sequence

import numpy as np

MAX_PROFILES = 16
MAX_GEOMS = 32
MAX_GPUS = 16
MAX_OTHER = 12

KIND_MIG, KIND_MPS = 0, 1
KIND_BY_NAME = {"mig": KIND_MIG, "mps": KIND_MPS}
FILTER_NONE, FILTER_REJECT_ALL, FILTER_RESOURCES_FIT = 0, 1, 2

# pkg/constant/constants.go:50-52, pkg/gpu/slicing/profile.go:30-31
_RE_MIG_RES = re.compile(r"nvidia\.com\/mig-\d+g\.\d+gb")
_RE_MPS_RES = re.compile(r"nvidia\.com/gpu-\d+gb")
_RE_GB = re.compile(r"\d+gb")
_RE_GI = re.compile(r"\d+g")
MIG_PREFIX = "nvidia.com/mig-"
MPS_PREFIX = "nvidia.com/gpu-"
CPU, MEMORY, EPHEMERAL, PODS = "cpu", "memory", "ephemeral-storage", "pods"

STATUS_PREFIX = "nos.nebuly.com/status-gpu"
SPEC_PREFIX = "nos.nebuly.com/spec-gpu"
LABEL_PARTITIONING = "nos.nebuly.com/gpu-partitioning"
LABEL_PRODUCT = "nvidia.com/gpu.product"
LABEL_COUNT = "nvidia.com/gpu.count"
LABEL_MEMORY = "nvidia.com/gpu.memory"


class PackError(ValueError):
    pass


def slice_profile_of(kind: int, resource: str) -> Optional[str]:
    """mig.ExtractProfileName (pkg/gpu/mig/util.go:46-52) / slicing.ExtractProfileName
    (pkg/gpu/slicing/util.go:34-40): unanchored regex match, then TrimPrefix."""
    if kind == KIND_MIG:
        if _RE_MIG_RES.search(resource) is None:
            return None
        return resource[len(MIG_PREFIX):] if resource.startswith(MIG_PREFIX) else resource
    if _RE_MPS_RES.search(resource) is None:
        return None
    return resource[len(MPS_PREFIX):] if resource.startswith(MPS_PREFIX) else resource


def resource_of_profile(kind: int, profile: str) -> str:
    return (MIG_PREFIX if kind == KIND_MIG else MPS_PREFIX) + profile


def mig_mem(profile: str) -> int:
    m = _RE_GB.search(profile)
    return int(m.group(0)[:-2]) if m else 0


def mig_gi(profile: str) -> int:
    m = _RE_GI.search(profile)
    return int(m.group(0)[:-1]) if m else 0


def mps_size(profile: str) -> int:  # slicing.ProfileName.GetMemorySizeGB (profile.go:52-59)
    t = profile[len(MPS_PREFIX):] if profile.startswith(MPS_PREFIX) else profile
    t = t[:-2] if t.endswith("gb") else t
    return int(t) if re.fullmatch(r"[+-]?\d+", t) else 0


_RE_MIG_PROFILE = re.compile(r"\d+g\.\d+gb")      # pkg/constant/constants.go:51, mig.ProfileName.isValid (profile.go:57-59)


def validate_known_geometries(configs: Dict[str, List[Dict[str, int]]]) -> None:
    """mig.ValidateConfigs (pkg/gpu/mig/known_configs.go:159-181): raises PackError where the reference returns an error."""
    if len(configs) == 0:
        raise PackError("no known configs provided")
    for geometries in configs.values():
        for geometry in geometries:
            for profile, quantity in geometry.items():
                if _RE_MIG_PROFILE.search(profile) is None:
                    raise PackError(f"invalid profile {profile}")
                if quantity < 1:
                    raise PackError(f"invalid quantity {quantity} for profile {profile}")


def group_by_model(allowed: Sequence[dict]) -> Dict[str, List[Dict[str, int]]]:
    """AllowedMigGeometriesList.GroupByModel (pkg/gpu/mig/allowed_geometries.go): the configuration file's list of
    {"models": [...], "allowedGeometries": [...]} entries -> model -> geometries, the shape `known_geometries` takes here and
    mig.SetKnownGeometries takes in the reference (cmd/gpupartitioner/gpupartitioner.go loads it at start-up)."""
    out: Dict[str, List[Dict[str, int]]] = {}
    for entry in allowed:
        for model in entry.get("models", []):
            out[model] = [dict(g) for g in entry.get("allowedGeometries", [])]
    return out


def compute_pod_request(pod: dict) -> Dict[str, int]:
    """resource.ComputePodRequest (pkg/resource/resource.go:127-146); the overhead sum is discarded there."""
    if "request" in pod:
        return dict(pod["request"])
    cont: Dict[str, int] = {}
    for c in pod.get("containers", []):
        for k, v in c.items():
            cont[k] = cont.get(k, 0) + v
    init: Dict[str, int] = {}
    for c in pod.get("init_containers", []):
        for k in set(init) | set(c):
            init[k] = max(init.get(k, 0), c.get(k, 0))
    return {k: max(cont.get(k, 0), init.get(k, 0)) for k in list(cont) + [k for k in init if k not in cont]}


# ------------------------------------------------------------------------------------------
# v1.Node (labels / annotations) -> logical node: mig.NewNode (pkg/gpu/mig/node.go:40-100),
# slicing.NewNode (pkg/gpu/slicing/node.go:32-105), TakeSnapshot (internal/partitioning/*/snapshot_taker.go)
# ------------------------------------------------------------------------------------------
def _atoi(s):
    return int(s) if isinstance(s, str) and re.fullmatch(r"[+-]?\d+", s) else None


def _status_annotations(annotations: Dict[str, str]):
    by_gpu: Dict[int, list] = {}
    for k, v in annotations.items():
        if k.startswith(SPEC_PREFIX) and len(k.split("-")) == 4 and _atoi(v) is not None and _atoi(k.split("-")[2]) is not None:
            continue                                   # parsed as a spec annotation first (pkg/gpu/annotation.go:91-94)
        if not k.startswith(STATUS_PREFIX):
            continue
        parts = k.split("-")
        if len(parts) != 5:
            continue
        q, idx = _atoi(v), _atoi(parts[2])
        st = parts[4].lower()
        if q is None or idx is None or st not in ("free", "used", "unknown"):
            continue
        by_gpu.setdefault(idx, []).append((parts[3], st, q))
    return by_gpu


def nodes_from_v1(kind: int, v1_nodes: Sequence[dict], known_geometries: Optional[dict] = None) -> List[dict]:
    """Keeps only nodes labelled for `kind`; raises PackError where NewNode returns an error."""
    kind_name = "mig" if kind == KIND_MIG else "mps"
    out = []
    for n in v1_nodes:
        labels = n.get("labels", {})
        if labels.get(LABEL_PARTITIONING) != kind_name:
            continue
        if LABEL_PRODUCT not in labels:
            raise PackError(f"node {n['name']}: missing label {LABEL_PRODUCT}")
        model = labels[LABEL_PRODUCT]
        count = _atoi(labels.get(LABEL_COUNT))
        if count is None:
            raise PackError(f"node {n['name']}: missing/invalid label {LABEL_COUNT}")
        mem = 0
        if kind == KIND_MPS:
            mb = _atoi(labels.get(LABEL_MEMORY))
            if mb is None:
                raise PackError(f"node {n['name']}: missing/invalid label {LABEL_MEMORY}")
            mem = int(math.ceil(float(mb) / 1000))     # pkg/gpu/util.go:71
        elif known_geometries is not None and model not in known_geometries:
            raise PackError(f"model {model!r} is not associated with any known GPU")
        gpus = []
        by_gpu = _status_annotations(n.get("annotations", {}))
        for idx in sorted(by_gpu):                     # canonical: ascending annotated GPU index
            used, free = {}, {}
            for prof, st, q in by_gpu[idx]:
                if st == "used":
                    used[prof] = q
                if st == "free":
                    free[prof] = q
            gpus.append({"model": model, "index": idx, "memory_gb": mem, "used": used, "free": free})
        for i in range(len(gpus), count):
            gpus.append({"model": model, "index": i, "memory_gb": mem, "used": {}, "free": {}})
        out.append({"name": n["name"], "gpus": gpus, "allocatable": dict(n.get("allocatable", {})),
                    "requested": dict(n.get("requested", {})), "num_pods": int(n.get("num_pods", 0))})
    return out


# ------------------------------------------------------------------------------------------
# packed containers
# ------------------------------------------------------------------------------------------
@dataclass
class PackedSnapshot:
    kind: int
    profiles: List[str]
    profile_mem_gb: np.ndarray
    profile_gi: np.ndarray
    models: List[str]
    model_geom_offset: np.ndarray
    geom_counts: np.ndarray
    node_names: List[str]
    gpus_per_node: int
    node_gpu_count: np.ndarray
    gpu_model: np.ndarray
    gpu_index: np.ndarray
    gpu_memory_gb: np.ndarray
    gpu_used: np.ndarray
    gpu_free: np.ndarray
    gpu_used_mask: np.ndarray
    gpu_free_mask: np.ndarray
    other_resources: List[str]
    node_alloc: np.ndarray
    node_requested: np.ndarray
    node_allowed_pods: np.ndarray
    node_num_pods: np.ndarray

    @property
    def n_nodes(self) -> int:
        return len(self.node_gpu_count)

    @property
    def n_profiles(self) -> int:
        return len(self.profiles)

    @property
    def n_res(self) -> int:
        return 3 + len(self.profiles) + len(self.other_resources)

    def column_of(self, resource: str) -> Optional[int]:
        if resource == CPU:
            return 0
        if resource == MEMORY:
            return 1
        if resource == EPHEMERAL:
            return 2
        p = slice_profile_of(self.kind, resource)
        if p is not None:
            return 3 + self.profiles.index(p) if p in self.profiles else None
        return 3 + len(self.profiles) + self.other_resources.index(resource) if resource in self.other_resources else None


@dataclass
class PackedPods:
    keys: List[str]
    priority: np.ndarray
    tracker_key: np.ndarray
    base: np.ndarray
    scalar_ptr: np.ndarray
    scalar_col: np.ndarray
    scalar_qty: np.ndarray

    @property
    def n_pods(self) -> int:
        return len(self.priority)


def _profile_order(kind: int, names) -> List[str]:
    if kind == KIND_MIG:
        return sorted(names)
    return sorted(names, key=lambda p: (mps_size(p), p))


def pack(kind: int, nodes: Sequence[dict], pods: Sequence[dict],
         known_geometries: Optional[Dict[str, List[Dict[str, int]]]] = None):
    """Returns (PackedSnapshot, PackedPods).  Nodes are sorted by byte-wise name; pods keep their order."""
    nodes = sorted(nodes, key=lambda n: n["name"].encode())
    known = known_geometries or {}
    pod_reqs = [compute_pod_request(p) for p in pods]

    # ---- dictionaries --------------------------------------------------------------------
    prof_set, other_set, model_list = set(), [], []
    for n in nodes:
        for g in n["gpus"]:
            prof_set.update(g["used"]); prof_set.update(g["free"])
            if g["model"] not in model_list:
                model_list.append(g["model"])
        for rl in (n.get("allocatable", {}), n.get("requested", {})):
            for r in rl:
                p = slice_profile_of(kind, r)
                if p is not None:
                    prof_set.add(p)
                elif r not in (CPU, MEMORY, EPHEMERAL, PODS) and r not in other_set:
                    other_set.append(r)
    for rq in pod_reqs:
        for r in rq:
            p = slice_profile_of(kind, r)
            if p is not None:
                prof_set.add(p)
            elif r not in (CPU, MEMORY, EPHEMERAL, PODS) and r not in other_set:
                other_set.append(r)
    if kind == KIND_MIG:
        for m in model_list:
            for geom in known.get(m, []):
                prof_set.update(geom)
    else:
        for p in prof_set:
            if p != f"{mps_size(p)}gb":
                raise PackError(f"non-canonical MPS profile name {p!r} (slicing.NewProfile writes '%dgb')")
    profiles = _profile_order(kind, prof_set)
    other = sorted(other_set)
    if len(profiles) > MAX_PROFILES:
        raise PackError(f"{len(profiles)} distinct profiles > {MAX_PROFILES}")
    if len(other) > MAX_OTHER:
        raise PackError(f"{len(other)} non-slice scalar resources > {MAX_OTHER}")
    pid = {p: i for i, p in enumerate(profiles)}
    K, R = len(profiles), 3 + len(profiles) + len(other)
    oid = {r: 3 + K + i for i, r in enumerate(other)}

    if kind == KIND_MIG:
        pmem = np.array([mig_mem(p) for p in profiles], np.int32)
        pgi = np.array([mig_gi(p) for p in profiles], np.int32)
    else:
        pmem = np.array([mps_size(p) for p in profiles], np.int32)
        pgi = np.zeros(K, np.int32)

    # ---- geometry tables -----------------------------------------------------------------
    offs, rows = [0], []
    if kind == KIND_MIG:
        for m in model_list:
            if m not in known:
                raise PackError(f"model {m!r} is not associated with any known GPU")   # mig.NewGPU (gpu.go:44-47)
            if len(known[m]) > MAX_GEOMS:
                raise PackError(f"model {m!r}: {len(known[m])} geometries > {MAX_GEOMS}")
            for geom in known[m]:
                row = np.zeros(16, np.uint8)
                for p, q in geom.items():
                    if q < 1 or q > 255:
                        raise PackError("geometry quantity out of range")          # mig.ValidateConfigs (known_configs.go:159-181)
                    row[pid[p]] = q
                rows.append(row)
            offs.append(len(rows))
    else:
        offs = [0] * (len(model_list) + 1)
    geom_counts = np.stack(rows) if rows else np.zeros((0, 16), np.uint8)

    # ---- nodes -----------------------------------------------------------------------------
    N = len(nodes)
    G = max([len(n["gpus"]) for n in nodes], default=0)
    G = max(G, 1)
    if G > MAX_GPUS:
        raise PackError(f"{G} GPUs per node > {MAX_GPUS}")
    gcount = np.zeros(N, np.int32)
    gmodel = np.zeros(N * G, np.int32); gindex = np.zeros(N * G, np.int32); gmem = np.zeros(N * G, np.int32)
    gused = np.zeros((N * G, 16), np.uint8); gfree = np.zeros((N * G, 16), np.uint8)
    gumask = np.zeros(N * G, np.uint16); gfmask = np.zeros(N * G, np.uint16)
    alloc = np.zeros((N, R), np.int64); req = np.zeros((N, R), np.int64)
    apods = np.zeros(N, np.int32); npods = np.zeros(N, np.int32)

    def col(r):
        if r == CPU: return 0
        if r == MEMORY: return 1
        if r == EPHEMERAL: return 2
        p = slice_profile_of(kind, r)
        return 3 + pid[p] if p is not None else oid[r]

    for i, n in enumerate(nodes):
        gcount[i] = len(n["gpus"])
        for j, g in enumerate(n["gpus"]):
            x = i * G + j
            gmodel[x] = model_list.index(g["model"]); gindex[x] = g["index"]; gmem[x] = g.get("memory_gb", 0)
            for p, q in g["used"].items():
                if not 0 <= q <= 255:
                    raise PackError("device count out of range")
                gused[x, pid[p]] = q; gumask[x] |= np.uint16(1 << pid[p])
            for p, q in g["free"].items():
                if not 0 <= q <= 255:
                    raise PackError("device count out of range")
                gfree[x, pid[p]] = q; gfmask[x] |= np.uint16(1 << pid[p])
        for r, v in n.get("allocatable", {}).items():
            if r == PODS:
                apods[i] += v
            else:
                alloc[i, col(r)] += v
        for r, v in n.get("requested", {}).items():
            if r != PODS:
                req[i, col(r)] += v
        npods[i] = n.get("num_pods", 0)

    snap = PackedSnapshot(kind, profiles, pmem, pgi, model_list, np.array(offs, np.int32), geom_counts,
                          [n["name"] for n in nodes], G, gcount, gmodel, gindex, gmem, gused, gfree, gumask, gfmask,
                          other, alloc, req, apods, npods)

    # ---- pods ------------------------------------------------------------------------------
    P = len(pods)
    prio = np.array([p.get("priority") or 0 for p in pods], np.int32).reshape(P)
    key_ids: Dict[str, int] = {}
    keys = [f"{p['namespace']}/{p['name']}" for p in pods]
    tkey = np.array([key_ids.setdefault(k, len(key_ids)) for k in keys], np.int32).reshape(P)
    base = np.zeros((P, 3), np.int64)
    ptr = np.zeros(P + 1, np.int32)
    cols, qtys = [], []
    for i, rq in enumerate(pod_reqs):
        ent = []
        for r, v in rq.items():
            if r == CPU: base[i, 0] += v
            elif r == MEMORY: base[i, 1] += v
            elif r == EPHEMERAL: base[i, 2] += v
            elif r == PODS: pass                       # framework.Resource.Add puts it in AllowedPodNumber: never read on this path
            else: ent.append((col(r), v))
        ent.sort()
        cols.extend(c for c, _ in ent); qtys.extend(v for _, v in ent)
        ptr[i + 1] = len(cols)
    ppods = PackedPods(keys, prio, tkey, base, ptr, np.array(cols, np.int32), np.array(qtys, np.int64))
    return snap, ppods


def unpack_plan(snap: PackedSnapshot, gpu_count: np.ndarray, gpu_mask: np.ndarray) -> Dict[str, list]:
    """-> {node name: [(GPUIndex, {resource name: count})]} == state.PartitioningState
    (internal/partitioning/state/partitioning.go:24-40)."""
    out: Dict[str, list] = {}
    G = snap.gpus_per_node
    gpu_count = np.asarray(gpu_count).reshape(-1, 16)
    for i, name in enumerate(snap.node_names):
        gl = []
        for j in range(int(snap.node_gpu_count[i])):
            x = i * G + j
            res = {}
            m = int(gpu_mask[x])
            for k, p in enumerate(snap.profiles):
                if m >> k & 1:
                    res[resource_of_profile(snap.kind, p)] = int(gpu_count[x, k])
            gl.append((int(snap.gpu_index[x]), res))
        out[name] = gl
    return out
