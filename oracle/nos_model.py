"""TEST INFRASTRUCTURE ONLY -- CPU restatement of the nos partition-planning hot path.

This file is the *faithful* half of the oracle: a pure-Python, dict-and-string model that
follows the Go reference statement by statement (Go maps -> Python dicts, Go pointers ->
Python object references, so the snapshot aliasing rules fall out of the object graph the same
way they do in Go).  It is used for small cases only: the golden vectors ported from the
reference's own tests (tests/golden/) and randomised cross-checks of the packed C++ oracle
(oracle/nos_oracle.cpp).  Nothing under nos_b200/ may import it.

All `path:line` citations are relative to /root/reference (WindowsXp-Beta/nos, v0.1.2).

Canonical resolution of the reference's map-iteration non-determinism:
  * geometry arg-max (pkg/gpu/mig/gpu.go:195-201): lowest index in the model's allowed-geometry
    list among the max scorers;
  * GPU order inside a node (pkg/gpu/mig/node.go:70): the order of the `gpus` list given;
  * multi-profile pods (internal/partitioning/core/util.go:59-64, pkg/gpu/mig/gpu.go:234):
    profiles are visited in ascending name order (MIG) / ascending (size, name) order (MPS);
  * MPS missing-slice sort ties and restore order (pkg/gpu/slicing/gpu.go:178-182,214-216):
    ascending (size, name).
Third-party arithmetic (k8s.io/kubernetes v1.25.4 framework.Resource / NodeInfo / NodeResourcesFit,
Go 1.18 sort.SliceStable) is restated from the published algorithms; each lives in one named
function below so it can be corrected in one place.
"""
from __future__ import annotations

import copy
import math
import re
from dataclasses import dataclass, field
from typing import Callable, Dict, List, Optional, Tuple

# --------------------------------------------------------------------------------------
# constants  (pkg/constant/constants.go:47-107)
# --------------------------------------------------------------------------------------
REGEX_MIG_RESOURCE = re.compile(r"nvidia\.com\/mig-\d+g\.\d+gb")      # constants.go:50
REGEX_MIG_PROFILE = re.compile(r"\d+g\.\d+gb")                         # constants.go:51
REGEX_MIG_MEMORY = re.compile(r"\d+gb")                                # constants.go:52, mig/profile.go:51
REGEX_MIG_GI = re.compile(r"\d+g")                                     # mig/profile.go:50
REGEX_MPS_RESOURCE = re.compile(r"nvidia\.com/gpu-\d+gb")              # slicing/profile.go:31
MIG_RESOURCE_PREFIX = "nvidia.com/mig-"                                # constants.go:58
MPS_RESOURCE_PREFIX = "nvidia.com/gpu-"                                # slicing/profile.go:30
RESOURCE_NVIDIA_GPU = "nvidia.com/gpu"                                 # constants.go:65
RESOURCE_GPU_MEMORY = "nos.nebuly.com/gpu-memory"                      # pkg/api/.../v1alpha1 ResourceGPUMemory
MIN_SLICE_MEMORY_GB = 1                                                # slicing/constant.go:24

CPU, MEMORY, EPHEMERAL, PODS = "cpu", "memory", "ephemeral-storage", "pods"

KIND_MIG, KIND_MPS = "mig", "mps"


# --------------------------------------------------------------------------------------
# framework.Resource restated (k8s.io/kubernetes v1.25.4 pkg/scheduler/framework/types.go)
# cpu is carried in milli-units everywhere, every other quantity as an integer Value().
# --------------------------------------------------------------------------------------
@dataclass
class Resource:
    milli_cpu: int = 0
    memory: int = 0
    ephemeral: int = 0
    allowed_pods: int = 0
    scalars: Dict[str, int] = field(default_factory=dict)

    def clone(self) -> "Resource":
        return Resource(self.milli_cpu, self.memory, self.ephemeral, self.allowed_pods, dict(self.scalars))

    def add_list(self, rl: Dict[str, int]) -> None:
        """framework.Resource.Add(v1.ResourceList)."""
        for name, q in rl.items():
            if name == CPU:
                self.milli_cpu += q
            elif name == MEMORY:
                self.memory += q
            elif name == PODS:
                self.allowed_pods += q
            elif name == EPHEMERAL:
                self.ephemeral += q
            else:
                self.scalars[name] = self.scalars.get(name, 0) + q


def resource_from_list(rl: Dict[str, int]) -> Resource:
    """framework.NewResource / resource.FromListToFramework (pkg/resource/resource.go:53-56)."""
    r = Resource()
    r.add_list(rl)
    return r


def resource_to_list(r: Resource) -> Dict[str, int]:
    """resource.FromFrameworkToList (pkg/resource/resource.go:35-50): cpu/memory/pods/ephemeral keys always set."""
    out = {CPU: r.milli_cpu, MEMORY: r.memory, PODS: r.allowed_pods, EPHEMERAL: r.ephemeral}
    out.update(r.scalars)
    return out


def _keys(*maps):  # pkg/util/util.go:106-120 (order irrelevant at every call site)
    s = []
    seen = set()
    for m in maps:
        for k in m:
            if k not in seen:
                seen.add(k)
                s.append(k)
    return s


def res_sum(a: Resource, b: Resource) -> Resource:  # pkg/resource/resource.go:59-72
    out = Resource(a.milli_cpu + b.milli_cpu, a.memory + b.memory, a.ephemeral + b.ephemeral,
                   a.allowed_pods + b.allowed_pods)
    for k in _keys(a.scalars, b.scalars):
        out.scalars[k] = a.scalars.get(k, 0) + b.scalars.get(k, 0)
    return out


def res_subtract(a: Resource, b: Resource) -> Resource:  # pkg/resource/resource.go:92-103
    out = Resource(a.milli_cpu - b.milli_cpu, a.memory - b.memory, a.ephemeral - b.ephemeral,
                   a.allowed_pods - b.allowed_pods)
    for k in _keys(a.scalars, b.scalars):
        out.scalars[k] = a.scalars.get(k, 0) - b.scalars.get(k, 0)
    return out


def res_subtract_non_negative(a: Resource, b: Resource) -> Resource:  # pkg/resource/resource.go:76-88
    out = res_subtract(a, b)
    out.milli_cpu = max(0, out.milli_cpu)
    out.memory = max(0, out.memory)
    out.allowed_pods = max(0, out.allowed_pods)
    out.ephemeral = max(0, out.ephemeral)
    for k in list(out.scalars):
        out.scalars[k] = max(0, out.scalars[k])
    return out


def res_abs(r: Resource) -> Resource:  # pkg/resource/resource.go:105-125
    out = r.clone()
    out.milli_cpu, out.memory = abs(out.milli_cpu), abs(out.memory)
    out.allowed_pods, out.ephemeral = abs(out.allowed_pods), abs(out.ephemeral)
    for k in list(out.scalars):
        out.scalars[k] = abs(out.scalars[k])
    return out


# --------------------------------------------------------------------------------------
# pods
# --------------------------------------------------------------------------------------
@dataclass
class Pod:
    namespace: str
    name: str
    priority: Optional[int] = None          # *Spec.Priority; corev1.PodPriority -> 0 when nil
    containers: List[Dict[str, int]] = field(default_factory=list)       # Resources.Requests per container
    init_containers: List[Dict[str, int]] = field(default_factory=list)
    labels: Dict[str, str] = field(default_factory=dict)
    creation_ts: int = 0
    node_name: str = ""

    def key(self) -> str:  # util.GetNamespacedName(&pod).String()
        return f"{self.namespace}/{self.name}"


def pod_priority(pod: Pod) -> int:  # k8s.io/component-helpers scheduling/corev1.PodPriority
    return pod.priority if pod.priority is not None else 0


def compute_pod_request(pod: Pod) -> Dict[str, int]:
    """resource.ComputePodRequest (pkg/resource/resource.go:127-146): sum of containers, max with each
    init container; the overhead sum is computed and discarded (:140-142), so it is not modelled."""
    containers: Dict[str, int] = {}
    for c in pod.containers:                       # quota.Add: union of keys
        for k, v in c.items():
            containers[k] = containers.get(k, 0) + v
    init: Dict[str, int] = {}
    for c in pod.init_containers:                  # quota.Max: union of keys, max value (missing == 0)
        for k in _keys(init, c):
            init[k] = max(init.get(k, 0), c.get(k, 0))
    out: Dict[str, int] = {}
    for k in _keys(containers, init):
        out[k] = max(containers.get(k, 0), init.get(k, 0))
    return out


# --------------------------------------------------------------------------------------
# MIG profiles (pkg/gpu/mig/profile.go, util.go)
# --------------------------------------------------------------------------------------
def mig_memory_slices(p: str) -> int:  # profile.go:70-75
    m = REGEX_MIG_MEMORY.search(p)
    return int(m.group(0)[:-2]) if m else 0


def mig_gi_slices(p: str) -> int:  # profile.go:77-82
    m = REGEX_MIG_GI.search(p)
    return int(m.group(0)[:-1]) if m else 0


def mig_smaller_than(a: str, b: str) -> bool:  # profile.go:84-96  (NOT a strict weak order)
    if mig_memory_slices(a) < mig_memory_slices(b):
        return True
    if mig_gi_slices(a) < mig_gi_slices(b):
        return True
    return False


def is_mig_resource(name: str) -> bool:  # mig/util.go:36-38 (unanchored MatchString)
    return REGEX_MIG_RESOURCE.search(name) is not None


def mig_extract_profile(name: str) -> Optional[str]:  # mig/util.go:46-52
    if not is_mig_resource(name):
        return None
    return name[len(MIG_RESOURCE_PREFIX):] if name.startswith(MIG_RESOURCE_PREFIX) else name


def mig_memory_gb_from_resource(name: str) -> int:  # mig/util.go:68-86 (error -> 0 as used by callers)
    if not is_mig_resource(name):
        return 0
    matches = REGEX_MIG_MEMORY.findall(name)
    if len(matches) != 1:
        return 0
    return int(re.search(r"\d+", matches[0]).group(0))


def mig_requested_profiles(pod: Pod) -> Dict[str, int]:  # mig/util.go:87-95
    res: Dict[str, int] = {}
    for r, q in compute_pod_request(pod).items():
        p = mig_extract_profile(r)
        if p is not None:
            res[p] = res.get(p, 0) + q
    return res


# --------------------------------------------------------------------------------------
# MPS profiles (pkg/gpu/slicing/profile.go, util.go)
# --------------------------------------------------------------------------------------
def mps_memory_gb(p: str) -> int:  # slicing/profile.go:52-59
    t = p[len(MPS_RESOURCE_PREFIX):] if p.startswith(MPS_RESOURCE_PREFIX) else p
    t = t[:-2] if t.endswith("gb") else t
    try:
        if not re.fullmatch(r"[+-]?\d+", t):
            return 0
        return int(t)
    except ValueError:
        return 0


def mps_new_profile(size_gb: int) -> str:  # slicing/profile.go:48-50
    return f"{size_gb}gb"


def mps_smaller_than(a: str, b: str) -> bool:  # slicing/profile.go:36-42
    return mps_memory_gb(a) < mps_memory_gb(b)


def is_mps_resource(name: str) -> bool:  # slicing/util.go:69-71
    return REGEX_MPS_RESOURCE.search(name) is not None


def mps_extract_profile(name: str) -> Optional[str]:  # slicing/util.go:34-40
    if not is_mps_resource(name):
        return None
    return name[len(MPS_RESOURCE_PREFIX):] if name.startswith(MPS_RESOURCE_PREFIX) else name


def mps_requested_profiles(pod: Pod) -> Dict[str, int]:  # slicing/util.go:59-67
    res: Dict[str, int] = {}
    for r, q in compute_pod_request(pod).items():
        p = mps_extract_profile(r)
        if p is not None:
            res[p] = res.get(p, 0) + q
    return res


def _mps_order(profiles):  # canonical (size, name) ascending
    return sorted(profiles, key=lambda p: (mps_memory_gb(p), p))


# --------------------------------------------------------------------------------------
# known MIG geometries (pkg/gpu/mig/known_configs.go:24-142) -- data, reproduced as given
# (including the 3g.20gb / 2g.10gb rows inside the A100-80GB table, :118,:121)
# --------------------------------------------------------------------------------------
GPU_MODEL_A30 = "NVIDIA-A30"
GPU_MODEL_A100_SXM4_40GB = "NVIDIA-A100-40GB-SXM4"
GPU_MODEL_A100_PCIE_80GB = "NVIDIA-A100-80GB-PCIe"

DEFAULT_KNOWN_GEOMETRIES: Dict[str, List[Dict[str, int]]] = {
    GPU_MODEL_A30: [
        {"4g.24gb": 1},
        {"2g.12gb": 2},
        {"2g.12gb": 1, "1g.6gb": 2},
        {"1g.6gb": 4},
    ],
    GPU_MODEL_A100_SXM4_40GB: [
        {"7g.40gb": 1},
        {"4g.20gb": 1, "2g.10gb": 1, "1g.5gb": 1},
        {"4g.20gb": 1, "1g.5gb": 3},
        {"3g.20gb": 2},
        {"3g.20gb": 1, "2g.10gb": 1, "1g.5gb": 1},
        {"3g.20gb": 1, "1g.5gb": 3},
        {"2g.10gb": 2, "3g.20gb": 1},
        {"2g.10gb": 1, "1g.5gb": 2, "3g.20gb": 1},
        {"2g.10gb": 3, "1g.5gb": 1},
        {"2g.10gb": 2, "1g.5gb": 3},
        {"2g.10gb": 1, "1g.5gb": 5},
        {"1g.5gb": 7},
    ],
    GPU_MODEL_A100_PCIE_80GB: [
        {"7g.79gb": 1},
        {"4g.40gb": 1, "2g.20gb": 1, "1g.10gb": 1},
        {"4g.40gb": 1, "1g.10gb": 3},
        {"3g.40gb": 2},
        {"3g.40gb": 1, "2g.20gb": 1, "1g.10gb": 1},
        {"3g.40gb": 1, "1g.10gb": 3},
        {"2g.20gb": 2, "3g.20gb": 1},
        {"2g.10gb": 1, "1g.10gb": 2, "3g.40gb": 1},
        {"2g.20gb": 3, "1g.10gb": 1},
        {"2g.20gb": 2, "1g.10gb": 3},
        {"2g.20gb": 1, "1g.10gb": 5},
        {"1g.10gb": 7},
    ],
}


# --------------------------------------------------------------------------------------
# NodeInfo restated (framework.NodeInfo: only what the planner path touches)
# --------------------------------------------------------------------------------------
@dataclass
class NodeInfo:
    allocatable: Resource = field(default_factory=Resource)
    requested: Resource = field(default_factory=Resource)
    num_pods: int = 0                      # len(Pods)

    def clone(self) -> "NodeInfo":
        return NodeInfo(self.allocatable.clone(), self.requested.clone(), self.num_pods)

    def value_copy(self) -> "NodeInfo":
        """`n.NodeInfo()` returns the struct by value: Allocatable / Requested are pointers and stay
        shared, the Pods slice header is copied (mig/node.go:119-121)."""
        return NodeInfo(self.allocatable, self.requested, self.num_pods)

    def add_pod(self, pod: Pod) -> None:
        """framework.NodeInfo.AddPod -> calculateResource: containers summed, max with init containers.
        (Upstream also adds Spec.Overhead; pods in this model carry none.)"""
        req = resource_from_list(compute_pod_request(pod))
        self.requested.milli_cpu += req.milli_cpu
        self.requested.memory += req.memory
        self.requested.ephemeral += req.ephemeral
        for k, v in req.scalars.items():
            self.requested.scalars[k] = self.requested.scalars.get(k, 0) + v
        self.num_pods += 1


# --------------------------------------------------------------------------------------
# mig.GPU / mig.Node  (pkg/gpu/mig/gpu.go, node.go)
# --------------------------------------------------------------------------------------
class MigGPU:
    def __init__(self, model: str, index: int, used: Dict[str, int], free: Dict[str, int],
                 allowed: List[Dict[str, int]]):
        self.model, self.index = model, index
        self.allowed = allowed            # shared, never mutated (gpu.go:61)
        self.used, self.free = used, free

    def clone(self) -> "MigGPU":  # gpu.go:57-72
        return MigGPU(self.model, self.index, dict(self.used), dict(self.free), self.allowed)

    def geometry(self) -> Dict[str, int]:  # gpu.go:82-93
        res: Dict[str, int] = {}
        for p, q in self.used.items():
            res[p] = res.get(p, 0) + q
        for p, q in self.free.items():
            res[p] = res.get(p, 0) + q
        return res

    def allows_geometry(self, g: Dict[str, int]) -> bool:  # gpu.go:215-222 (cmp.Equal: same keys, same values)
        return any(g == a for a in self.allowed)

    def can_apply_geometry(self, g: Dict[str, int]) -> bool:  # gpu.go:97-110
        if not self.allows_geometry(g):
            return False
        for p, q in self.used.items():
            if g.get(p, 0) < q:
                return False
        return True

    def apply_geometry(self, g: Dict[str, int]) -> bool:  # gpu.go:132-152
        if not self.can_apply_geometry(g):
            return False
        for p, q in g.items():
            self.free[p] = q - self.used.get(p, 0)      # zero kept (:141)
        for p in list(self.free):
            if p not in g:
                del self.free[p]
        return True

    def update_geometry_for(self, required: Dict[str, int]) -> Tuple[bool, int]:  # gpu.go:158-212
        """Returns (updated, geometry_evaluations)."""
        scores: List[int] = [0] * len(self.allowed)
        hit: List[bool] = [False] * len(self.allowed)
        evals = 0
        for gi, cand in enumerate(self.allowed):
            for p, rq in required.items():
                evals += 1
                if self.free.get(p, 0) >= rq:
                    continue
                n = min(cand.get(p, 0) - self.used.get(p, 0), rq)
                if n <= 0:
                    continue
                if not self.can_apply_geometry(cand):
                    continue
                scores[gi] += n
                hit[gi] = True
        best, best_score = None, 0
        for gi in range(len(self.allowed)):       # canonical: table order, strict >
            if hit[gi] and scores[gi] > best_score:
                best_score, best = scores[gi], gi
        if best is None:
            return False, evals
        self.apply_geometry(self.allowed[best])
        return True, evals

    def add_pod(self, pod: Pod) -> bool:  # gpu.go:233-247 (partial mutation on failure is reference behaviour)
        req = mig_requested_profiles(pod)
        for r in sorted(req):                     # canonical: ascending name
            q = req[r]
            if self.free.get(r, 0) < q:
                return False
            self.free[r] = self.free.get(r, 0) - q
            self.used[r] = self.used.get(r, 0) + q
        return True

    def has_free(self) -> bool:  # gpu.go:249-251
        return len(self.free) > 0


class MigNode:
    kind = KIND_MIG

    def __init__(self, name: str, gpus: List[MigGPU], node_info: NodeInfo):
        self.name, self.gpus, self.node_info = name, gpus, node_info

    def clone(self) -> "MigNode":  # node.go:212-222
        return MigNode(self.name, [g.clone() for g in self.gpus], self.node_info.clone())

    def geometry(self) -> Dict[str, int]:  # node.go:108-117
        res: Dict[str, int] = {}
        for g in self.gpus:
            for p, q in g.geometry().items():
                res[p] = res.get(p, 0) + q
        return res

    def has_free_capacity(self) -> bool:  # node.go:124-139
        if not self.gpus:
            return False
        for g in self.gpus:
            if g.has_free():
                return True
            if not g.allows_geometry(g.geometry()):
                return True
        return False

    def update_geometry_for(self, slices: Dict[str, int]) -> Tuple[bool, int]:  # node.go:145-177
        if not self.gpus or not slices:
            return False, 0
        required = dict(slices)
        any_updated, evals = False, 0
        for g in self.gpus:
            upd, e = g.update_geometry_for(required)
            evals += e
            any_updated = any_updated or upd
            for p, q in g.free.items():
                required[p] = required.get(p, 0) - q
                if required[p] <= 0:
                    del required[p]
        # node.go:172-174 + computeScalarResources :179-195 -- done even when nothing was updated
        scalars = {r: v for r, v in self.node_info.allocatable.scalars.items() if not is_mig_resource(r)}
        for p, v in self.geometry().items():
            scalars[MIG_RESOURCE_PREFIX + p] = v
        self.node_info.allocatable.scalars = scalars
        return any_updated, evals

    def add_pod(self, pod: Pod) -> bool:  # node.go:201-210
        for g in self.gpus:
            if g.add_pod(pod):
                self.node_info.value_copy().add_pod(pod)
                return True
        return False

    def partitioning(self) -> List[Tuple[int, Dict[str, int]]]:
        """PartitionCalculator.GetPartitioning (internal/partitioning/mig/partitition_calculator.go:30-46)."""
        return [(g.index, {MIG_RESOURCE_PREFIX + p: q for p, q in g.geometry().items()}) for g in self.gpus]

    def requested_slices(self, pod: Pod) -> Dict[str, int]:  # mig/slice_calculator.go:30-37
        return mig_requested_profiles(pod)


# --------------------------------------------------------------------------------------
# slicing.GPU / slicing.Node  (pkg/gpu/slicing/gpu.go, node.go)
# --------------------------------------------------------------------------------------
class MpsGPU:
    def __init__(self, model: str, index: int, memory_gb: int, used: Dict[str, int], free: Dict[str, int]):
        self.model, self.index, self.memory_gb = model, index, memory_gb
        self.used, self.free = used, free

    def validate(self) -> Optional[str]:  # gpu.go:66-97
        total = 0
        for m in (self.used, self.free):
            for p, q in m.items():
                mem = mps_memory_gb(p)
                if mem < MIN_SLICE_MEMORY_GB:
                    return "min slice size"
                total += mem * q
        if total > self.memory_gb:
            return "total memory exceeds GPU memory"
        return None

    def clone(self) -> "MpsGPU":  # gpu.go:110-129
        return MpsGPU(self.model, self.index, self.memory_gb, dict(self.used), dict(self.free))

    def geometry(self) -> Dict[str, int]:  # gpu.go:99-108
        res: Dict[str, int] = {}
        for p, q in self.used.items():
            res[p] = res.get(p, 0) + q
        for p, q in self.free.items():
            res[p] = res.get(p, 0) + q
        return res

    def _tot_slices_memory(self) -> int:  # gpu.go:256-265
        return sum(mps_memory_gb(p) * q for p, q in self.used.items()) + \
            sum(mps_memory_gb(p) * q for p, q in self.free.items())

    def _can_create_more(self) -> bool:  # gpu.go:250-254
        return self.memory_gb - self._tot_slices_memory() >= MIN_SLICE_MEMORY_GB

    def _create_slices(self, size_gb: int, num: int) -> bool:  # gpu.go:238-247
        spare = self.memory_gb - self._tot_slices_memory()
        if spare < size_gb * num:
            return False
        p = mps_new_profile(size_gb)
        self.free[p] = self.free.get(p, 0) + num
        return True

    def has_free_capacity(self) -> bool:  # gpu.go:131-136
        if len(self.free) > 0:
            return True
        return self._can_create_more()

    def add_pod(self, pod: Pod) -> bool:  # gpu.go:142-156
        req = mps_requested_profiles(pod)
        for r in _mps_order(req):
            q = req[r]
            if self.free.get(r, 0) < q:
                return False
            self.free[r] = self.free.get(r, 0) - q
            self.used[r] = self.used.get(r, 0) + q
        return True

    def update_geometry_for(self, slices: Dict[str, int]) -> bool:  # gpu.go:162-232
        missing: Dict[str, int] = {}
        for p, rq in slices.items():                       # getMissingSlices :222-232
            d = rq - self.free.get(p, 0)
            if d > 0:
                missing[p] = d
        if not missing:
            return False
        updated = False
        original_free = dict(self.free)
        for mp in _mps_order(missing):                     # :173-182 (canonical tie order)
            size = mps_memory_gb(mp)
            if self._can_create_more():                    # :188-197 spare capacity first
                q = missing[mp]
                for _ in range(q):
                    if not self._create_slices(size, 1):
                        break
                    missing[mp] -= 1
                    updated = True
            q = missing[mp]                                # :199-211 delete ALL originally-free keys, then create
            for k in original_free:
                self.free.pop(k, None)
            for _ in range(q):
                if not self._can_create_more():
                    break
                if not self._create_slices(size, 1):
                    break
                missing[mp] -= 1
                updated = True
            for k in _mps_order(original_free):            # :213-216 restore, all-or-nothing per profile
                self._create_slices(mps_memory_gb(k), original_free[k])
        return updated


class MpsNode:
    kind = KIND_MPS

    def __init__(self, name: str, gpus: List[MpsGPU], node_info: NodeInfo):
        self.name, self.gpus, self.node_info = name, gpus, node_info

    def clone(self) -> "MpsNode":  # node.go:107-118
        return MpsNode(self.name, [g.clone() for g in self.gpus], self.node_info.clone())

    def geometry(self) -> Dict[str, int]:  # node.go:175-183
        res: Dict[str, int] = {}
        for g in self.gpus:
            for p, q in g.geometry().items():
                res[p] = res.get(p, 0) + q
        return res

    def has_free_capacity(self) -> bool:  # node.go:208-215
        return any(g.has_free_capacity() for g in self.gpus)

    def update_geometry_for(self, slices: Dict[str, int]) -> Tuple[bool, int]:  # node.go:120-152
        if not self.gpus or not slices:
            return False, 0
        required = dict(slices)
        any_updated, evals = False, 0
        for g in self.gpus:
            evals += len(required)
            upd = g.update_geometry_for(required)
            any_updated = any_updated or upd
            for p, q in g.free.items():
                required[p] = required.get(p, 0) - q
                if required[p] <= 0:
                    del required[p]
        scalars = {r: v for r, v in self.node_info.allocatable.scalars.items() if not is_mps_resource(r)}
        for p, v in self.geometry().items():               # node.go:154-170
            scalars[MPS_RESOURCE_PREFIX + p] = v
        self.node_info.allocatable.scalars = scalars
        return any_updated, evals

    def add_pod(self, pod: Pod) -> bool:  # node.go:193-202
        for g in self.gpus:
            if g.add_pod(pod):
                self.node_info.value_copy().add_pod(pod)
                return True
        return False

    def partitioning(self) -> List[Tuple[int, Dict[str, int]]]:  # internal/partitioning/mps/partition_calculator.go:30-46
        return [(g.index, {MPS_RESOURCE_PREFIX + p: q for p, q in g.geometry().items()}) for g in self.gpus]

    def requested_slices(self, pod: Pod) -> Dict[str, int]:  # mps/slice_calculator.go:30-37
        return mps_requested_profiles(pod)


# --------------------------------------------------------------------------------------
# Go 1.18 sort.SliceStable restated (sort/zsortfunc.go: stable_func, insertionSort_func,
# symMerge_func, rotate_func, swapRange_func) -- needed because the pod comparator is not a
# strict weak order in general (core/util.go:34-73, mig/profile.go:84-96)
# --------------------------------------------------------------------------------------
def go_stable_sort(items: list, less: Callable[[object, object], bool]) -> list:
    data = list(items)
    n = len(data)

    def lt(i, j):
        return less(data[i], data[j])

    def swap(i, j):
        data[i], data[j] = data[j], data[i]

    def insertion(a, b):
        for i in range(a + 1, b):
            j = i
            while j > a and lt(j, j - 1):
                swap(j, j - 1)
                j -= 1

    def swap_range(a, b, cnt):
        for i in range(cnt):
            swap(a + i, b + i)

    def rotate(a, m, b):
        i, j = m - a, b - m
        while i != j:
            if i > j:
                swap_range(m - i, m, j)
                i -= j
            else:
                swap_range(m - i, m + j - i, i)
                j -= i
        swap_range(m - i, m, i)

    def sym_merge(a, m, b):
        if m - a == 1:
            i, j = m, b
            while i < j:
                h = (i + j) >> 1
                if lt(h, a):
                    i = h + 1
                else:
                    j = h
            for k in range(a, i - 1):
                swap(k, k + 1)
            return
        if b - m == 1:
            i, j = a, m
            while i < j:
                h = (i + j) >> 1
                if not lt(m, h):
                    i = h + 1
                else:
                    j = h
            for k in range(m, i, -1):
                swap(k, k - 1)
            return
        mid = (a + b) >> 1
        nn = mid + m
        if m > mid:
            start, r = nn - b, mid
        else:
            start, r = a, m
        p = nn - 1
        while start < r:
            c = (start + r) >> 1
            if not lt(p - c, c):
                start = c + 1
            else:
                r = c
        end = nn - start
        if start < m and m < end:
            rotate(start, m, end)
        if a < start and start < mid:
            sym_merge(a, start, mid)
        if mid < end and end < b:
            sym_merge(mid, end, b)

    block = 20
    a, b = 0, block
    while b <= n:
        insertion(a, b)
        a = b
        b += block
    insertion(a, n)
    while block < n:
        a, b = 0, 2 * block
        while b <= n:
            sym_merge(a, a + block, b)
            a = b
            b += 2 * block
        m = a + block
        if m < n:
            sym_merge(a, m, n)
        block *= 2
    return data


# --------------------------------------------------------------------------------------
# core: sorter, snapshot, tracker, planner (internal/partitioning/core)
# --------------------------------------------------------------------------------------
def _first_profile(kind: str, slices: Dict[str, int]) -> str:
    """core/util.go:59-64 takes the first key of a Go map; canonical: first in profile order."""
    return sorted(slices)[0] if kind == KIND_MIG else _mps_order(slices)[0]


def sort_pods(kind: str, pods: List[Pod]) -> List[Pod]:
    """NewPodSorter (core/util.go:34-73)."""
    req = mig_requested_profiles if kind == KIND_MIG else mps_requested_profiles
    smaller = mig_smaller_than if kind == KIND_MIG else mps_smaller_than

    def less(a: Pod, b: Pod) -> bool:
        pa, pb = pod_priority(a), pod_priority(b)
        if pa != pb:
            return pa > pb
        ra = req(a)
        if not ra:
            return False
        rb = req(b)
        if not rb:
            return False
        return smaller(_first_profile(kind, ra), _first_profile(kind, rb))

    return go_stable_sort(pods, less)


FILTER_NONE = "none"                  # the reference's own test setup: mocked framework, always Success
FILTER_RESOURCES_FIT = "resources_fit"


def node_resources_fit(pod_req: Resource, ni: NodeInfo) -> bool:
    """NodeResourcesFit.Filter -> fitsRequest (k8s.io/kubernetes v1.25.4 plugins/noderesources/fit.go),
    restated from the published source; parity unpinned by the reference (its planner tests mock it)."""
    ok = True
    if ni.num_pods + 1 > ni.allocatable.allowed_pods:
        ok = False
    if pod_req.milli_cpu == 0 and pod_req.memory == 0 and pod_req.ephemeral == 0 and len(pod_req.scalars) == 0:
        return ok
    if pod_req.milli_cpu > ni.allocatable.milli_cpu - ni.requested.milli_cpu:
        ok = False
    if pod_req.memory > ni.allocatable.memory - ni.requested.memory:
        ok = False
    if pod_req.ephemeral > ni.allocatable.ephemeral - ni.requested.ephemeral:
        ok = False
    for r, q in pod_req.scalars.items():
        if q > ni.allocatable.scalars.get(r, 0) - ni.requested.scalars.get(r, 0):
            ok = False
    return ok


class ClusterSnapshot:  # core/snapshot.go:43-190
    def __init__(self, kind: str, nodes: Dict[str, object]):
        self.kind = kind
        self.data: Dict[str, object] = nodes
        self.forked: Optional[Dict[str, object]] = None
        self.sum_calls = 0            # resource.Sum invocations, for the faithful-cost counter

    def _get(self) -> Dict[str, object]:
        return self.forked if self.forked is not None else self.data

    def clone(self) -> "ClusterSnapshot":  # :93-106
        c = ClusterSnapshot(self.kind, {k: v.clone() for k, v in self.data.items()})
        if self.forked is not None:
            c.forked = {k: v.clone() for k, v in self.forked.items()}
        return c

    def fork(self) -> None:  # :85-91
        if self.forked is not None:
            raise RuntimeError("snapshot already forked")
        self.forked = {k: v.clone() for k, v in self._get().items()}

    def commit(self) -> None:  # :108-113
        if self.forked is not None:
            self.data, self.forked = self.forked, None

    def revert(self) -> None:  # :115-117
        self.forked = None

    def get_nodes(self) -> Dict[str, object]:
        return self._get()

    def get_node(self, name: str):
        return self._get().get(name)

    def set_node(self, node) -> None:  # :177-179
        self._get()[node.name] = node

    def add_pod(self, node_name: str, pod: Pod) -> bool:  # :181-190
        node = self._get().get(node_name)
        if node is None:
            return False
        return node.add_pod(pod)

    def partitioning_state(self) -> Dict[str, list]:  # :66-72
        return {name: n.partitioning() for name, n in self.get_nodes().items()}

    def candidate_nodes(self) -> list:  # :119-130 (byte-wise name order)
        res = [n for n in self._get().values() if n.has_free_capacity()]
        return sorted(res, key=lambda n: n.name.encode())

    def lacking_resources(self, pod: Pod) -> Resource:  # :132-165
        pod_req = compute_pod_request(pod)
        tot_a, tot_r = Resource(), Resource()
        for n in self.get_nodes().values():
            tot_a = res_sum(tot_a, n.node_info.allocatable)
            tot_r = res_sum(tot_r, n.node_info.requested)
            self.sum_calls += 2
        available = res_subtract_non_negative(tot_a, tot_r)
        diff = res_subtract(available, resource_from_list(pod_req))
        res = Resource()
        if diff.milli_cpu < 0:
            res.milli_cpu = diff.milli_cpu
        if diff.memory < 0:
            res.memory = diff.memory
        if diff.ephemeral < 0:
            res.ephemeral = diff.ephemeral
        if diff.allowed_pods < 0:
            res.allowed_pods = diff.allowed_pods
        for k, v in diff.scalars.items():
            if v < 0:
                res.scalars[k] = v
        return res_abs(res)

    def lacking_slices(self, pod: Pod) -> Dict[str, int]:  # :74-76 + SliceFilter.ExtractSlices
        scal = self.lacking_resources(pod).scalars
        out: Dict[str, int] = {}
        extract = mig_extract_profile if self.kind == KIND_MIG else mps_extract_profile
        for r, q in scal.items():
            p = extract(r)
            if p is not None:
                out[p] = out.get(p, 0) + int(q)
        return out


class SliceTracker:  # core/tracker.go:26-88
    def __init__(self, snapshot: ClusterSnapshot, pods: List[Pod]):
        self.requested: Dict[str, int] = {}
        self.lacking: Dict[str, int] = {}
        self.lookup: Dict[str, Dict[str, int]] = {}
        self._req = mig_requested_profiles if snapshot.kind == KIND_MIG else mps_requested_profiles
        for pod in pods:
            ent = self.lookup.setdefault(pod.key(), {})
            for s, q in snapshot.lacking_slices(pod).items():
                self.lacking[s] = self.lacking.get(s, 0) + q
                ent[s] = ent.get(s, 0) + q
            for s, q in self._req(pod).items():
                self.requested[s] = self.requested.get(s, 0) + q

    def remove(self, pod: Pod) -> None:  # :67-88
        for s, q in self._req(pod).items():
            self.requested[s] = self.requested.get(s, 0) - q
            if self.requested[s] <= 0:
                del self.requested[s]
        ent = self.lookup.get(pod.key())
        if ent is not None:
            for s, q in list(ent.items()):
                self.lacking[s] = self.lacking.get(s, 0) - q
                ent[s] -= q
                if ent[s] <= 0:
                    del ent[s]
                if self.lacking[s] <= 0:
                    del self.lacking[s]


@dataclass
class PlanStats:
    candidates: int = 0            # (visited candidate node, pod) pairs == logical tryAddPod calls
    visited_nodes: int = 0
    geometry_evals: int = 0
    placements: int = 0
    sum_calls: int = 0


def plan(snapshot: ClusterSnapshot, pods: List[Pod], prefilter_ok: Optional[List[bool]] = None,
         filter_mode: str = FILTER_NONE, filter_const: bool = True) -> Tuple[Dict[str, list], PlanStats]:
    """planner.Plan (core/planner.go:67-153).  `prefilter_ok[i]` is the PreFilter verdict for pods[i]
    (a per-pod constant during one Plan()); `filter_mode` selects the Filter simulation:
    FILTER_NONE with `filter_const` reproduces the reference tests' mocked framework."""
    stats = PlanStats()
    state = snapshot.partitioning_state()
    tracker = SliceTracker(snapshot, pods)
    if not tracker.lacking:
        stats.sum_calls = snapshot.sum_calls
        return state, stats
    ok_by_id = {id(p): (True if prefilter_ok is None else bool(prefilter_ok[i])) for i, p in enumerate(pods)}
    sorted_pods = sort_pods(snapshot.kind, pods)
    cands = snapshot.candidate_nodes()
    for n in cands:
        if not tracker.lacking:
            break
        snapshot.fork()
        stats.visited_nodes += 1
        upd, evals = n.update_geometry_for(tracker.lacking)
        stats.geometry_evals += evals
        if upd:
            snapshot.set_node(n)
        added = 0
        for pod in sorted_pods:
            stats.candidates += 1
            # tryAddPod (core/planner.go:155-175)
            if snapshot.lacking_slices(pod):
                continue
            node = snapshot.get_node(n.name)
            if node is None:
                continue
            # canSchedulePod (:178-207)
            if not ok_by_id[id(pod)]:
                continue
            if filter_mode == FILTER_NONE:
                if not filter_const:
                    continue
            else:
                if not node_resources_fit(resource_from_list(compute_pod_request(pod)), node.node_info.value_copy()):
                    continue
            if not snapshot.add_pod(n.name, pod):
                continue
            state[n.name] = n.partitioning()
            tracker.remove(pod)
            added += 1
            stats.placements += 1
        if added == 0:
            snapshot.revert()
        else:
            snapshot.commit()
    stats.sum_calls = snapshot.sum_calls
    return state, stats


# --------------------------------------------------------------------------------------
# snapshot taking: v1.Node labels / annotations -> mig.Node / slicing.Node
# (pkg/gpu/annotation.go:59-101, pkg/gpu/util.go:30-73, mig/node.go:40-100, slicing/node.go:32-105,
#  internal/partitioning/{mig,mps}/snapshot_taker.go:31-53)
# --------------------------------------------------------------------------------------
ANNOTATION_STATUS_PREFIX = "nos.nebuly.com/status-gpu"     # pkg/api/nos.nebuly.com/v1alpha1/annotations.go:23
ANNOTATION_SPEC_PREFIX = "nos.nebuly.com/spec-gpu"         # :22
LABEL_GPU_PARTITIONING = "nos.nebuly.com/gpu-partitioning"  # labels.go:23
LABEL_NVIDIA_PRODUCT = "nvidia.com/gpu.product"             # pkg/constant/constants.go:81
LABEL_NVIDIA_COUNT = "nvidia.com/gpu.count"                 # :84
LABEL_NVIDIA_MEMORY = "nvidia.com/gpu.memory"               # :87


def _atoi(s: str) -> Optional[int]:
    return int(s) if re.fullmatch(r"[+-]?\d+", s) else None


def parse_status_annotation(key: str, value: str):
    """gpu.ParseStatusAnnotation (pkg/gpu/annotation.go:59-85) -> (index, profile, status, quantity) or None."""
    if not key.startswith(ANNOTATION_STATUS_PREFIX):
        return None
    parts = key.split("-")
    if len(parts) != 5:
        return None
    q, idx = _atoi(value), _atoi(parts[2])
    if q is None or idx is None:
        return None
    status = parts[4].lower()
    if status not in ("free", "used", "unknown"):
        return None
    return idx, parts[3], status, q


def parse_spec_annotation(key: str, value: str):  # pkg/gpu/annotation.go:29-57
    if not key.startswith(ANNOTATION_SPEC_PREFIX):
        return None
    parts = key.split("-")
    if len(parts) != 4:
        return None
    q, idx = _atoi(value), _atoi(parts[2])
    if q is None or idx is None:
        return None
    return idx, parts[3], q


def _status_by_gpu(annotations: Dict[str, str]) -> Dict[int, list]:
    by_gpu: Dict[int, list] = {}
    for k, v in annotations.items():
        if parse_spec_annotation(k, v) is not None:     # ParseNodeAnnotations tries spec first (:91-94)
            continue
        st = parse_status_annotation(k, v)
        if st is not None:
            by_gpu.setdefault(st[0], []).append(st)
    return by_gpu


def node_info_from(node: dict) -> NodeInfo:
    ni = NodeInfo(resource_from_list(node.get("allocatable", {})), resource_from_list(node.get("requested", {})),
                  int(node.get("num_pods", 0)))
    return ni


def new_mig_node(node: dict, known: Dict[str, List[Dict[str, int]]]) -> MigNode:  # mig/node.go:40-100
    labels = node.get("labels", {})
    if LABEL_NVIDIA_PRODUCT not in labels:
        raise ValueError("missing product label")
    model = labels[LABEL_NVIDIA_PRODUCT]
    count = _atoi(labels.get(LABEL_NVIDIA_COUNT, "x"))
    if count is None:
        raise ValueError("missing/invalid count label")
    if model not in known:
        raise ValueError("model not associated with any known GPU")
    gpus: List[MigGPU] = []
    by_gpu = _status_by_gpu(node.get("annotations", {}))
    for idx in sorted(by_gpu):                       # canonical: ascending GPU index
        used, free = {}, {}
        for (_, prof, status, q) in by_gpu[idx]:
            if status == "used":
                used[prof] = q
            if status == "free":
                free[prof] = q
        gpus.append(MigGPU(model, idx, used, free, known[model]))
    for i in range(len(gpus), count):
        gpus.append(MigGPU(model, i, {}, {}, known[model]))
    return MigNode(node["name"], gpus, node_info_from(node))


def gpu_memory_gb(labels: Dict[str, str]) -> int:  # pkg/gpu/util.go:58-73
    mb = _atoi(labels.get(LABEL_NVIDIA_MEMORY, "x"))
    if mb is None:
        raise ValueError("missing/invalid memory label")
    return int(math.ceil(float(mb) / 1000))


def new_mps_node(node: dict) -> MpsNode:  # slicing/node.go:32-105
    labels = node.get("labels", {})
    if LABEL_NVIDIA_PRODUCT not in labels:
        raise ValueError("missing product label")
    model = labels[LABEL_NVIDIA_PRODUCT]
    count = _atoi(labels.get(LABEL_NVIDIA_COUNT, "x"))
    if count is None:
        raise ValueError("missing/invalid count label")
    mem = gpu_memory_gb(labels)
    gpus: List[MpsGPU] = []
    by_gpu = _status_by_gpu(node.get("annotations", {}))
    for idx in sorted(by_gpu):
        used, free = {}, {}
        for (_, prof, status, q) in by_gpu[idx]:
            if status == "used":
                used[prof] = q
            if status == "free":
                free[prof] = q
        g = MpsGPU(model, idx, mem, used, free)
        err = g.validate()
        if err:
            raise ValueError(err)
        gpus.append(g)
    for i in range(len(gpus), count):
        gpus.append(MpsGPU(model, i, mem, {}, {}))
    return MpsNode(node["name"], gpus, node_info_from(node))


def take_snapshot(kind: str, nodes: List[dict], known: Optional[Dict[str, List[Dict[str, int]]]] = None) -> ClusterSnapshot:
    known = DEFAULT_KNOWN_GEOMETRIES if known is None else known
    out: Dict[str, object] = {}
    for n in nodes:
        if n.get("labels", {}).get(LABEL_GPU_PARTITIONING) != kind:
            continue
        out[n["name"]] = new_mig_node(n, known) if kind == KIND_MIG else new_mps_node(n)
    return ClusterSnapshot(kind, out)


def pod_from(d: dict) -> Pod:
    return Pod(d["namespace"], d["name"], d.get("priority"), [dict(c) for c in d.get("containers", [])],
               [dict(c) for c in d.get("init_containers", [])], dict(d.get("labels", {})), int(d.get("creation_ts", 0)))
