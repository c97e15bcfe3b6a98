"""TEST INFRASTRUCTURE ONLY -- faithful Python restatement of the capacity-scheduling ElasticQuota accounting
(pkg/scheduler/plugins/capacityscheduling/{elasticquotainfo,capacity_scheduling}.go), of the quota-aware pod request
(pkg/gpu/util/resource.go) and of the operator's in-quota / over-quota labelling
(internal/controllers/elasticquota/elasticquota.go).  Citations relative to /root/reference.  Used only by tests/.

Third-party arithmetic restated here: apimachinery v0.25.4 resource.Quantity.{Value, AsApproximateFloat64} on the
quantities FromFrameworkToList builds (cpu = NewMilliQuantity -> base * math.Pow10(-3); everything else scale 0),
apiserver v0.25.4 quota.{Add, LessThanOrEqual, Equals}.
"""
from __future__ import annotations

import math
from dataclasses import dataclass, field
from typing import Dict, List, Optional, Tuple

from nos_model import (CPU, EPHEMERAL, MEMORY, PODS, RESOURCE_GPU_MEMORY, RESOURCE_NVIDIA_GPU, Pod, Resource,
                       compute_pod_request, is_mig_resource, mig_memory_gb_from_resource, pod_priority, res_subtract_non_negative,
                       res_sum, resource_from_list, resource_to_list)

INT64_MIN = -(1 << 63)


def go_int64_of_float(x: float) -> int:
    """int64(f) as compiled for amd64 (CVTTSD2SQ): out-of-range / NaN -> 0x8000000000000000."""
    if x != x or x >= 9.223372036854775808e18 or x < -9.223372036854775808e18:
        return INT64_MIN
    return int(x)


def quota_pod_request(pod: Pod, nvidia_gpu_memory_gb: int) -> Dict[str, int]:
    """gpu_util.ResourceCalculator.ComputePodRequest (pkg/gpu/util/resource.go:60-86): the plain request plus the
    scalar nos.nebuly.com/gpu-memory = memGB * #nvidia.com/gpu + sum(migGB * q); MPS slices contribute 0; key always set."""
    res = compute_pod_request(pod)
    total = 0
    for name, q in res.items():
        if name == RESOURCE_NVIDIA_GPU:
            total += nvidia_gpu_memory_gb * q
        elif is_mig_resource(name):
            total += mig_memory_gb_from_resource(name) * q
    res[RESOURCE_GPU_MEMORY] = total
    return res


@dataclass
class ElasticQuotaInfo:  # elasticquotainfo.go:177-192
    name: str = ""
    namespaces: List[str] = field(default_factory=list)
    min: Optional[Resource] = None
    max: Optional[Resource] = None
    used: Optional[Resource] = None
    max_enforced: bool = False
    pods: set = field(default_factory=set)

    def reserve(self, req: Resource) -> None:  # :194-200 (Memory, MilliCPU, scalars only)
        self.used.memory += req.memory
        self.used.milli_cpu += req.milli_cpu
        for k, v in req.scalars.items():
            self.used.scalars[k] = self.used.scalars.get(k, 0) + v

    def unreserve(self, req: Resource) -> None:  # :202-208
        self.used.memory -= req.memory
        self.used.milli_cpu -= req.milli_cpu
        for k, v in req.scalars.items():
            self.used.scalars[k] = self.used.scalars.get(k, 0) - v

    def used_over_min_with(self, req: Resource) -> bool:  # :210-212
        return sum_greater_than(req, self.used, self.min)

    def used_over_max_with(self, req: Resource) -> bool:  # :214-219
        return sum_greater_than(req, self.used, self.max) if self.max_enforced else False

    def used_over_min(self) -> bool:  # :222-224
        return greater_than(self.used, self.min)

    def used_over(self, r: Resource) -> bool:  # :227-229
        return greater_than(self.used, r)

    def used_lte_with(self, r: Resource, req: Resource) -> bool:  # :237-239
        return sum_less_than_equal(req, self.used, r)

    def add_pod_if_not_present(self, uid: str, req: Resource) -> None:  # :276-292
        if uid in self.pods:
            return
        self.pods.add(uid)
        self.reserve(req)

    def delete_pod_if_present(self, uid: str, req: Resource) -> None:  # :294-310
        if uid not in self.pods:
            return
        self.pods.discard(uid)
        self.unreserve(req)


def greater_than(x: Resource, y: Resource) -> bool:  # :313-315
    return sum_greater_than(x, Resource(), y)


def sum_greater_than(x1: Resource, x2: Resource, y: Resource) -> bool:  # :319-339 (scalars only if present in y)
    if x1.milli_cpu + x2.milli_cpu > y.milli_cpu:
        return True
    if x1.memory + x2.memory > y.memory:
        return True
    for r in set(x1.scalars) | set(x2.scalars) | set(y.scalars):
        if r in y.scalars and x1.scalars.get(r, 0) + x2.scalars.get(r, 0) > y.scalars[r]:
            return True
    return False


def sum_less_than_equal(x1: Resource, x2: Resource, y: Resource) -> bool:  # :343-361
    if x1.milli_cpu + x2.milli_cpu > y.milli_cpu:
        return False
    if x1.memory + x2.memory > y.memory:
        return False
    for r in set(x1.scalars) | set(x2.scalars) | set(y.scalars):
        if r in y.scalars and x1.scalars.get(r, 0) + x2.scalars.get(r, 0) > y.scalars[r]:
            return False
    return True


def _approx_float(name: str, v: int) -> float:
    """Quantity.AsApproximateFloat64 of the quantity FromFrameworkToList builds for `name` (resource.go:35-50)."""
    if name == CPU:
        return float(v) * 1e-3          # NewMilliQuantity: base * math.Pow10(-3)
    return float(v)


def _quantity_value(name: str, v: int) -> int:
    """Quantity.Value(): milli quantities round up."""
    if name == CPU:
        return -((-v) // 1000)
    return v


class ElasticQuotaInfos(dict):  # namespace -> *ElasticQuotaInfo (a composite quota sits under several keys) :31-174
    def add(self, e: "ElasticQuotaInfo") -> None:  # :68-72
        for ns in sorted(e.namespaces):
            self[ns] = e

    def update(self, old: "ElasticQuotaInfo", new: "ElasticQuotaInfo") -> None:  # :51-66 (dict.update is not used on this class)
        for ns in sorted(new.namespaces):
            cur = self.get(ns)
            if cur is not None:
                new.pods, new.used = cur.pods, cur.used
            self[ns] = new
        for ns in sorted(old.namespaces):
            if ns not in new.namespaces:
                self.pop(ns, None)

    def delete(self, e: "ElasticQuotaInfo") -> None:  # :45-49
        for ns in sorted(e.namespaces):
            self.pop(ns, None)

    def aggregated_min(self) -> Resource:  # :154-163
        tot = Resource()
        for e in self.values():
            if e.min is not None:
                tot = res_sum(tot, e.min)
        return tot

    def aggregated_used(self) -> Resource:  # :165-174
        tot = Resource()
        for e in self.values():
            if e.used is not None:
                tot = res_sum(tot, e.used)
        return tot

    def aggregated_overquotas(self) -> Resource:  # :145-152
        tot = Resource()
        for e in self.values():
            tot = res_sum(tot, res_subtract_non_negative(e.min, e.used))
        return tot

    def aggregated_used_over_min_with(self, req: Resource) -> bool:  # :74-79
        mn = self.aggregated_min()
        used = self.aggregated_used()
        used.add_list(resource_to_list(req))
        return greater_than(used, mn)

    def guaranteed_overquota_percentages(self, e: ElasticQuotaInfo) -> Dict[str, float]:  # :101-119
        out: Dict[str, float] = {}
        if e.min is None:
            return out
        total = resource_to_list(self.aggregated_min())
        for r, m in resource_to_list(e.min).items():
            t = total.get(r, 0)
            p = 0.0
            if _quantity_value(r, t) > 0:
                p = _approx_float(r, m) / _approx_float(r, t)
            out[r] = p
        return out

    def guaranteed_overquotas(self, ns: str) -> Optional[Resource]:  # :81-99
        e = self.get(ns)
        if e is None:
            return None
        pct = self.guaranteed_overquota_percentages(e)
        agg = self.aggregated_overquotas()
        res = Resource()
        res.milli_cpu = go_int64_of_float(math.floor(float(agg.milli_cpu) * pct.get(CPU, 0.0)))
        res.memory = go_int64_of_float(math.floor(float(agg.memory) * pct.get(MEMORY, 0.0)))
        res.allowed_pods = go_int64_of_float(math.floor(float(agg.allowed_pods) * pct.get(PODS, 0.0)))
        res.ephemeral = go_int64_of_float(math.floor(float(agg.ephemeral) * pct.get(EPHEMERAL, 0.0)))
        for r, v in agg.scalars.items():
            res.scalars[r] = go_int64_of_float(math.floor(float(v) * pct.get(r, 0.0)))
        return res


PREFILTER_SUCCESS, PREFILTER_OVER_MAX, PREFILTER_OVER_MIN = 0, 1, 2


def prefilter(infos: ElasticQuotaInfos, pod: Pod, nvidia_gpu_memory_gb: int) -> int:
    """CapacityScheduling.PreFilter (capacity_scheduling.go:190-278) as it runs inside the gpupartitioner: the shared
    lister holds no nodes (cmd/gpupartitioner/gpupartitioner.go:316), so the nominated-pod sums reduce to the pod's own
    request."""
    req = resource_from_list(quota_pod_request(pod, nvidia_gpu_memory_gb))
    eq = infos.get(pod.namespace)
    if eq is None:
        return PREFILTER_SUCCESS
    nominated = Resource()
    nominated.add_list(resource_to_list(req))
    if eq.used_over_max_with(nominated):
        return PREFILTER_OVER_MAX
    if infos.aggregated_used_over_min_with(nominated):
        return PREFILTER_OVER_MIN
    return PREFILTER_SUCCESS


# ------------------------------------------------------------------------------------------------------------
# operator side: internal/controllers/elasticquota/elasticquota.go:38-118
# ------------------------------------------------------------------------------------------------------------
def _lte_on_keys_of_b(a: Dict[str, int], b: Dict[str, int]) -> bool:  # quota.LessThanOrEqual
    return all(a[k] <= v for k, v in b.items() if k in a)


def operator_sort(pods: List[Pod], nvidia_gpu_memory_gb: int) -> List[Pod]:
    """sortPodListForFindingOverQuotaPods (:77-104).  The reference uses sort.Slice (unstable) with a third key that is
    only a partial order; the canonical regime has unique (creationTimestamp, priority) pairs inside one quota, where any
    correct sort agrees.  Here: Python's stable sort on the full comparator."""
    import functools

    def cmp(a: Pod, b: Pod) -> int:
        def less(x: Pod, y: Pod) -> bool:
            if x.creation_ts != y.creation_ts:
                return x.creation_ts < y.creation_ts
            px, py = x.priority or 0, y.priority or 0
            if px != py:
                return px < py
            rx, ry = quota_pod_request(x, nvidia_gpu_memory_gb), quota_pod_request(y, nvidia_gpu_memory_gb)
            if rx != ry:
                return _lte_on_keys_of_b(rx, ry)
            return x.name < y.name
        if less(a, b):
            return -1
        if less(b, a):
            return 1
        return 0

    return sorted(pods, key=functools.cmp_to_key(cmp))


def patch_pods_and_compute_used(pods: List[Pod], quota_min: Dict[str, int], quota_max: Dict[str, int],
                                nvidia_gpu_memory_gb: int) -> Tuple[Dict[str, int], List[Tuple[str, bool]]]:
    """PatchPodsAndComputeUsedQuota (:38-72) -> (status.used, [(pod key, over_quota)] in sorted order)."""
    ordered = operator_sort(pods, nvidia_gpu_memory_gb)
    used: Dict[str, int] = {k: 0 for k in list(quota_min) + list(quota_max)}      # newZeroUsed :107-118
    labels = []
    for pod in ordered:
        for k, v in quota_pod_request(pod, nvidia_gpu_memory_gb).items():          # quota.Add: union of keys
            used[k] = used.get(k, 0) + v
        in_quota = _lte_on_keys_of_b(used, quota_min)
        labels.append((pod.key(), not in_quota))
    return {k: v for k, v in used.items() if k in quota_min}, labels


# ------------------------------------------------------------------------------------------------------------
# preemption, first pass of SelectVictimsOnNode (capacity_scheduling.go:511-596): which pods of a node become
# potential victims.  Every node works on its own clone of the quota snapshot; removing a victim un-reserves its request
# there (RemovePod extension :311-321 -> deletePodIfPresent), so later pods of the same node see the updated Used.
# `node_pods` must already be in the order the reference's sort gives (least important first, :516).
# ------------------------------------------------------------------------------------------------------------
def capacity_info_patches(pods: List[Pod], quota_min: Dict[str, int], quota_max: Dict[str, int],
                          nvidia_gpu_memory_gb: int) -> List[Tuple[str, str]]:
    """The Patch calls one PatchPodsAndComputeUsedQuota issues (patchCapacityInfoIfDifferent :120-149): (pod key, desired
    label) for every pod whose nos.nebuly.com/capacity label differs from the desired value, in the order they are patched."""
    _, labels = patch_pods_and_compute_used(pods, quota_min, quota_max, nvidia_gpu_memory_gb)
    by_key = {p.key(): p for p in pods}
    out = []
    for key, over in labels:
        want = "over-quota" if over else "in-quota"
        if (by_key[key].labels or {}).get("nos.nebuly.com/capacity", "") != want:
            out.append((key, want))
    return out


def clone_infos(infos: ElasticQuotaInfos) -> ElasticQuotaInfos:
    # elasticquotainfo.go:37-43 clones every map entry separately, so the namespaces of a composite quota become
    # DISTINCT objects in the snapshot: un-reserving a victim only touches the copy of the victim's own namespace.
    out = ElasticQuotaInfos()
    for ns, e in infos.items():
        out[ns] = ElasticQuotaInfo(e.name, list(e.namespaces), e.min.clone() if e.min else None, e.max.clone() if e.max else None,
                                   e.used.clone() if e.used else None, e.max_enforced, set(e.pods))
    return out


def over_quota_label(pod: Pod) -> bool:  # pkg/util/pod/pod.go:31-36
    return pod.labels.get("nos.nebuly.com/capacity") == "over-quota"


def select_potential_victims(infos: ElasticQuotaInfos, preemptor: Pod, nominated_in_eq: Resource, node_pods: List[Pod],
                             nvidia_gpu_memory_gb: int) -> List[bool]:
    snap = clone_infos(infos)
    prio = pod_priority(preemptor)
    pre_eq = snap.get(preemptor.namespace)
    victims = [False] * len(node_pods)

    def remove(pv: Pod):
        e = snap.get(pv.namespace)
        if e is not None:
            e.unreserve(resource_from_list(quota_pod_request(pv, nvidia_gpu_memory_gb)))   # pod assumed accounted (present)

    if pre_eq is not None:
        more_than_min = pre_eq.used_over_min_with(nominated_in_eq)
        for i, pv in enumerate(node_pods):
            pv_eq = snap.get(pv.namespace)
            if pv_eq is None:
                continue
            if more_than_min:
                if pv.namespace == preemptor.namespace:
                    if pod_priority(pv) < prio:
                        victims[i] = True
                        remove(pv)
                    continue
                if not over_quota_label(pv):
                    continue
                g = snap.guaranteed_overquotas(preemptor.namespace)
                min_plus_g = res_sum(g, pre_eq.min)
                if pre_eq.used_lte_with(min_plus_g, nominated_in_eq):
                    pg = snap.guaranteed_overquotas(pv.namespace)
                    pv_min_plus_g = res_sum(pg, pv_eq.min)
                    if pv_eq.used_over(pv_min_plus_g):
                        victims[i] = True
                        remove(pv)
            else:
                if pv.namespace != preemptor.namespace and pv_eq.used_over_min() and over_quota_label(pv):
                    victims[i] = True
                    remove(pv)
    else:
        for i, pv in enumerate(node_pods):
            if pv.namespace in snap:
                continue
            if pod_priority(pv) < prio:
                victims[i] = True
    return victims


# ------------------------------------------------------------------------------------------------------------
# preemption dry run, the whole of preemptor.SelectVictimsOnNode (capacity_scheduling.go:468-675) for one node:
# first pass as above, Filter with the potential victims gone, quota checks, the PDB split, and the reprieve loops.
# Third-party pieces restated from k8s.io/kubernetes v1.25.4 (not in /root/reference; the reference's TestDryRunPreemption
# asserts the candidate count only -- its three fixtures are tests/golden/dryrun_preemption.json): NodeInfo.RemovePod / AddPodInfo (Requested and len(Pods) bookkeeping, an
# error when the pod is not on the node), NodeResourcesFit as the only Filter, no nominated pods on the node.
# ------------------------------------------------------------------------------------------------------------
VICTIMS_OK, VICTIMS_NONE, VICTIMS_FILTER, VICTIMS_MAX_QUOTA, VICTIMS_MIN_QUOTA, VICTIMS_ERROR = 0, 1, 2, 3, 4, 5


@dataclass
class PreemptionNode:
    allocatable: Resource
    requested: Resource            # of every pod in `pods`
    pods: List[Pod]                # least important first (the order sort.Slice with !MoreImportantPod leaves, :516)
    pdbs_of_pod: List[List[int]]   # per pod: ids of the PDBs that match it and do not list it in DisruptedPods (:864-883)


def select_victims_on_node(infos: ElasticQuotaInfos, preemptor: Pod, pod_req: Resource, nominated_in_eq: Resource, nominated: Resource,
                           node: PreemptionNode, pdb_allowed: List[int], nvidia_gpu_memory_gb: int):
    """-> (status, victims as indices into node.pods in the order the reference appends them, numViolatingVictim)"""
    from nos_model import NodeInfo, node_resources_fit
    snap = clone_infos(infos)
    prio = pod_priority(preemptor)
    pre_eq = snap.get(preemptor.namespace)
    ni = NodeInfo(node.allocatable.clone(), node.requested.clone(), len(node.pods))
    on_node = set(range(len(node.pods)))
    fit_req = resource_from_list(compute_pod_request(preemptor))

    def node_req(pv: Pod) -> Resource:
        return resource_from_list(compute_pod_request(pv))

    def quota_req(pv: Pod) -> Resource:
        return resource_from_list(quota_pod_request(pv, nvidia_gpu_memory_gb))

    def remove_pod(i: int) -> bool:                      # removePod closure (:492-501); False = NodeInfo.RemovePod error
        if i not in on_node:
            return False
        on_node.discard(i)
        r = node_req(node.pods[i])
        ni.requested.milli_cpu -= r.milli_cpu; ni.requested.memory -= r.memory; ni.requested.ephemeral -= r.ephemeral
        for k, v in r.scalars.items():
            ni.requested.scalars[k] = ni.requested.scalars.get(k, 0) - v
        ni.num_pods -= 1
        e = snap.get(node.pods[i].namespace)
        if e is not None:
            e.delete_pod_if_present(("pod", i), quota_req(node.pods[i]))
        return True

    def add_pod(i: int) -> None:                         # addPod closure (:502-509)
        on_node.add(i)
        ni.add_pod(node.pods[i])
        e = snap.get(node.pods[i].namespace)
        if e is not None:
            e.add_pod_if_not_present(("pod", i), quota_req(node.pods[i]))

    for i, pv in enumerate(node.pods):                   # every pod of the node is accounted in its quota (Reserve / informer)
        e = snap.get(pv.namespace)
        if e is not None:
            e.pods.add(("pod", i))

    potential: List[int] = []
    if pre_eq is not None:
        more_than_min = pre_eq.used_over_min_with(nominated_in_eq)
        for i, pv in enumerate(node.pods):
            pv_eq = snap.get(pv.namespace)
            if pv_eq is None:
                continue
            if more_than_min:
                if pv.namespace == preemptor.namespace:
                    if pod_priority(pv) < prio:
                        potential.append(i); remove_pod(i)
                    continue
                if not over_quota_label(pv):
                    continue
                g = snap.guaranteed_overquotas(preemptor.namespace)
                if pre_eq.used_lte_with(res_sum(g, pre_eq.min), nominated_in_eq):
                    pg = snap.guaranteed_overquotas(pv.namespace)
                    if pv_eq.used_over(res_sum(pg, pv_eq.min)):
                        potential.append(i); remove_pod(i)
            else:
                if pv.namespace != preemptor.namespace and pv_eq.used_over_min() and over_quota_label(pv):
                    potential.append(i); remove_pod(i)
    else:
        for i, pv in enumerate(node.pods):
            if pv.namespace in snap:
                continue
            if pod_priority(pv) < prio:
                potential.append(i); remove_pod(i)

    if not potential:                                    # :598-602
        return VICTIMS_NONE, [], 0
    if not node_resources_fit(fit_req, ni):              # :610-612
        return VICTIMS_FILTER, [], 0
    if pre_eq is not None:                               # :617-624
        if pre_eq.used_over_max_with(pod_req):
            return VICTIMS_MAX_QUOTA, [], 0
        if snap.aggregated_used_over_min_with(pod_req):
            return VICTIMS_MIN_QUOTA, [], 0

    potential = list(reversed(potential))                # sort by MoreImportantPod (:628-630): most important first
    allowed = list(pdb_allowed)                          # filterPodsWithPDBViolation (:850-895)
    violating, non_violating = [], []
    for i in potential:
        bad = False
        for b in node.pdbs_of_pod[i]:
            allowed[b] -= 1
            if allowed[b] < 0:
                bad = True
        (violating if bad else non_violating).append(i)

    victims: List[int] = []

    def reprieve(i: int):                                # :636-660 -> (fits, ok)
        add_pod(i)
        fits = node_resources_fit(fit_req, ni)
        if not fits:
            if not remove_pod(i):
                return fits, False
            victims.append(i)
        if pre_eq is not None and (pre_eq.used_over_max_with(nominated_in_eq) or snap.aggregated_used_over_min_with(nominated)):
            if not remove_pod(i):                        # the pod is already gone when it did not fit: RemovePod errors (:653-655)
                return fits, False
            victims.append(i)
        return fits, True

    num_violating = 0
    for i in violating:
        fits, ok = reprieve(i)
        if not ok:
            return VICTIMS_ERROR, [], 0
        if not fits:
            num_violating += 1
    for i in non_violating:
        _, ok = reprieve(i)
        if not ok:
            return VICTIMS_ERROR, [], 0
    return VICTIMS_OK, victims, num_violating
