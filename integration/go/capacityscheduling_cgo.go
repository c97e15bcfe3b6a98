// capacityscheduling_cgo.go -- the quota side: glue between the CapacityScheduling plugin
// (pkg/scheduler/plugins/capacityscheduling/capacity_scheduling.go) and the resident quota state of libnosplan.so.
//
// UNCOMPILED HERE (no Go toolchain in this image); integration/go/check_arity.py checks every C call against
// include/nosplan.h.  The plugin keeps its ElasticQuotaInfos exactly as today for everything that needs the object model
// (informers, PostFilter bookkeeping); this type mirrors them on the GPU so that the arithmetic -- Used, the aggregates,
// GetGuaranteedOverquotas and the PreFilter verdicts of a batch of pending pods -- comes from the engine:
//
//   capacityscheduling.New (capacity_scheduling.go:113)            -> newQuotaMirror + Begin(every ElasticQuotaInfo)
//   addPod / updatePod / deletePod, Reserve / Unreserve            -> Apply(events): AddPodIfNotPresent / DeletePodIfPresent
//     (capacity_scheduling.go:311-321, 381-445; elasticquotainfo.go:276-310)
//   PreFilter's quota part (capacity_scheduling.go:190-278)        -> PreFilter(pending pods): one verdict per pod
package capacityscheduling

/*
#cgo LDFLAGS: -lnosplan -lcudart
#include <stdlib.h>
#include "nosplan.h"
*/
import "C"

import (
	"fmt"
	"sort"
	"unsafe"

	v1 "k8s.io/api/core/v1"
	"k8s.io/kubernetes/pkg/scheduler/framework"
)

// quotaMirror interns namespaces, quota rows and resource columns once (Begin) and afterwards only ships events.
type quotaMirror struct {
	h       *C.nosplan_t
	cols    []v1.ResourceName // column 0 cpu (milli), 1 memory, 2 ephemeral-storage, 3 pods, 4.. scalar resources ascending
	colID   map[v1.ResourceName]int
	nsID    map[string]int32
	podID   map[string]int32 // pod UID -> dense id < capacity
	cap     int32
	gpuMem  int // column of nos.nebuly.com/gpu-memory, -1 = not tracked
}

const nonNil = uint32(1) << 31 // NOSPLAN_QUOTA_NONNIL: "the *framework.Resource is not nil"

func (m *quotaMirror) row(r *framework.Resource, dst []int64) uint32 {
	if r == nil {
		return 0
	}
	mask := nonNil
	dst[0], dst[1], dst[2], dst[3] = r.MilliCPU, r.Memory, r.EphemeralStorage, int64(r.AllowedPodNumber)
	for name, q := range r.ScalarResources {
		c := m.colID[name]
		dst[c] = q
		mask |= 1 << uint(c)
	}
	return mask
}

// Begin uploads the quota tables: one row per distinct *ElasticQuotaInfo, referenced by every namespace it is stored under
// (a CompositeElasticQuota is the same pointer under several keys, elasticquotainfo.go:68-72).
func newQuotaMirror(h *C.nosplan_t, infos ElasticQuotaInfos, gpuMemoryGB int64, migGB map[v1.ResourceName]int64, podCapacity int32) (*quotaMirror, error) {
	m := &quotaMirror{h: h, colID: map[v1.ResourceName]int{}, nsID: map[string]int32{}, podID: map[string]int32{}, cap: podCapacity, gpuMem: -1}
	scalars := map[v1.ResourceName]bool{}
	rows := map[*ElasticQuotaInfo]int32{}
	order := []*ElasticQuotaInfo{}
	nss := make([]string, 0, len(infos))
	for ns := range infos {
		nss = append(nss, ns)
	}
	sort.Strings(nss)
	for _, ns := range nss {
		e := infos[ns]
		if _, ok := rows[e]; !ok {
			rows[e] = int32(len(order))
			order = append(order, e)
		}
		for _, r := range []*framework.Resource{e.Min, e.Max, e.Used} {
			if r != nil {
				for name := range r.ScalarResources {
					scalars[name] = true
				}
			}
		}
	}
	m.cols = []v1.ResourceName{v1.ResourceCPU, v1.ResourceMemory, v1.ResourceEphemeralStorage, v1.ResourcePods}
	names := make([]string, 0, len(scalars))
	for s := range scalars {
		names = append(names, string(s))
	}
	sort.Strings(names)
	for _, s := range names {
		m.cols = append(m.cols, v1.ResourceName(s))
	}
	for c, name := range m.cols {
		m.colID[name] = c
		if name == "nos.nebuly.com/gpu-memory" {
			m.gpuMem = c
		}
	}
	E, R := len(order), len(m.cols)
	nsQuota := make([]int32, len(nss))
	for i, ns := range nss {
		m.nsID[ns] = int32(i)
		nsQuota[i] = rows[infos[ns]]
	}
	min, max, used := make([]int64, E*R), make([]int64, E*R), make([]int64, E*R)
	minMask, maxMask, usedMask, enforced := make([]uint32, E), make([]uint32, E), make([]uint32, E), make([]uint8, E)
	for e, info := range order {
		minMask[e] = m.row(info.Min, min[e*R:(e+1)*R])
		maxMask[e] = m.row(info.Max, max[e*R:(e+1)*R])
		usedMask[e] = m.row(info.Used, used[e*R:(e+1)*R])
		if info.MaxEnforced {
			enforced[e] = 1
		}
	}
	coeff := make([]int64, R) // pkg/gpu/util/resource.go:70-86: GB of gpu-memory one unit of the column stands for
	for c, name := range m.cols {
		if name == "nvidia.com/gpu" {
			coeff[c] = gpuMemoryGB
		} else if gb, ok := migGB[name]; ok {
			coeff[c] = gb
		}
	}
	q := C.nosplan_quotas{
		n_quotas: C.int32_t(E), n_res: C.int32_t(R), n_namespaces: C.int32_t(len(nss)), gpu_memory_col: C.int32_t(m.gpuMem),
		ns_quota: (*C.int32_t)(unsafe.Pointer(&nsQuota[0])),
		min: (*C.int64_t)(unsafe.Pointer(&min[0])), max: (*C.int64_t)(unsafe.Pointer(&max[0])), used_init: (*C.int64_t)(unsafe.Pointer(&used[0])),
		min_mask: (*C.uint32_t)(unsafe.Pointer(&minMask[0])), max_mask: (*C.uint32_t)(unsafe.Pointer(&maxMask[0])),
		used_mask: (*C.uint32_t)(unsafe.Pointer(&usedMask[0])), max_enforced: (*C.uint8_t)(unsafe.Pointer(&enforced[0])),
		gpu_memory_coeff: (*C.int64_t)(unsafe.Pointer(&coeff[0])),
	}
	if rc := C.nosplan_quota_state_begin(h, &q, C.int32_t(podCapacity)); rc != 0 {
		return nil, fmt.Errorf("nosplan_quota_state_begin: %s", C.GoString(C.nosplan_last_error(h)))
	}
	return m, nil
}

// podEvent is one AddPodIfNotPresent (add == true) or DeletePodIfPresent with the request the pod has NOW
// (the reference unreserves what the pod requests at the time of the call, elasticquotainfo.go:294-310).
type podEvent struct {
	add     bool
	pod     *v1.Pod
	request v1.ResourceList // computePodResourceRequest(pod), capacity_scheduling.go:752-770
}

func (m *quotaMirror) requestRow(rl v1.ResourceList, dst []int64) uint32 {
	var mask uint32
	for name, q := range rl {
		switch name {
		case v1.ResourceCPU:
			dst[0] = q.MilliValue()
		case v1.ResourceMemory:
			dst[1] = q.Value()
		case v1.ResourceEphemeralStorage:
			dst[2] = q.Value()
		case v1.ResourcePods:
			dst[3] = q.Value()
		default:
			if c, ok := m.colID[name]; ok {
				dst[c] = q.Value()
				mask |= 1 << uint(c) // every requested key becomes present in Used, quantity 0 included (:197-199)
			}
		}
	}
	return mask
}

// Apply sends one batch of events (a pod may occur once per batch) and returns Used of every quota row afterwards.
func (m *quotaMirror) Apply(events []podEvent) (used []int64, err error) {
	n, R := len(events), len(m.cols)
	if n == 0 {
		return nil, nil
	}
	op, pod, ns, mask := make([]uint8, n), make([]int32, n), make([]int32, n), make([]uint32, n)
	req := make([]int64, n*R)
	for i, ev := range events {
		if !ev.add {
			op[i] = C.NOSPLAN_QUOTA_EVENT_DELETE
		}
		uid := string(ev.pod.UID)
		id, ok := m.podID[uid]
		if !ok {
			id = int32(len(m.podID))
			if id >= m.cap {
				return nil, fmt.Errorf("more than %d pods: begin a larger state", m.cap)
			}
			m.podID[uid] = id
		}
		pod[i] = id
		nsid, ok := m.nsID[ev.pod.Namespace]
		if !ok {
			nsid = -1 // not subject to any quota: the event changes nothing
		}
		ns[i] = nsid
		mask[i] = m.requestRow(ev.request, req[i*R:(i+1)*R])
	}
	evs := C.nosplan_quota_events{
		n_events: C.int32_t(n), op: (*C.uint8_t)(unsafe.Pointer(&op[0])), pod_id: (*C.int32_t)(unsafe.Pointer(&pod[0])),
		ns: (*C.int32_t)(unsafe.Pointer(&ns[0])), request: (*C.int64_t)(unsafe.Pointer(&req[0])),
		request_mask: (*C.uint32_t)(unsafe.Pointer(&mask[0])),
	}
	E := 0
	for range m.nsID {
		E++ // upper bound of the quota rows; the engine writes E rows
	}
	used = make([]int64, E*R)
	out := C.nosplan_quota_out{used: (*C.int64_t)(unsafe.Pointer(&used[0]))}
	var applied C.int32_t
	var st C.nosplan_stats
	if rc := C.nosplan_quota_state_apply(m.h, &evs, &out, &applied, &st); rc != 0 {
		return nil, fmt.Errorf("nosplan_quota_state_apply: %s", C.GoString(C.nosplan_last_error(m.h)))
	}
	return used, nil
}

// PreFilter is the quota part of CapacityScheduling.PreFilter (capacity_scheduling.go:190-278) for a batch of pending pods
// against the state as the events applied so far left it: NOSPLAN_PREFILTER_SUCCESS / _OVER_MAX / _OVER_MIN per pod.
func (m *quotaMirror) PreFilter(pods []*v1.Pod, requests []v1.ResourceList) ([]uint8, error) {
	n, R := len(pods), len(m.cols)
	if n == 0 {
		return nil, nil
	}
	ns, mask, req, verdict := make([]int32, n), make([]uint32, n), make([]int64, n*R), make([]uint8, n)
	for i, p := range pods {
		nsid, ok := m.nsID[p.Namespace]
		if !ok {
			nsid = -1
		}
		ns[i] = nsid
		mask[i] = m.requestRow(requests[i], req[i*R:(i+1)*R])
	}
	pend := C.nosplan_quota_pods{
		n_pods: C.int32_t(n), ns: (*C.int32_t)(unsafe.Pointer(&ns[0])), request: (*C.int64_t)(unsafe.Pointer(&req[0])),
		request_mask: (*C.uint32_t)(unsafe.Pointer(&mask[0])),
	}
	var st C.nosplan_stats
	if rc := C.nosplan_quota_state_prefilter(m.h, &pend, (*C.uint8_t)(unsafe.Pointer(&verdict[0])), &st); rc != 0 {
		return nil, fmt.Errorf("nosplan_quota_state_prefilter: %s", C.GoString(C.nosplan_last_error(m.h)))
	}
	return verdict, nil
}
