// planner_cgo.go -- the cgo shim a maintainer drops into internal/partitioning/core/ of nos to plan on libnosplan.so.
//
// UNCOMPILED HERE: the image this repository is built in has no Go toolchain and no module cache, so this file has never
// been through `go build`.  What IS checked mechanically (integration/go/check_arity.py, run by the CPU test-suite) is
// that every C.nosplan_* call below exists in include/nosplan.h with the number of arguments used here, and that every
// C.nosplan_<struct> field named in a composite literal is a field of that struct.
//
// It implements core.Planner (internal/partitioning/core/interface.go:27-29) and is returned by the two factories
// mig.NewPlanner / mps.NewPlanner (internal/partitioning/mig/factory.go:31-37, internal/partitioning/mps/factory.go:44-50)
// in place of core.NewPlanner(...); nothing else in the gpupartitioner changes.
package core

/*
#cgo LDFLAGS: -lnosplan -lcudart
#include <stdlib.h>
#include <cuda_runtime_api.h>
#include "nosplan.h"
*/
import "C"

import (
	"context"
	"fmt"
	"regexp"
	"sort"
	"strconv"
	"unsafe"

	"github.com/nebuly-ai/nos/internal/partitioning/state"
	"github.com/nebuly-ai/nos/pkg/gpu"
	"github.com/nebuly-ai/nos/pkg/gpu/mig"
	"github.com/nebuly-ai/nos/pkg/gpu/slicing"
	"github.com/nebuly-ai/nos/pkg/resource"
	v1 "k8s.io/api/core/v1"
	"k8s.io/component-helpers/scheduling/corev1"
	"k8s.io/kubernetes/pkg/scheduler/framework"
)

// pinned is a page-locked host buffer (cudaHostAlloc) viewed as Go slices.  The engine reads page-locked caller arrays in
// place with the copy engine (nosplan_upload checks cudaPointerGetAttributes) and writes the plan straight into
// page-locked output buffers; pageable Go memory works too, it just passes through the engine's own staging buffer.
// Buffers are grown, never shrunk, and live as long as the planner: cgo's pointer rules are met because they are C memory.
type pinned struct {
	p   unsafe.Pointer
	cap int
}

func (b *pinned) bytes(n int) []byte {
	if n > b.cap {
		if b.p != nil {
			C.cudaFreeHost(b.p)
		}
		c := n + n/4 + 4096
		if rc := C.cudaHostAlloc(&b.p, C.size_t(c), C.cudaHostAllocDefault); rc != C.cudaSuccess {
			panic(fmt.Sprintf("cudaHostAlloc(%d): %d", c, int(rc)))
		}
		b.cap = c
	}
	if n == 0 {
		return nil
	}
	return unsafe.Slice((*byte)(b.p), n)
}
func i32s(b *pinned, n int) []int32   { x := b.bytes(4 * n); if n == 0 { return nil }; return unsafe.Slice((*int32)(unsafe.Pointer(&x[0])), n) }
func i64s(b *pinned, n int) []int64   { x := b.bytes(8 * n); if n == 0 { return nil }; return unsafe.Slice((*int64)(unsafe.Pointer(&x[0])), n) }
func u16s(b *pinned, n int) []uint16  { x := b.bytes(2 * n); if n == 0 { return nil }; return unsafe.Slice((*uint16)(unsafe.Pointer(&x[0])), n) }
func u8s(b *pinned, n int) []uint8    { return b.bytes(n) }
func p32(s []int32) *C.int32_t        { if len(s) == 0 { return nil }; return (*C.int32_t)(unsafe.Pointer(&s[0])) }
func p64(s []int64) *C.int64_t        { if len(s) == 0 { return nil }; return (*C.int64_t)(unsafe.Pointer(&s[0])) }
func pu16(s []uint16) *C.uint16_t     { if len(s) == 0 { return nil }; return (*C.uint16_t)(unsafe.Pointer(&s[0])) }
func pu8(s []uint8) *C.uint8_t        { if len(s) == 0 { return nil }; return (*C.uint8_t)(unsafe.Pointer(&s[0])) }

// packer holds the SoA form of one (snapshot, pending pods) pair: the Go twin of nos_b200/packing.py:pack, which is the
// executable specification the parity tests drive (tests/test_new_node_golden.py, tests/test_host_mirror.py).
type packer struct {
	kind     C.int32_t
	profiles []string // canonical order: MIG ascending name, MPS ascending (size GB, name)
	profID   map[string]int
	others   []string // non-slice scalar resource names, ascending
	otherID  map[string]int
	models   []gpu.Model
	modelID  map[gpu.Model]int
	names    []string // node names, byte-wise ascending (core/snapshot.go:126-128)
	G        int

	// buffers (page-locked, reused across Plan() calls)
	bProfMem, bProfGI, bModelOff, bGeom, bGPUCount, bGPUModel, bGPUIndex, bGPUMem, bUsed, bFree, bUMask, bFMask,
	bAlloc, bReq, bAllowed, bNPods, bPrio, bKey, bBase, bSPtr, bSCol, bSQty, bOutCount, bOutMask pinned

	profMem, profGI, modelOff, gpuCount, gpuModel, gpuIndex, gpuMem, allowed, nPods, prio, key, sPtr, sCol []int32
	geom, used, free, outCount                                                                            []uint8
	uMask, fMask, outMask                                                                                 []uint16
	alloc, req, base, sQty                                                                                []int64
}

var (
	migSizeRe = regexp.MustCompile(`(\d+)gb`) // pkg/gpu/mig/profile.go:70-75 getMemorySlices
	migGiRe   = regexp.MustCompile(`(\d+)g\.`) // pkg/gpu/mig/profile.go:77-82 getGiSlices
)

func atoiOr0(s string) int32 { n, _ := strconv.Atoi(s); return int32(n) }

// sliceProfileOf classifies a resource name with the reference's own predicates: "" = not a slice resource of this kind.
func (pk *packer) sliceProfileOf(r v1.ResourceName) string {
	if pk.kind == C.NOSPLAN_KIND_MIG {
		if p, err := mig.ExtractProfileName(r); err == nil { // pkg/gpu/mig/util.go:46-58
			return string(p)
		}
		return ""
	}
	if p, err := slicing.ExtractProfileName(r); err == nil { // pkg/gpu/slicing/util.go:34-57
		return string(p)
	}
	return ""
}

func (pk *packer) resourceName(prof string) v1.ResourceName {
	if pk.kind == C.NOSPLAN_KIND_MIG {
		return mig.ProfileName(prof).AsResourceName()
	}
	return slicing.ProfileName(prof).AsResourceName()
}

// podRequest is resource.ComputePodRequest (pkg/resource/resource.go:127-146) split into the engine's columns.
type podRequest struct {
	cpuMilli, memory, ephemeral int64
	scalars                     map[v1.ResourceName]int64 // every key present in the request, zero quantities included
}

func computePodRequest(pod *v1.Pod) podRequest {
	rl := resource.ComputePodRequest(*pod)
	pr := podRequest{scalars: map[v1.ResourceName]int64{}}
	for name, q := range rl {
		switch name {
		case v1.ResourceCPU:
			pr.cpuMilli = q.MilliValue()
		case v1.ResourceMemory:
			pr.memory = q.Value()
		case v1.ResourceEphemeralStorage:
			pr.ephemeral = q.Value()
		case v1.ResourcePods:
		default:
			pr.scalars[name] = q.Value()
		}
	}
	return pr
}

// newPacker interns names and sizes every buffer: pass 1 over the snapshot and the pods collects the profile names (from
// the GPUs' maps, the geometry tables, the nodes' scalar resources and the pods' requests), the other scalar resource
// names and the GPU models; pass 2 fills the arrays.
func (pk *packer) pack(kind C.int32_t, snapshot Snapshot, pods []v1.Pod) error {
	pk.kind = kind
	nodes := snapshot.GetNodes()
	pk.names = pk.names[:0]
	for n := range nodes {
		pk.names = append(pk.names, n)
	}
	sort.Strings(pk.names)

	profSet, otherSet, modelSet := map[string]bool{}, map[string]bool{}, map[gpu.Model]bool{}
	scan := func(rl map[v1.ResourceName]int64) {
		for r := range rl {
			if p := pk.sliceProfileOf(r); p != "" {
				profSet[p] = true
			} else {
				otherSet[string(r)] = true
			}
		}
	}
	pk.G = 1
	known := mig.GetKnownGeometries() // pkg/gpu/mig/known_configs.go:152-157: runtime data, handed to the engine per call
	for _, name := range pk.names {
		ni := nodes[name].NodeInfo()
		scan(ni.Allocatable.ScalarResources)
		scan(ni.Requested.ScalarResources)
		switch n := nodes[name].(type) {
		case *mig.Node:
			if len(n.GPUs) > pk.G {
				pk.G = len(n.GPUs)
			}
			for i := range n.GPUs {
				modelSet[n.GPUs[i].GetModel()] = true
				for p := range n.GPUs[i].GetUsedMigDevices() {
					profSet[string(p)] = true
				}
				for p := range n.GPUs[i].GetFreeMigDevices() {
					profSet[string(p)] = true
				}
			}
		case *slicing.Node:
			if len(n.GPUs) > pk.G {
				pk.G = len(n.GPUs)
			}
			for i := range n.GPUs {
				for p := range n.GPUs[i].UsedProfiles {
					profSet[string(p)] = true
				}
				for p := range n.GPUs[i].FreeProfiles {
					profSet[string(p)] = true
				}
			}
		default:
			return fmt.Errorf("node %q: neither *mig.Node nor *slicing.Node", name)
		}
	}
	if kind == C.NOSPLAN_KIND_MIG {
		for m := range modelSet {
			for _, g := range known[m] {
				for s := range g {
					profSet[string(s.(mig.ProfileName))] = true
				}
			}
		}
	}
	reqs := make([]podRequest, len(pods))
	for i := range pods {
		reqs[i] = computePodRequest(&pods[i])
		scan(reqs[i].scalars)
	}
	pk.profiles = pk.profiles[:0]
	for p := range profSet {
		pk.profiles = append(pk.profiles, p)
	}
	if kind == C.NOSPLAN_KIND_MIG {
		sort.Strings(pk.profiles)
	} else {
		sort.Slice(pk.profiles, func(a, b int) bool {
			sa, sb := slicing.ProfileName(pk.profiles[a]).GetMemorySizeGB(), slicing.ProfileName(pk.profiles[b]).GetMemorySizeGB()
			if sa != sb {
				return sa < sb
			}
			return pk.profiles[a] < pk.profiles[b]
		})
	}
	if len(pk.profiles) > int(C.NOSPLAN_MAX_PROFILES) {
		return fmt.Errorf("%d slice profiles > NOSPLAN_MAX_PROFILES", len(pk.profiles))
	}
	pk.others = pk.others[:0]
	for o := range otherSet {
		pk.others = append(pk.others, o)
	}
	sort.Strings(pk.others)
	pk.models = pk.models[:0]
	for m := range modelSet {
		pk.models = append(pk.models, m)
	}
	sort.Slice(pk.models, func(a, b int) bool { return pk.models[a] < pk.models[b] })
	pk.profID, pk.otherID, pk.modelID = map[string]int{}, map[string]int{}, map[gpu.Model]int{}
	for i, p := range pk.profiles {
		pk.profID[p] = i
	}
	for i, o := range pk.others {
		pk.otherID[o] = i
	}
	for i, m := range pk.models {
		pk.modelID[m] = i
	}

	K, N, G, P := len(pk.profiles), len(pk.names), pk.G, len(pods)
	R := 3 + K + len(pk.others)
	// ---- profiles and geometry tables
	pk.profMem, pk.profGI = i32s(&pk.bProfMem, K), i32s(&pk.bProfGI, K)
	for i, p := range pk.profiles {
		if kind == C.NOSPLAN_KIND_MIG {
			if m := migSizeRe.FindStringSubmatch(p); m != nil {
				pk.profMem[i] = atoiOr0(m[1])
			}
			if m := migGiRe.FindStringSubmatch(p); m != nil {
				pk.profGI[i] = atoiOr0(m[1])
			}
		} else {
			pk.profMem[i], pk.profGI[i] = int32(slicing.ProfileName(p).GetMemorySizeGB()), 0
		}
	}
	rows := 0
	pk.modelOff = i32s(&pk.bModelOff, len(pk.models)+1)
	for i, m := range pk.models {
		pk.modelOff[i] = int32(rows)
		rows += len(known[m])
	}
	pk.modelOff[len(pk.models)] = int32(rows)
	pk.geom = u8s(&pk.bGeom, 16*rows)
	for i := range pk.geom {
		pk.geom[i] = 0
	}
	row := 0
	for _, m := range pk.models {
		for _, g := range known[m] { // table order is kept: it is the canonical tie-break of the geometry arg-max
			for s, q := range g {
				pk.geom[16*row+pk.profID[string(s.(mig.ProfileName))]] = uint8(q)
			}
			row++
		}
	}
	// ---- nodes
	pk.gpuCount, pk.allowed, pk.nPods = i32s(&pk.bGPUCount, N), i32s(&pk.bAllowed, N), i32s(&pk.bNPods, N)
	pk.gpuModel, pk.gpuIndex, pk.gpuMem = i32s(&pk.bGPUModel, N*G), i32s(&pk.bGPUIndex, N*G), i32s(&pk.bGPUMem, N*G)
	pk.used, pk.free = u8s(&pk.bUsed, 16*N*G), u8s(&pk.bFree, 16*N*G)
	pk.uMask, pk.fMask = u16s(&pk.bUMask, N*G), u16s(&pk.bFMask, N*G)
	pk.alloc, pk.req = i64s(&pk.bAlloc, N*R), i64s(&pk.bReq, N*R)
	for i := range pk.used {
		pk.used[i], pk.free[i] = 0, 0
	}
	for i := range pk.uMask {
		pk.uMask[i], pk.fMask[i], pk.gpuModel[i], pk.gpuIndex[i], pk.gpuMem[i] = 0, 0, 0, 0, 0
	}
	fillRes := func(dst []int64, r *framework.Resource) {
		for c := range dst {
			dst[c] = 0
		}
		dst[0], dst[1], dst[2] = r.MilliCPU, r.Memory, r.EphemeralStorage
		for name, q := range r.ScalarResources {
			if p := pk.sliceProfileOf(name); p != "" {
				dst[3+pk.profID[p]] = q
			} else {
				dst[3+K+pk.otherID[string(name)]] = q
			}
		}
	}
	putMap := func(cnt []uint8, mask *uint16, p string, q int) { // a Go map is (count, key present): zero-valued keys matter
		id := pk.profID[p]
		cnt[id] = uint8(q)
		*mask |= 1 << uint(id)
	}
	for i, name := range pk.names {
		ni := nodes[name].NodeInfo()
		fillRes(pk.alloc[i*R:(i+1)*R], ni.Allocatable)
		fillRes(pk.req[i*R:(i+1)*R], ni.Requested)
		pk.allowed[i], pk.nPods[i] = int32(ni.Allocatable.AllowedPodNumber), int32(len(ni.Pods))
		switch n := nodes[name].(type) {
		case *mig.Node:
			pk.gpuCount[i] = int32(len(n.GPUs))
			for j := range n.GPUs { // the order of Node.GPUs is the order the reference iterates them in
				x := i*G + j
				pk.gpuModel[x], pk.gpuIndex[x] = int32(pk.modelID[n.GPUs[j].GetModel()]), int32(n.GPUs[j].GetIndex())
				for p, q := range n.GPUs[j].GetUsedMigDevices() {
					putMap(pk.used[16*x:16*x+16], &pk.uMask[x], string(p), q)
				}
				for p, q := range n.GPUs[j].GetFreeMigDevices() {
					putMap(pk.free[16*x:16*x+16], &pk.fMask[x], string(p), q)
				}
			}
		case *slicing.Node:
			pk.gpuCount[i] = int32(len(n.GPUs))
			for j := range n.GPUs {
				x := i*G + j
				pk.gpuIndex[x], pk.gpuMem[x] = int32(n.GPUs[j].Index), int32(n.GPUs[j].MemoryGB)
				for p, q := range n.GPUs[j].UsedProfiles {
					putMap(pk.used[16*x:16*x+16], &pk.uMask[x], string(p), q)
				}
				for p, q := range n.GPUs[j].FreeProfiles {
					putMap(pk.free[16*x:16*x+16], &pk.fMask[x], string(p), q)
				}
			}
		}
	}
	// ---- pods, in the caller's order (the engine applies NewPodSorter itself, core/util.go:34-73)
	pk.prio, pk.key, pk.base, pk.sPtr = i32s(&pk.bPrio, P), i32s(&pk.bKey, P), i64s(&pk.bBase, 3*P), i32s(&pk.bSPtr, P+1)
	nnz := 0
	for i := range reqs {
		nnz += len(reqs[i].scalars)
	}
	pk.sCol, pk.sQty = i32s(&pk.bSCol, nnz), i64s(&pk.bSQty, nnz)
	keyID := map[string]int32{} // duplicates of namespace/name share one SliceTracker entry (core/tracker.go:38-47)
	w := 0
	for i := range pods {
		pk.prio[i] = corev1.PodPriority(&pods[i]) // core/util.go:41-42
		k := pods[i].Namespace + "/" + pods[i].Name
		id, ok := keyID[k]
		if !ok {
			id = int32(len(keyID))
			keyID[k] = id
		}
		pk.key[i] = id
		pk.base[3*i], pk.base[3*i+1], pk.base[3*i+2] = reqs[i].cpuMilli, reqs[i].memory, reqs[i].ephemeral
		pk.sPtr[i] = int32(w)
		cols := make([]int, 0, len(reqs[i].scalars))
		qty := map[int]int64{}
		for name, q := range reqs[i].scalars {
			c := 0
			if p := pk.sliceProfileOf(name); p != "" {
				c = 3 + pk.profID[p]
			} else {
				c = 3 + K + pk.otherID[string(name)]
			}
			cols = append(cols, c)
			qty[c] = q
		}
		sort.Ints(cols) // ascending columns within a pod
		for _, c := range cols {
			pk.sCol[w], pk.sQty[w] = int32(c), qty[c]
			w++
		}
	}
	pk.sPtr[P] = int32(w)
	pk.outCount, pk.outMask = u8s(&pk.bOutCount, 16*N*G), u16s(&pk.bOutMask, N*G)
	return nil
}

// records is the node side in the compact form of nosplan_snapshot_compact (include/nosplan.h): one record per node, one DMA,
// expanded on the GPU -- about half the bytes of the SoA arrays.  Written here from the arrays pack() has just filled, for
// brevity; a packer that targets only this form writes the records directly while it walks the nodes.
func (pk *packer) records(b *pinned) ([]byte, int) {
	N, G, K, nOther := len(pk.names), pk.G, len(pk.profiles), len(pk.others)
	R := 3 + K + nOther
	rb := int(C.nosplan_compact_record_bytes(C.int32_t(G), C.int32_t(nOther)))
	rec := b.bytes(N * rb)
	for i := range rec {
		rec[i] = 0
	}
	for i := 0; i < N; i++ {
		r := rec[i*rb : (i+1)*rb]
		head := unsafe.Slice((*int32)(unsafe.Pointer(&r[0])), 4)
		head[0], head[1], head[2] = pk.gpuCount[i], pk.allowed[i], pk.nPods[i]
		base := unsafe.Slice((*int64)(unsafe.Pointer(&r[16])), 6)
		copy(base[0:3], pk.alloc[i*R:i*R+3])
		copy(base[3:6], pk.req[i*R:i*R+3])
		sl := unsafe.Slice((*int32)(unsafe.Pointer(&r[64])), 16)
		for k := 0; k < K; k++ {
			sl[k], sl[8+k] = int32(pk.alloc[i*R+3+k]), int32(pk.req[i*R+3+k])
		}
		if nOther > 0 {
			ot := unsafe.Slice((*int64)(unsafe.Pointer(&r[128])), 2*nOther)
			copy(ot[0:nOther], pk.alloc[i*R+3+K:(i+1)*R])
			copy(ot[nOther:2*nOther], pk.req[i*R+3+K:(i+1)*R])
		}
		for j := 0; j < G; j++ {
			x, g := i*G+j, r[128+16*nOther+24*j:]
			copy(g[0:8], pk.used[16*x:16*x+8])
			copy(g[8:16], pk.free[16*x:16*x+8])
			g[16], g[17], g[18] = uint8(pk.uMask[x]), uint8(pk.fMask[x]), uint8(pk.gpuModel[x])
			g[20], g[21] = uint8(pk.gpuMem[x]), uint8(pk.gpuMem[x]>>8)
		}
	}
	return rec, rb
}

// cudaPlanner implements core.Planner on the engine.
type cudaPlanner struct {
	h         *C.nosplan_t
	kind      C.int32_t // NOSPLAN_KIND_MIG | NOSPLAN_KIND_MPS
	framework framework.Framework
	pk        packer
	pre, rec  pinned
}

// NewCudaPlanner is what mig.NewPlanner / mps.NewPlanner return instead of core.NewPlanner(partitionCalculator, sliceCalculator,
// scheduler).  There is no CPU fallback: without a CUDA device construction fails and the controller does not start.
func NewCudaPlanner(kind gpu.PartitioningKind, fw framework.Framework, cudaDevice int) (Planner, error) {
	var h *C.nosplan_t
	if rc := C.nosplan_create(C.int(cudaDevice), &h); rc != 0 {
		return nil, fmt.Errorf("nosplan_create(device %d): %d", cudaDevice, int(rc))
	}
	k := C.int32_t(C.NOSPLAN_KIND_MIG)
	if kind == gpu.PartitioningKindMps {
		k = C.NOSPLAN_KIND_MPS
	}
	return &cudaPlanner{h: h, kind: k, framework: fw}, nil
}

func (p *cudaPlanner) Close() { C.nosplan_destroy(p.h) }

// Plan implements core.Planner.  Error behaviour follows planner.go:100-108: an error is returned only for an engine
// failure (the reconciler requeues); a pod that does not fit is never an error.
func (p *cudaPlanner) Plan(ctx context.Context, snapshot Snapshot, pods []v1.Pod) (PartitioningPlan, error) {
	pk := &p.pk
	if err := pk.pack(p.kind, snapshot, pods); err != nil {
		return PartitioningPlan{}, err
	}
	// PreFilter verdict per pod: a per-pod constant during one Plan() (no Reserve happens inside Plan, so quota Used is fixed);
	// the quota part can also come from nosplan_quota_state_prefilter (capacityscheduling_cgo.go)
	pre := u8s(&p.pre, len(pods))
	for i := range pods {
		pre[i] = 0
		if _, st := p.framework.RunPreFilterPlugins(ctx, framework.NewCycleState(), &pods[i]); st.IsSuccess() {
			pre[i] = 1
		}
	}
	snap := C.nosplan_snapshot{
		kind: p.kind, n_profiles: C.int32_t(len(pk.profiles)), profile_mem_gb: p32(pk.profMem), profile_gi: p32(pk.profGI),
		n_models: C.int32_t(len(pk.models)), model_geom_offset: p32(pk.modelOff), geom_counts: pu8(pk.geom),
		n_nodes: C.int32_t(len(pk.names)), gpus_per_node: C.int32_t(pk.G), node_gpu_count: p32(pk.gpuCount),
		gpu_model: p32(pk.gpuModel), gpu_index: p32(pk.gpuIndex), gpu_memory_gb: p32(pk.gpuMem),
		gpu_used: pu8(pk.used), gpu_free: pu8(pk.free), gpu_used_mask: pu16(pk.uMask), gpu_free_mask: pu16(pk.fMask),
		n_other: C.int32_t(len(pk.others)), node_alloc: p64(pk.alloc), node_requested: p64(pk.req),
		node_allowed_pods: p32(pk.allowed), node_num_pods: p32(pk.nPods),
	}
	cp := C.nosplan_pods{
		n_pods: C.int32_t(len(pods)), priority: p32(pk.prio), tracker_key: p32(pk.key), base: p64(pk.base),
		scalar_ptr: p32(pk.sPtr), scalar_col: p32(pk.sCol), scalar_qty: p64(pk.sQty),
	}
	out := C.nosplan_partitioning{gpu_count: pu8(pk.outCount), gpu_mask: pu16(pk.outMask)}
	var st C.nosplan_stats
	if len(pk.profiles) <= 8 && len(pk.models) <= 256 {
		// the compact form: same plan, about half the bytes over PCIe
		rec, rb := pk.records(&p.rec)
		cs := C.nosplan_snapshot_compact{
			kind: p.kind, n_profiles: C.int32_t(len(pk.profiles)), profile_mem_gb: p32(pk.profMem), profile_gi: p32(pk.profGI),
			n_models: C.int32_t(len(pk.models)), model_geom_offset: p32(pk.modelOff), geom_counts: pu8(pk.geom),
			n_nodes: C.int32_t(len(pk.names)), gpus_per_node: C.int32_t(pk.G), n_other: C.int32_t(len(pk.others)),
			record_bytes: C.int32_t(rb), records: pu8(rec),
		}
		if rc := C.nosplan_plan_compact(p.h, &cs, &cp, pu8(pre), C.NOSPLAN_FILTER_RESOURCES_FIT, &out, &st); rc != 0 {
			return PartitioningPlan{}, fmt.Errorf("nosplan_plan_compact: %s", C.GoString(C.nosplan_last_error(p.h)))
		}
	} else if rc := C.nosplan_plan(p.h, &snap, &cp, pu8(pre), C.NOSPLAN_FILTER_RESOURCES_FIT, &out, &st); rc != 0 {
		return PartitioningPlan{}, fmt.Errorf("nosplan_plan: %s", C.GoString(C.nosplan_last_error(p.h)))
	}
	// state.PartitioningState (internal/partitioning/state/partitioning.go:24-40): zero-count keys are kept (mask bit set)
	desired := make(state.PartitioningState, len(pk.names))
	for i, name := range pk.names {
		np := state.NodePartitioning{GPUs: make([]state.GPUPartitioning, 0, pk.gpuCount[i])}
		for j := 0; j < int(pk.gpuCount[i]); j++ {
			x := i*pk.G + j
			res := make(map[v1.ResourceName]int)
			for k, prof := range pk.profiles {
				if pk.outMask[x]>>uint(k)&1 == 1 {
					res[pk.resourceName(prof)] = int(pk.outCount[16*x+k])
				}
			}
			np.GPUs = append(np.GPUs, state.GPUPartitioning{GPUIndex: int(pk.gpuIndex[x]), Resources: res})
		}
		desired[name] = np
	}
	return NewPartitioningPlan(desired), nil
}

// ChangedNodes is the per-node form of the comparison Actuator.Apply makes as a whole (core/actuator.go:39-66,
// state/partitioning.go:33-56): the names of the nodes whose plan differs from their current partitioning, in name order,
// computed on the rows the last Plan() left on the GPU.
func (p *cudaPlanner) ChangedNodes() ([]string, error) {
	idx := make([]int32, len(p.pk.names))
	var n C.int32_t
	if rc := C.nosplan_changed_nodes(p.h, p32(idx), C.int32_t(len(idx)), &n); rc != 0 {
		return nil, fmt.Errorf("nosplan_changed_nodes: %s", C.GoString(C.nosplan_last_error(p.h)))
	}
	out := make([]string, 0, int(n))
	for _, i := range idx[:int(n)] {
		out = append(out, p.pk.names[i])
	}
	return out, nil
}
