// nos/host.hpp -- C++ host side above the C-ABI of libnosplan.so.
//
// The reference (WindowsXp-Beta/nos v0.1.2) is Go and this image has no Go toolchain, so the types a
// maintainer's cgo shim would keep on the Go side are mirrored here in C++: same names, same argument
// meaning, same error behaviour (a Go `error` return is a thrown nos::Error).  Nothing in this layer plans:
// it parses v1.Node labels/annotations, interns names, packs the SoA of include/nosplan.h, calls
// nosplan_plan on the GPU and rebuilds state.PartitioningState.  Paths are relative to the reference root.
#pragma once
#include <cstdint>
#include <functional>
#include <map>
#include <memory>
#include <optional>
#include <stdexcept>
#include <string>
#include <vector>

#include "nosplan.h"

namespace nos {

struct Error : std::runtime_error { using std::runtime_error::runtime_error; };

// v1.ResourceList with quantities already reduced the way the path reads them: cpu = MilliValue(), everything
// else Value() (pkg/resource/resource.go:37-47 FromListToFramework).
using ResourceList = std::map<std::string, int64_t>;

namespace v1 {
struct Container { ResourceList requests; };
struct Pod {                                      // k8s.io/api/core/v1.Pod, the fields Plan() reads
    std::string ns, name;
    std::optional<int32_t> priority;              // *Spec.Priority; corev1.PodPriority -> 0 when unset
    std::vector<Container> containers, initContainers;
};
struct Node {                                     // v1.Node
    std::string name;
    std::map<std::string, std::string> labels, annotations;
    ResourceList allocatable;                     // Status.Allocatable
};
}  // namespace v1

namespace framework {
struct NodeInfo {                                 // k8s framework.NodeInfo, the fields Plan() reads
    std::optional<v1::Node> node;                 // Node() == nil when unset
    ResourceList requested;                       // Requested
    int32_t numPods = 0;                          // len(Pods)
};
}  // namespace framework

namespace resource {
// resource.ComputePodRequest (pkg/resource/resource.go:127-146): sum of containers, max with every init container;
// the pod overhead is computed and discarded by the reference, so it is not an input here.
ResourceList ComputePodRequest(const v1::Pod& pod);
}  // namespace resource

namespace gpu {
enum class Status { Free, Used, Unknown };
struct StatusAnnotation { std::string ProfileName; int Index; Status status; int Quantity; };
struct SpecAnnotation { std::string ProfileName; int Index; int Quantity; };
// pkg/gpu/annotation.go:29-55 / :57-85; nullopt where the reference returns an error
std::optional<SpecAnnotation> ParseSpecAnnotation(const std::string& key, const std::string& value);
std::optional<StatusAnnotation> ParseStatusAnnotation(const std::string& key, const std::string& value);
// pkg/gpu/annotation.go:87-104 (spec parse is tried first)
void ParseNodeAnnotations(const v1::Node& node, std::vector<StatusAnnotation>& status, std::vector<SpecAnnotation>& spec);
std::string GetModel(const v1::Node& node);       // pkg/gpu/util.go:30-40
int GetCount(const v1::Node& node);               // :43-56
int GetMemoryGB(const v1::Node& node);            // :59-73, ceil(MB / 1000)
using Geometry = std::map<std::string, int>;      // profile name -> quantity
}  // namespace gpu

// One partitionable GPU: mig.GPU (pkg/gpu/mig/gpu.go:27-33) / slicing.GPU (pkg/gpu/slicing/gpu.go:30-38)
struct GPU {
    std::string model;
    int index = 0;
    int memoryGB = 0;                             // slicing only
    std::map<std::string, int> used, free;        // profile -> devices; zero-valued keys are kept
    int GetIndex() const { return index; }
    const std::string& GetModel() const { return model; }
    gpu::Geometry GetGeometry() const;            // used + free per profile (mig/gpu.go:82-95)
};

// mig.Node (pkg/gpu/mig/node.go:30-38) / slicing.Node (pkg/gpu/slicing/node.go:26-30)
struct Node {
    std::string Name;
    std::vector<GPU> GPUs;
    framework::NodeInfo nodeInfo;
    const std::string& GetName() const { return Name; }
    const framework::NodeInfo& NodeInfo() const { return nodeInfo; }
};

namespace mig {
using KnownGeometries = std::map<std::string, std::vector<gpu::Geometry>>;   // gpu.Model -> allowed geometries
// mig.SetKnownGeometries / GetKnownGeometries (pkg/gpu/mig/known_configs.go:143-157): process-wide runtime data.
// ValidateConfigs (:159-181) rejects empty lists and quantities < 1.
void SetKnownGeometries(const KnownGeometries& g);
const KnownGeometries& GetKnownGeometries();
const KnownGeometries& DefaultKnownGeometries();  // the built-in table (:24-142)
Node NewNode(const framework::NodeInfo& n);       // pkg/gpu/mig/node.go:40-100
}  // namespace mig

namespace slicing {
Node NewNode(const framework::NodeInfo& n);       // pkg/gpu/slicing/node.go:32-105
}  // namespace slicing

namespace state {
struct GPUPartitioning {                          // internal/partitioning/state/partitioning.go:24-27
    int GPUIndex = 0;
    std::map<std::string, int> Resources;
    bool operator==(const GPUPartitioning& o) const { return GPUIndex == o.GPUIndex && Resources == o.Resources; }
};
struct NodePartitioning {                         // :29-38, Equal ignores the order of the GPUs (util.UnorderedEqual)
    std::vector<GPUPartitioning> GPUs;
    bool Equal(const NodePartitioning& other) const;
};
struct PartitioningState : std::map<std::string, NodePartitioning> {   // :40-56
    bool IsEmpty() const { return empty(); }
    bool Equal(const PartitioningState& other) const;
};
using ClusterState = std::map<std::string, framework::NodeInfo>;       // state.ClusterState.GetNodes()
}  // namespace state

namespace mig {
// getGPUSpecAnnotationList (internal/partitioning/mig/partitioner.go:77-94); GPUs in the order given, resources of a GPU in
// name order (the reference iterates a Go map; the annotations end up in a map, so the order is not observable).
std::vector<gpu::SpecAnnotation> GetGPUSpecAnnotationList(const state::NodePartitioning& p);
std::string SpecAnnotationKey(const gpu::SpecAnnotation& a);   // SpecAnnotation.String (pkg/gpu/annotation.go:152-154)
}  // namespace mig

namespace mps {
struct MPSResource {                                           // nvidiav1.MPSResource as ToPluginConfig fills it
    std::string Name, Rename;
    int MemoryGB = 0;
    std::vector<std::string> Devices;
    int Replicas = 0;
};
struct PluginConfig {                                          // nvidiav1.Config, the fields ToPluginConfig sets
    std::string MigStrategy = "none";
    bool FailRequestsGreaterThanOne = true;
    std::vector<MPSResource> Resources;
};
PluginConfig ToPluginConfig(const state::NodePartitioning& p); // internal/partitioning/mps/partitioner.go:123-157
}  // namespace mps

// The quota side of the C-ABI: what a cgo shim inside pkg/scheduler/plugins/capacityscheduling would build from its
// ElasticQuotaInfos and pod lists before nosplan_quota_recompute / nosplan_quota_label_changes / nosplan_preempt_dryrun.
namespace capacityscheduling {
struct ElasticQuotaInfo {                         // elasticquotainfo.go:177-192, the fields the accounting reads
    std::string ResourceName;
    std::vector<std::string> Namespaces;          // one entry: ElasticQuota; several: CompositeElasticQuota
    std::optional<ResourceList> Min, Max, Used;   // nil *framework.Resource == nullopt
    bool MaxEnforced = false;
};
struct QuotaPod {                                 // a pod as the accounting sees it
    std::string ns, name;
    int32_t priority = 0;                         // *Spec.Priority, 0 when unset
    int64_t creationTimestamp = 0;                // unix seconds
    ResourceList request;                         // resource.ComputePodRequest, without nos.nebuly.com/gpu-memory
    bool overQuota = false;                       // label nos.nebuly.com/capacity == over-quota (pkg/util/pod/pod.go:31-36)
};
struct PackedQuotaPods {
    std::vector<std::string> keys;                // namespace/name
    std::vector<int32_t> ns, priority, nameRank;
    std::vector<int64_t> request, creationTs;     // request: [P][Rq]
    std::vector<uint32_t> requestMask;
    nosplan_quota_pods c() const;
};
struct PackedQuotas {
    std::vector<std::string> names, namespaces, columns;
    int32_t gpuMemoryCol = -1;
    std::vector<int32_t> nsQuota;
    std::vector<int64_t> min, max, usedInit, gpuMemoryCoeff;
    std::vector<uint32_t> minMask, maxMask, usedMask;
    std::vector<uint8_t> maxEnforced;
    std::vector<PackedQuotaPods> podSets;         // one per pod list given to PackQuotas, same order
    int R() const { return (int)columns.size(); }
    nosplan_quotas c() const;
};
// Interns namespaces (quotas first, in the order given, then the pods'), resource names (cpu, memory, ephemeral-storage, pods,
// then every scalar seen, sorted, nos.nebuly.com/gpu-memory always among them) and quota rows; a namespace listed by two
// quotas points at the later one, as ElasticQuotaInfos.Add overwrites (elasticquotainfo.go:68-72).  Throws nos::Error beyond
// NOSPLAN_QUOTA_MAX_RES columns.
PackedQuotas PackQuotas(const std::vector<ElasticQuotaInfo>& quotas, const std::vector<std::vector<QuotaPod>>& podSets,
                        int nvidiaGpuMemoryGB);
// nosplan_preempt_query for one preemptor over the candidate nodes (preemptor.SelectVictimsOnNode, capacity_scheduling.go:468-675).
struct PreemptNode {
    ResourceList allocatable;                     // NodeInfo.Allocatable, "pods" = allowed pod number
    std::vector<QuotaPod> pods;                   // NodeInfo.Pods, least important first (the order of :516)
    std::vector<std::vector<int32_t>> pdbsOfPod;  // per pod: budgets that match it and do not list it in DisruptedPods (:856-877)
};
struct PackedPreemptQuery {
    int32_t preemptorNs = -1, preemptorPriority = 0;
    uint32_t fitMask = 0;
    std::vector<int64_t> nominatedInEQ, podReq, nominated, fitRequest, request, nodeAlloc, nodeRequested;
    std::vector<int32_t> nodePtr, ns, priority, pdbAllowed, podPdbPtr, podPdb;
    std::vector<uint8_t> overQuota;
    std::vector<uint32_t> requestMask;
    nosplan_preempt_query c() const;
};
PackedPreemptQuery PackPreemptQuery(const PackedQuotas& pq, const QuotaPod& preemptor, const ResourceList& nominatedInEQ,
                                    const ResourceList& nominated, const std::vector<PreemptNode>& nodes,
                                    const std::vector<int32_t>& pdbAllowed);
}  // namespace capacityscheduling

namespace core {

enum class Kind { MIG = NOSPLAN_KIND_MIG, MPS = NOSPLAN_KIND_MPS };

// core.Snapshot (internal/partitioning/core/interface.go:53-65), the read side Plan() needs.
class Snapshot {
public:
    Snapshot(Kind kind, std::map<std::string, Node> nodes) : kind_(kind), nodes_(std::move(nodes)) {}
    Kind kind() const { return kind_; }
    const std::map<std::string, Node>& GetNodes() const { return nodes_; }
    const Node* GetNode(const std::string& name) const;
    state::PartitioningState GetPartitioningState() const;            // core/snapshot.go:66-72 + PartitionCalculator
private:
    Kind kind_;
    std::map<std::string, Node> nodes_;
};

// SnapshotTaker.TakeSnapshot (internal/partitioning/mig/snapshot_taker.go:31-53, mps twin): keeps the nodes labelled
// nos.nebuly.com/gpu-partitioning=<kind>, fails where NewNode fails.
Snapshot TakeSnapshot(Kind kind, const state::ClusterState& cluster);

struct PartitioningPlan {                         // core/planner.go:30-49
    state::PartitioningState DesiredState;
    std::string id;                               // unix seconds; not part of parity
    const std::string& GetId() const { return id; }
};
PartitioningPlan NewPartitioningPlan(state::PartitioningState s);

// Stand-in for the framework.Framework the planner simulates scheduling with (core/planner.go:178-207).
struct SchedulerFramework {
    int filterMode = NOSPLAN_FILTER_RESOURCES_FIT;             // NOSPLAN_FILTER_*
    std::function<bool(const v1::Pod&)> preFilter;             // RunPreFilterPlugins verdict; empty = Success
};

// The flat buffers handed to nosplan_plan; exposed so tests can compare the packing with other packers.
struct Packed {
    std::vector<std::string> profiles, models, others, nodeNames, podKeys;
    std::vector<int32_t> profileMemGB, profileGi, modelGeomOffset;
    std::vector<uint8_t> geomCounts;
    int32_t G = 1;
    std::vector<int32_t> nodeGpuCount, gpuModel, gpuIndex, gpuMemoryGB;
    std::vector<uint8_t> gpuUsed, gpuFree;
    std::vector<uint16_t> gpuUsedMask, gpuFreeMask;
    std::vector<int64_t> nodeAlloc, nodeRequested;
    std::vector<int32_t> nodeAllowedPods, nodeNumPods;
    std::vector<int32_t> podPriority, podTrackerKey, scalarPtr, scalarCol;
    std::vector<int64_t> podBase, scalarQty;
    int R() const { return 3 + (int)profiles.size() + (int)others.size(); }
    nosplan_snapshot snapshot(Kind kind) const;
    nosplan_pods pods() const;
};
Packed Pack(const Snapshot& snapshot, const std::vector<v1::Pod>& pods);

// core.Planner (interface.go:27-29).  mig.NewPlanner / mps.NewPlanner (internal/partitioning/{mig,mps}/factory.go)
// become NewPlanner(kind, scheduler).  Owns one nosplan engine bound to one CUDA device; construction fails
// without one -- there is no CPU planning path.
class Planner {
public:
    Planner(Kind kind, SchedulerFramework scheduler, int cudaDevice = 0);
    ~Planner();
    Planner(const Planner&) = delete;
    Planner& operator=(const Planner&) = delete;
    // Plan(ctx, snapshot, pendingPods) (core/planner.go:67-153).  The snapshot is not modified (the reference
    // plans on snapshot.Clone(), partitioner_controller.go:185).
    PartitioningPlan Plan(const Snapshot& snapshot, const std::vector<v1::Pod>& pendingPods);
    const nosplan_stats& LastStats() const { return stats_; }
private:
    Kind kind_;
    SchedulerFramework scheduler_;
    nosplan_t* h_ = nullptr;
    nosplan_stats stats_{};
};

// What Actuator.Apply (core/actuator.go:39-66) writes, as data: per node of the desired state, in map order, the node
// annotations after mig Partitioner.ApplyPartitioning (internal/partitioning/mig/partitioner.go:43-74: old
// nos.nebuly.com/spec-gpu* keys dropped, one spec annotation per (GPU, profile), the plan id) or, for MPS, the device-plugin
// ConfigMap entry and node label of mps Partitioner.ApplyPartitioning (internal/partitioning/mps/partitioner.go:60-112).
// Throws Error where the reference returns one: a node of the plan that is not in the cluster ("failed to get node"), a
// resource that is not a slice of the kind.
struct NodePatch {
    std::string node;
    std::map<std::string, std::string> annotations;            // MIG: the node's full annotation map after the patch
    std::string configKey;                                     // MPS: "<node>-<plan id>" (DevicePluginConfigKeyFormat, :38)
    mps::PluginConfig config;                                  // MPS: ToPluginConfig(partitioning)
};
struct ApplyResult { bool applied = false; std::vector<NodePatch> patches; };
ApplyResult Apply(Kind kind, const state::ClusterState& cluster, const Snapshot& snapshot, const PartitioningPlan& plan);

// Actuator.Apply's decision (core/actuator.go:39-66): false when the desired state equals the current one or is empty;
// otherwise the node names that Apply would hand to the Partitioner.  changedOnly = skip nodes whose partitioning
// already equals the desired one (the reference patches every node of the plan).
std::vector<std::string> NodesToPartition(const Snapshot& snapshot, const PartitioningPlan& plan, bool changedOnly);

}  // namespace core
}  // namespace nos
