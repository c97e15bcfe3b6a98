/*
 * nosplan.h -- C-ABI of the B200-native partition-planning engine (libnosplan.so).
 *
 * This is the boundary a cgo shim binds.  The reference (WindowsXp-Beta/nos,
 * v0.1.2) is 100 % Go and has no FFI of its own, so every entry point below names the Go interface
 * it stands behind (paths relative to the reference root).
 *
 * Conventions
 *   - plain pointers and sizes only; every buffer is caller-allocated HOST memory unless a function
 *     says otherwise; nothing is retained after a call returns (cgo pointer rules);
 *   - every function returns 0 on success or a negative NOSPLAN_E* code; nosplan_last_error(h)
 *     returns a message for the last failure on that handle;
 *   - a handle owns its CUDA stream and a device arena reused across calls; one call at a time per
 *     handle, distinct handles may be used from different threads (the MIG and MPS controllers of
 *     the reference run up to two Plan() calls at once, cmd/gpupartitioner/gpupartitioner.go:215-250);
 *   - there is no CPU fallback: without a CUDA device nosplan_create fails with NOSPLAN_E_CUDA.
 *
 * Packed snapshot ("SoA") layout
 *   Profiles are interned by the caller into ids 0..K-1, K <= NOSPLAN_MAX_PROFILES, in canonical
 *   order: MIG ascending profile name, MPS ascending (size GB, name).  Wherever the reference
 *   iterates a Go map of profiles and the order is observable, the engine visits ids ascending.
 *   Nodes are given in byte-wise ascending name order (core/snapshot.go:126-128); GPUs of a node in
 *   the order of the reference's Node.GPUs slice.  A Go map is a (count, key-present) pair here:
 *   counts in a [16] uint8 row, presence in a 16-bit mask (zero-valued keys matter:
 *   pkg/gpu/mig/gpu.go:141,243,249-251).
 *   Node resources are int64 rows of R = 3 + K + n_other columns:
 *     0 cpu (milli)  1 memory  2 ephemeral-storage  3..3+K-1 slice resource of profile id  3+K.. other scalars
 */
#ifndef NOSPLAN_H
#define NOSPLAN_H

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define NOSPLAN_MAX_PROFILES 16
#define NOSPLAN_MAX_GEOMS    32   /* allowed geometries per GPU model */
#define NOSPLAN_MAX_GPUS     16   /* GPUs per node */
#define NOSPLAN_MAX_OTHER    12   /* non-slice scalar resource columns */

#define NOSPLAN_KIND_MIG 0
#define NOSPLAN_KIND_MPS 1

/* Filter simulation (core/planner.go:178-207 canSchedulePod).  NONE / REJECT_ALL reproduce the
 * constant statuses of the reference's mocked framework (core/planner_test.go:467-480);
 * RESOURCES_FIT is NodeResourcesFit of k8s.io/kubernetes v1.25.4 restated. */
#define NOSPLAN_FILTER_NONE          0
#define NOSPLAN_FILTER_REJECT_ALL    1
#define NOSPLAN_FILTER_RESOURCES_FIT 2

#define NOSPLAN_OK          0
#define NOSPLAN_E_INVALID  -1   /* malformed input (sizes, ids out of range) */
#define NOSPLAN_E_CUDA     -2   /* CUDA runtime failure, message in nosplan_last_error */
#define NOSPLAN_E_LIMIT    -3   /* input exceeds a compiled-in limit */
#define NOSPLAN_E_NOMEM    -4

typedef struct nosplan nosplan_t;

/* Replaces: the inputs core.Planner.Plan reads through core.Snapshot (internal/partitioning/core/
 * interface.go:53-65), i.e. map[string]*mig.Node / *slicing.Node with their framework.NodeInfo, and
 * mig.GetKnownGeometries() (pkg/gpu/mig/known_configs.go:152-157). */
typedef struct nosplan_snapshot {
    int32_t kind;                      /* NOSPLAN_KIND_* */
    int32_t n_profiles;                /* K */
    const int32_t* profile_mem_gb;     /* [K] MIG: digits of `\d+gb` (mig/profile.go:70-75); MPS: slice size GB */
    const int32_t* profile_gi;         /* [K] MIG: digits of `\d+g`  (mig/profile.go:77-82); MPS: 0 */
    /* MIG allowed geometries per GPU model (runtime data, never compiled in); unused for MPS */
    int32_t n_models;
    const int32_t* model_geom_offset;  /* [n_models+1] row range of each model in geom_counts */
    const uint8_t* geom_counts;        /* [rows][16] count per profile id, 0 = profile not in the geometry */
    /* nodes */
    int32_t n_nodes;                   /* N */
    int32_t gpus_per_node;             /* G: row stride of the per-GPU arrays (max GPUs of any node) */
    const int32_t* node_gpu_count;     /* [N] <= G */
    const int32_t* gpu_model;          /* [N*G] model id (MIG) */
    const int32_t* gpu_index;          /* [N*G] GPUIndex reported back in the plan */
    const int32_t* gpu_memory_gb;      /* [N*G] MPS: GPU memory GB (pkg/gpu/util.go:58-73) */
    const uint8_t* gpu_used;           /* [N*G][16] */
    const uint8_t* gpu_free;           /* [N*G][16] */
    const uint16_t* gpu_used_mask;     /* [N*G] key presence of the used map */
    const uint16_t* gpu_free_mask;     /* [N*G] key presence of the free map */
    int32_t n_other;                   /* non-slice scalar columns */
    const int64_t* node_alloc;         /* [N][R] NodeInfo.Allocatable */
    const int64_t* node_requested;     /* [N][R] NodeInfo.Requested */
    const int32_t* node_allowed_pods;  /* [N] Allocatable.AllowedPodNumber */
    const int32_t* node_num_pods;      /* [N] len(NodeInfo.Pods) */
} nosplan_snapshot;

/* Replaces: candidatePods []v1.Pod of Plan (core/planner.go:67) after resource.ComputePodRequest
 * (pkg/resource/resource.go:127-146).  Pods are given in the caller's ORIGINAL order; the engine
 * applies NewPodSorter (core/util.go:34-73) itself. */
typedef struct nosplan_pods {
    int32_t n_pods;                    /* P */
    const int32_t* priority;           /* [P] corev1.PodPriority */
    const int32_t* tracker_key;        /* [P] dense id of namespace/name; duplicates share one SliceTracker entry (core/tracker.go:38-47) */
    const int64_t* base;               /* [P][3] cpu milli, memory, ephemeral */
    const int32_t* scalar_ptr;         /* [P+1] CSR over the scalar entries of each pod */
    const int32_t* scalar_col;         /* [nnz] resource column (>= 3); ascending within a pod */
    const int64_t* scalar_qty;         /* [nnz] quantity (a present key may be 0) */
} nosplan_pods;

/* Replaces: state.PartitioningState inside core.PartitioningPlan (internal/partitioning/state/
 * partitioning.go:24-40): per GPU the resource->count map of used+free devices; zero-count keys kept. */
typedef struct nosplan_partitioning {
    uint8_t* gpu_count;                /* [N*G][16] */
    uint16_t* gpu_mask;                /* [N*G] */
} nosplan_partitioning;

typedef struct nosplan_stats {
    int64_t candidates;                /* (visited candidate node, pod) pairs == logical tryAddPod calls */
    int64_t visited_nodes;
    int64_t geometry_evals;            /* (geometry, required profile) pairs scored, mig/gpu.go:164-165 */
    int64_t placements;                /* successful tryAddPod */
    int64_t candidate_nodes;           /* len(GetCandidateNodes()) */
    int32_t lacking_at_start;          /* number of lacking profiles after NewSliceTracker */
    int32_t lacking_at_end;
    float   ms_h2d, ms_kernels, ms_d2h;/* CUDA-event times of the last call (ms_kernels = whole run()) */
    float   ms_prepass, ms_tracker, ms_parallel, ms_sweep; /* per-kernel CUDA-event times inside ms_kernels */
    int32_t waves;                     /* path 0 only: > 0 = the sweep ran as that many parallel waves (exact lacking vector) */
    int32_t kernel_launches;
    int32_t path;                      /* 1: node-parallel pass (saturated lacking vector), 0: sequential sweep */
    int64_t h2d_bytes, d2h_bytes;      /* bytes copied by upload() / download() */
    int32_t fused;                     /* 1: the plan is the single-launch kernel's (path == 1), 0: the classic kernel sequence's */
    int32_t fused_reject;              /* != 0: the single-launch kernel ran first and rejected its own hypothesis (reason bits) */
    int32_t gate_replayed;             /* some node's accounts disagree with its GPUs: nodes were planned in parallel without the
                                          cluster-wide gate (core/snapshot.go:132-165), the gate was replayed in name order, and this
                                          many nodes (>= 0) had to be planned again with it; -1: no replay took place */
    int32_t reserved0;
} nosplan_stats;

/* Create an engine bound to one CUDA device.  Replaces: mig.NewPlanner / mps.NewPlanner
 * (internal/partitioning/mig/factory.go:31-37, mps/factory.go:44-50). */
int nosplan_create(int cuda_device, nosplan_t** out);
void nosplan_destroy(nosplan_t* h);
const char* nosplan_last_error(const nosplan_t* h);

/* One Plan() call, host buffers in, host buffers out (H2D + kernels + D2H).
 * Replaces: core.Planner.Plan (internal/partitioning/core/interface.go:27-29, planner.go:67-153).
 * prefilter_ok: [P] PreFilter verdict per pod in original order, NULL = all Success
 * (the verdict is a per-pod constant during one Plan(), see nosplan_quota_prefilter). */
int nosplan_plan(nosplan_t* h, const nosplan_snapshot* snap, const nosplan_pods* pods,
                 const uint8_t* prefilter_ok, int32_t filter_mode,
                 nosplan_partitioning* out, nosplan_stats* stats);

/* Plan() against a snapshot that STAYS in HBM between calls -- the device-side twin of state.ClusterState, the cache the
 * reference's informers keep current and every TakeSnapshot reads (internal/partitioning/state/cluster_state.go,
 * internal/partitioning/mig/snapshot_taker.go:31-53).  nosplan_plan_resident plans new pending pods against the node arrays
 * of the last nosplan_plan / nosplan_upload on this handle without copying them again (Plan() never modifies the snapshot
 * it is given: partitioner_controller.go:185 plans on a clone).  nosplan_update_nodes overwrites nodes
 * [first, first + rows->n_nodes) of the resident snapshot from a snapshot struct that holds just those rows (same kind,
 * profiles, gpus_per_node and n_other; its tables are ignored); the snapshot's shape is fixed by the full upload. */
int nosplan_plan_resident(nosplan_t* h, const nosplan_pods* pods, const uint8_t* prefilter_ok, int32_t filter_mode,
                          nosplan_partitioning* out, nosplan_stats* stats);
int nosplan_update_nodes(nosplan_t* h, int32_t first, const nosplan_snapshot* rows);
/* nosplan_upload for the pods alone, against the resident snapshot: what nosplan_plan_resident does before it runs the
 * kernels, for the callers that drive the run themselves (nosplan_run, the sliced and sharded protocols: every rank keeps
 * its slice of the nodes resident). */
int nosplan_upload_pods(nosplan_t* h, const nosplan_pods* pods, const uint8_t* prefilter_ok, int32_t filter_mode);

/* Change detection for the actuator.  Actuator.Apply (core/actuator.go:39-66) compares the
 * desired state with the current one as a whole (PartitioningState.Equal, state/partitioning.go:33-56) and then patches EVERY
 * node of the plan; this call names the nodes whose desired partitioning differs from the snapshot's current one -- compared
 * as NodePartitioning.Equal compares them, zero-count keys included -- so that only those are patched.  After a Plan() on this
 * handle (any of nosplan_plan / plan_resident / run); indices ascending (name order), of the handle's own nodes in slice
 * mode.  More than `capacity` changes: the first `capacity` are returned with NOSPLAN_E_LIMIT, *n_changed holds the number. */
int nosplan_changed_nodes(nosplan_t* h, int32_t* node_idx, int32_t capacity, int32_t* n_changed);

/* Host-only: the pod order the engine derives before uploading.  Replaces: NewPodSorter(...).Sort (core/util.go:34-73):
 * stable by priority descending, then by SmallerThan of the first requested profile.  order_out[k] = index of the k-th pod.
 * Where that comparator is a strict weak order on the pods present, a counting sort is used (every stable sort agrees);
 * otherwise Go 1.18's sort.SliceStable (insertion sort of 20 + symMerge) is executed step for step.  Needs no GPU. */
int nosplan_host_pod_order(const nosplan_snapshot* snap, const nosplan_pods* pods, int32_t* order_out);

/* Host-only: the pod structure the single-launch kernel reads (runs of identical requests in NewPodSorter order, (cpu, memory)
 * pairs in padded positions, minima per 8 / 64 / 512 positions, suffix minima, cumulative sums), staged exactly as nosplan_plan
 * stages it, into the caller's buffer -- so that the staging can be checked and timed without a GPU.  meta[0..10] = byte offsets
 * of the sections (geometry key masks, model ranges, run records, runs per profile, then the five levels, the pods, the end),
 * meta[11..15] = runs, padded positions, cpu shift, memory shift, requested-profile set.  Returns the size in bytes (also when
 * `blob` is NULL or too small: nothing is written then), 0 when these pods take the general path, < 0 on error. */
int64_t nosplan_host_stage_pods(const nosplan_snapshot* snap, const nosplan_pods* pods, const uint8_t* prefilter_ok,
                                uint8_t* blob, int64_t capacity, int32_t* meta);

/* nosplan_plan with the node side in COMPACT form: one record per node, one contiguous array, one DMA -- 128 + 24*G (+ 16 per
 * other scalar resource) bytes per node against ~490 in the SoA form of nosplan_snapshot (G = 8, K = 7), expanded on the GPU
 * into the arrays the kernels read (128 + 24*G rounded up to 16).  Same semantics, errors and outputs as nosplan_plan; the snapshot stays resident for
 * nosplan_plan_resident / nosplan_update_nodes exactly as after nosplan_plan.  Replaces the same reference inputs as
 * nosplan_snapshot (internal/partitioning/core/interface.go:53-65); the cgo packer fills records instead of twelve slices.
 * Limits: K <= 8 profiles, model ids < 256, slice-column values < 2^31.  Record layout (little-endian):
 *     0  int32  gpu_count, allowed_pods, num_pods, 0
 *    16  int64  allocatable cpu (milli), memory, ephemeral-storage; requested cpu, memory, ephemeral-storage
 *    64  int32  allocatable of slice columns 0..7, requested of slice columns 0..7  (columns >= K: 0)
 *   128  int64  allocatable of the n_other scalar columns, then requested of them
 *   128 + 16*n_other, for each GPU slot g < G (24 bytes):
 *        uint8 used[8], uint8 free[8], uint8 used key mask, uint8 free key mask, uint8 model id, uint8 0, uint16 memory GB, uint16 0
 *   padded with zeros to a multiple of 16 bytes.  The per-node checks of nosplan_plan (GPU count, model ids) are made on the
 *   GPU: a malformed record is reported by the same call, after its kernels have run, and nothing of that call is kept. */
typedef struct nosplan_snapshot_compact {
    int32_t kind, n_profiles;
    const int32_t* profile_mem_gb;     /* [K] as in nosplan_snapshot */
    const int32_t* profile_gi;         /* [K] */
    int32_t n_models;
    const int32_t* model_geom_offset;  /* [n_models+1] */
    const uint8_t* geom_counts;        /* [rows][16] */
    int32_t n_nodes, gpus_per_node, n_other;
    int32_t record_bytes;              /* must equal nosplan_compact_record_bytes(gpus_per_node, n_other) */
    const uint8_t* records;            /* [N][record_bytes], nodes in byte-wise name order */
} nosplan_snapshot_compact;
int32_t nosplan_compact_record_bytes(int32_t gpus_per_node, int32_t n_other);
int nosplan_plan_compact(nosplan_t* h, const nosplan_snapshot_compact* snap, const nosplan_pods* pods,
                         const uint8_t* prefilter_ok, int32_t filter_mode,
                         nosplan_partitioning* out, nosplan_stats* stats);

/* The same call split in three so the planning kernels can be timed with the snapshot resident in
 * HBM: upload() packs and copies, run() launches the kernels on the handle's stream (repeatable),
 * download() copies the plan back.  cuda_stream of the handle is exposed for CUDA-event timing. */
int nosplan_upload(nosplan_t* h, const nosplan_snapshot* snap, const nosplan_pods* pods,
                   const uint8_t* prefilter_ok, int32_t filter_mode);
int nosplan_run(nosplan_t* h);
int nosplan_download(nosplan_t* h, nosplan_partitioning* out, nosplan_stats* stats);
void* nosplan_stream(nosplan_t* h);    /* cudaStream_t */
/* CUDA events between the kernels of a run feed the per-kernel times of nosplan_stats (default on); switching them off
 * removes their few microseconds from a Plan(). */
int nosplan_set_kernel_events(nosplan_t* h, int32_t on);

/* Node-sharded Plan() over `world` ranks -- one handle per rank (one process per GPU), the SAME snapshot uploaded on
 * every rank (the reference's Plan() is single-process, core/planner.go:67-153, so there is no
 * counterpart call).  Each rank runs the node-parallel pass on its slice [lo, hi), chunk = ceil(N/world), lo = min(N, rank*chunk),
 * hi = min(N, lo+chunk) (equal-sized chunks, the last possibly short or empty, so that the rows all-gather), of the
 * name-ordered node list.  The only data-path exchange is one all-gather of nosplan_shard_exchange_bytes() per rank:
 * 4 counters and the bitmap of tracker keys its nodes placed.
 *   begin(): cluster totals and tracker initialisation (replicated: cheaper than a collective), node-parallel pass on
 *            the slice; the rank's block is written to the DEVICE buffer d_exchange (caller-owned, e.g. a torch tensor);
 *   the caller all-gathers the blocks (NCCL) into d_gathered [world][exchange_bytes] on the handle's stream;
 *   end():   OR / sum of the blocks, verification of the saturation hypothesis; if it fails, or the pass is not
 *            applicable, every rank runs the full sequential sweep (identical result on every rank).
 * After end(), nosplan_download works as usual; rows of nodes outside the rank's slice hold the initial state unless the
 * sequential sweep ran (stats.path == 0), in which case the plan is complete on every rank.  nosplan_device_outputs
 * exposes the device rows so the slices can be all-gathered without a host round trip. */
int64_t nosplan_shard_exchange_bytes(nosplan_t* h);
int nosplan_shard_begin(nosplan_t* h, int32_t rank, int32_t world, void* d_exchange, int64_t exchange_bytes);
int nosplan_shard_end(nosplan_t* h, const void* d_gathered, int32_t world);
int nosplan_device_outputs(nosplan_t* h, void** d_gpu_count /* [N*G][16] u8 */, void** d_gpu_mask /* [N*G] u16 */);

/* The same protocol with the collectives issued by the library itself (NCCL resolved with dlopen at run time): rank 0
 * obtains an id with nosplan_nccl_unique_id and hands its 128 bytes to the other ranks by any means; every rank calls
 * nosplan_shard_init once, then per Plan(): nosplan_upload, nosplan_run_sharded (begin + ONE grouped NCCL launch with the
 * exchange block and, if gather_plan, the plan rows + end), nosplan_download_sharded (complete plan on every rank). */
int nosplan_nccl_unique_id(uint8_t* out128);
int nosplan_shard_init(nosplan_t* h, int32_t rank, int32_t world, const uint8_t* id128);
int nosplan_run_sharded(nosplan_t* h, int32_t gather_plan);
int nosplan_download_sharded(nosplan_t* h, nosplan_partitioning* out, nosplan_stats* stats);

/* Slice mode: every rank uploads ONLY its own contiguous slice of the name-ordered node list (rank order == name order;
 * the same pods, profiles and geometry tables everywhere), so H2D, k_prepass and the plan rows are per-rank work and the
 * job scales weakly.  Two small exchanges per Plan(): the slice's cluster totals (one NOSPLAN_SLICE_TOTALS_BYTES block per
 * rank: sums of allocatable - requested per slice resource, largest free counts, consistency flags) between k_prepass
 * and the tracker, then the exchange block of the replicated protocol above.
 *   slice_begin(): totals of the slice -> d_totals (device, caller-owned);   caller all-gathers into [world][256];
 *   slice_mid():   folds the gathered totals, tracker initialisation, node-parallel pass on the slice -> d_exchange;
 *   caller all-gathers the exchange blocks, then nosplan_shard_end() as above; nosplan_download() returns the slice's rows
 *   and the job-wide counters.
 * The sequential sweep needs the whole snapshot on one GPU: where the node-parallel pass does not apply (or its
 * hypothesis fails a posteriori) nosplan_download returns NOSPLAN_SLICED_NOT_APPLICABLE on EVERY rank (same data, same
 * verdict) and the caller plans that problem unsharded (nosplan_plan on the whole snapshot, or the replicated protocol).
 * nosplan_run_sliced does the three steps with the library's own NCCL communicator (nosplan_shard_init). */
#define NOSPLAN_SLICE_TOTALS_BYTES 256
#define NOSPLAN_SLICED_NOT_APPLICABLE 1
int nosplan_slice_begin(nosplan_t* h, void* d_totals);
int nosplan_slice_mid(nosplan_t* h, int32_t rank, int32_t world, const void* d_totals_gathered, void* d_exchange, int64_t exchange_bytes);

/* Slice mode, problems the node-parallel pass does not cover (nosplan_download reported NOSPLAN_SLICED_NOT_APPLICABLE, the
 * same verdict on every rank): the reference's sequential sweep (core/planner.go:86-141) run rank after rank in name order.
 * Rank 0 calls with d_state_in = NULL; every other rank calls once its predecessor's d_state_out block has reached it (any
 * transport; nosplan_slice_chain_bytes bytes of device memory: lacking vector, running totals, counters, removed tracker
 * keys).  Every rank must take its turn, also after the lacking vector has emptied (its turn then costs one launch).
 * Afterwards nosplan_download gives this rank's rows; its stats are cumulative up to this rank, i.e. final on the last one. */
int64_t nosplan_slice_chain_bytes(nosplan_t* h);
int nosplan_slice_chain(nosplan_t* h, const void* d_state_in, void* d_state_out);
int nosplan_run_sliced(nosplan_t* h);
/* 1: nosplan_run_sliced moves its two blocks over peer memory -- every rank stores its block straight into a mailbox in each
 * peer's HBM (CUDA IPC mapping, NVLink) and raises a flag there; the consuming kernel waits on the flags of its own mailbox,
 * so no collective is launched on the path.  0: NCCL all-gathers (no peer access between the GPUs, more than 8 ranks,
 * NOSPLAN_NO_PEER set, or an exchange block above 64 KiB).  Decided once, unanimously, in nosplan_shard_init. */
int nosplan_shard_transport(nosplan_t* h);

/* No counterpart in the reference.  The tracker initialisation of Plan() -- lacking[p] summed over the pods, and the
 * number of distinct pod keys that lack any slice -- for callers that may not allocate device memory: processes
 * sharing the GPU with latency-sensitive streams (cudaMalloc/cudaFree synchronise the whole device) or running inside a
 * fixed device-memory reservation.  All pointers are DEVICE pointers owned by the caller; pods in the nosplan_pods
 * layout (tracker_key, scalar_ptr/col/qty), d_totals = [16] int64 sum over nodes of (allocatable - requested) per slice
 * column, outputs d_lack = [16] int32 and d_lacking_entries = [1] int32.  The per-key working table lives in a static
 * device array of NOSPLAN_STATIC_MAX_KEYS rows (1 MiB per GPU, kept until the process exits): n_keys above that is
 * refused with NOSPLAN_E_LIMIT.  The table is shared by every call on that GPU -- overlapping calls race -- and is
 * cleared per call.  Work is enqueued on cuda_stream, no synchronisation, no allocation; the first call on a GPU
 * resolves the table's device address and should be made outside stream capture.  errbuf (optional) gets a message.
 * The pods are device data the host cannot check: a pod whose tracker key lies outside [0, n_keys) or whose scalar
 * column is below 3 is skipped on the GPU and *d_lacking_entries comes back negative (sign bit set; d_lack incomplete). */
#define NOSPLAN_STATIC_MAX_KEYS 16384
int nosplan_lacking_static(int cuda_device, void* cuda_stream, int32_t n_pods, int32_t n_keys, int32_t n_profiles,
                           const int32_t* d_tracker_key, const int32_t* d_scalar_ptr, const int32_t* d_scalar_col,
                           const int64_t* d_scalar_qty, const int64_t* d_totals, int32_t* d_lack,
                           int32_t* d_lacking_entries, char* errbuf, int32_t errbuf_len);

/* ------------------------------------------------------------------------------------------------------------
 * ElasticQuota accounting (pkg/scheduler/plugins/capacityscheduling/elasticquotainfo.go, capacity_scheduling.go:190-278,
 * pkg/gpu/util/resource.go:60-86, internal/controllers/elasticquota/elasticquota.go:38-118).
 * Quota resource columns: 0 cpu (milli)  1 memory  2 ephemeral-storage  3 pods  4.. scalar resources, Rq <= 16.
 * Key presence of a scalar in a framework.Resource / ResourceList is a bit in a 32-bit mask (bit c = column c;
 * bit 31 = the whole Resource is non-nil). */
#define NOSPLAN_QUOTA_MAX_RES 16
#define NOSPLAN_QUOTA_NONNIL  0x80000000u

#define NOSPLAN_PREFILTER_SUCCESS  0
#define NOSPLAN_PREFILTER_OVER_MAX 1   /* Unschedulable: quota used + request > Max (capacity_scheduling.go:257-266) */
#define NOSPLAN_PREFILTER_OVER_MIN 2   /* Unschedulable: aggregated used + request > aggregated Min (:268-275) */

/* Replaces: ElasticQuotaInfos, the namespace -> *ElasticQuotaInfo map (elasticquotainfo.go:31-72).  A
 * CompositeElasticQuota is one quota row referenced by several namespaces and is counted once per namespace by the
 * aggregates, exactly like the same pointer stored under several map keys (:68-72). */
typedef struct nosplan_quotas {
    int32_t n_quotas;                  /* E */
    int32_t n_res;                     /* Rq */
    int32_t n_namespaces;              /* NS */
    int32_t gpu_memory_col;            /* column of nos.nebuly.com/gpu-memory (>= 4), -1 = not tracked */
    const int32_t* ns_quota;           /* [NS] quota row of the namespace, -1 = not subject to any quota */
    const int64_t* min;                /* [E][Rq] */
    const int64_t* max;                /* [E][Rq] */
    const int64_t* used_init;          /* [E][Rq] Used before the pods below are reserved; NULL = zeros */
    const uint32_t* min_mask;          /* [E] */
    const uint32_t* max_mask;          /* [E] */
    const uint32_t* used_mask;         /* [E] scalar keys present in used_init */
    const uint8_t* max_enforced;       /* [E] ElasticQuotaInfo.MaxEnforced */
    const int64_t* gpu_memory_coeff;   /* [Rq] GB of gpu-memory one unit of column c stands for: NvidiaGPUDeviceMemoryGB for
                                          nvidia.com/gpu, the profile's GB for nvidia.com/mig-*, else 0 (pkg/gpu/util/resource.go:70-86) */
} nosplan_quotas;

/* Replaces: []*v1.Pod after resource.ComputePodRequest; the engine adds the gpu-memory scalar itself (Q2). */
typedef struct nosplan_quota_pods {
    int32_t n_pods;
    const int32_t* ns;                 /* [P] namespace id */
    const int64_t* request;            /* [P][Rq] plain request (column gpu_memory_col is ignored on input) */
    const uint32_t* request_mask;      /* [P] scalar keys present in the request; NULL = the non-zero ones.  reserveResource
                                          makes every requested key present in Used, even with quantity 0 (:197-199) */
    const int32_t* priority;           /* [P]  *Spec.Priority                (operator labelling only, may be NULL otherwise) */
    const int64_t* creation_ts;        /* [P]  CreationTimestamp, unix s     (operator labelling only) */
    const int32_t* name_rank;          /* [P]  rank of the pod name in byte order (operator labelling only) */
} nosplan_quota_pods;

typedef struct nosplan_quota_out {     /* every pointer may be NULL = not wanted */
    int64_t* used;                     /* [E][Rq] Used after reserveResource of every running pod (:194-200, 276-292) */
    uint32_t* used_mask;               /* [E] scalar keys present in that Used */
    int64_t* agg_min;                  /* [Rq] getAggregatedMin (:154-163) */
    int64_t* agg_used;                 /* [Rq] getAggregatedUsed (:165-174) */
    int64_t* agg_overquota;            /* [Rq] getAggregatedOverquotas (:145-152) */
    uint32_t* agg_masks;               /* [3]  scalar key sets of the three aggregates */
    int64_t* guaranteed;               /* [E][Rq] GetGuaranteedOverquotas of every quota (:81-119), float64 like the reference */
    uint8_t* prefilter;                /* [P_pending] NOSPLAN_PREFILTER_* (capacity_scheduling.go:190-278) */
    uint8_t* over_quota;               /* [P_running] operator label: 1 = over-quota (elasticquota.go:38-72) */
    int64_t* operator_used;            /* [E][Rq] status.used as the operator computes it (restricted to the keys of Min) */
    int32_t* ambiguous_order;          /* [1] != 0: two pods of one quota tie on (creationTimestamp, priority) with requests that
                                          the reference's partial-order comparator cannot rank consistently */
} nosplan_quota_out;

/* Recompute the quota state from scratch for E quotas, the running pods (reserved into Used) and the pending pods
 * (PreFilter verdict each).  Host buffers in and out.  Replaces, as one batch: addPod/reserveResource event handling,
 * the aggregate getters, GetGuaranteedOverquotas, the quota part of PreFilter and PatchPodsAndComputeUsedQuota. */
int nosplan_quota_recompute(nosplan_t* h, const nosplan_quotas* quotas, const nosplan_quota_pods* running,
                            const nosplan_quota_pods* pending, nosplan_quota_out* out, nosplan_stats* stats);

/* The same with the pods split between the ranks of a multi-GPU job (after nosplan_shard_init; SURVEY.md section 8e: "quota
 * sums reduced the same way").  Every rank passes the SAME quota tables and ITS OWN share of the running and pending pods;
 * each rank reserves its pods into a Used that starts from zero (used_init counts on rank 0 only), the [E][Rq] int64 rows are
 * summed over the ranks with one NCCL all-reduce and the key masks OR-ed (all-gather + fold) inside the call, and aggregates,
 * guaranteed over-quotas and the PreFilter verdicts of the rank's own pending pods follow from the reduced rows -- identical,
 * bit for bit, to one nosplan_quota_recompute over all pods (int64 sums and ORs are order-independent).  No counterpart in
 * the reference (its plugin is one process); the rows replaced are elasticquotainfo.go:194-208,74-174.  over_quota /
 * operator_used are computed from the rank's own pods only: route every quota's pods to one rank if they are wanted. */
int nosplan_quota_recompute_sharded(nosplan_t* h, const nosplan_quotas* quotas, const nosplan_quota_pods* running,
                                    const nosplan_quota_pods* pending, nosplan_quota_out* out, nosplan_stats* stats);

/* Event-driven accounting on a quota state that stays in HBM -- the way the plugin itself keeps its ElasticQuotaInfos current:
 * the pod informer handlers and Reserve / Unreserve call AddPodIfNotPresent / DeletePodIfPresent
 * (pkg/scheduler/plugins/capacityscheduling/elasticquotainfo.go:276-310, capacity_scheduling.go:311-321,381-445), which
 * reserve / unreserve the pod's request unless the pod is already / not accounted in its quota (the `pods` set).
 * nosplan_quota_state_begin uploads the quota tables (Used = used_init, no pod accounted); nosplan_quota_state_apply applies
 * one batch of events and returns what `out` asks for of used, used_mask, agg_*, guaranteed (NULL pointers = not wanted).
 * pod_id: dense id the caller assigns to a pod UID, < pod_capacity.  The events of a batch are applied in no particular
 * order, so a pod may occur at most once per batch (checked; two batches for an add and a delete of the same pod). */
#define NOSPLAN_QUOTA_EVENT_ADD 0      /* AddPodIfNotPresent (:276-292) */
#define NOSPLAN_QUOTA_EVENT_DELETE 1   /* DeletePodIfPresent (:294-310) */
typedef struct nosplan_quota_events {
    int32_t n_events;
    const uint8_t* op;                 /* [n] NOSPLAN_QUOTA_EVENT_* */
    const int32_t* pod_id;             /* [n] */
    const int32_t* ns;                 /* [n] namespace id */
    const int64_t* request;            /* [n][Rq] plain request at the time of the event (the engine adds gpu-memory) */
    const uint32_t* request_mask;      /* [n] or NULL, as in nosplan_quota_pods */
} nosplan_quota_events;

int nosplan_quota_state_begin(nosplan_t* h, const nosplan_quotas* quotas, int32_t pod_capacity);
int nosplan_quota_state_apply(nosplan_t* h, const nosplan_quota_events* events, nosplan_quota_out* out,
                              int32_t* n_applied /* [1] out, may be NULL: events that changed the state */, nosplan_stats* stats);

/* The quota part of CapacityScheduling.PreFilter (capacity_scheduling.go:190-278) for a batch of pending pods against the
 * resident state as the events applied so far left it: verdict[i] = NOSPLAN_PREFILTER_*. */
int nosplan_quota_state_prefilter(nosplan_t* h, const nosplan_quota_pods* pending, uint8_t* verdict /* [n_pods] */, nosplan_stats* stats);

/* The operator pass returning only what has to be patched.  The reference walks the pods of a
 * quota in the order of sortPodListForFindingOverQuotaPods and calls patchCapacityInfoIfDifferent on each
 * (internal/controllers/elasticquota/elasticquota.go:44-60, :120-149): a pod is patched iff its label
 * nos.nebuly.com/capacity differs from the desired value.  Here every quota is reconciled in one call and the pods to patch
 * come back as one list: by quota row, and inside a quota in the reference's patch order.  Pods of namespaces without a
 * quota are never reconciled and never listed. */
#define NOSPLAN_LABEL_OTHER 0          /* label absent, or any value but the two below */
#define NOSPLAN_LABEL_IN_QUOTA 1       /* constant.CapacityInfoInQuota   "in-quota"   */
#define NOSPLAN_LABEL_OVER_QUOTA 2     /* constant.CapacityInfoOverQuota "over-quota" */
typedef struct nosplan_label_changes {
    const uint8_t* current;            /* [P_running] in: NOSPLAN_LABEL_* now on the pod */
    int32_t capacity;                  /* in: entries available in pod / desired */
    int32_t* pod;                      /* [capacity] out: index of the running pod to patch */
    uint8_t* desired;                  /* [capacity] out: NOSPLAN_LABEL_IN_QUOTA / NOSPLAN_LABEL_OVER_QUOTA */
    int32_t* n_changes;                /* [1] out: how many pods need a patch; when it exceeds capacity the first `capacity`
                                          are returned and the call reports NOSPLAN_E_LIMIT */
} nosplan_label_changes;

/* `out` may be NULL or carry any of the nosplan_quota_out outputs computed from the running pods on the way
 * (operator_used = the status.used each reconcile would write, over_quota, ambiguous_order, used ...). */
int nosplan_quota_label_changes(nosplan_t* h, const nosplan_quotas* quotas, const nosplan_quota_pods* running,
                                const nosplan_label_changes* changes, nosplan_quota_out* out, nosplan_stats* stats);

/* First pass of preemptor.SelectVictimsOnNode (capacity_scheduling.go:511-596) for ONE preemptor over MANY nodes: which
 * pods of each node become potential victims under the quota predicates.  Upstream fans SelectVictimsOnNode out per
 * node, each on its own clone of the quota snapshot; removing a victim un-reserves its request in that clone
 * (RemovePod extension :311-321), so the pods of one node are evaluated in order and the nodes independently.
 * Pods are grouped by node (node_ptr CSR) and, inside a node, given in the order of the reference's sort (:516, least
 * important first).  quotas->used_init / used_mask hold the CURRENT Used of every quota (e.g. nosplan_quota_out.used).
 * The Filter re-check and the reprieve loop (:598-673) stay on the caller's side. */
typedef struct nosplan_victim_query {
    int32_t preemptor_ns;              /* namespace id of the preemptor pod */
    int32_t preemptor_priority;
    const int64_t* nominated_in_eq;    /* [Rq] PreFilterState.nominatedPodsReqInEQWithPodReq (the pod's own quota request
                                          included, gpu-memory column already filled) */
    int32_t n_nodes;
    const int32_t* node_ptr;           /* [n_nodes+1] */
    int32_t n_pods;
    const int32_t* ns;                 /* [n_pods] */
    const int32_t* priority;           /* [n_pods] */
    const uint8_t* over_quota;         /* [n_pods] pod carries the label nos.nebuly.com/capacity=over-quota (pkg/util/pod/pod.go:31-36) */
    const int64_t* request;            /* [n_pods][Rq] plain request (gpu-memory added by the engine) */
    const uint32_t* request_mask;      /* [n_pods] or NULL */
} nosplan_victim_query;

int nosplan_quota_victims(nosplan_t* h, const nosplan_quotas* quotas, const nosplan_victim_query* query,
                          uint8_t* victim /* [n_pods] out: 1 = potential victim */, nosplan_stats* stats);

/* The WHOLE of preemptor.SelectVictimsOnNode (capacity_scheduling.go:468-675) for ONE preemptor over MANY nodes (SURVEY.md
 * section 8f row 2): the first pass above, then "no victims" (:598-602), the Filter with the potential victims gone
 * (:610-612), max quota / aggregated min quota with the pod's request (:617-624), the PodDisruptionBudget split
 * (filterPodsWithPDBViolation :850-895) and the two reprieve loops (:636-673).  Nodes are independent (one clone of the
 * quota snapshot and one NodeInfo each); the pods of a node are sequential.
 * The Filter is NodeResourcesFit on the node account given here (k8s.io/kubernetes v1.25.4, not part of the reference tree;
 * no nominated pods on the node).  Quota columns double as node resource columns: 0 cpu milli, 1 memory, 2 ephemeral
 * storage, 3 pods (node_alloc = allowed pod number, node_requested = len(Pods)), >= 4 scalars; the gpu-memory column is
 * ignored on the node side.  The PDB match (namespace, selector, DisruptedPods: :856-877) is string work done by the caller,
 * who passes per pod the ids of the budgets that would be decremented. */
#define NOSPLAN_VICTIMS_OK 0           /* framework.Success: victim[] / num_violating are valid for the node */
#define NOSPLAN_VICTIMS_NONE 1         /* UnschedulableAndUnresolvable "No victims found on node" (:598-602) */
#define NOSPLAN_VICTIMS_FILTER 2       /* the Filter still rejects the pod with every potential victim gone (:610-612) */
#define NOSPLAN_VICTIMS_MAX_QUOTA 3    /* Unschedulable "max quota exceeded" (:618-620) */
#define NOSPLAN_VICTIMS_MIN_QUOTA 4    /* Unschedulable "total min quota exceeded" (:621-623) */
#define NOSPLAN_VICTIMS_ERROR 5        /* framework.AsStatus(err): removePod of a pod that already left the node (:653-655) */
#define NOSPLAN_PREEMPT_MAX_PDBS 64    /* PodDisruptionBudgets per query */

typedef struct nosplan_preempt_query {
    nosplan_victim_query v;            /* preemptor, nominated_in_eq and the pods grouped by node, as for nosplan_quota_victims */
    const int64_t* pod_req;            /* [Rq] PreFilterState.podReq: the preemptor's quota request (gpu-memory filled) */
    const int64_t* nominated;          /* [Rq] PreFilterState.nominatedPodsReqWithPodReq */
    const int64_t* fit_request;        /* [Rq] the preemptor's request as NodeResourcesFit computes it (column 3 unused) */
    uint32_t fit_mask;                 /* scalar columns present in fit_request */
    int32_t n_pdbs;                    /* <= NOSPLAN_PREEMPT_MAX_PDBS */
    const int64_t* node_alloc;         /* [n_nodes][Rq] NodeInfo.Allocatable */
    const int64_t* node_requested;     /* [n_nodes][Rq] NodeInfo.Requested */
    const int32_t* pdb_allowed;        /* [n_pdbs] pdb.Status.DisruptionsAllowed */
    const int32_t* pod_pdb_ptr;        /* [n_pods+1] CSR */
    const int32_t* pod_pdb;            /* budgets matching each pod and not listing it in DisruptedPods, ascending id */
} nosplan_preempt_query;

int nosplan_preempt_dryrun(nosplan_t* h, const nosplan_quotas* quotas, const nosplan_preempt_query* query,
                           uint8_t* victim /* [n_pods] out: 1 = victim (meaningful where node_status is OK) */,
                           int32_t* node_status /* [n_nodes] out: NOSPLAN_VICTIMS_* */,
                           int32_t* num_violating /* [n_nodes] out: numViolatingVictim */, nosplan_stats* stats);

/* The next preemptor against the SAME cluster state: quotas, nodes, pods and budgets of the last nosplan_preempt_dryrun on this
 * handle stay in HBM (upstream's DryRunPreemption runs once per unschedulable pod of a scheduling cycle, each time over the
 * same snapshot); only the preemptor's fields of `preemptor` are read -- v.preemptor_ns, v.preemptor_priority,
 * v.nominated_in_eq, pod_req, nominated, fit_request, fit_mask -- and the outputs are sized as in that call. */
int nosplan_preempt_dryrun_again(nosplan_t* h, const nosplan_preempt_query* preemptor, uint8_t* victim, int32_t* node_status,
                                 int32_t* num_violating, nosplan_stats* stats);

/* ------------------------------------------------------------------------------------------------------------
 * Snapshot packer: the status annotations of N nodes -> the per-GPU arrays of nosplan_snapshot.
 * Replaces: gpu.ParseNodeAnnotations (pkg/gpu/annotation.go:29-104), StatusAnnotationList.GroupByGpuIndex (:186-196) and
 * the GPU construction of mig.NewNode / slicing.NewNode (pkg/gpu/mig/node.go:65-100, pkg/gpu/slicing/node.go:48-105).
 * The caller flattens the node.Annotations maps into byte buffers (every annotation of every node, related or not; the
 * engine applies the reference's own acceptance rules: prefix, strings.Split(key, "-") part count, strconv.Atoi of value
 * and GPU index, case-insensitive status word) and supplies the profile dictionary in the id order of the snapshot it is
 * building.  GPUs come out in ascending annotated index (the canonical resolution of the reference's map iteration),
 * followed by the GPUs the nvidia.com/gpu.count label adds (index = position). */
typedef struct nosplan_annotations {
    int32_t n_nodes;                   /* N */
    const int32_t* node_ptr;           /* [N+1] CSR over the annotations of each node */
    const int32_t* key_off;            /* [A+1] byte offsets into key_bytes (A = node_ptr[N]) */
    const char* key_bytes;
    const int32_t* val_off;            /* [A+1] */
    const char* val_bytes;
    const int32_t* gpu_count;          /* [N] gpu.GetCount: label nvidia.com/gpu.count (pkg/gpu/util.go:43-56) */
    int32_t n_profiles;                /* K */
    const int32_t* profile_off;        /* [K+1] */
    const char* profile_bytes;         /* profile names, id order */
    int32_t gpus_per_node;             /* G: row stride of the outputs */
} nosplan_annotations;

typedef struct nosplan_parsed_gpus {   /* every pointer may be NULL = not wanted; layouts as in nosplan_snapshot */
    uint8_t* gpu_used;                 /* [N*G][16] */
    uint8_t* gpu_free;                 /* [N*G][16] */
    uint16_t* gpu_used_mask;           /* [N*G] */
    uint16_t* gpu_free_mask;           /* [N*G] */
    int32_t* gpu_index;                /* [N*G] */
    int32_t* node_gpu_count;           /* [N] */
    int32_t* errors;                   /* [4] annotations naming a profile outside the dictionary; device counts outside [0,255];
                                          nodes with more than G GPUs; GPU indices beyond 32 bits.  Any non-zero entry also
                                          fails the call (NOSPLAN_E_INVALID / NOSPLAN_E_LIMIT). */
} nosplan_parsed_gpus;

int nosplan_parse_annotations(nosplan_t* h, const nosplan_annotations* in, nosplan_parsed_gpus* out, nosplan_stats* stats);

#ifdef __cplusplus
}
#endif
#endif /* NOSPLAN_H */
