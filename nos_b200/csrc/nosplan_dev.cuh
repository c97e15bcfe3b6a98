// Device-side problem description shared by the planning kernels (sm_100a).
// Layout notes are in DESIGN.md section "Data layout in HBM".
#pragma once
#include <cstdint>
#include <cuda_runtime.h>

#include "../../include/nosplan.h"

namespace nsp {

constexpr int KP = NOSPLAN_MAX_PROFILES;                                   // 16 profile slots
constexpr int RMAX = 3 + NOSPLAN_MAX_PROFILES + NOSPLAN_MAX_OTHER;         // resource columns
constexpr int MAX_ROWS = 256;                                              // geometry rows cached in smem (all models)
constexpr int MAX_MODELS = 32;
constexpr int RUN_CACHE = 64;                                              // runs whose first chunk of pods lives in smem

// per-pod flags (sorted order)
constexpr uint32_t PF_PRE_OK = 1u;      // PreFilter verdict (capacity_scheduling.go:190-278 is a per-pod constant)
constexpr uint32_t PF_HAS_OTHER = 2u;   // requests a non-slice scalar resource
constexpr uint32_t PF_ZERO_REQ = 4u;    // all-zero request (NodeResourcesFit short-circuit)

// A run = maximal stretch of consecutive pods (in NewPodSorter order) with the same slice request.
struct __align__(16) RunRec {
    int32_t start, len;
    uint16_t q[KP];        // requested quantity per profile id, 0 = not requested
    uint16_t n_ent;        // number of requested profiles
    uint16_t p_single;     // the profile id when n_ent == 1
    uint16_t pad[2];
};
static_assert(sizeof(RunRec) == 48, "RunRec layout");

// ---- single-launch Plan() (fused.cu) ---------------------------------------------------------------------------------
// Pods of the fused path live in a PADDED position space: every run starts at a multiple of 64 and is padded to a multiple
// of 64 with ineligible dummies, so that the minima over 8 and over 64 positions never straddle a run (the ones over 512
// positions may: a minimum over more pods only skips less).  Everything the fused kernel reads besides the node rows is ONE
// blob -- [geometry rows | their key masks | row range per model | run records | runs per profile] as a header, then
// (cpu, memory) int32 pairs [minima per 512 | minima per 64 | minima per 8 | pods], every section a multiple of 16 bytes --
// so that a prefix of it is copied into shared memory in one piece.  A pod is its two NodeResourcesFit costs as non-negative int32 in common power-of-two units (cpu milli >> cpu_shift,
// memory bytes >> mem_shift; exact because every pod request is a multiple of the unit); F_NONE marks a pod that can never be
// placed (PreFilter verdict false, or padding) and, in the minima, a group without an eligible pod.
constexpr int FRUN_MAX = 64;                 // runs the fused kernel keeps in shared memory
// sections of the blob behind the header (indices into DevProblem::fp_off; FP_END is the total size).  Per group of eight
// positions there are, besides the minima: FP_SFX, the minima over the rest of the run from this group on ("nothing from here
// to the end of the run fits any more" is one test), and FP_CUM, {cpu, memory, count} summed over the eligible pods from the
// head of the run through this group plus the group's eligibility mask ("the whole head of the run fits" is one test per
// group, no pod is looked at); sums saturate at F_NONE.
constexpr int FP_L3 = 4, FP_L2 = 5, FP_L1 = 6, FP_SFX = 7, FP_CUM = 8, FP_PODS = 9, FP_END = 10, FP_NOFF = 11;
constexpr int F_NONE = 0x7fffffff;
constexpr int F_REM_MAX = 0x7ffffffe;        // largest remaining capacity representable; larger node values reject the fused plan
constexpr int F_COST_LIMIT = 1 << 28;        // pod costs stay below this so that eight of them sum without overflow
struct __align__(16) FRun {
    int32_t pos0;                            // first padded position (multiple of 64)
    int32_t nblk;                            // 64-pod blocks
    int32_t n_real;                          // pods of the run, eligible or not (SliceTracker counts them all)
    int32_t p;                               // requested profile id, -1: no slice request
    int32_t q;                               // requested quantity
    int32_t min_c, min_m;                    // minima over the run's eligible pods
    int32_t pad;
};
static_assert(sizeof(FRun) == 32, "FRun layout");
// reasons the fused plan is rejected (DevProblem::wave[5]); the host then runs the classic kernel sequence
constexpr int FREJ_COUNTS = 1;               // a used / free count above 31: the byte-wise sums over a node's GPUs could wrap
constexpr int FREJ_REM = 2;                  // a node's remaining cpu / memory does not fit the int32 units
constexpr int FREJ_HYPOTHESIS = 4;           // lacking set != requested profiles, or not saturated to the end
constexpr int FREJ_ACCOUNTS = 8;             // a node's Allocatable / Requested disagree with its GPUs

// State handed from rank to rank by the chained sweep: lackingSlices, the running sum(alloc - requested) per slice column,
// counters (visited nodes, geometry evaluations, placements, candidate nodes), committed flag, lacking profiles at the start,
// then the bitmap of tracker keys removed so far.
struct ChainHead {
    int32_t lack[16];
    long long D[16];
    long long counters[4];
    int32_t committed, lack_start, pad[2];
};

struct DevStats {
    long long candidates, visited_nodes, geometry_evals, placements, candidate_nodes;
    int lacking_at_start, lacking_at_end;
    int path, pad;             // 1: node-parallel pass stood, 0: sequential sweep
    int gate_again, pad2;      // nodes processed a second time, with the cluster-wide gate, after the gate replay (sweep.cu)
};

// ---- exchange over peer memory (slice mode on NVLink-connected GPUs, one process per GPU) ------------------------------
// Every rank owns one mailbox in its own HBM, mapped into the other ranks through CUDA IPC.  A rank PUSHES its block into
// every mailbox (remote stores over NVLink), fences, and raises its flag there; the consumer kernel on each rank spins on
// the W flags of its own mailbox (local loads) and reads W blocks from local memory.  Flags carry the Plan() epoch and
// slots alternate by epoch parity, so nothing is ever reset and a rank at most one step ahead never overwrites unread data.
constexpr int PEER_MAX = 8;
struct PeerBox { unsigned char* base[PEER_MAX]; int world, rank; unsigned long long xb_max; };
struct PeerWait { const unsigned* flags; int world; unsigned epoch; };      // flags == nullptr: nothing to wait for
__host__ __device__ inline size_t box_flags(int which, int parity) { return (size_t)(which - 1) * 64 + (size_t)parity * 32; }
__host__ __device__ inline size_t box_tot(int parity, int r) { return 256 + ((size_t)parity * PEER_MAX + r) * 256; }
__host__ __device__ inline size_t box_x(int parity, int r, size_t xb_max) { return 256 + 2 * PEER_MAX * 256 + ((size_t)parity * PEER_MAX + r) * xb_max; }
// a second set of slots for blocks in flag-in-word form (fused.cu): kept apart from the plain blocks so that a stale plain word
// can never be taken for a word of the current epoch
__host__ __device__ inline size_t box_ll(int parity, int r, size_t xb_max) { return 256 + 2 * PEER_MAX * 256 + 2 * PEER_MAX * xb_max + ((size_t)parity * PEER_MAX + r) * xb_max; }
inline size_t box_bytes(size_t xb_max) { return 256 + 2 * PEER_MAX * 256 + 4 * PEER_MAX * xb_max; }

// all threads of the block call it; false = a peer did not show up within several seconds (the caller records the failure
// instead of hanging the GPU)
__device__ __forceinline__ bool peer_wait(const PeerWait& w) {
    bool ok = true;
    if (w.flags && (int)threadIdx.x < w.world) {
        const volatile unsigned* f = w.flags + threadIdx.x;
        unsigned spins = 0;
        while ((int)(*f - w.epoch) < 0) { __nanosleep(500); if (++spins > (1u << 24)) { ok = false; break; } }
    }
    ok = __syncthreads_and(ok) != 0;
    __threadfence_system();
    return ok;
}

struct DevProblem {
    int kind, K, R, n_other;
    int N, G, P, n_keys, n_runs, filter_mode;
    int n_models, n_rows;
    int node_lo, node_hi;            // node range of the node-parallel pass on this rank (whole snapshot: 0, N)
    int32_t pmem[KP], pgi[KP];
    // geometry tables
    const int32_t* model_off;        // [n_models+1]
    const uint4* geom;               // [n_rows] 16 counts
    // snapshot
    const int32_t* node_gpu_count;   // [N]
    const int32_t* gpu_model;        // [N*G]
    const int32_t* gpu_mem;          // [N*G]
    const uint4* gpu_used;           // [N*G]
    const uint4* gpu_free;           // [N*G]
    const uint16_t* gpu_umask;       // [N*G]
    const uint16_t* gpu_fmask;       // [N*G]
    const long long* node_alloc;     // [N][R]
    const long long* node_req;       // [N][R]
    const int32_t* node_allowed_pods;
    const int32_t* node_num_pods;
    // pods in sorted order
    const long long* pod_cpu;        // [P]
    const long long* pod_mem;
    const long long* pod_eph;
    const int32_t* pod_key;
    const uint32_t* pod_flags;
    const int32_t* pod_sptr;         // [P+1] CSR of scalar entries
    const int32_t* pod_scol;
    const long long* pod_sqty;
    const RunRec* runs;              // [n_runs]
    const int32_t* run_chunk_off;    // [n_runs+1] first 32-pod chunk of each run in the chunk-min arrays
    const long long* cmin_cpu;       // [n_chunks] min request over the PreFilter-ok pods of the chunk (LLONG_MAX: none)
    const long long* cmin_mem;
    const long long* cmin_eph;
    int n_chunks;
    // node-parallel fast path (see sweep.cu): host-side eligibility and saturation bound per profile
    int no_eph;                      // no pod requests ephemeral storage (the kernels then skip that column's sums)
    int fast_eligible;               // every pod class requests at most one slice profile
    int32_t sat_m[KP];               // MIG: M_p, most devices of profile p one GPU can hold (allowed-geometry rows); MPS: from gflags[1]
    int slice_mode;                  // this rank holds only its own slice of the node list: no sequential sweep is possible here
    // slice mode, the sequential sweep rank after rank (nosplan_slice_chain): k_plan_sweep runs the chain over this rank's
    // slice, starting from the state its predecessor left (chain_in; NULL = the first rank) and leaving its own in chain_out
    int chain_mode; const unsigned char* chain_in; unsigned char* chain_out;
    // slice mode with the two exchanges fused into the kernels (peer-memory transport): k_prepass's last CTA pushes the totals,
    // k_tracker_init waits for and folds them, k_plan_parallel's last CTA pushes the exchange block, k_plan_sweep waits and folds
    int xf_on; unsigned xf_epoch;
    PeerBox xf_box;
    unsigned char* xf_exchange;      // this rank's exchange block [4 x u64 counters | placed-key bitmap]
    unsigned long long xf_xb;        // its size in bytes (8-byte granular)
    // work area.  The small fields exist twice; k_prepass of run k clears the copy of run k + 1 (zero_next) and, for its own
    // run, the tracker tables (placed, entry_lack), which nothing touches before k_tracker_init
    unsigned long long* zero_next; int zero_next_n8;
    long long* tot;                  // [KP] sum over nodes of (alloc - requested) per slice column
    int32_t* lack;                   // [KP] SliceTracker.lackingSlices
    int32_t* entry_lack;             // [n_keys][KP] lackingSlicesLookup
    uint8_t* cand_flag;              // [N] HasFreeCapacity
    int32_t* maxfree;                // [KP] max free count of a profile on any GPU at t0 (work area)
    int32_t* gflags;                 // [4]  [0] some node's Allocatable/Requested disagree with its GPUs' used/free
                                     //      [1] largest GPU memory GB of any GPU (MPS: bounds the devices one GPU can hold)
                                     //      [3] a peer exchange timed out on this rank (never exchanged)
    unsigned* placed;                // [(n_keys+31)/32] tracker keys placed by the node-parallel pass
    unsigned long long* acc;         // [4]  placements, geometry evals, visited, candidate nodes of the node-parallel pass
    int32_t* lack_dec;               // [KP] lacking quantities removed by the node-parallel pass (single-rank runs)
    int32_t* next_node;              // [4]  [0] work counter of the node-parallel pass, [1] / [2] CTAs of k_prepass / k_plan_parallel done
    int32_t* wave;                   // [8]  [0] k_plan_sweep hands the plan to k_plan_waves  [1],[2] first node of the wave that changes
                                     //      the lacking vector (by wave parity)  [3] candidate nodes  [4] waves run
    // gate replay (sweep.cu, gate_scan): what a node processed WITHOUT the cluster-wide gate did to the slice columns, kept when
    // some node's accounts disagree with its GPUs, so that the gate can be checked afterwards in name order
    long long* rec_dw;               // [N][KP] Allocatable from the geometry minus the snapshot's (mig/node.go:172-174)
    int2* rec_pn;                    // [N][KP] .x devices the node's new pods took, per profile; .y most devices the gate would have had to
                                     //         grant at any GPU.AddPod attempt of the node (>= .x: failed attempts of multi-profile pods count)
    int32_t* rec_fl;                 // [N]     1 | updated << 1 | pods added << 2
    long long* rec_chunk;            // [N/32 + 1][KP][4] the same per 32 nodes: change of D and smallest slack, with / without an earlier Commit
    int32_t* rec_chunk_fl;           // [N/32 + 1] some node of the chunk added a pod
    long long* gate_state;           // [KP + 2] waves: sum(alloc - requested) per slice column, `committed`, first node that fails the gate
    // single-launch Plan() (fused.cu): padded pod costs, their minima per 8 and per 64 positions, runs, units
    const unsigned char* fp_blob;
    int fp_off[FP_NOFF];             // byte offsets in the blob: key masks, model ranges, runs, runs per profile (geometry rows
                                     // at 0), minima per 512 / 64 / 8 positions, suffix minima, cumulative sums, pods, total
    int fp_n_runs, fp_pad_pods, fp_cpu_shift, fp_mem_shift;
    unsigned fp_rset;                // profiles some pod requests: the lacking set the fused kernel speculates on
    int whole_snapshot;              // the node-parallel pass covers every node on this rank (no exchange needed)
    int verify_inline;               // single GPU: the last CTA of k_plan_parallel verifies the pass itself (else k_plan_sweep does)
    // outputs
    uint4* out_count;                // [N*G]
    uint16_t* out_mask;              // [N*G]
    DevStats* stats;
};

// Programmatic dependent launch: a kernel launched with this helper may become resident while its predecessor on the stream
// is still running; it must call cudaGridDependencySynchronize() before touching anything the predecessor writes.
// Predecessors call cudaTriggerProgrammaticLaunchCompletion() early to allow it.  Saves the launch latency between the
// short kernels of one Plan().
template <class... KArgs, class... Args>
inline cudaError_t launch_pdl(void (*kernel)(KArgs...), dim3 grid, dim3 block, size_t smem, cudaStream_t st, Args... args) {
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = grid; cfg.blockDim = block; cfg.dynamicSmemBytes = smem; cfg.stream = st;
    cudaLaunchAttribute at[1];
    at[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
    at[0].val.programmaticStreamSerializationAllowed = 1;
    cfg.attrs = at; cfg.numAttrs = 1;
    return cudaLaunchKernelEx(&cfg, kernel, KArgs(args)...);
}

// last-CTA-done: true in exactly one CTA of the grid, after every other CTA's global writes and atomics are visible to it
__device__ __forceinline__ bool last_block_done(int32_t* counter) {
    __shared__ int s_last;
    __threadfence();
    __syncthreads();
    if (threadIdx.x == 0) s_last = atomicAdd(counter, 1) == (int)gridDim.x - 1;
    __syncthreads();
    if (s_last) __threadfence();
    return s_last != 0;
}

// one CTA stores `bytes` (8-byte granular, read through L2) into slot `which` of every rank's mailbox and raises its flag
__device__ __forceinline__ void peer_push(const PeerBox& box, int which, unsigned epoch, const unsigned char* src, unsigned long long bytes) {
    const int par = epoch & 1;
    const unsigned long long n8 = bytes >> 3;
    for (int r = 0; r < box.world; ++r) {
        unsigned long long* dst = reinterpret_cast<unsigned long long*>(
            box.base[r] + (which == 1 ? box_tot(par, box.rank) : box_x(par, box.rank, box.xb_max)));
        for (unsigned long long i = threadIdx.x; i < n8; i += blockDim.x) dst[i] = __ldcg(reinterpret_cast<const unsigned long long*>(src) + i);
    }
    __threadfence_system();
    __syncthreads();
    if ((int)threadIdx.x < box.world)
        *reinterpret_cast<volatile unsigned*>(box.base[threadIdx.x] + box_flags(which, par) + 4 * box.rank) = epoch;
}

__device__ __forceinline__ unsigned byte_of(const uint4& v, int p) {
    unsigned w = (p < 8) ? ((p < 4) ? v.x : v.y) : ((p < 12) ? v.z : v.w);
    return (w >> ((p & 3) * 8)) & 0xffu;
}

// bit p set <=> byte p of v is non-zero.  Per word: 0xff where the byte is non-zero, then the four byte MSBs are gathered
// into the low nibble by one multiplication (bits 7,15,23,31 -> 28..31).
__device__ __forceinline__ unsigned nz_nibble(unsigned w) {
    return ((__vcmpne4(w, 0u) & 0x80808080u) * 0x00204081u) >> 28;
}
__device__ __forceinline__ uint16_t nz_mask(const uint4& v) {
    return (uint16_t)(nz_nibble(v.x) | (nz_nibble(v.y) << 4) | (nz_nibble(v.z) << 8) | (nz_nibble(v.w) << 12));
}

__device__ __forceinline__ uint4 add_bytes(const uint4& a, const uint4& b) {
    return make_uint4(__vadd4(a.x, b.x), __vadd4(a.y, b.y), __vadd4(a.z, b.z), __vadd4(a.w, b.w));
}

}  // namespace nsp
