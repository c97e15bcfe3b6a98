// The node sweep of planner.Plan (internal/partitioning/core/planner.go:92-150) on sm_100a.
//
// process_node<>  everything the reference does for ONE candidate node, executed by one warp:
//   1. Node.UpdateGeometryFor (pkg/gpu/mig/node.go:145-177 / pkg/gpu/slicing/node.go:120-152): GPUs in order;
//      MIG: lanes = allowed geometries of the GPU model (table staged in shared memory), each lane scores its
//      geometry (pkg/gpu/mig/gpu.go:164-191), warp arg-max with the canonical tie-break (lowest table index),
//      ApplyGeometry by lanes = profiles.  MPS: lanes = profiles, closed-form create / delete / restore
//      (pkg/gpu/slicing/gpu.go:162-232).
//   2. the pod loop (planner.go:116-141 tryAddPod) over RUNS of identical slice requests instead of single pods.
//      Gate (core/snapshot.go:132-165), NodeResourcesFit on the slice columns and GPU.AddPod feasibility are
//      class-level and only ever get tighter while one node is processed, so a run whose class fails is skipped
//      whole; inside a feasible run, per-chunk minima of the pods' cpu/memory/ephemeral requests skip 32-pod
//      chunks that cannot hold a fitting pod, and a loaded chunk is resolved with warp prefix sums: every lane
//      tests its pod against the node state it would see if all earlier lanes were placed, the first lane that
//      fails under that hypothesis splits the step -- exactly the sequential first-fit.
//   3. Commit / Revert bookkeeping including the aliasing rules of the reference's object graph (state rules
//      written out in oracle/nos_oracle.cpp plan_incremental and DESIGN.md).
//
// Two drivers share it:
//   k_plan_parallel  node-parallel pass, one warp per candidate node over the whole GPU.  Exact only under three
//                    conditions it checks itself: every pod class requests at most one profile; every node's
//                    Allocatable/Requested agree with its GPUs (then the cluster-wide gate is implied by AddPod
//                    feasibility and the running totals drop out); and the lacking-slices vector is SATURATED,
//                    lack[p] > (G+2)*M_p, so that min(cand-used, lack) and the lack>free tests cannot see its exact
//                    value.  Then no state is carried from node to node except the set of tracker keys already
//                    removed, which is collected in a bitmap and turned into the final lacking vector afterwards.
//   k_plan_sweep     verifies the saturation hypothesis a posteriori (lack only decreases, so checking its final value
//                    is enough); if it held, the parallel result stands.  Otherwise -- and whenever the fast path is
//                    not applicable -- it discards it and walks the candidate nodes sequentially with one warp and the
//                    full loop-carried state (lacking vector, cluster totals, removed entries, "a Commit happened")
//                    in registers / shared memory, prefetching the next node while the current one is processed.
#include <climits>

#include <cooperative_groups.h>

#include "nosplan_dev.cuh"

namespace nsp {

namespace {

constexpr unsigned FULL = 0xffffffffu;
constexpr long long CAP_INF = (1ll << 40);
constexpr int SATURATED = INT_MAX / 2;

struct SNode {
    uint4 used[NOSPLAN_MAX_GPUS];
    uint4 free_[NOSPLAN_MAX_GPUS];
    uint16_t umask[NOSPLAN_MAX_GPUS], fmask[NOSPLAN_MAX_GPUS];
    int model[NOSPLAN_MAX_GPUS], mem[NOSPLAN_MAX_GPUS];
    long long alloc[RMAX + 1], req[RMAX + 1];
    int ng, allowed, npods, idx;
};

struct WarpScratch {       // per-warp shared memory
    SNode nd;
    uint4 pack[2];         // min(required, 255) and the required >= 256 mask, one byte per profile
    long long Dg[KP], allocF[KP], reqF[KP];
    int need[KP];          // gate-less modes: most devices of a profile the gate would have had to grant a multi-profile pod (see gate_scan)
};

struct Tables {            // CTA-wide shared memory, read-only after the prologue
    uint4 geom[MAX_ROWS];
    uint16_t gmask[MAX_ROWS];
    int off[MAX_MODELS + 1];
    int max_rows;          // most geometries of any model
};

struct PodCache {          // run records of the first RUN_CACHE runs; first 32 pods of each (sequential driver only)
    long long (*cpu)[32];
    long long (*mem)[32];
    long long (*eph)[32];
    int (*key)[32];
    unsigned (*flag)[32];
    const RunRec* runs;    // shared-memory copy of the first n_cached run records
    const int* chunk_off;  // shared-memory copy of run_chunk_off for those runs
    int n_cached;
    // bit r of prof_runs[p]: cached run r requests profile p and nothing else; prof_runs[KP]: slice-less or multi-profile runs
    const unsigned long long* prof_runs;
};

// built by the whole CTA once the run records are in shared memory (RUN_CACHE == 64 runs == one 64-bit word per profile)
__device__ __forceinline__ void build_run_masks(const RunRec* s_runs, int n_cached, unsigned long long* s_masks, int tid, int nthreads) {
    for (int p = tid; p <= KP; p += nthreads) s_masks[p] = 0;
    __syncthreads();
    for (int r = tid; r < n_cached; r += nthreads)
        atomicOr(&s_masks[s_runs[r].n_ent == 1 ? s_runs[r].p_single : KP], 1ull << r);
    __syncthreads();
}
static_assert(RUN_CACHE == 64, "one 64-bit word of runs per profile");

struct NodeRegs {          // one node spread over the lanes of a warp
    uint4 used, fre;
    int um, fm, model, mem;
    long long alloc, req;
    int ng, allowed, npods;
};

__device__ __forceinline__ void load_node(const DevProblem& pb, int i, int lane, NodeRegs& r) {
    r.ng = pb.node_gpu_count[i];
    r.allowed = pb.node_allowed_pods[i];
    r.npods = pb.node_num_pods[i];
    r.used = make_uint4(0, 0, 0, 0); r.fre = r.used; r.um = r.fm = r.model = r.mem = 0;
    if (lane < r.ng) {
        const int x = i * pb.G + lane;
        r.used = pb.gpu_used[x]; r.fre = pb.gpu_free[x];
        r.um = pb.gpu_umask[x]; r.fm = pb.gpu_fmask[x];
        r.model = pb.gpu_model[x]; r.mem = pb.gpu_mem[x];
    }
    r.alloc = r.req = 0;
    if (lane < pb.R) {
        r.alloc = pb.node_alloc[(size_t)i * pb.R + lane];
        r.req = pb.node_req[(size_t)i * pb.R + lane];
    }
}

__device__ __forceinline__ void store_node(SNode& nd, const NodeRegs& r, int i, int lane) {
    if (lane < NOSPLAN_MAX_GPUS) {
        nd.used[lane] = r.used; nd.free_[lane] = r.fre;
        nd.umask[lane] = (uint16_t)r.um; nd.fmask[lane] = (uint16_t)r.fm;
        nd.model[lane] = r.model; nd.mem[lane] = r.mem;
    }
    nd.alloc[lane] = r.alloc; nd.req[lane] = r.req;     // lane < 32 == RMAX+1
    if (lane == 0) { nd.ng = r.ng; nd.allowed = r.allowed; nd.npods = r.npods; nd.idx = i; }
}

__device__ __forceinline__ uint8_t* bytes(uint4& v) { return reinterpret_cast<uint8_t*>(&v); }

__device__ __forceinline__ long long warp_sum_ll(long long v) {
#pragma unroll
    for (int o = 16; o; o >>= 1) v += __shfl_xor_sync(FULL, v, o);
    return v;
}
__device__ __forceinline__ int warp_sum_int(int v) { return __reduce_add_sync(FULL, v); }
// sum of 32 NON-NEGATIVE int64 with the 32-bit warp-reduce unit: three 21-bit limbs, each limb sum < 2^26
__device__ __forceinline__ long long warp_sum_nonneg_ll(long long v) {
    const unsigned long long u = (unsigned long long)v;
    const unsigned s0 = __reduce_add_sync(FULL, (unsigned)(u & 0x1fffffu));
    const unsigned s1 = __reduce_add_sync(FULL, (unsigned)((u >> 21) & 0x1fffffu));
    const unsigned s2 = __reduce_add_sync(FULL, (unsigned)(u >> 42));
    return (long long)((unsigned long long)s0 + ((unsigned long long)s1 << 21) + ((unsigned long long)s2 << 42));
}

__device__ __forceinline__ long long warp_excl_ll(long long v, int lane) {
    long long s = v;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
        long long t = __shfl_up_sync(FULL, s, o);
        if (lane >= o) s += t;
    }
    return s - v;
}
__device__ __forceinline__ int warp_excl_int(int v, int lane) {
    int s = v;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
        int t = __shfl_up_sync(FULL, s, o);
        if (lane >= o) s += t;
    }
    return s - v;
}

// ---------------------------------------------------------------- MIG GPU.UpdateGeometryFor (mig/gpu.go:158-212)
// Scoring of one allowed geometry against one GPU, 4 profiles per 32-bit word with byte-SIMD integer ops:
//   score = sum_p [required[p] > 0 and free[p] < required[p]] * max(0, min(row[p] - used[p], required[p]))
// `rq` holds min(required, 255) per byte (exact: row - used <= 255), `big` 0xff where required >= 256 (then
// free < required always).  CanApplyGeometry (:97-110) for an allowed row is row[p] >= used[p] for every p.
struct ReqPack { uint4 rq, big; int n_req; };

__device__ __forceinline__ unsigned score_word(unsigned row, unsigned used, unsigned fre, unsigned rq, unsigned big) {
    const unsigned m = __vminu4(__vsubus4(row, used), rq);
    return __vsadu4(m & (__vcmpltu4(fre, rq) | big), 0u);
}

__device__ __forceinline__ unsigned score_row(const uint4& row, const uint4& used, const uint4& fre, const ReqPack& rp, int K) {
    // bytes of profiles >= K are zero in row, used and free, so only the words below K matter
    unsigned can = __vcmpgeu4(row.x, used.x);
    unsigned score = score_word(row.x, used.x, fre.x, rp.rq.x, rp.big.x);
    if (K > 4) { can &= __vcmpgeu4(row.y, used.y); score += score_word(row.y, used.y, fre.y, rp.rq.y, rp.big.y); }
    if (K > 8) {
        can &= __vcmpgeu4(row.z, used.z) & __vcmpgeu4(row.w, used.w);
        score += score_word(row.z, used.z, fre.z, rp.rq.z, rp.big.z) + score_word(row.w, used.w, fre.w, rp.rq.w, rp.big.w);
    }
    return can == 0xffffffffu ? score : 0u;
}

// lane p (< 16) holds required[p]; every lane gets the packed form through shared memory
__device__ __forceinline__ ReqPack pack_required(int required, int lane, uint4* s_pack) {
    uint8_t* b = reinterpret_cast<uint8_t*>(s_pack);
    if (lane < KP) { b[lane] = (uint8_t)min(max(required, 0), 255); b[16 + lane] = required >= 256 ? 0xff : 0; }
    __syncwarp();
    ReqPack rp;
    rp.rq = s_pack[0]; rp.big = s_pack[1];
    rp.n_req = __popc(__ballot_sync(FULL, required > 0));
    __syncwarp();
    return rp;
}

// ApplyGeometry (:132-152) of table row `br` on GPU g by the 16 lanes `pl` = 0..15 of the caller's segment
__device__ __forceinline__ void apply_row(SNode& nd, int g, int br, int pl, const Tables& tb) {
    const unsigned rw = byte_of(tb.geom[br], pl), us = byte_of(nd.used[g], pl);
    bytes(nd.free_[g])[pl] = (uint8_t)(rw ? rw - us : 0);                     // zero kept (:141)
    if (pl == 0) nd.fmask[g] = tb.gmask[br];
}

// One GPU: lanes = geometries of the GPU's model.  Returns updated (warp-uniform).
__device__ __forceinline__ bool mig_gpu_update(SNode& nd, int g, const ReqPack& rp, int lane, int K, const Tables& tb,
                                               long long& evals) {
    const int m = nd.model[g];
    const int r0 = tb.off[m], nrows = tb.off[m + 1] - r0;
    evals += (long long)nrows * rp.n_req;
    unsigned key = 0;
    if (lane < nrows) {
        const unsigned score = score_row(tb.geom[r0 + lane], nd.used[g], nd.free_[g], rp, K);
        if (score) key = (score << 5) | (unsigned)(31 - lane);
    }
    const unsigned best = __reduce_max_sync(FULL, key);
    if (best == 0) return false;
    const int br = r0 + 31 - (int)(best & 31u);            // strict > in table order == lowest index among max
    if (lane < KP) apply_row(nd, g, br, lane, tb);
    __syncwarp();
    return true;
}

// All GPUs of a node at once -- valid only when `required` is not consumed from GPU to GPU (saturated lacking
// vector, node-parallel pass).  Lanes = (GPU, geometry) pairs, S lanes per GPU, 32/S GPUs per step.
__device__ __forceinline__ bool mig_node_update_all(SNode& nd, int ng, const ReqPack& rp, int lane, int K, const Tables& tb,
                                                    long long& evals) {
    const int S = tb.max_rows <= 16 ? 16 : 32;
    const int sub = lane / S, r = lane - sub * S;
    const unsigned segmask = S == 32 ? FULL : (0xffffu << (16 * sub));
    bool upd = false;
    for (int g0 = 0; g0 < ng; g0 += 32 / S) {
        const int g = g0 + sub;
        unsigned key = 0;
        int r0 = 0;
        if (g < ng) {
            const int m = nd.model[g];
            r0 = tb.off[m];
            const int nrows = tb.off[m + 1] - r0;
            if (r == 0) evals += (long long)nrows * rp.n_req;
            if (r < nrows) {
                const unsigned score = score_row(tb.geom[r0 + r], nd.used[g], nd.free_[g], rp, K);
                if (score) key = (score << 5) | (unsigned)(31 - r);
            }
        }
        const unsigned best = __reduce_max_sync(segmask, key);
        if (best != 0 && r < KP) apply_row(nd, g, r0 + 31 - (int)(best & 31u), r, tb);
        upd |= __any_sync(FULL, best != 0);
        __syncwarp();
    }
    evals = warp_sum_ll(evals);                                // per-segment partial counts -> warp total
    return upd;
}

// ---------------------------------------------------------------- MPS GPU.UpdateGeometryFor (slicing/gpu.go:162-232)
// lanes = profiles (ids ascend by (size, name)).  Returns updated (warp-uniform).
__device__ __forceinline__ bool mps_gpu_update(SNode& nd, int g, int required, int lane, const DevProblem& pb) {
    const int size = lane < pb.K ? pb.pmem[lane] : 0;
    const int used = lane < KP ? (int)bytes(nd.used[g])[lane] : 0;
    int fre = lane < KP ? (int)bytes(nd.free_[g])[lane] : 0;
    unsigned fmask = nd.fmask[g];
    const int mem = nd.mem[g];
    int missing = 0;
    if (lane < pb.K && required > 0) { const int d = required - fre; if (d > 0) missing = d; }
    unsigned miss_mask = __ballot_sync(FULL, missing > 0);
    if (miss_mask == 0) return false;
    bool updated = false;
    const int ofree = fre;
    const unsigned omask = fmask;
    while (miss_mask) {                                            // ascending profile id == (size, name) order
        const int mp = __ffs(miss_mask) - 1;
        miss_mask &= miss_mask - 1;
        const int msize = __shfl_sync(FULL, size, mp);
        int mq = __shfl_sync(FULL, missing, mp);
        int spare = mem - warp_sum_int(size * (used + fre));
        if (spare >= 1) {                                          // spare capacity first (:188-197)
            const int c = msize > 0 ? min(mq, spare / msize) : mq;
            if (c > 0) { if (lane == mp) fre += c; fmask |= 1u << mp; mq -= c; updated = true; }
        }
        if (lane < KP && (omask >> lane & 1)) fre = 0;             // delete ALL originally-free keys (:199-202)
        fmask &= ~omask;
        spare = mem - warp_sum_int(size * (used + fre));
        {                                                          // :203-211
            int c = 0;
            if (spare >= 1) c = msize > 0 ? min(mq, spare / msize) : mq;
            if (c > 0) { if (lane == mp) fre += c; fmask |= 1u << mp; mq -= c; updated = true; spare -= c * msize; }
        }
        unsigned rest = omask;                                     // restore, all-or-nothing per profile (:213-216)
        while (rest) {
            const int k = __ffs(rest) - 1;
            rest &= rest - 1;
            const int need = __shfl_sync(FULL, size * ofree, k);
            if (spare >= need) { if (lane == k) fre += ofree; fmask |= 1u << k; spare -= need; }
        }
    }
    if (lane < KP) bytes(nd.free_[g])[lane] = (uint8_t)fre;
    if (lane == 0) nd.fmask[g] = (uint16_t)fmask;
    __syncwarp();
    return updated;
}

// SliceTracker.Remove (core/tracker.go:67-88).
//   sequential driver: `removed` is a shared-memory bitmap of emptied entries; lanes = profiles subtract the entry.
//   parallel driver:   only the global "placed" bitmap is set; the lacking vector is rebuilt from it afterwards.
// MODE: how one node is processed
//   M_SEQ     sequential sweep: exact lacking vector, gate, shared-memory removed-set
//   M_PAR     node-parallel pass: saturated view, geometry of all GPUs at once, global placed bitmap
//   M_SPEC    wave, speculative: exact view, nothing written; reports whether the node would change the lacking vector
//   M_COMMIT  wave, the one node of a wave that does change it: exact view, updates the global removed-set and vector
constexpr int M_SEQ = 0, M_PAR = 1, M_SPEC = 2, M_COMMIT = 3;

template <int MODE>
__device__ __forceinline__ void tracker_remove(const DevProblem& pb, int key, unsigned* removed, int& lack, int lane, int& dirty) {
    constexpr bool PAR = MODE == M_PAR;
    if (MODE == M_SPEC) {
        // would this placement change the lacking vector?  (the entry's one and only removal, tracker.go:78-86)
        const unsigned w = __ldcg(&removed[key >> 5]);       // written by other SMs in earlier waves: read at L2
        if (w >> (key & 31) & 1) return;
        const int e = (lane < KP) ? pb.entry_lack[(size_t)key * KP + lane] : 0;
        if (__ballot_sync(FULL, e > 0)) dirty = 1;
        return;
    }
    if (PAR) {
        if (lane == 0) {
            const unsigned bit = 1u << (key & 31);
            if (!(removed[key >> 5] & bit) && !(atomicOr(&removed[key >> 5], bit) & bit) && pb.whole_snapshot)
                for (int pp = 0; pp < pb.K; ++pp) {
                    const int e = pb.entry_lack[(size_t)key * KP + pp];
                    if (e > 0) atomicAdd(&pb.lack_dec[pp], e);
                }
        }
        return;
    }
    const unsigned w = MODE == M_COMMIT ? __ldcg(&removed[key >> 5]) : removed[key >> 5];
    if (w >> (key & 31) & 1) return;
    __syncwarp();
    if (lane == 0) removed[key >> 5] = w | (1u << (key & 31));
    if (lane < KP) {
        const int e = pb.entry_lack[(size_t)key * KP + lane];
        if (e > 0) lack -= e;
    }
    __syncwarp();
}

// One candidate node, start to finish.  `ws.nd` holds the node.  Returns the number of pods added.
//   lack      lane p: SliceTracker.lackingSlices[p] (PAR: SATURATED on the lacking profiles, never modified)
//   D         lane p: cluster-wide sum(Allocatable - Requested) of slice column p (unused when PAR)
//   GATE      the cluster-wide gate (core/snapshot.go:132-165) is evaluated.  Always in the sequential sweep; in the other
//             modes it is vacuous when every node's accounts agree with its GPUs, and where they do not the node is first
//             processed without it, leaving a record (write_gate_record) from which gate_scan decides afterwards, in name order, whether the
//             gate would have made a difference -- only then is the node processed again with GATE set
template <int MODE, bool GATE = (MODE == M_SEQ)>
__device__ __forceinline__ int process_node(const DevProblem& pb, WarpScratch& ws, const Tables& tb, const PodCache* pc,
                                            unsigned* removed, int& lack, long long& D, bool& committed,
                                            long long& st_evals, long long& st_place, int lane, int& dirty, bool& upd_out) {
    constexpr bool PAR = MODE == M_PAR;        // saturated view: geometry of all GPUs at once, global placed bitmap
    SNode& nd = ws.nd;
    const bool mig = pb.kind == NOSPLAN_KIND_MIG;
    const int K = pb.K, fm = pb.filter_mode;
    const int ng = nd.ng, node_i = nd.idx;
    const long long oldA = (lane < K) ? nd.alloc[3 + lane] : 0;
    const long long oldR = (lane < K) ? nd.req[3 + lane] : 0;

    // ---- 1. Node.UpdateGeometryFor
    int required = lack;
    bool upd = false;
    if (PAR && mig) {
        const ReqPack rp = pack_required(required, lane, ws.pack);
        long long ev = 0;
        upd = mig_node_update_all(nd, ng, rp, lane, K, tb, ev);
        st_evals += ev;
    } else {
        for (int g = 0; g < ng; ++g) {
            bool u;
            if (mig) {
                const ReqPack rp = pack_required(required, lane, ws.pack);
                u = mig_gpu_update(nd, g, rp, lane, K, tb, st_evals);
            } else {
                st_evals += __popc(__ballot_sync(FULL, required > 0));
                u = mps_gpu_update(nd, g, required, lane, pb);
            }
            upd |= u;
            if (!PAR && lane < KP) {
                const int f = (int)bytes(nd.free_[g])[lane];
                if ((nd.fmask[g] >> lane & 1) && required > 0) required = max(0, required - f);
            }
        }
    }
    long long newA = 0;                                        // computeScalarResources (mig/node.go:179-195)
    if (lane < K)
        for (int g = 0; g < ng; ++g) newA += (long long)bytes(nd.used[g])[lane] + bytes(nd.free_[g])[lane];
    const long long allocF = upd ? newA : oldA;                // Allocatable of the forked entry the pods are added to
    if (lane < KP) { ws.allocF[lane] = allocF; ws.reqF[lane] = oldR; ws.Dg[lane] = GATE ? D - oldA + allocF : 0; ws.need[lane] = 0; }
    long long remc = nd.alloc[0] - nd.req[0], remm = nd.alloc[1] - nd.req[1], reme = nd.alloc[2] - nd.req[2];
    const bool podcount_ok = nd.npods + 1 <= nd.allowed;
    // no pod requests ephemeral storage and the node is not over-committed on it: every `0 <= reme` test holds
    const bool use_eph = !(pb.no_eph && reme >= 0);
    const bool fit = fm == NOSPLAN_FILTER_RESOURCES_FIT;
    __syncwarp();
    // profiles with at least one free device on some GPU of the node: a single-profile class outside this set cannot place
    // a pod (GPU.AddPod needs free >= q >= 1), and the set only shrinks while the node is processed
    uint4 any_free = make_uint4(0, 0, 0, 0);
    for (int g = 0; g < ng; ++g) { const uint4 f = nd.free_[g]; any_free.x |= f.x; any_free.y |= f.y; any_free.z |= f.z; any_free.w |= f.w; }
    const unsigned feas = nz_mask(any_free);

    // ---- 2. the pod loop, run by run
    int added = 0;
    const bool all_fail = (fm == NOSPLAN_FILTER_REJECT_ALL) || (fit && !podcount_ok);
    // cached runs: only those whose single profile has a free device somewhere on the node (plus slice-less / multi-profile ones)
    unsigned long long live = pc->prof_runs[KP];
    for (unsigned f = feas; f; f &= f - 1) live |= pc->prof_runs[__ffs(f) - 1];
    int next_uncached = pc->n_cached;
    while (!all_fail) {
        int ri;
        if (live) { ri = __ffsll((long long)live) - 1; live &= live - 1; }
        else if (next_uncached < pb.n_runs) ri = next_uncached++;
        else break;
        const bool rcached = ri < pc->n_cached;
        const bool cached = rcached && pc->cpu != nullptr;
        const RunRec* rr = rcached ? &pc->runs[ri] : &pb.runs[ri];
        const int n_ent = rr->n_ent;
        if (!rcached && n_ent == 1 && !(feas >> rr->p_single & 1)) continue;
        const int rstart = rr->start, rend = rr->start + rr->len;
        const int q_l = (lane < KP) ? (int)rr->q[lane] : 0;
        // class-level tests; all of them only get tighter while this node is processed
        bool ok_l = true;
        if (q_l > 0) {
            if (GATE) ok_l = ws.Dg[lane] >= q_l;                                            // gate, snapshot.go:132-165
            if (fit) ok_l = ok_l && (q_l <= ws.allocF[lane] - ws.reqF[lane]);
        }
        if (!__all_sync(FULL, ok_l)) continue;

        if (n_ent <= 1) {
            // ---------- single-profile (or slice-less) class: bulk first-fit
            const unsigned emask = __ballot_sync(FULL, q_l > 0);
            const int p = n_ent ? __ffs(emask) - 1 : 0;
            const int q = n_ent ? __shfl_sync(FULL, q_l, p) : 0;
            long long cap = CAP_INF;
            if (n_ent) {
                const int fr0 = (lane < ng) ? (int)bytes(nd.free_[lane])[p] : 0;
                const int cg = q == 1 ? fr0 : fr0 / q;                                        // GPU.AddPod feasibility
                cap = warp_sum_int(cg);
                if (GATE) cap = min(cap, ws.Dg[p] / q);
                if (fit) cap = min(cap, (ws.allocF[p] - ws.reqF[p]) / q);
            }
            if (cap <= 0) continue;
            const int c0 = rcached ? pc->chunk_off[ri] : pb.run_chunk_off[ri];
            const int nch = (rr->len + 31) >> 5;
            for (int cb = 0; cb < nch && cap > 0; cb += 32) {
                // chunk minima of 32 chunks at a time: a chunk whose cheapest eligible pod does not fit is skipped
                long long mc = LLONG_MAX, mm = LLONG_MAX, me = LLONG_MAX;
                if (cb + lane < nch) {
                    mc = pb.cmin_cpu[c0 + cb + lane]; mm = pb.cmin_mem[c0 + cb + lane];
                    me = use_eph ? pb.cmin_eph[c0 + cb + lane] : 0;        // !use_eph: reme >= 0, the test below holds
                }
                int cdone = 0;
                while (cap > 0) {
                    bool pass = mc != LLONG_MAX && lane >= cdone;
                    if (fit) pass = pass && mc <= remc && mm <= remm && me <= reme;
                    const unsigned pm = __ballot_sync(FULL, pass);
                    if (pm == 0) break;
                    const int cl = __ffs(pm) - 1;
                    cdone = cl + 1;
                    const int cursor = rstart + (cb + cl) * 32;
                    const int s = cursor + lane;
                    const bool in = s < rend;
                    long long cpu = 0, mem = 0, eph = 0; int key = 0; unsigned fl = 0;
                    if (in) {
                        if (cached && cursor == rstart) {
                            cpu = pc->cpu[ri][lane]; mem = pc->mem[ri][lane]; eph = pc->eph[ri][lane];
                            key = pc->key[ri][lane]; fl = pc->flag[ri][lane];
                        } else {
                            cpu = pb.pod_cpu[s]; mem = pb.pod_mem[s]; if (use_eph) eph = pb.pod_eph[s];
                            key = pb.pod_key[s]; fl = pb.pod_flags[s];
                        }
                    }
                    const bool slow = in && fit && (fl & PF_HAS_OTHER);
                    const bool zero = (fl & PF_ZERO_REQ) != 0;
                    int from = 0;
                    while (from < 32 && cap > 0) {
                        // pods with other scalar resources take the exact per-pod path; they split the step
                        const unsigned slow_m = __ballot_sync(FULL, slow && lane >= from);
                        const int stop = slow_m ? __ffs(slow_m) - 1 : 32;
                        bool valid = in && lane >= from && lane < stop && (fl & PF_PRE_OK);
                        if (fit) valid = valid && (zero || (cpu <= remc && mem <= remm && eph <= reme));   // can never pass later
                        const unsigned vm = __ballot_sync(FULL, valid);
                        if (vm == 0 && stop == 32) break;
                        const int rank = __popc(vm & ((1u << lane) - 1));
                        const bool cand = valid && rank < cap;
                        // totals of the candidates first: if they all fit together (the common case) no prefix sums are needed
                        long long sc = 0, sm = 0, se = 0;
                        int firstfail = 32;
                        if (fit) {
                            sc = warp_sum_nonneg_ll(cand ? cpu : 0); sm = warp_sum_nonneg_ll(cand ? mem : 0);
                            se = use_eph ? warp_sum_nonneg_ll(cand ? eph : 0) : 0;
                            if (sc > remc || sm > remm || se > reme) {
                                const long long ec = warp_excl_ll(cand ? cpu : 0, lane), em = warp_excl_ll(cand ? mem : 0, lane),
                                                ee = use_eph ? warp_excl_ll(cand ? eph : 0, lane) : 0;
                                const bool test = cand && (zero || (cpu <= remc - ec && mem <= remm - em && eph <= reme - ee));
                                const unsigned failm = __ballot_sync(FULL, cand && !test);
                                if (failm) {
                                    firstfail = __ffs(failm) - 1;
                                    sc = __shfl_sync(FULL, ec, firstfail); sm = __shfl_sync(FULL, em, firstfail); se = __shfl_sync(FULL, ee, firstfail);
                                }
                            }
                        }
                        const bool acc = cand && lane < firstfail;
                        const unsigned accm = __ballot_sync(FULL, acc);
                        const int n = __popc(accm);
                        if (n > 0) {
                            remc -= sc; remm -= sm; reme -= se;
                            if (n_ent) {                                   // first-fit over GPUs (mig/node.go:201-210)
                                const int fr1 = (lane < ng) ? (int)bytes(nd.free_[lane])[p] : 0;
                                const int cg = q == 1 ? fr1 : fr1 / q;
                                const int ex = warp_excl_int(cg, lane);
                                const int take = max(0, min(cg, n - ex));
                                if (lane < ng && take > 0) {
                                    bytes(nd.free_[lane])[p] -= (uint8_t)(take * q);
                                    bytes(nd.used[lane])[p] += (uint8_t)(take * q);
                                    nd.umask[lane] |= (uint16_t)(1u << p);
                                }
                                if (lane == 0) { ws.reqF[p] += (long long)n * q; ws.Dg[p] -= (long long)n * q; }
                                cap -= n;
                            }
                            if (PAR) {                                     // tracker.Remove per placed pod
                                if (acc) {
                                    const unsigned bit = 1u << (key & 31);
                                    if (!(removed[key >> 5] & bit) && !(atomicOr(&removed[key >> 5], bit) & bit) && pb.whole_snapshot) {
                                        // this lane set the bit: it owns the entry's one and only removal (tracker.go:78-86)
                                        for (int pp = 0; pp < K; ++pp) {
                                            const int e = pb.entry_lack[(size_t)key * KP + pp];
                                            if (e > 0) atomicAdd(&pb.lack_dec[pp], e);
                                        }
                                    }
                                }
                            } else {
                                unsigned todo = accm;
                                while (todo) {
                                    const int l = __ffs(todo) - 1;
                                    todo &= todo - 1;
                                    tracker_remove<MODE>(pb, __shfl_sync(FULL, key, l), removed, lack, lane, dirty);
                                }
                            }
                            added += n; st_place += n;
                            __syncwarp();
                        }
                        if (firstfail < 32) { from = firstfail + 1; continue; }
                        if (stop < 32 && cap > 0) {
                            // exact per-pod path for the pod at lane `stop` (has non-slice scalar requests)
                            const int ps = cursor + stop;
                            const long long pcpu = __shfl_sync(FULL, cpu, stop), pmem_ = __shfl_sync(FULL, mem, stop), peph = __shfl_sync(FULL, eph, stop);
                            const unsigned pfl = __shfl_sync(FULL, fl, stop);
                            const int pkey = __shfl_sync(FULL, key, stop);
                            bool ok = (pfl & PF_PRE_OK) != 0;
                            ok = ok && pcpu <= remc && pmem_ <= remm && peph <= reme;
                            if (ok && lane == 0) {
                                for (int e = pb.pod_sptr[ps]; e < pb.pod_sptr[ps + 1]; ++e) {
                                    const int col = pb.pod_scol[e];
                                    if (col >= 3 + K && pb.pod_sqty[e] > nd.alloc[col] - nd.req[col]) ok = false;
                                }
                            }
                            ok = __shfl_sync(FULL, (int)ok, 0) != 0;
                            if (ok) {
                                remc -= pcpu; remm -= pmem_; reme -= peph;
                                if (lane == 0)
                                    for (int e = pb.pod_sptr[ps]; e < pb.pod_sptr[ps + 1]; ++e) {
                                        const int col = pb.pod_scol[e];
                                        if (col >= 3 + K) nd.req[col] += pb.pod_sqty[e];
                                    }
                                if (n_ent) {
                                    const unsigned gm = __ballot_sync(FULL, lane < ng && (int)bytes(nd.free_[lane < ng ? lane : 0])[p] >= q);
                                    const int gsel = __ffs(gm) - 1;        // cap > 0 guarantees one exists
                                    if (lane == gsel) {
                                        bytes(nd.free_[lane])[p] -= (uint8_t)q; bytes(nd.used[lane])[p] += (uint8_t)q;
                                        nd.umask[lane] |= (uint16_t)(1u << p);
                                    }
                                    if (lane == 0) { ws.reqF[p] += q; ws.Dg[p] -= q; }
                                    cap -= 1;
                                }
                                tracker_remove<MODE>(pb, pkey, removed, lack, lane, dirty);
                                ++added; ++st_place;
                                __syncwarp();
                            }
                            from = stop + 1;
                            continue;
                        }
                        break;
                    }
                }
            }
        } else {
            // ---------- multi-profile class: exact per-pod path (GPU.AddPod keeps partial moves, mig/gpu.go:234-245)
            for (int s = rstart; s < rend; ++s) {
                const unsigned fl = pb.pod_flags[s];
                const long long cpu = pb.pod_cpu[s], mem = pb.pod_mem[s], eph = pb.pod_eph[s];
                bool okc = true;                                   // gate + class-level fit again (state moves inside the run)
                if (q_l > 0) {
                    if (GATE) okc = ws.Dg[lane] >= q_l;
                    if (fit) okc = okc && (q_l <= ws.allocF[lane] - ws.reqF[lane]);
                }
                if (!__all_sync(FULL, okc)) break;                 // every later pod of the run fails the same way
                bool ok = (fl & PF_PRE_OK) != 0;
                if (fit) {
                    ok = ok && cpu <= remc && mem <= remm && eph <= reme;
                    if (ok && (fl & PF_HAS_OTHER)) {
                        if (lane == 0)
                            for (int e = pb.pod_sptr[s]; e < pb.pod_sptr[s + 1]; ++e) {
                                const int col = pb.pod_scol[e];
                                if (col >= 3 + K && pb.pod_sqty[e] > nd.alloc[col] - nd.req[col]) ok = false;
                            }
                        ok = __shfl_sync(FULL, (int)ok, 0) != 0;
                    }
                }
                if (!ok) continue;
                // GPU.AddPod is attempted, and a failed attempt may leave devices moved: without the gate, note what it would
                // have had to grant for this attempt (devices of the profile taken so far on this node + this pod's)
                if (!GATE && q_l > 0) ws.need[lane] = max(ws.need[lane], (int)(ws.reqF[lane] - oldR) + q_l);
                bool placed = false, moved = false;                // Node.AddPod: GPUs in order, profiles ascending
                for (int g = 0; g < ng && !placed; ++g) {
                    const int fr = (lane < KP) ? (int)bytes(nd.free_[g])[lane] : 0;
                    const unsigned bad = __ballot_sync(FULL, q_l > 0 && fr < q_l);
                    const unsigned want = __ballot_sync(FULL, q_l > 0);
                    const unsigned mv = bad ? (want & ((1u << (__ffs(bad) - 1)) - 1)) : want;   // moved before the first failure
                    if (mv >> lane & 1) {
                        bytes(nd.free_[g])[lane] -= (uint8_t)q_l; bytes(nd.used[g])[lane] += (uint8_t)q_l;
                    }
                    if (lane == 0 && mv) nd.umask[g] |= (uint16_t)mv;
                    moved |= mv != 0;
                    placed = bad == 0;
                    __syncwarp();
                }
                if (!placed) { if (!moved) break; else continue; }   // a clean failure repeats for the rest of the run
                remc -= cpu; remm -= mem; reme -= eph;
                if (q_l > 0) { ws.reqF[lane] += q_l; ws.Dg[lane] -= q_l; }
                if (lane == 0 && (fl & PF_HAS_OTHER))
                    for (int e = pb.pod_sptr[s]; e < pb.pod_sptr[s + 1]; ++e) {
                        const int col = pb.pod_scol[e];
                        if (col >= 3 + K) nd.req[col] += pb.pod_sqty[e];
                    }
                tracker_remove<MODE>(pb, pb.pod_key[s], removed, lack, lane, dirty);
                ++added; ++st_place;
                __syncwarp();
            }
        }
    }

    // ---- 3. Commit / Revert (planner.go:144-149) with the object-graph aliasing rules
    if (added > 0) {
        if (GATE && lane < KP) D = ws.Dg[lane];
        committed = true;
        if (MODE != M_SPEC && upd && lane < ng) {                                // partitioningState[n] = GetPartitioning(n) (planner.go:138)
            const int x = node_i * pb.G + lane;
            pb.out_count[x] = add_bytes(nd.used[lane], nd.free_[lane]);
            pb.out_mask[x] = nd.umask[lane] | nd.fmask[lane];
        }
    } else if (GATE && !committed) {
        D += newA - oldA;                                      // the mutated original is still what the snapshot holds
    }
    upd_out = upd;
    __syncwarp();
    return added;
}

// What gate_scan replays for a node that process_node has just taken WITHOUT the gate: Allocatable from the geometry minus the
// snapshot's (devices only move from free to used while pods are added, so the sum is still the one UpdateGeometryFor left),
// the devices its new pods took, and the outcome.
__device__ __forceinline__ void write_gate_record(const DevProblem& pb, const WarpScratch& ws, int lane, int added, bool upd) {
    const SNode& nd = ws.nd;
    if (lane < KP) {
        long long newA = 0;
        if (lane < pb.K)
            for (int g = 0; g < nd.ng; ++g)
                newA += (long long)reinterpret_cast<const uint8_t*>(&nd.used[g])[lane] + reinterpret_cast<const uint8_t*>(&nd.free_[g])[lane];
        const long long oldA = (lane < pb.K) ? nd.alloc[3 + lane] : 0, oldR = (lane < pb.K) ? nd.req[3 + lane] : 0;
        const int pl = (int)(ws.reqF[lane] - oldR);
        pb.rec_dw[(size_t)nd.idx * KP + lane] = newA - oldA;
        pb.rec_pn[(size_t)nd.idx * KP + lane] = make_int2(pl, max(pl, ws.need[lane]));
    }
    if (lane == 0) pb.rec_fl[nd.idx] = 1 | (upd ? 2 : 0) | (added << 2);
}

__device__ __forceinline__ void load_tables(const DevProblem& pb, Tables& tb, int tid, int nthreads) {
    if (pb.kind != NOSPLAN_KIND_MIG) return;
    for (int r = tid; r < pb.n_rows; r += nthreads) { uint4 row = pb.geom[r]; tb.geom[r] = row; tb.gmask[r] = nz_mask(row); }
    for (int m = tid; m <= pb.n_models; m += nthreads) tb.off[m] = pb.model_off[m];
    if (tid == 0) {
        int mr = 0;
        for (int m = 0; m < pb.n_models; ++m) mr = max(mr, pb.model_off[m + 1] - pb.model_off[m]);
        tb.max_rows = mr;
    }
}

// Is the node-parallel pass applicable?  lane p evaluates profile p; warp-uniform result.
// Saturation bound: one GPU holds at most M_p devices of profile p (allowed geometries / memory, or its t0 free
// count), a node has at most G GPUs, so lack[p] > (G+2)*M_p keeps every comparison against `required` one-sided.
// any_accounts: also when some node's accounts disagree with its GPUs or a pod class asks for several profiles (the pass then
// leaves records for the gate replay)
__device__ __forceinline__ bool fast_applicable(const DevProblem& pb, int lack_p, int lane, bool any_accounts = false) {
    bool ok = true;
    if (lane < KP && lack_p > 0) {
        int sat = pb.sat_m[lane];
        if (pb.kind == NOSPLAN_KIND_MPS) sat = pb.pmem[lane] > 0 ? pb.gflags[1] / pb.pmem[lane] : -1;   // devices of this size a GPU can hold
        const long long m = max(sat, pb.maxfree[lane]);
        ok = sat >= 0 && (long long)lack_p > (long long)(pb.G + 2) * m;
    }
    const bool any_lack = __ballot_sync(FULL, lane < KP && lack_p > 0) != 0;
    return __all_sync(FULL, ok) && any_lack && (any_accounts || (pb.fast_eligible && pb.gflags[0] == 0));
}

// ---- gate replay.  Nodes [a, b) were processed without the cluster-wide gate (core/snapshot.go:132-165) and left records.
// The gate compares, per requested profile, the running cluster total D[p] = sum(Allocatable - Requested) with the pod's
// quantity, in front of GPU.AddPod.  While a node is processed the total only shrinks (by the devices its pods take), so every
// attempt of the node passes the gate iff D + dA - need >= 0, need = the largest (devices taken so far + the pod's own) over
// the attempts that matter: the placements (a pod GPU.AddPod refuses cleanly is skipped whatever the gate says) and, for pods
// asking for several profiles, every attempt -- a refused one may leave devices moved (mig/gpu.go:234-245), which the gate
// would have prevented.  Walks the candidates in name order, 32 per step (lane = node, warp scans of the per-node change of
// D under the Commit / Revert rules of process_node), and stops at the first node with an attempt that would not have passed:
// returns it (b: none) with D (lane p: profile p) and `committed` as they are in front of that node.
template <bool COHERENT>
__device__ __forceinline__ int gate_scan(const DevProblem& pb, int a, int b, unsigned profs, long long& D, bool& committed, int lane) {
    // COHERENT: the records were written by other SMs inside this kernel (waves): read them at L2
    auto ld_dw = [&](int i, int p) { const long long* q = &pb.rec_dw[(size_t)i * KP + p]; return COHERENT ? __ldcg(q) : *q; };
    auto ld_pn = [&](int i, int p) { const int2* q = &pb.rec_pn[(size_t)i * KP + p]; return COHERENT ? __ldcg(q) : *q; };
    const int n_prof = __popc(profs);
    for (int base = a; base < b; base += 32) {
        const int i = base + lane;
        const bool on = i < b && __ldcg(reinterpret_cast<const unsigned char*>(pb.cand_flag) + i) != 0;
        const int fl = on ? (COHERENT ? __ldcg(&pb.rec_fl[i]) : pb.rec_fl[i]) : 0;
        const bool added = (fl >> 2) > 0, upd = (fl & 2) != 0;
        const unsigned cb = __ballot_sync(FULL, added);
        const bool comm_before = committed || (cb & ((1u << lane) - 1)) != 0;
        // one profile: the node's change of D[p] under the Commit / Revert rules, its inclusive scan over the 32 nodes, and
        // whether some attempt of the node would have failed the gate
        auto step = [&](int p, long long dw, int2 pn, long long& incl, bool& bad) {
            const long long g = upd ? dw : 0;                                  // Allocatable of the forked entry minus the snapshot's
            const long long delta = !on ? 0 : added ? g - pn.x : (comm_before ? 0 : dw);
            incl = delta;
#pragma unroll
            for (int o = 1; o < 32; o <<= 1) { const long long t = __shfl_up_sync(FULL, incl, o); if (lane >= o) incl += t; }
            const long long Dp = __shfl_sync(FULL, D, p);
            bad = on && pn.y > 0 && Dp + incl - delta + g - pn.y < 0;
        };
        auto apply = [&](int p, long long incl, int v) {
            const long long tot = v == 0 ? 0 : __shfl_sync(FULL, incl, v - 1);
            if (lane == p) D += tot;
        };
        int v = 32;
        if (n_prof <= 4) {                                      // the usual case: every load of the step in flight at once
            int pp[4]; long long dw[4], incl[4]; int2 pn[4];
            unsigned m = profs;
#pragma unroll
            for (int k = 0; k < 4; ++k) { pp[k] = m ? __ffs(m) - 1 : -1; m &= m - 1; }
#pragma unroll
            for (int k = 0; k < 4; ++k) { const bool ld = on && pp[k] >= 0; dw[k] = ld ? ld_dw(i, pp[k]) : 0; pn[k] = ld ? ld_pn(i, pp[k]) : make_int2(0, 0); }
            bool viol = false;
#pragma unroll
            for (int k = 0; k < 4; ++k) if (pp[k] >= 0) { bool bad; step(pp[k], dw[k], pn[k], incl[k], bad); viol = viol || bad; }
            const unsigned vm = __ballot_sync(FULL, viol);
            v = vm ? __ffs(vm) - 1 : 32;
#pragma unroll
            for (int k = 0; k < 4; ++k) if (pp[k] >= 0) apply(pp[k], incl[k], v);
        } else {
            bool viol = false;
            for (unsigned m = profs; m; m &= m - 1) {
                const int p = __ffs(m) - 1;
                long long incl; bool bad;
                step(p, on ? ld_dw(i, p) : 0, on ? ld_pn(i, p) : make_int2(0, 0), incl, bad);
                viol = viol || bad;
            }
            const unsigned vm = __ballot_sync(FULL, viol);
            v = vm ? __ffs(vm) - 1 : 32;
            for (unsigned m = profs; m; m &= m - 1) {
                const int p = __ffs(m) - 1;
                long long incl; bool bad;
                step(p, on ? ld_dw(i, p) : 0, on ? ld_pn(i, p) : make_int2(0, 0), incl, bad);
                apply(p, incl, v);
            }
        }
        committed = committed || (cb & (v == 32 ? FULL : ((1u << v) - 1))) != 0;
        if (v < 32) return base + v;
    }
    return b;
}

// ---- the same, two levels.  gate_chunks (every warp of the grid, one chunk of 32 nodes at a time) condenses the records of a
// chunk, per requested profile, into the chunk's total change of D and the smallest slack any attempt of its nodes leaves
// relative to D at the chunk's start -- once for a Commit having happened before the chunk (1) and once for none (0: nodes in
// front of the chunk's first adding node then leave their re-partitioned self behind, planner.go:144-149 with the aliasing of
// DESIGN.md section 4).  chunk_scan (one warp, lane = chunk) walks 32 chunks per step exactly as gate_scan walks nodes and
// stops at the first chunk with a node that would not have passed; gate_find puts the two together.
constexpr long long GATE_INF = (1ll << 62);

__device__ __forceinline__ long long warp_min_ll(long long v) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) { const long long t = __shfl_xor_sync(FULL, v, o); v = t < v ? t : v; }
    return v;
}

template <bool COHERENT>
__device__ __forceinline__ void gate_chunks(const DevProblem& pb, int a, int b, unsigned profs, int lane, int gw, int total_warps) {
    const int n_chunks = (b - a + 31) >> 5;
    for (int c = gw; c < n_chunks; c += total_warps) {
        const int i = a + 32 * c + lane;
        const bool on = i < b && __ldcg(reinterpret_cast<const unsigned char*>(pb.cand_flag) + i) != 0;
        const int fl = on ? (COHERENT ? __ldcg(&pb.rec_fl[i]) : pb.rec_fl[i]) : 0;
        const bool added = (fl >> 2) > 0, upd = (fl & 2) != 0;
        const unsigned cb = __ballot_sync(FULL, added);
        const bool before_first = (cb & ((1u << lane) - 1)) == 0;      // no node of the chunk in front of this one added a pod
        for (unsigned m = profs; m; m &= m - 1) {
            const int p = __ffs(m) - 1;
            long long dw = 0; int2 pn = make_int2(0, 0);
            if (on) {
                const long long* qd = &pb.rec_dw[(size_t)i * KP + p]; const int2* qp = &pb.rec_pn[(size_t)i * KP + p];
                dw = COHERENT ? __ldcg(qd) : *qd; pn = COHERENT ? __ldcg(qp) : *qp;
            }
            const long long g = upd ? dw : 0;
            const long long d1 = !on ? 0 : added ? g - pn.x : 0;
            const long long d0 = !on ? 0 : added ? g - pn.x : (before_first ? dw : 0);
            long long i0 = d0, i1 = d1;
#pragma unroll
            for (int o = 1; o < 32; o <<= 1) {
                const long long t0 = __shfl_up_sync(FULL, i0, o), t1 = __shfl_up_sync(FULL, i1, o);
                if (lane >= o) { i0 += t0; i1 += t1; }
            }
            const bool chk = on && pn.y > 0;
            const long long m0 = warp_min_ll(chk ? i0 - d0 + g - pn.y : GATE_INF), m1 = warp_min_ll(chk ? i1 - d1 + g - pn.y : GATE_INF);
            if (lane == 31) {
                long long* q = pb.rec_chunk + ((size_t)c * KP + p) * 4;
                q[0] = i0; q[1] = i1; q[2] = m0; q[3] = m1;
            }
        }
        if (lane == 0) pb.rec_chunk_fl[c] = cb != 0;
    }
}

// chunks [c_lo, c_hi), at most 32: first chunk with a node that fails the gate (c_hi: none); D and `committed` advanced to its start
__device__ __forceinline__ int chunk_scan(const DevProblem& pb, int c_lo, int c_hi, unsigned profs, long long& D, bool& committed, int lane) {
    const int c = c_lo + lane;
    const bool on = c < c_hi;
    const unsigned cb = __ballot_sync(FULL, on && __ldcg(&pb.rec_chunk_fl[on ? c : c_lo]) != 0);
    const bool comm_before = committed || (cb & ((1u << lane) - 1)) != 0;
    auto load = [&](int p, long long& T, long long& M) {
        T = 0; M = GATE_INF;
        if (on) {
            const long long* q = pb.rec_chunk + ((size_t)c * KP + p) * 4;
            T = __ldcg(q + (comm_before ? 1 : 0)); M = __ldcg(q + (comm_before ? 3 : 2));
        }
    };
    auto step = [&](int p, long long T, long long M, long long& incl, bool& bad) {
        incl = T;
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) { const long long t = __shfl_up_sync(FULL, incl, o); if (lane >= o) incl += t; }
        const long long Dp = __shfl_sync(FULL, D, p);
        bad = on && M < GATE_INF && Dp + incl - T + M < 0;
    };
    auto apply = [&](int p, long long incl, int u) {
        const long long tot = u == 0 ? 0 : __shfl_sync(FULL, incl, u - 1);
        if (lane == p) D += tot;
    };
    int u = 32;
    if (__popc(profs) <= 4) {                                   // the usual case: every load of the step in flight at once
        int pp[4]; long long T[4], M[4], incl[4];
        unsigned m = profs;
#pragma unroll
        for (int k = 0; k < 4; ++k) { pp[k] = m ? __ffs(m) - 1 : -1; m &= m - 1; }
#pragma unroll
        for (int k = 0; k < 4; ++k) { T[k] = 0; M[k] = GATE_INF; if (pp[k] >= 0) load(pp[k], T[k], M[k]); }
        bool viol = false;
#pragma unroll
        for (int k = 0; k < 4; ++k) if (pp[k] >= 0) { bool bad; step(pp[k], T[k], M[k], incl[k], bad); viol = viol || bad; }
        const unsigned vm = __ballot_sync(FULL, viol);
        u = vm ? __ffs(vm) - 1 : 32;
#pragma unroll
        for (int k = 0; k < 4; ++k) if (pp[k] >= 0) apply(pp[k], incl[k], u);
    } else {
        bool viol = false;
        for (unsigned m = profs; m; m &= m - 1) {
            const int p = __ffs(m) - 1;
            long long T, M, incl; bool bad;
            load(p, T, M); step(p, T, M, incl, bad);
            viol = viol || bad;
        }
        const unsigned vm = __ballot_sync(FULL, viol);
        u = vm ? __ffs(vm) - 1 : 32;
        for (unsigned m = profs; m; m &= m - 1) {
            const int p = __ffs(m) - 1;
            long long T, M, incl; bool bad;
            load(p, T, M); step(p, T, M, incl, bad);
            apply(p, incl, u);
        }
    }
    committed = committed || (cb & (u == 32 ? FULL : ((1u << u) - 1))) != 0;
    return min(c_lo + u, c_hi);
}

// first node of [from, b) that fails the gate (b: none), given the chunk summaries of [a, b) (chunks count from a); D and
// `committed` are those in front of `from` on entry and in front of the returned node on exit
template <bool COHERENT>
__device__ __forceinline__ int gate_find(const DevProblem& pb, int a, int b, int from, unsigned profs, long long& D, bool& committed, int lane) {
    const int n_chunks = (b - a + 31) >> 5;
    while (from < b) {
        const int c = (from - a) >> 5;
        if ((from - a) & 31) {                                  // the rest of a chunk, node by node
            const int end = min(b, a + 32 * (c + 1));
            const int v = gate_scan<COHERENT>(pb, from, end, profs, D, committed, lane);
            if (v < end) return v;
            from = end;
            continue;
        }
        const int c_hi = min(c + 32, n_chunks);
        const int u = chunk_scan(pb, c, c_hi, profs, D, committed, lane);
        from = min(b, a + 32 * u);
        if (u < c_hi) {
            const int end = min(b, from + 32);
            const int v = gate_scan<COHERENT>(pb, from, end, profs, D, committed, lane);
            if (v < end) return v;
            from = end;
        }
    }
    return b;
}

// profiles some pending pod requests: the only slice columns the gate ever looks at
__device__ __forceinline__ unsigned requested_profiles(const DevProblem& pb, int lane) {
    unsigned m = 0;
    for (int r = lane; r < pb.n_runs; r += 32) {
        const RunRec& rr = pb.runs[r];
        if (rr.n_ent == 1) m |= 1u << rr.p_single;
        else if (rr.n_ent > 1)
            for (int p = 0; p < KP; ++p) if (rr.q[p]) m |= 1u << p;
    }
    return __reduce_or_sync(FULL, m);
}

}  // namespace

// ------------------------------------------------------------------------------------------------------------
// node-parallel pass: grid-stride, one warp per candidate node
#ifndef PAR_MIN_BLOCKS
#define PAR_MIN_BLOCKS 6
#endif
__global__ void __launch_bounds__(128, PAR_MIN_BLOCKS) k_plan_parallel(DevProblem pb) {
    __shared__ Tables tb;
    __shared__ WarpScratch ws_all[4];
    __shared__ RunRec s_runs[RUN_CACHE];
    __shared__ int s_choff[RUN_CACHE];
    __shared__ unsigned long long s_masks[KP + 1];
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    cudaTriggerProgrammaticLaunchCompletion();
    // prologue on uploaded data only; the results of k_prepass / k_tracker_init are read after the dependency wait
    load_tables(pb, tb, tid, blockDim.x);
    PodCache pc;
    pc.cpu = nullptr; pc.mem = nullptr; pc.eph = nullptr; pc.key = nullptr; pc.flag = nullptr;
    pc.runs = s_runs; pc.chunk_off = s_choff; pc.n_cached = min(pb.n_runs, RUN_CACHE); pc.prof_runs = s_masks;
    for (int r = tid; r < pc.n_cached; r += blockDim.x) { s_runs[r] = pb.runs[r]; s_choff[r] = pb.run_chunk_off[r]; }
    __syncthreads();
    build_run_masks(s_runs, pc.n_cached, s_masks, tid, blockDim.x);
    cudaGridDependencySynchronize();
    const int lack0 = (lane < KP) ? pb.lack[lane] : 0;
    // single GPU, some node's accounts disagree with its GPUs or a pod class asks for several profiles (GPU.AddPod may then fail
    // half-way and leave devices moved, which the gate would have prevented): the pass runs without the gate and leaves records;
    // k_plan_waves replays the gate over them in name order (mode 3) and processes again the few nodes where it matters
    const bool gate_spec = pb.verify_inline && !pb.xf_on && (pb.gflags[0] != 0 || !pb.fast_eligible);
    const bool applicable = fast_applicable(pb, lack0, lane, gate_spec);
    if (!applicable) {
        if (pb.verify_inline && blockIdx.x == 0 && tid == 0) pb.wave[0] = 2;     // k_plan_waves takes the plan from here
        if (!pb.xf_on) return;
    }

    WarpScratch& ws = ws_all[warp];
    int lack = lack0 > 0 ? SATURATED : 0;
    long long D = 0, st_evals = 0, st_place = 0, visited = 0;
    bool committed = true, upd_unused = false;
    int dirty_unused = 0;
    // dynamic distribution: nodes differ a lot in cost (how many runs are feasible), so warps pull the next node.  Pulling
    // ahead of need (to hide the counter's round trip or to prefetch rows) was measured slower: with ~3 nodes per warp the
    // reserved nodes wait behind a busy warp while other warps idle.
    while (applicable) {
        int i = 0;
        if (lane == 0) i = pb.node_lo + atomicAdd(pb.next_node, 1);
        i = __shfl_sync(FULL, i, 0);
        if (i >= pb.node_hi) break;
        if (!pb.cand_flag[i]) continue;
        NodeRegs nr;
        load_node(pb, i, lane, nr);
        store_node(ws.nd, nr, i, lane);
        __syncwarp();
        ++visited;
        const int added = process_node<M_PAR>(pb, ws, tb, &pc, pb.placed, lack, D, committed, st_evals, st_place, lane, dirty_unused, upd_unused);
        if (gate_spec) write_gate_record(pb, ws, lane, added, upd_unused);
    }
    if (lane == 0 && visited) {
        atomicAdd(&pb.acc[0], (unsigned long long)st_place);
        atomicAdd(&pb.acc[1], (unsigned long long)st_evals);
        atomicAdd(&pb.acc[2], (unsigned long long)visited);
    }
    // single GPU: the CTA that finishes last verifies the saturation hypothesis on the final lacking vector (what
    // k_plan_sweep does for the sharded drivers) and either publishes the counters or hands the plan to k_plan_waves
    if (pb.verify_inline && applicable && last_block_done(&pb.next_node[2]) && warp == 0) {
        if (gate_spec) { if (lane == 0) pb.wave[0] = 3; return; }
        const int le = lack0 - ((lane < KP) ? __ldcg(&pb.lack_dec[lane]) : 0);
        const bool ok = fast_applicable(pb, lack0 > 0 ? le : 0, lane) && __all_sync(FULL, !(lane < KP && lack0 > 0) || le > 0);
        const int lack_start = __popc(__ballot_sync(FULL, lack0 > 0)), lack_end = __popc(__ballot_sync(FULL, le > 0));
        if (lane == 0) {
            if (ok) {
                DevStats* st = pb.stats;
                const long long vis = (long long)__ldcg(&pb.acc[2]);
                st->visited_nodes = vis;
                st->candidates = vis * pb.P;
                st->geometry_evals = (long long)__ldcg(&pb.acc[1]);
                st->placements = (long long)__ldcg(&pb.acc[0]);
                st->candidate_nodes = vis;
                st->lacking_at_start = lack_start;
                st->lacking_at_end = lack_end;
                st->path = 1;
            } else pb.wave[0] = 2;
        }
    }
    // slice mode, fused exchange 2: the CTA that finishes last pushes this rank's counters and placed-key bitmap (pb.acc and
    // pb.placed point into the exchange block) into every rank's mailbox -- also when the pass did not apply (all zero)
    if (pb.xf_on && last_block_done(&pb.next_node[2])) peer_push(pb.xf_box, 2, pb.xf_epoch, pb.xf_exchange, pb.xf_xb);
}

// verification of the node-parallel pass, or the sequential sweep
__global__ void __launch_bounds__(256, 1) k_plan_sweep(DevProblem pb, int32_t* cand_idx) {
    extern __shared__ __align__(16) unsigned char s_dyn[];
    __shared__ Tables tb;
    __shared__ WarpScratch ws2[2];
    __shared__ RunRec s_runs[RUN_CACHE];
    __shared__ int s_choff[RUN_CACHE];
    __shared__ int s_scan[256];
    __shared__ unsigned long long s_masks[KP + 1];
    __shared__ int s_ncand, s_fast_ok;
    __shared__ int s_lackend[KP];

    typedef long long (*chunk_ll)[32];
    typedef int (*chunk_i)[32];
    typedef unsigned (*chunk_u)[32];
    PodCache pc;
    pc.cpu = reinterpret_cast<chunk_ll>(s_dyn);
    pc.mem = pc.cpu + RUN_CACHE;
    pc.eph = pc.mem + RUN_CACHE;
    pc.key = reinterpret_cast<chunk_i>(pc.eph + RUN_CACHE);
    pc.flag = reinterpret_cast<chunk_u>(pc.key + RUN_CACHE);
    pc.runs = s_runs;
    pc.chunk_off = s_choff;
    pc.n_cached = min(pb.n_runs, RUN_CACHE);
    pc.prof_runs = s_masks;
    unsigned* s_removed = reinterpret_cast<unsigned*>(pc.flag + RUN_CACHE);      // [(n_keys+31)/32]

    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    cudaGridDependencySynchronize();
    if (pb.xf_on) {
        // slice mode, fused exchange 2: wait for every rank's block in this rank's mailbox; OR the bitmaps, sum the counters
        const int par = pb.xf_epoch & 1, W = pb.xf_box.world;
        const unsigned char* me = pb.xf_box.base[pb.xf_box.rank];
        if (!peer_wait(PeerWait{reinterpret_cast<const unsigned*>(me + box_flags(2, par)), W, pb.xf_epoch}) && tid == 0) pb.gflags[3] = 1;
        const int words = (pb.n_keys + 31) / 32;
        for (int x = tid; x < words; x += blockDim.x) {
            unsigned w = 0;
            for (int r = 0; r < W; ++r) w |= __ldcg(reinterpret_cast<const unsigned*>(me + box_x(par, r, pb.xf_box.xb_max) + 32) + x);
            pb.placed[x] = w;
        }
        if (tid < 4) {
            unsigned long long a = 0;
            for (int r = 0; r < W; ++r) a += __ldcg(reinterpret_cast<const unsigned long long*>(me + box_x(par, r, pb.xf_box.xb_max)) + tid);
            pb.acc[tid] = a;
        }
        __threadfence_block();
        __syncthreads();
    }
    if (pb.gflags[3]) {                  // a peer never delivered its block: nothing below would be meaningful
        if (tid == 0) { pb.stats->path = 0; pb.stats->pad = 2; }
        return;
    }
    const ChainHead* cin = pb.chain_mode ? reinterpret_cast<const ChainHead*>(pb.chain_in) : nullptr;
    const int lack0 = (lane < KP) ? (cin ? cin->lack[lane] : pb.lack[lane]) : 0;
    const int lack_start = cin ? cin->lack_start : __popc(__ballot_sync(FULL, lack0 > 0));

    // ---------------- did the node-parallel pass run, and did its hypothesis hold to the end?
    if (!pb.chain_mode && fast_applicable(pb, lack0, lane)) {
        if (tid < KP) s_lackend[tid] = pb.lack[tid] - (pb.whole_snapshot ? pb.lack_dec[tid] : 0);
        __syncthreads();
        for (int k = tid; !pb.whole_snapshot && k < pb.n_keys; k += blockDim.x)
            if (pb.placed[k >> 5] >> (k & 31) & 1)
                for (int p = 0; p < pb.K; ++p) {
                    const int e = pb.entry_lack[(size_t)k * KP + p];
                    if (e > 0) atomicSub(&s_lackend[p], e);
                }
        __syncthreads();
        if (warp == 0) {
            const int le = (lane < KP) ? s_lackend[lane] : 0;
            const bool ok = fast_applicable(pb, lack0 > 0 ? le : 0, lane) &&
                            __all_sync(FULL, !(lane < KP && lack0 > 0) || le > 0);
            if (lane == 0) s_fast_ok = ok;
            if (ok) {
                const int lack_end = __popc(__ballot_sync(FULL, le > 0));
                if (lane == 0) {
                    DevStats* st = pb.stats;
                    st->visited_nodes = (long long)pb.acc[2];
                    st->candidates = (long long)pb.acc[2] * pb.P;
                    st->geometry_evals = (long long)pb.acc[1];
                    st->placements = (long long)pb.acc[0];
                    st->candidate_nodes = (long long)pb.acc[2];
                    st->lacking_at_start = lack_start;
                    st->lacking_at_end = lack_end;
                    st->path = 1;
                }
            }
        }
        __syncthreads();
        if (s_fast_ok) return;
        // hypothesis failed: discard the speculative plan (restore GetPartitioningState) and fall through
        for (int x = tid; x < pb.N * pb.G; x += blockDim.x) {
            const int i = x / pb.G, j = x - i * pb.G;
            if (j < pb.node_gpu_count[i]) {
                pb.out_count[x] = add_bytes(pb.gpu_used[x], pb.gpu_free[x]);
                pb.out_mask[x] = pb.gpu_umask[x] | pb.gpu_fmask[x];
            }
        }
        __syncthreads();
    }

    // The sweep must follow the EXACT lacking vector.  It changes only when a node places a pod key nobody placed before
    // (at most sum(lack) times), and between two changes the nodes are independent of each other: k_plan_waves processes
    // them in parallel waves.  That needs what the node-parallel pass needs except saturation -- single-profile classes and
    // consistent node accounts, which make the gate vacuous -- and pays off when few changes are possible.
    if (!pb.slice_mode && pb.fast_eligible && pb.gflags[0] == 0 && pb.N > 0 && lack_start > 0) {
        long long total = 0;
        for (int p = 0; p < KP; ++p) total += max(0, __shfl_sync(FULL, lack0, p));
        if (total * 45 < (long long)pb.N * 17) {            // ~45 us per change against ~17 us per node of the chain below
            for (int w = tid; w < (pb.n_keys + 31) / 32; w += blockDim.x) pb.placed[w] = 0;
            if (tid < 4) pb.acc[tid] = 0;
            if (tid == 0) { pb.stats->lacking_at_start = lack_start; pb.wave[0] = 1; }
            return;
        }
    }
    if (pb.slice_mode && !pb.chain_mode) {
        // this rank holds only its slice of the node list: the sequential chain cannot run here.  Report it (path 0, no
        // visits); the caller plans this problem unsharded (nosplan_download_sliced returns NOSPLAN_SLICED_NOT_APPLICABLE).
        if (tid == 0) {
            DevStats* st = pb.stats;
            st->visited_nodes = 0; st->candidates = 0; st->geometry_evals = 0; st->placements = 0; st->candidate_nodes = 0;
            st->lacking_at_start = lack_start; st->lacking_at_end = lack_start; st->path = 0;
            st->pad = lack_start > 0;       // nothing lacking: the unchanged state IS the plan (planner.go:80-83)
        }
        return;
    }
    // ---------------- sequential sweep.  prologue (whole CTA): tables, run cache, candidate compaction in name order
    load_tables(pb, tb, tid, blockDim.x);
    for (int w = tid; w < (pb.n_keys + 31) / 32; w += blockDim.x)
        s_removed[w] = cin ? reinterpret_cast<const unsigned*>(pb.chain_in + sizeof(ChainHead))[w] : 0u;
    for (int r = tid; r < pc.n_cached; r += blockDim.x) { s_runs[r] = pb.runs[r]; s_choff[r] = pb.run_chunk_off[r]; }
    __syncthreads();
    build_run_masks(s_runs, pc.n_cached, s_masks, tid, blockDim.x);
    for (int t = tid; t < pc.n_cached * 32; t += blockDim.x) {
        const int r = t >> 5, l = t & 31;
        if (l < s_runs[r].len) {
            const int s = s_runs[r].start + l;
            pc.cpu[r][l] = pb.pod_cpu[s]; pc.mem[r][l] = pb.pod_mem[s]; pc.eph[r][l] = pb.pod_eph[s];
            pc.key[r][l] = pb.pod_key[s]; pc.flag[r][l] = pb.pod_flags[s];
        }
    }
    // ordered compaction of candidate flags -> cand_idx (GetCandidateNodes, core/snapshot.go:119-130): 1024 nodes per step, four
    // flag bytes per thread in one load, warp scan + eight warp totals.  Not reached on the reference's early return
    // (nothing lacking, planner.go:80-83), where the list is never asked for.
    if (lack_start > 0) {
        int base = 0;
        int* s_wsum = s_scan;
        for (int c0 = 0; c0 < pb.N; c0 += 4 * (int)blockDim.x) {
            const int i0 = c0 + 4 * tid;
            unsigned f4 = 0;
            if (i0 + 3 < pb.N) f4 = *reinterpret_cast<const unsigned*>(pb.cand_flag + i0);
            else for (int k = 0; k < 4; ++k) if (i0 + k < pb.N && pb.cand_flag[i0 + k]) f4 |= 1u << (8 * k);
            const int cnt = ((f4 & 0xffu) != 0) + ((f4 & 0xff00u) != 0) + ((f4 & 0xff0000u) != 0) + ((f4 & 0xff000000u) != 0);
            int incl = cnt;
#pragma unroll
            for (int o = 1; o < 32; o <<= 1) { const int t = __shfl_up_sync(FULL, incl, o); if (lane >= o) incl += t; }
            if (lane == 31) s_wsum[warp] = incl;
            __syncthreads();
            int woff = 0, total = 0;
            for (int w = 0; w < (int)(blockDim.x >> 5); ++w) { const int t = s_wsum[w]; if (w < warp) woff += t; total += t; }
            int pos = base + woff + incl - cnt;
#pragma unroll
            for (int k = 0; k < 4; ++k) if (f4 >> (8 * k) & 0xffu) cand_idx[pos++] = i0 + k;
            base += total;
            __syncthreads();
        }
        if (tid == 0) s_ncand = base;
    } else if (tid == 0) s_ncand = 0;
    __syncthreads();
    if (warp != 0) return;

    // ---------------- the chain (warp 0)
    const int n_cand = s_ncand;
    int lack = lack0;                                              // lane p: lackingSlices[p]
    long long D = (lane < KP) ? (cin ? cin->D[lane] : pb.tot[lane]) : 0;   // lane p: sum(alloc - requested) of slice column p
    bool committed = cin ? cin->committed != 0 : false, upd_unused = false;
    int dirty_unused = 0;
    long long st_visited = cin ? cin->counters[0] : 0, st_evals = cin ? cin->counters[1] : 0, st_place = cin ? cin->counters[2] : 0;
    const long long cand_before = cin ? cin->counters[3] : 0;

    NodeRegs nr;
    if (n_cand > 0 && lack_start > 0) {
        load_node(pb, cand_idx[0], lane, nr);
        store_node(ws2[0].nd, nr, cand_idx[0], lane);
    }
    __syncwarp();
    for (int c = 0; c < n_cand && lack_start > 0; ++c) {
        if (__ballot_sync(FULL, lack > 0) == 0) break;             // planner.go:94-97
        const bool have_next = c + 1 < n_cand;
        int next_i = 0;
        if (have_next) { next_i = cand_idx[c + 1]; load_node(pb, next_i, lane, nr); }   // prefetch, consumed below
        ++st_visited;
        process_node<M_SEQ>(pb, ws2[c & 1], tb, &pc, s_removed, lack, D, committed, st_evals, st_place, lane, dirty_unused, upd_unused);
        if (have_next) store_node(ws2[(c + 1) & 1].nd, nr, next_i, lane);
        __syncwarp();
    }
    const int lack_end = __popc(__ballot_sync(FULL, lack > 0));
    if (lane == 0) {
        DevStats* st = pb.stats;
        st->visited_nodes = st_visited;
        st->candidates = st_visited * pb.P;
        st->geometry_evals = st_evals;
        st->placements = st_place;
        st->candidate_nodes = cand_before + (lack_start > 0 ? n_cand : 0);   // GetCandidateNodes is not reached on the early return (planner.go:80-83)
        st->lacking_at_start = lack_start;
        st->lacking_at_end = lack_end;
        st->path = 0;
        if (pb.chain_mode) st->pad = 0;
    }
    if (pb.chain_mode && pb.chain_out) {                              // what the next rank starts from
        ChainHead* co = reinterpret_cast<ChainHead*>(pb.chain_out);
        if (lane < 16) { co->lack[lane] = lane < KP ? lack : 0; co->D[lane] = lane < KP ? D : 0; }
        if (lane == 0) {
            co->counters[0] = st_visited; co->counters[1] = st_evals; co->counters[2] = st_place;
            co->counters[3] = cand_before + (lack_start > 0 ? n_cand : 0);
            co->committed = committed ? 1 : 0; co->lack_start = lack_start; co->pad[0] = co->pad[1] = 0;
        }
        __syncwarp();
        unsigned* ro = reinterpret_cast<unsigned*>(pb.chain_out + sizeof(ChainHead));
        for (int w = lane; w < (pb.n_keys + 31) / 32; w += 32) ro[w] = s_removed[w];
    }
}

constexpr size_t SWEEP_DYN_MAX = 160 * 1024;
constexpr size_t SWEEP_CHUNK_BYTES = (size_t)RUN_CACHE * 32 * (8 + 8 + 8 + 4 + 4);

// ------------------------------------------------------------------------------------------------------------
// The sweep as parallel waves (cooperative launch, one warp per node, grid = one resident wave of CTAs).
//   wave over nodes [a, a + warps):  phase 1  every warp processes its node against the exact lacking vector without
//                                             writing anything (M_SPEC) and reports whether the node would change the vector;
//                                    grid sync
//                                    phase 2  f = first such node.  Nodes a..f saw the vector they would have seen in the
//                                             sequential sweep: the clean ones write their rows, node f is processed again
//                                             on the real state (M_COMMIT: removed-set and vector updated); nodes > f are
//                                             dropped and start the next wave.
//                                    grid sync
// Exact by construction: a node's decisions depend on earlier nodes only through the lacking vector (the gate is vacuous
// for consistent node accounts, see fast_applicable), and the reference's loop exit (planner.go:94-97) is the wave loop's.
constexpr int WAVE_PRELUDE = 3;            // candidates the chain takes before the first wave
__global__ void __launch_bounds__(128, 4) k_plan_waves(DevProblem pb) {
    __shared__ Tables tb;
    __shared__ WarpScratch ws_all[4];
    __shared__ RunRec s_runs[RUN_CACHE];
    __shared__ int s_choff[RUN_CACHE];
    __shared__ unsigned long long s_masks[KP + 1];
    __shared__ int s_cnt[4];
    const int mode = __ldcg(&pb.wave[0]);
    if (mode == 0) return;                                    // the node-parallel pass stands (same verdict in every CTA)
    cooperative_groups::grid_group grid = cooperative_groups::this_grid();
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    load_tables(pb, tb, tid, blockDim.x);
    PodCache pc;
    pc.cpu = nullptr; pc.mem = nullptr; pc.eph = nullptr; pc.key = nullptr; pc.flag = nullptr;
    pc.runs = s_runs; pc.chunk_off = s_choff; pc.n_cached = min(pb.n_runs, RUN_CACHE); pc.prof_runs = s_masks;
    for (int r = tid; r < pc.n_cached; r += blockDim.x) { s_runs[r] = pb.runs[r]; s_choff[r] = pb.run_chunk_off[r]; }
    __syncthreads();
    build_run_masks(s_runs, pc.n_cached, s_masks, tid, blockDim.x);
    const int lack0 = (lane < KP) ? pb.lack[lane] : 0;        // the vector at t0: nothing on this path has touched it yet
    const int lack_start = __popc(__ballot_sync(FULL, lack0 > 0));
    bool as_waves = true;
    // some node's accounts disagree with its GPUs, or a pod class asks for several profiles: the cluster-wide gate is not implied
    // by GPU.AddPod, or has a say in what a failed GPU.AddPod leaves behind; it is replayed over the records the gate-less passes
    // leave (gate_scan)
    const bool gate_on = pb.gflags[0] != 0 || !pb.fast_eligible;
    int chain_cands = -1;                                      // >= 0: mode 3 fell through to the chain with this many candidates
    if (mode == 3) {
        // ---- the node-parallel pass ran without the gate.  One warp replays the gate in name order; a node whose pods would
        //      not all have passed it is processed again, with the gate, on the exact running totals.
        const unsigned profs = requested_profiles(pb, lane);
        gate_chunks<false>(pb, 0, pb.N, profs, lane, blockIdx.x * (blockDim.x >> 5) + warp, gridDim.x * (blockDim.x >> 5));
        grid.sync();
        if (blockIdx.x != 0 || warp != 0) return;
        WarpScratch& cw = ws_all[0];
        int lack = lack0 > 0 ? SATURATED : 0, dirty_unused = 0, again = 0;
        long long D = (lane < KP) ? pb.tot[lane] : 0, ev_unused = 0, pl_diff = 0;
        bool committed = false, upd_unused = false;
        for (int from = 0;;) {
            const int v = gate_find<false>(pb, 0, pb.N, from, profs, D, committed, lane);
            if (v >= pb.N) break;
            const int spec_added = pb.rec_fl[v] >> 2;
            NodeRegs nr;
            load_node(pb, v, lane, nr);
            store_node(cw.nd, nr, v, lane);
            if (lane < nr.ng) {                                 // GetPartitioningState again: the gate-less pass may have written the row
                const int x = v * pb.G + lane;
                pb.out_count[x] = add_bytes(pb.gpu_used[x], pb.gpu_free[x]);
                pb.out_mask[x] = pb.gpu_umask[x] | pb.gpu_fmask[x];
            }
            __syncwarp();
            long long pl = 0;
            process_node<M_PAR, true>(pb, cw, tb, &pc, pb.placed, lack, D, committed, ev_unused, pl, lane, dirty_unused, upd_unused);
            pl_diff += pl - spec_added;
            ++again;
            from = v + 1;
        }
        // the saturation hypothesis, as the last CTA of k_plan_parallel checks it.  Keys only the gate-less run of a node placed
        // stay in the bitmap: the final vector can come out smaller than it is, never larger, so a plan that stands here stands.
        const int le = lack0 - ((lane < KP) ? __ldcg(&pb.lack_dec[lane]) : 0);
        const bool ok = fast_applicable(pb, lack0 > 0 ? le : 0, lane, true) && __all_sync(FULL, !(lane < KP && lack0 > 0) || le > 0);
        const int lack_end = __popc(__ballot_sync(FULL, le > 0));
        if (ok) {
            if (lane == 0) {
                DevStats* st = pb.stats;
                const long long vis = (long long)__ldcg(&pb.acc[2]);
                st->visited_nodes = vis; st->candidates = vis * pb.P; st->candidate_nodes = vis;
                st->geometry_evals = (long long)__ldcg(&pb.acc[1]);
                st->placements = (long long)__ldcg(&pb.acc[0]) + pl_diff;
                st->lacking_at_start = lack_start; st->lacking_at_end = lack_end;
                st->path = 1; st->pad = 0; st->gate_again = again + 1;
            }
            return;
        }
        // saturation lost on the way: drop the speculative plan; the chain below runs it from the untouched t0 state
        int c = 0;
        for (int i = lane; i < pb.N; i += 32) c += pb.cand_flag[i] != 0;
        chain_cands = __reduce_add_sync(FULL, c);
        for (int x = lane; x < pb.N * pb.G; x += 32) {
            const int i = x / pb.G, j = x - i * pb.G;
            if (j < pb.node_gpu_count[i]) {
                pb.out_count[x] = add_bytes(pb.gpu_used[x], pb.gpu_free[x]);
                pb.out_mask[x] = pb.gpu_umask[x] | pb.gpu_fmask[x];
            }
        }
        for (int w = lane; w < (pb.n_keys + 31) / 32; w += 32) pb.placed[w] = 0;
        if (lane == 0) pb.stats->lacking_at_start = lack_start;
        __syncwarp();
        as_waves = false;
    }
    if (mode == 2) {
        // the node-parallel pass was rejected or did not apply (single-GPU driver): drop whatever it wrote, then choose
        // between waves and the single-warp chain exactly as k_plan_sweep does for the sharded drivers
        const int gtid = blockIdx.x * blockDim.x + tid, gthreads = gridDim.x * blockDim.x;
        for (int x = gtid; x < pb.N * pb.G; x += gthreads) {
            const int i = x / pb.G, j = x - i * pb.G;
            if (j < pb.node_gpu_count[i]) {
                pb.out_count[x] = add_bytes(pb.gpu_used[x], pb.gpu_free[x]);
                pb.out_mask[x] = pb.gpu_umask[x] | pb.gpu_fmask[x];
            }
        }
        for (int w = gtid; w < (pb.n_keys + 31) / 32; w += gthreads) pb.placed[w] = 0;
        if (gtid < 4) pb.acc[gtid] = 0;
        long long total = 0;
        for (int p = 0; p < KP; ++p) total += max(0, __shfl_sync(FULL, lack0, p));
        as_waves = lack_start > 0 && total * 45 < (long long)pb.N * 17;
    }
    if (chain_cands < 0) {   // len(GetCandidateNodes()) for the stats; first-change slots start at "none"
        int c = 0;
        for (int i = blockIdx.x * blockDim.x + tid; i < pb.N; i += gridDim.x * blockDim.x) c += pb.cand_flag[i] != 0;
        c = __reduce_add_sync(FULL, c);
        if (lane == 0) s_cnt[warp] = c;
        __syncthreads();
        if (tid == 0) {
            atomicAdd(&pb.wave[3], s_cnt[0] + s_cnt[1] + s_cnt[2] + s_cnt[3]);
            if (blockIdx.x == 0) { pb.wave[1] = INT_MAX; pb.wave[2] = INT_MAX; pb.stats->lacking_at_start = lack_start; }
        }
        grid.sync();
    }

    if (!as_waves) {
        // ---- the chain: one warp, nodes in name order, exact lacking vector, gate and removed-set as in k_plan_sweep (the
        //      removed-set lives in global memory here)
        if (blockIdx.x != 0 || warp != 0) return;
        WarpScratch& cw = ws_all[0];
        int lack = lack0, dirty_unused = 0;
        long long D = (lane < KP) ? pb.tot[lane] : 0, vis = 0, ev = 0, pl = 0;
        bool committed = false, upd_unused = false;
        for (int i = 0; i < pb.N && lack_start > 0; ++i) {
            if (!pb.cand_flag[i]) continue;
            if (__ballot_sync(FULL, lack > 0) == 0) break;         // planner.go:94-97
            NodeRegs nr;
            load_node(pb, i, lane, nr);
            store_node(cw.nd, nr, i, lane);
            __syncwarp();
            ++vis;
            process_node<M_SEQ>(pb, cw, tb, &pc, pb.placed, lack, D, committed, ev, pl, lane, dirty_unused, upd_unused);
        }
        const int lack_end = __popc(__ballot_sync(FULL, lack > 0));
        if (lane == 0) {
            DevStats* st = pb.stats;
            st->visited_nodes = vis; st->candidates = vis * pb.P; st->geometry_evals = ev; st->placements = pl;
            st->candidate_nodes = lack_start > 0 ? (chain_cands >= 0 ? chain_cands : __ldcg(&pb.wave[3])) : 0;   // not reached on the early return (planner.go:80-83)
            st->lacking_at_start = lack_start; st->lacking_at_end = lack_end; st->path = 0;
        }
        return;
    }

    WarpScratch& ws = ws_all[warp];
    const int total_warps = gridDim.x * (blockDim.x >> 5), gw = blockIdx.x * (blockDim.x >> 5) + warp;
    long long st_evals = 0, st_place = 0, st_visited = 0, D = 0;
    bool committed = true;
    // ---- a short chain first: a handful of pending pods is usually satisfied by the first candidate or two, and a node of the
    //      chain costs less than a wave with its two grid barriers.  One warp takes the first WAVE_PRELUDE candidates in name
    //      order exactly as the chain does (removed-set in global memory) and leaves the vector, the counters and the node to
    //      go on from where the waves read them; if the vector is empty by then the wave loop ends before its first wave.
    if (gw == 0) {
        int lack = lack0, dirty_unused = 0, i = 0, seen = 0;
        long long Dc = (lane < KP) ? pb.tot[lane] : 0, vis = 0, ev = 0, pl = 0;
        bool comm = false, upd_unused = false;
        for (; i < pb.N && seen < WAVE_PRELUDE; ++i) {
            if (!pb.cand_flag[i]) continue;
            if (__ballot_sync(FULL, lack > 0) == 0) break;         // planner.go:94-97
            NodeRegs nr;
            load_node(pb, i, lane, nr);
            store_node(ws.nd, nr, i, lane);
            __syncwarp();
            ++vis; ++seen;
            process_node<M_SEQ>(pb, ws, tb, &pc, pb.placed, lack, Dc, comm, ev, pl, lane, dirty_unused, upd_unused);
        }
        if (lane < KP) { pb.lack[lane] = lack; pb.gate_state[lane] = Dc; }
        if (lane == 0) {
            pb.acc[0] += (unsigned long long)pl; pb.acc[1] += (unsigned long long)ev; pb.acc[2] += (unsigned long long)vis;
            pb.wave[4] = i;
            pb.gate_state[KP] = comm ? 1 : 0;
        }
        __threadfence();
    }
    grid.sync();
    int a = __ldcg(&pb.wave[4]), n_waves = 0, n_again = 0;
    const unsigned profs = gate_on ? requested_profiles(pb, lane) : 0u;
    for (;;) {
        const int lack = (lane < KP) ? __ldcg(&pb.lack[lane]) : 0;
        if (a >= pb.N || __ballot_sync(FULL, lack > 0) == 0) break;          // planner.go:94-97
        int32_t* first_change = &pb.wave[1 + (n_waves & 1)];
        // ---- phase 1
        const int i = a + gw;
        const bool have = i < pb.N && pb.cand_flag[i];
        int dirty = 0, added = 0, lk = lack;
        bool upd = false;
        long long ev = 0, pl = 0;
        NodeRegs nr;
        if (have) {
            load_node(pb, i, lane, nr);
            store_node(ws.nd, nr, i, lane);
            __syncwarp();
            added = process_node<M_SPEC>(pb, ws, tb, &pc, pb.placed, lk, D, committed, ev, pl, lane, dirty, upd);
            if (gate_on) write_gate_record(pb, ws, lane, added, upd);
            if (dirty && lane == 0) atomicMin(first_change, i);
        }
        grid.sync();
        // ---- phase 2
        int f = __ldcg(first_change);
        int last_valid = min(f, a + total_warps - 1);
        if (gate_on) {
            // the gate, replayed over the nodes in front of f (or over the whole wave): the first node whose pods would not all
            // have passed it ends the wave in f's place and is processed again with the gate, on the exact running totals
            const int scan_end = min(f <= last_valid ? f : last_valid + 1, pb.N);
            gate_chunks<true>(pb, a, scan_end, profs, lane, gw, total_warps);
            grid.sync();
            if (gw == 0) {
                long long Dg = (lane < KP) ? __ldcg(&pb.gate_state[lane]) : 0;
                bool cm = __ldcg(&pb.gate_state[KP]) != 0;
                const int g = gate_find<true>(pb, a, scan_end, a, profs, Dg, cm, lane);
                if (lane < KP) pb.gate_state[lane] = Dg;
                if (lane == 0) { pb.gate_state[KP] = cm ? 1 : 0; pb.gate_state[KP + 1] = g; }
                __threadfence();
            }
            grid.sync();
            const int g = (int)__ldcg(&pb.gate_state[KP + 1]);
            if (g < scan_end) { f = g; last_valid = g; if (i == g) ++n_again; }
        }
        if (have && i <= last_valid) {
            if (i == f) {
                store_node(ws.nd, nr, i, lane);                              // phase 1 moved devices around in the copy
                __syncwarp();
                ev = 0; pl = 0; lk = lack;
                int d2 = 0;
                if (gate_on) {
                    long long Dg = (lane < KP) ? __ldcg(&pb.gate_state[lane]) : 0;
                    bool cm = __ldcg(&pb.gate_state[KP]) != 0;
                    process_node<M_COMMIT, true>(pb, ws, tb, &pc, pb.placed, lk, Dg, cm, ev, pl, lane, d2, upd);
                    if (lane < KP) pb.gate_state[lane] = Dg;
                    if (lane == 0) pb.gate_state[KP] = cm ? 1 : 0;
                } else
                    process_node<M_COMMIT>(pb, ws, tb, &pc, pb.placed, lk, D, committed, ev, pl, lane, d2, upd);
                if (lane < KP) pb.lack[lane] = lk;
            } else if (added > 0 && upd && lane < ws.nd.ng) {                // partitioningState[n] = GetPartitioning(n) (planner.go:138)
                const int x = i * pb.G + lane;
                pb.out_count[x] = add_bytes(ws.nd.used[lane], ws.nd.free_[lane]);
                pb.out_mask[x] = ws.nd.umask[lane] | ws.nd.fmask[lane];
            }
            ++st_visited; st_evals += ev; st_place += pl;
        }
        if (gw == 0 && lane == 0) pb.wave[1 + ((n_waves + 1) & 1)] = INT_MAX; // the other slot, last read one wave ago
        a = last_valid + 1;
        ++n_waves;
        grid.sync();
    }
    if (lane == 0 && st_visited) {
        atomicAdd(&pb.acc[0], (unsigned long long)st_place);
        atomicAdd(&pb.acc[1], (unsigned long long)st_evals);
        atomicAdd(&pb.acc[2], (unsigned long long)st_visited);
        if (n_again) atomicAdd(&pb.stats->gate_again, n_again);
    }
    grid.sync();
    if (gw == 0) {
        const int le = (lane < KP) ? __ldcg(&pb.lack[lane]) : 0;
        const int lack_end = __popc(__ballot_sync(FULL, le > 0));
        if (lane == 0) {
            DevStats* st = pb.stats;
            st->visited_nodes = (long long)__ldcg(&pb.acc[2]);
            st->candidates = st->visited_nodes * pb.P;
            st->geometry_evals = (long long)__ldcg(&pb.acc[1]);
            st->placements = (long long)__ldcg(&pb.acc[0]);
            st->candidate_nodes = __ldcg(&pb.wave[3]);
            st->lacking_at_end = lack_end;
            st->path = 0;
            st->pad = n_waves;
            if (gate_on) atomicAdd(&st->gate_again, 1);       // count + 1: 0 means no replay took place
        }
    }
}

size_t sweep_dyn_smem(const DevProblem& pb) { return SWEEP_CHUNK_BYTES + (size_t)((pb.n_keys + 31) / 32) * 4 + 16; }

// largest number of tracker keys (distinct namespace/name of pending pods) the shared-memory bitmap can hold
int sweep_max_keys() { return (int)((SWEEP_DYN_MAX - SWEEP_CHUNK_BYTES - 16) / 4) * 32; }

// grid sizes of one resident wave on the CURRENT device (the handle's): kept in the handle, so that handles on GPUs with
// different SM counts, or under an SM-limited context, never share them
cudaError_t sweep_prepare_device(int* par_blocks, int* wave_blocks) {
    int dev = 0, sms = 148, per_sm = 4;
    if (cudaGetDevice(&dev) == cudaSuccess) cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
    if (cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, k_plan_parallel, 128, 0) != cudaSuccess || per_sm < 1) per_sm = 4;
    *par_blocks = sms * per_sm;                                // one resident wave, grid-stride over the nodes
    int per_sm_w = 4;
    if (cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm_w, k_plan_waves, 128, 0) != cudaSuccess || per_sm_w < 1) per_sm_w = 2;
    *wave_blocks = sms * per_sm_w;                             // cooperative launch: every CTA must be resident
    return cudaFuncSetAttribute(k_plan_sweep, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)SWEEP_DYN_MAX);
}

cudaError_t launch_parallel(const DevProblem& pb, int max_blocks, cudaStream_t st) {
    const bool work = (pb.fast_eligible || pb.verify_inline) && pb.N > 0 && pb.P > 0 && pb.node_hi > pb.node_lo;
    if (!work && !pb.xf_on) return cudaSuccess;            // fused slice mode: the kernel still pushes this rank's (empty) block
    int blocks = work ? (pb.node_hi - pb.node_lo + 3) / 4 : 1;
    if (blocks > max_blocks) blocks = max_blocks;
    return launch_pdl(k_plan_parallel, dim3(blocks), dim3(128), 0, st, pb);
}

// Node-sharded Plan(): every rank ran the node-parallel pass on its own node range and all-gathered its
// [acc (4 x u64) | placed bitmap] block; OR the bitmaps and sum the counters into this rank's work area.
__global__ void __launch_bounds__(256) k_shard_combine(DevProblem pb, const unsigned char* gathered, int world, size_t stride, PeerWait wait) {
    if (!peer_wait(wait) && threadIdx.x == 0) pb.gflags[3] = 1;
    const int words = (pb.n_keys + 31) / 32;
    const int x = blockIdx.x * blockDim.x + threadIdx.x;
    if (x < words) {
        unsigned w = 0;
        for (int r = 0; r < world; ++r) w |= reinterpret_cast<const unsigned*>(gathered + r * stride + 32)[x];
        pb.placed[x] = w;
    }
    if (x < 4) {
        unsigned long long a = 0;
        for (int r = 0; r < world; ++r) a += reinterpret_cast<const unsigned long long*>(gathered + r * stride)[x];
        pb.acc[x] = a;
    }
}

void launch_shard_combine(const DevProblem& pb, const void* gathered, int world, size_t stride, const PeerWait& wait, cudaStream_t st) {
    const int words = (pb.n_keys + 31) / 32;
    const int n = words > 4 ? words : 4;
    k_shard_combine<<<(n + 255) / 256, 256, 0, st>>>(pb, static_cast<const unsigned char*>(gathered), world, stride, wait);
}

// always queued behind k_plan_sweep for problems the waves can take; returns at once unless k_plan_sweep handed over
cudaError_t launch_waves(const DevProblem& pb, int max_blocks, cudaStream_t st) {
    if ((!pb.fast_eligible && !pb.verify_inline) || pb.N == 0 || pb.P == 0 || pb.slice_mode) return cudaSuccess;
    int blocks = (pb.N + 3) / 4;
    if (blocks > max_blocks) blocks = max_blocks;
    DevProblem arg = pb;
    void* args[] = {&arg};
    return cudaLaunchCooperativeKernel((const void*)k_plan_waves, dim3(blocks), dim3(128), args, 0, st);
}

cudaError_t launch_sweep(const DevProblem& pb, int32_t* cand_idx, cudaStream_t st) {
    const size_t dyn = sweep_dyn_smem(pb);
    if (dyn > SWEEP_DYN_MAX) return cudaErrorInvalidValue;
    return launch_pdl(k_plan_sweep, dim3(1), dim3(256), dyn, st, pb, cand_idx);
}

}  // namespace nsp
