// k_plan_sweep -- the sequential node sweep of planner.Plan (internal/partitioning/core/planner.go:92-150)
// as ONE persistent kernel: the loop-carried state (lacking-slices vector, cluster-wide free totals,
// tracker entries already removed, "a Commit happened") never leaves the SM.
//
// One warp walks the candidate nodes in name order.  Per node:
//   1. Node.UpdateGeometryFor (pkg/gpu/mig/node.go:145-177 / pkg/gpu/slicing/node.go:120-152): GPUs in order;
//      MIG: lanes = allowed geometries of the GPU model (table staged in shared memory), each lane scores its
//      geometry (pkg/gpu/mig/gpu.go:164-191), warp arg-max with the canonical tie-break (lowest table index),
//      ApplyGeometry by lanes = profiles.  MPS: lanes = profiles, closed-form create / delete / restore
//      (pkg/gpu/slicing/gpu.go:162-232).
//   2. the pod loop (planner.go:116-141 tryAddPod) over RUNS of identical slice requests instead of single pods.
//      Gate (core/snapshot.go:132-165), NodeResourcesFit on the slice columns and GPU.AddPod feasibility are
//      class-level and only ever get tighter while one node is processed, so a run whose class fails is skipped
//      whole, and inside a feasible run 32 pods are tested per step: warp prefix sums of cpu/memory/ephemeral
//      requests give every lane the node state it would see if all earlier lanes were placed; the first lane that
//      fails under that hypothesis splits the step, so the result is exactly the sequential first-fit.
//   3. Commit / Revert bookkeeping including the aliasing rules of the reference's object graph (see the state
//      rules in oracle/nos_oracle.cpp plan_incremental and DESIGN.md).
// The next candidate node is prefetched into registers while the current one is processed.
#include "nosplan_dev.cuh"

namespace nsp {

namespace {

constexpr unsigned FULL = 0xffffffffu;
constexpr long long CAP_INF = (1ll << 40);

struct SNode {
    uint4 used[NOSPLAN_MAX_GPUS];
    uint4 free_[NOSPLAN_MAX_GPUS];
    uint16_t umask[NOSPLAN_MAX_GPUS], fmask[NOSPLAN_MAX_GPUS];
    int model[NOSPLAN_MAX_GPUS], mem[NOSPLAN_MAX_GPUS];
    long long alloc[RMAX + 1], req[RMAX + 1];
    int ng, allowed, npods, idx;
};

struct NodeRegs {          // one node spread over the lanes of a warp (prefetch registers)
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
    nd.alloc[lane] = r.alloc; nd.req[lane] = r.req;     // lane < 32 <= RMAX+1
    if (lane == 0) { nd.ng = r.ng; nd.allowed = r.allowed; nd.npods = r.npods; nd.idx = i; }
}

__device__ __forceinline__ uint8_t* bytes(uint4& v) { return reinterpret_cast<uint8_t*>(&v); }

__device__ __forceinline__ long long warp_sum_ll(long long v) {
#pragma unroll
    for (int o = 16; o; o >>= 1) v += __shfl_xor_sync(FULL, v, o);
    return v;
}
__device__ __forceinline__ int warp_sum_int(int v) { return __reduce_add_sync(FULL, v); }

// exclusive prefix sum over the lanes of a warp
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
// lanes = geometries of the GPU's model.  `required` lives in lane p (p < 16).  Returns updated (warp-uniform).
__device__ __forceinline__ bool mig_gpu_update(SNode& nd, int g, int required, int lane, int K,
                                               const uint4* s_geom, const uint16_t* s_gmask, const int* s_off,
                                               int* s_req, long long& evals) {
    if (lane < KP) s_req[lane] = required;
    __syncwarp();
    const int m = nd.model[g];
    const int r0 = s_off[m], nrows = s_off[m + 1] - r0;
    const int n_req = __popc(__ballot_sync(FULL, required > 0));
    evals += (long long)nrows * n_req;
    unsigned key = 0;
    if (lane < nrows) {
        const uint4 row = s_geom[r0 + lane];
        const uint4 used = nd.used[g], fre = nd.free_[g];
        const unsigned um = nd.umask[g];
        bool can = true;                                   // CanApplyGeometry (:97-110) for an allowed row
        int score = 0;
#pragma unroll
        for (int p = 0; p < KP; ++p) {
            const int rw = (int)byte_of(row, p), us = (int)byte_of(used, p);
            if ((um >> p & 1) && rw < us) can = false;
        }
#pragma unroll
        for (int p = 0; p < KP; ++p) {
            if (p < K) {
                const int rq = s_req[p];
                const int rw = (int)byte_of(row, p), us = (int)byte_of(used, p), fr = (int)byte_of(fre, p);
                if (rq > 0 && fr < rq) {
                    const int n = min(rw - us, rq);
                    if (n > 0) score += n;
                }
            }
        }
        if (can && score > 0) key = ((unsigned)score << 5) | (unsigned)(31 - lane);
    }
    const unsigned best = __reduce_max_sync(FULL, key);
    if (best == 0) return false;
    const int br = r0 + 31 - (int)(best & 31u);            // strict > in table order == lowest index among max
    if (lane < KP) {                                        // ApplyGeometry (:132-152): lanes = profiles
        const unsigned rw = byte_of(s_geom[br], lane), us = byte_of(nd.used[g], lane);
        bytes(nd.free_[g])[lane] = (uint8_t)(rw ? rw - us : 0);
    }
    if (lane == 0) nd.fmask[g] = s_gmask[br];
    __syncwarp();
    return true;
}

// ---------------------------------------------------------------- MPS GPU.UpdateGeometryFor (slicing/gpu.go:162-232)
// lanes = profiles (ids ascend by (size, name)).  Returns updated (warp-uniform).
__device__ __forceinline__ bool mps_gpu_update(SNode& nd, int g, int required, int lane, const DevProblem& pb) {
    const int size = lane < pb.K ? pb.pmem[lane] : 0;
    int used = lane < KP ? (int)bytes(nd.used[g])[lane] : 0;
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

// SliceTracker.Remove (core/tracker.go:67-88): lanes = profiles.  s_removed caches "entry already emptied".
__device__ __forceinline__ void tracker_remove(const DevProblem& pb, int key, unsigned* s_removed, int& lack, int lane) {
    const unsigned w = s_removed[key >> 5];
    if (w >> (key & 31) & 1) return;
    __syncwarp();
    if (lane == 0) s_removed[key >> 5] = w | (1u << (key & 31));
    if (lane < KP) {
        const int e = pb.entry_lack[(size_t)key * KP + lane];
        if (e > 0) lack -= e;
    }
    __syncwarp();
}

}  // namespace

// ------------------------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(256, 1) k_plan_sweep(DevProblem pb, int32_t* cand_idx) {
    extern __shared__ __align__(16) unsigned char s_dyn[];
    __shared__ uint4 s_geom[MAX_ROWS];
    __shared__ uint16_t s_gmask[MAX_ROWS];
    __shared__ int s_off[MAX_MODELS + 1];
    __shared__ SNode s_node[2];
    __shared__ RunRec s_runs[RUN_CACHE];
    __shared__ int s_req[KP];
    __shared__ long long s_Dg[KP], s_allocF[KP], s_reqF[KP];
    __shared__ int s_scan[256];
    __shared__ int s_ncand;

    // dynamic shared memory: first chunk (32 pods) of every cached run, then the tracker "entry removed" bitmap
    typedef long long (*chunk_ll)[32];
    typedef int (*chunk_i)[32];
    typedef unsigned (*chunk_u)[32];
    chunk_ll s_ccpu = reinterpret_cast<chunk_ll>(s_dyn);
    chunk_ll s_cmem = s_ccpu + RUN_CACHE;
    chunk_ll s_ceph = s_cmem + RUN_CACHE;
    chunk_i s_ckey = reinterpret_cast<chunk_i>(s_ceph + RUN_CACHE);
    chunk_u s_cflag = reinterpret_cast<chunk_u>(s_ckey + RUN_CACHE);
    unsigned* s_removed = reinterpret_cast<unsigned*>(s_cflag + RUN_CACHE);      // [(n_keys+31)/32]
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const bool mig = pb.kind == NOSPLAN_KIND_MIG;
    const int K = pb.K;

    // ---------------- prologue (whole CTA): tables, run cache, candidate compaction in name order
    if (mig) {
        for (int r = tid; r < pb.n_rows; r += blockDim.x) { uint4 row = pb.geom[r]; s_geom[r] = row; s_gmask[r] = nz_mask(row); }
        for (int m = tid; m <= pb.n_models; m += blockDim.x) s_off[m] = pb.model_off[m];
    }
    for (int w = tid; w < (pb.n_keys + 31) / 32; w += blockDim.x) s_removed[w] = 0;
    const int n_cached = min(pb.n_runs, RUN_CACHE);
    for (int r = tid; r < n_cached; r += blockDim.x) s_runs[r] = pb.runs[r];
    __syncthreads();
    for (int t = tid; t < n_cached * 32; t += blockDim.x) {
        const int r = t >> 5, l = t & 31;
        if (l < s_runs[r].len) {
            const int s = s_runs[r].start + l;
            s_ccpu[r][l] = pb.pod_cpu[s]; s_cmem[r][l] = pb.pod_mem[s]; s_ceph[r][l] = pb.pod_eph[s];
            s_ckey[r][l] = pb.pod_key[s]; s_cflag[r][l] = pb.pod_flags[s];
        }
    }
    {   // ordered compaction of candidate flags -> cand_idx (GetCandidateNodes, core/snapshot.go:119-130)
        int base = 0;
        for (int c0 = 0; c0 < pb.N; c0 += blockDim.x) {
            const int i = c0 + tid;
            const int f = (i < pb.N) ? pb.cand_flag[i] : 0;
            s_scan[tid] = f;
            __syncthreads();
            for (int o = 1; o < (int)blockDim.x; o <<= 1) {
                int v = (tid >= o) ? s_scan[tid - o] : 0;
                __syncthreads();
                s_scan[tid] += v;
                __syncthreads();
            }
            if (f) cand_idx[base + s_scan[tid] - 1] = i;
            base += s_scan[blockDim.x - 1];
            __syncthreads();
        }
        if (tid == 0) s_ncand = base;
    }
    __threadfence_block();
    __syncthreads();
    if (warp != 0) return;

    // ---------------- the chain (warp 0)
    const int n_cand = s_ncand;
    int lack = (lane < KP) ? pb.lack[lane] : 0;                    // lane p: lackingSlices[p]
    long long D = (lane < KP) ? pb.tot[lane] : 0;                  // lane p: sum(alloc - requested) of slice column p
    bool committed = false;
    long long st_visited = 0, st_evals = 0, st_place = 0;
    const int lack_start = __popc(__ballot_sync(FULL, lack > 0));
    const int fm = pb.filter_mode;

    NodeRegs nr;
    if (n_cand > 0 && lack_start > 0) {
        load_node(pb, cand_idx[0], lane, nr);
        store_node(s_node[0], nr, cand_idx[0], lane);
    }
    __syncwarp();

    for (int c = 0; c < n_cand && lack_start > 0; ++c) {
        if (__ballot_sync(FULL, lack > 0) == 0) break;             // planner.go:94-97
        SNode& nd = s_node[c & 1];
        const bool have_next = c + 1 < n_cand;
        int next_i = 0;
        if (have_next) { next_i = cand_idx[c + 1]; load_node(pb, next_i, lane, nr); }   // prefetch (consumed at loop end)

        ++st_visited;
        const int ng = nd.ng, node_i = nd.idx;
        const long long oldA = (lane < K) ? nd.alloc[3 + lane] : 0;
        const long long oldR = (lane < K) ? nd.req[3 + lane] : 0;

        // ---- 1. Node.UpdateGeometryFor
        int required = lack;
        bool upd = false;
        for (int g = 0; g < ng; ++g) {
            bool u;
            if (mig) u = mig_gpu_update(nd, g, required, lane, K, s_geom, s_gmask, s_off, s_req, st_evals);
            else {
                st_evals += __popc(__ballot_sync(FULL, required > 0));
                u = mps_gpu_update(nd, g, required, lane, pb);
            }
            upd |= u;
            if (lane < KP) {
                const int f = (int)bytes(nd.free_[g])[lane];
                if ((nd.fmask[g] >> lane & 1) && required > 0) required = max(0, required - f);
            }
        }
        long long newA = 0;                                        // computeScalarResources (mig/node.go:179-195)
        if (lane < K)
            for (int g = 0; g < ng; ++g) newA += (long long)bytes(nd.used[g])[lane] + bytes(nd.free_[g])[lane];
        const long long allocF = upd ? newA : oldA;                // the forked entry the pods are added to
        if (lane < KP) { s_allocF[lane] = allocF; s_reqF[lane] = oldR; s_Dg[lane] = D - oldA + allocF; }
        long long remc = nd.alloc[0] - nd.req[0], remm = nd.alloc[1] - nd.req[1], reme = nd.alloc[2] - nd.req[2];
        const bool podcount_ok = nd.npods + 1 <= nd.allowed;
        __syncwarp();

        // ---- 2. the pod loop, run by run
        int added = 0;
        const bool all_fail = (fm == NOSPLAN_FILTER_REJECT_ALL) || (fm == NOSPLAN_FILTER_RESOURCES_FIT && !podcount_ok);
        for (int ri = 0; ri < pb.n_runs && !all_fail; ++ri) {
            const RunRec* rr = (ri < RUN_CACHE) ? &s_runs[ri] : &pb.runs[ri];
            const int rstart = rr->start, rend = rr->start + rr->len, n_ent = rr->n_ent;
            const int q_l = (lane < KP) ? (int)rr->q[lane] : 0;
            // class-level tests; all of them only get tighter while this node is processed
            bool ok_l = true;
            if (q_l > 0) {
                ok_l = s_Dg[lane] >= q_l;                                                       // gate, snapshot.go:132-165
                if (fm == NOSPLAN_FILTER_RESOURCES_FIT) ok_l = ok_l && (q_l <= s_allocF[lane] - s_reqF[lane]);
            }
            if (!__all_sync(FULL, ok_l)) continue;

            if (n_ent <= 1) {
                // ---------- single-profile (or slice-less) class: bulk first-fit
                const unsigned emask = __ballot_sync(FULL, q_l > 0);
                const int p = n_ent ? __ffs(emask) - 1 : 0;
                const int q = n_ent ? __shfl_sync(FULL, q_l, p) : 0;
                long long cap = CAP_INF;
                if (n_ent) {
                    const int cg = (lane < ng) ? (int)bytes(nd.free_[lane])[p] / q : 0;       // GPU.AddPod feasibility
                    cap = warp_sum_int(cg);
                    cap = min(cap, s_Dg[p] / q);
                    if (fm == NOSPLAN_FILTER_RESOURCES_FIT) cap = min(cap, (s_allocF[p] - s_reqF[p]) / q);
                }
                for (int cursor = rstart; cursor < rend && cap > 0; cursor += 32) {
                    const int s = cursor + lane;
                    const bool in = s < rend;
                    long long cpu = 0, mem = 0, eph = 0; int key = 0; unsigned fl = 0;
                    if (in) {
                        if (ri < RUN_CACHE && cursor == rstart) {
                            cpu = s_ccpu[ri][lane]; mem = s_cmem[ri][lane]; eph = s_ceph[ri][lane];
                            key = s_ckey[ri][lane]; fl = s_cflag[ri][lane];
                        } else {
                            cpu = pb.pod_cpu[s]; mem = pb.pod_mem[s]; eph = pb.pod_eph[s];
                            key = pb.pod_key[s]; fl = pb.pod_flags[s];
                        }
                    }
                    const bool fit = fm == NOSPLAN_FILTER_RESOURCES_FIT;
                    const bool slow = in && fit && (fl & PF_HAS_OTHER);
                    int from = 0;
                    while (from < 32 && cap > 0) {
                        // pods with other scalar resources take the exact per-pod path; they split the step
                        const unsigned slow_m = __ballot_sync(FULL, slow && lane >= from);
                        const int stop = slow_m ? __ffs(slow_m) - 1 : 32;
                        const bool zero = (fl & PF_ZERO_REQ) != 0;
                        bool valid = in && lane >= from && lane < stop && (fl & PF_PRE_OK);
                        if (fit) valid = valid && (zero || (cpu <= remc && mem <= remm && eph <= reme));  // can never pass later
                        const unsigned vm = __ballot_sync(FULL, valid);
                        if (vm == 0 && stop == 32) break;
                        const int rank = __popc(vm & ((1u << lane) - 1));
                        const bool cand = valid && rank < cap;
                        bool test = cand;
                        const long long ccpu = cand ? cpu : 0, cmem = cand ? mem : 0, ceph = cand ? eph : 0;
                        if (fit) {
                            const long long ec = warp_excl_ll(ccpu, lane), em = warp_excl_ll(cmem, lane), ee = warp_excl_ll(ceph, lane);
                            test = cand && (zero || (cpu <= remc - ec && mem <= remm - em && eph <= reme - ee));
                        }
                        const unsigned failm = __ballot_sync(FULL, cand && !test);
                        const int firstfail = failm ? __ffs(failm) - 1 : 32;
                        const bool acc = cand && lane < firstfail;
                        const unsigned accm = __ballot_sync(FULL, acc);
                        const int n = __popc(accm);
                        if (n > 0) {
                            const long long sc = warp_sum_ll(acc ? cpu : 0), sm = warp_sum_ll(acc ? mem : 0), se = warp_sum_ll(acc ? eph : 0);
                            remc -= sc; remm -= sm; reme -= se;
                            if (n_ent) {                                   // first-fit over GPUs (mig/node.go:201-210)
                                const int cg = (lane < ng) ? (int)bytes(nd.free_[lane])[p] / q : 0;
                                const int ex = warp_excl_int(cg, lane);
                                const int take = max(0, min(cg, n - ex));
                                if (lane < ng && take > 0) {
                                    bytes(nd.free_[lane])[p] -= (uint8_t)(take * q);
                                    bytes(nd.used[lane])[p] += (uint8_t)(take * q);
                                    nd.umask[lane] |= (uint16_t)(1u << p);
                                }
                                if (lane == 0) { s_reqF[p] += (long long)n * q; s_Dg[p] -= (long long)n * q; }
                                cap -= n;
                            }
                            unsigned todo = accm;                          // tracker.Remove per placed pod
                            while (todo) {
                                const int l = __ffs(todo) - 1;
                                todo &= todo - 1;
                                tracker_remove(pb, __shfl_sync(FULL, key, l), s_removed, lack, lane);
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
                                    if (lane == 0) { s_reqF[p] += q; s_Dg[p] -= q; }
                                    cap -= 1;
                                }
                                tracker_remove(pb, pkey, s_removed, lack, lane);
                                ++added; ++st_place;
                                __syncwarp();
                            }
                            from = stop + 1;
                            continue;
                        }
                        break;
                    }
                }
            } else {
                // ---------- multi-profile class: exact per-pod path (GPU.AddPod keeps partial moves, mig/gpu.go:234-245)
                for (int s = rstart; s < rend; ++s) {
                    const unsigned fl = pb.pod_flags[s];
                    const long long cpu = pb.pod_cpu[s], mem = pb.pod_mem[s], eph = pb.pod_eph[s];
                    // gate + class-level fit again (state may have changed inside this run)
                    bool okc = true;
                    if (q_l > 0) {
                        okc = s_Dg[lane] >= q_l;
                        if (fm == NOSPLAN_FILTER_RESOURCES_FIT) okc = okc && (q_l <= s_allocF[lane] - s_reqF[lane]);
                    }
                    if (!__all_sync(FULL, okc)) break;                     // every later pod of the run fails the same way
                    bool ok = (fl & PF_PRE_OK) != 0;
                    if (fm == NOSPLAN_FILTER_RESOURCES_FIT) {
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
                    // Node.AddPod: GPUs in order, profiles ascending, partial moves kept
                    bool placed = false, dirty = false;
                    for (int g = 0; g < ng && !placed; ++g) {
                        const int fr = (lane < KP) ? (int)bytes(nd.free_[g])[lane] : 0;
                        const unsigned bad = __ballot_sync(FULL, q_l > 0 && fr < q_l);
                        const unsigned want = __ballot_sync(FULL, q_l > 0);
                        const unsigned mv = bad ? (want & ((1u << (__ffs(bad) - 1)) - 1)) : want;   // profiles moved before the first failure
                        if (mv >> lane & 1) {
                            bytes(nd.free_[g])[lane] -= (uint8_t)q_l; bytes(nd.used[g])[lane] += (uint8_t)q_l;
                        }
                        if (lane == 0 && mv) nd.umask[g] |= (uint16_t)mv;
                        dirty |= mv != 0;
                        placed = bad == 0;
                        __syncwarp();
                    }
                    if (!placed) { if (!dirty) break; else continue; }     // clean failure repeats for the rest of the run
                    remc -= cpu; remm -= mem; reme -= eph;
                    if (q_l > 0) { s_reqF[lane] += q_l; s_Dg[lane] -= q_l; }
                    if (lane == 0 && (fl & PF_HAS_OTHER))
                        for (int e = pb.pod_sptr[s]; e < pb.pod_sptr[s + 1]; ++e) {
                            const int col = pb.pod_scol[e];
                            if (col >= 3 + K) nd.req[col] += pb.pod_sqty[e];
                        }
                    tracker_remove(pb, pb.pod_key[s], s_removed, lack, lane);
                    ++added; ++st_place;
                    __syncwarp();
                }
            }
        }

        // ---- 3. Commit / Revert (planner.go:144-149) with the object-graph aliasing rules
        if (added > 0) {
            if (lane < KP) D = s_Dg[lane];
            committed = true;
            if (upd && lane < ng) {                                // partitioningState[n] = GetPartitioning(n) (planner.go:138)
                const int x = node_i * pb.G + lane;
                pb.out_count[x] = add_bytes(nd.used[lane], nd.free_[lane]);
                pb.out_mask[x] = nd.umask[lane] | nd.fmask[lane];
            }
        } else if (!committed) {
            D += newA - oldA;                                      // the mutated original is still what the snapshot holds
        }
        __syncwarp();
        if (have_next) store_node(s_node[(c + 1) & 1], nr, next_i, lane);
        __syncwarp();
    }

    if (lane == 0) {
        DevStats* st = pb.stats;
        st->visited_nodes = st_visited;
        st->candidates = st_visited * pb.P;
        st->geometry_evals = st_evals;
        st->placements = st_place;
        st->candidate_nodes = lack_start > 0 ? n_cand : 0;   // GetCandidateNodes is not reached on the early return (planner.go:80-83)
        st->lacking_at_start = lack_start;
    }
    const int lack_end = __popc(__ballot_sync(FULL, lack > 0));
    if (lane == 0) pb.stats->lacking_at_end = lack_end;
}

constexpr size_t SWEEP_DYN_MAX = 200 * 1024;
constexpr size_t SWEEP_CHUNK_BYTES = (size_t)RUN_CACHE * 32 * (8 + 8 + 8 + 4 + 4);

size_t sweep_dyn_smem(const DevProblem& pb) { return SWEEP_CHUNK_BYTES + (size_t)((pb.n_keys + 31) / 32) * 4 + 16; }

// largest number of tracker keys (distinct namespace/name of pending pods) the shared-memory bitmap can hold
int sweep_max_keys() { return (int)((SWEEP_DYN_MAX - SWEEP_CHUNK_BYTES - 16) / 4) * 32; }

cudaError_t sweep_prepare_device() {
    return cudaFuncSetAttribute(k_plan_sweep, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)SWEEP_DYN_MAX);
}

cudaError_t launch_sweep(const DevProblem& pb, int32_t* cand_idx, cudaStream_t st) {
    const size_t dyn = sweep_dyn_smem(pb);
    if (dyn > SWEEP_DYN_MAX) return cudaErrorInvalidValue;
    k_plan_sweep<<<1, 256, dyn, st>>>(pb, cand_idx);
    return cudaGetLastError();
}

}  // namespace nsp
