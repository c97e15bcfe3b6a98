// Plan() in ONE launch (sm_100a): k_plan_fused reads every node row once and does, per node, everything k_prepass and the
// node-parallel pass of sweep.cu do in three launches -- initial partitioning state, HasFreeCapacity, cluster totals,
// Node.UpdateGeometryFor, the pod loop of planner.go:116-141 -- under a HYPOTHESIS that is verified by the CTA that finishes
// last:
//
//   every profile some pending pod requests is lacking, and the lacking vector stays saturated to the end
//   (lack[p] > (G+2)*M_p, the condition under which sweep.cu's k_plan_parallel is exact).
//
// Which profiles are lacking depends on the cluster totals, i.e. on every node; speculating on it removes the grid-wide
// dependency between "sum the totals" and "plan the nodes", so the node rows cross HBM once and the whole Plan() is one
// kernel.  If the hypothesis does not hold (or a node's accounts disagree with its GPUs, or a value does not fit the
// compact units below) the kernel reports it and the host runs the classic sequence k_prepass -> k_tracker_init ->
// k_plan_parallel -> ... on the untouched inputs; nothing the fused kernel wrote survives that.
//
// Mapping.  8 lanes per node (lane = GPU of the node), 4 nodes per warp, 16 per CTA; all group-level votes and reductions
// use the 8-lane mask.  reference -> here:
//   mig.GPU.UpdateGeometryFor (pkg/gpu/mig/gpu.go:158-212)   each lane walks the allowed geometries of ITS GPU: SWAR
//        "row >= used" on 4 profiles per word, score = dp4a(row - used, lacking mask); strict > in table order
//   mig.Node.AddPod / GPU.AddPod (mig/node.go:201-210, gpu.go:233-247)   capacity per run = sum over lanes of free[p] / q,
//        first-fit over the GPUs by an 8-lane exclusive prefix, once per run
//   tryAddPod over all pods (planner.go:116-141) + NodeResourcesFit   runs of identical slice requests in NewPodSorter order;
//        pods are (cpu, memory) int32 pairs in a padded position space with minima per 8 and per 64 pods, so that a group
//        descends only into blocks that can still hold a fitting pod; 8 pods per step, sequential first-fit reproduced with
//        an 8-lane prefix sum when the step's candidates do not all fit together
//   SliceTracker.Remove (core/tracker.go:67-88)   one bit per pod POSITION (tracker keys are unique on this path), OR-ed in
//        shared memory and flushed once per CTA; the final lacking vector is rebuilt from the bitmap by the last CTA
#include <climits>

#include "nosplan_dev.cuh"

namespace nsp {

namespace {

constexpr int FUSED_THREADS = 128;
constexpr int FUSED_NODES = FUSED_THREADS / 8;       // nodes per CTA step
constexpr unsigned HI = 0x80808080u;
constexpr int CAP_NONE = 1 << 30;                     // capacity of a run without slice request

__device__ __forceinline__ unsigned byte_at(unsigned lo, unsigned hi, int p) {   // byte p (< 8) of the pair
    return ((p < 4 ? lo : hi) >> ((p & 3) * 8)) & 0xffu;
}

template <bool WIDE>       // WIDE: profiles 8..15 in use (words z, w)
__device__ __forceinline__ void plan_nodes(const DevProblem& pb, const uint4* s_geom, const uint16_t* s_gmask, const int* s_off,
                                           const FRun* s_runs, const unsigned long long* s_prof_runs, unsigned* s_bits,
                                           unsigned long long* s_D, int* s_maxfree, unsigned long long* s_cnt, int* s_flags) {
    const int tid = threadIdx.x, lane = tid & 31, grp = lane >> 3, j = lane & 7;
    const unsigned gmask = 0xffu << (grp * 8);
    const int gsh = grp * 8;
    const int G = pb.G, K = pb.K, R = pb.R;
    const bool fit = pb.filter_mode == NOSPLAN_FILTER_RESOURCES_FIT;
    const bool reject_all = pb.filter_mode == NOSPLAN_FILTER_REJECT_ALL;
    const int n_req = __popc(pb.fp_rset);
    // lacking profiles as a 0/1 byte per profile (the dp4a weights)
    unsigned Lw[4];
#pragma unroll
    for (int w = 0; w < 4; ++w) {
        unsigned v = 0;
#pragma unroll
        for (int b = 0; b < 4; ++b) v |= ((pb.fp_rset >> (w * 4 + b)) & 1u) << (8 * b);
        Lw[w] = v;
    }
    long long accD0 = 0, accD1 = 0;                   // sum(alloc - requested) of slice columns j and j + 8
    uint4 mf = make_uint4(0, 0, 0, 0);                // byte-wise max of the free counts this lane has seen
    unsigned flags = 0;
    long long st_place = 0, st_evals = 0;
    int st_visited = 0;

    const int n_tiles = (pb.node_hi - pb.node_lo + FUSED_NODES - 1) / FUSED_NODES;
    for (int tile = blockIdx.x; tile < n_tiles; tile += gridDim.x) {
        const int i = pb.node_lo + tile * FUSED_NODES + (tid >> 3);
        const bool node_on = i < pb.node_hi;
        // ---- the node's rows: every load is issued before the first use
        int ng = 0, allowed_pods = 0, npods = 0;
        uint4 used = make_uint4(0, 0, 0, 0), fre = used;
        unsigned um = 0, fm = 0;
        int model = 0;
        long long ac = 0, am = 0, ae = 0, rc = 0, rm = 0, re = 0, a0 = 0, r0 = 0, a1 = 0, r1 = 0;
        if (node_on) {
            ng = min(max(pb.node_gpu_count[i], 0), G);
            allowed_pods = pb.node_allowed_pods[i]; npods = pb.node_num_pods[i];
            const long long* al = pb.node_alloc + (size_t)i * R;
            const long long* rq = pb.node_req + (size_t)i * R;
            ac = al[0]; am = al[1]; ae = al[2]; rc = rq[0]; rm = rq[1]; re = rq[2];
            if (j < K) { a0 = al[3 + j]; r0 = rq[3 + j]; }
            if (WIDE && j + 8 < K) { a1 = al[3 + 8 + j]; r1 = rq[3 + 8 + j]; }
            if (j < ng) {
                const int x = i * G + j;
                used = pb.gpu_used[x]; fre = pb.gpu_free[x];
                um = pb.gpu_umask[x]; fm = pb.gpu_fmask[x];
                model = min(max(pb.gpu_model[x], 0), max(pb.n_models - 1, 0));
            }
        }
        const bool has = j < ng;
        // ---- k_prepass duties: totals, largest free count, node accounts against the GPU maps
        accD0 += a0 - r0; accD1 += a1 - r1;
        if (((used.x | used.y | fre.x | fre.y | (WIDE ? (used.z | used.w | fre.z | fre.w) : 0u)) & 0xe0e0e0e0u) != 0) flags |= FREJ_COUNTS;
        mf.x = __vmaxu4(mf.x, fre.x); mf.y = __vmaxu4(mf.y, fre.y);
        if (WIDE) { mf.z = __vmaxu4(mf.z, fre.z); mf.w = __vmaxu4(mf.w, fre.w); }
        {
            unsigned fx = fre.x, fy = fre.y, ux = used.x, uy = used.y, fz = fre.z, fw = fre.w, uz = used.z, uw = used.w;
#pragma unroll
            for (int o = 1; o < 8; o <<= 1) {          // byte-wise sums over the node's GPUs (counts <= 31: no carry)
                fx += __shfl_xor_sync(gmask, fx, o); fy += __shfl_xor_sync(gmask, fy, o);
                ux += __shfl_xor_sync(gmask, ux, o); uy += __shfl_xor_sync(gmask, uy, o);
                if (WIDE) {
                    fz += __shfl_xor_sync(gmask, fz, o); fw += __shfl_xor_sync(gmask, fw, o);
                    uz += __shfl_xor_sync(gmask, uz, o); uw += __shfl_xor_sync(gmask, uw, o);
                }
            }
            bool bad = false;
            if (j < K) { const long long sf = byte_at(fx, fy, j), su = byte_at(ux, uy, j); bad = (a0 - r0 < sf) || (r0 > su); }
            if (WIDE && j + 8 < K) { const long long sf = byte_at(fz, fw, j), su = byte_at(uz, uw, j); bad |= (a1 - r1 < sf) || (r1 > su); }
            if (bad) flags |= FREJ_ACCOUNTS;
        }
        // ---- mig.GPU.UpdateGeometryFor for this lane's GPU + AllowsGeometry(current) for HasFreeCapacity
        const uint4 cnt0 = make_uint4(used.x + fre.x, used.y + fre.y, used.z + fre.z, used.w + fre.w);
        const unsigned gm0 = um | fm;
        unsigned best = 0;
        bool allowed = false;
        int nrows = 0, row0 = 0;
        if (has && pb.n_models > 0) { row0 = s_off[model]; nrows = s_off[model + 1] - row0; }
        for (int r = 0; r < nrows; ++r) {
            const uint4 row = s_geom[row0 + r];
            const unsigned tx = (row.x | HI) - used.x, ty = (row.y | HI) - used.y;   // bit 7 of a byte: row >= used
            unsigned ok = tx & ty;
            unsigned sc = __dp4a(tx & ~HI, Lw[0], 0u);
            sc = __dp4a(ty & ~HI, Lw[1], sc);
            bool same = row.x == cnt0.x && row.y == cnt0.y;
            if (WIDE) {
                const unsigned tz = (row.z | HI) - used.z, tw = (row.w | HI) - used.w;
                ok &= tz & tw;
                sc = __dp4a(tz & ~HI, Lw[2], sc);
                sc = __dp4a(tw & ~HI, Lw[3], sc);
                same = same && row.z == cnt0.z && row.w == cnt0.w;
            }
            if ((ok & HI) == HI && sc) best = max(best, (sc << 5) | (unsigned)(31 - r));   // strict > in table order
            allowed |= same && s_gmask[row0 + r] == gm0;
        }
        const bool has_free = has && (fm != 0 || !allowed);
        const bool candidate = (__ballot_sync(gmask, has_free) >> gsh & 0xffu) != 0;
        bool upd_g = false;
        if (candidate && best) {
            const int br = row0 + 31 - (int)(best & 31u);
            const uint4 row = s_geom[br];
            fre = make_uint4(row.x - used.x, row.y - used.y, WIDE ? row.z - used.z : 0u, WIDE ? row.w - used.w : 0u);   // zero kept (:141)
            fm = s_gmask[br];
            upd_g = true;
        }
        const bool upd = (__ballot_sync(gmask, upd_g) >> gsh & 0xffu) != 0;
        int added = 0;
        if (candidate) {
            if (j == 0) ++st_visited;
            if (has) st_evals += (long long)nrows * n_req;
            // ---- NodeResourcesFit state in the pods' units
            int remc = F_REM_MAX, remm = F_REM_MAX;
            bool all_fail = reject_all;
            if (fit) {
                const long long dc = ac - rc, dm = am - rm;
                const long long qc = dc >> pb.fp_cpu_shift, qm = dm >> pb.fp_mem_shift;
                if (qc > F_REM_MAX || qm > F_REM_MAX) flags |= FREJ_REM;
                remc = dc < 0 ? -1 : (int)min(qc, (long long)F_REM_MAX);
                remm = dm < 0 ? -1 : (int)min(qm, (long long)F_REM_MAX);
                if (ae - re < 0) remc = -1;                 // no pod requests ephemeral storage: 0 <= negative fails for every pod
                all_fail = all_fail || !(npods + 1 <= allowed_pods);
            }
            // profiles with a free device somewhere on the node
            unsigned nzm = nz_nibble(fre.x) | (nz_nibble(fre.y) << 4);
            if (WIDE) nzm |= (nz_nibble(fre.z) << 8) | (nz_nibble(fre.w) << 12);
            const unsigned feas = __reduce_or_sync(gmask, has ? nzm : 0u);
            unsigned long long live = 0;
            if (!all_fail) {
                live = s_prof_runs[KP];
                for (unsigned f = feas; f; f &= f - 1) live |= s_prof_runs[__ffs(f) - 1];
            }
            // ---- the pod loop, run by run
            while (live) {
                const int ri = __ffsll((long long)live) - 1;
                live &= live - 1;
                const FRun fr = s_runs[ri];
                if (fr.min_c > remc || fr.min_m > remm) continue;          // no pod of the run can pass any more
                const int p = fr.p, q = fr.q;
                int cg = 0, cap = CAP_NONE;
                if (p >= 0) {
                    const unsigned w = p < 8 ? (p < 4 ? fre.x : fre.y) : (p < 12 ? fre.z : fre.w);
                    const int f0 = (int)((w >> ((p & 3) * 8)) & 0xffu);
                    cg = has ? (q == 1 ? f0 : f0 / q) : 0;                 // GPU.AddPod feasibility
                    cap = (int)__reduce_add_sync(gmask, (unsigned)cg);
                    if (cap <= 0) continue;
                }
                const int cap0 = cap;
                const int blk0 = fr.pos0 >> 6;
                for (int b0 = 0; b0 < fr.nblk && cap > 0; b0 += 8) {
                    uint2 l2 = make_uint2(F_NONE, F_NONE);
                    if (b0 + j < fr.nblk) l2 = __ldg(&pb.fp_l2[blk0 + b0 + j]);
                    unsigned m2 = __ballot_sync(gmask, (int)l2.x <= remc && (int)l2.y <= remm) >> gsh & 0xffu;
                    while (m2 && cap > 0) {
                        const int blk = blk0 + b0 + __ffs(m2) - 1;
                        m2 &= m2 - 1;
                        const uint2 l1 = __ldg(&pb.fp_l1[blk * 8 + j]);
                        unsigned m1 = __ballot_sync(gmask, (int)l1.x <= remc && (int)l1.y <= remm) >> gsh & 0xffu;
                        while (m1 && cap > 0) {
                            const int sub = blk * 8 + __ffs(m1) - 1;
                            m1 &= m1 - 1;
                            const uint2 pod = __ldg(&pb.fp_pod[sub * 8 + j]);
                            const int c = (int)pod.x, m = (int)pod.y;
                            unsigned acc_bits = 0;
                            int from = 0;
                            for (;;) {
                                // a pod that does not fit now never fits later on this node: the state only gets tighter
                                const bool valid = j >= from && c <= remc && m <= remm;
                                const unsigned vm = __ballot_sync(gmask, valid) >> gsh & 0xffu;
                                if (!vm) break;
                                const bool cand = valid && __popc(vm & ((1u << j) - 1)) < cap;
                                unsigned accm = __ballot_sync(gmask, cand) >> gsh & 0xffu;
                                int next_from = 8;
                                if (fit) {
                                    int sc = (int)__reduce_add_sync(gmask, cand ? (unsigned)c : 0u);
                                    int sm = (int)__reduce_add_sync(gmask, cand ? (unsigned)m : 0u);
                                    if (sc > remc || sm > remm) {
                                        // the candidates do not all fit together: sequential first-fit by prefix sums; the first
                                        // pod that fails under "all earlier candidates placed" splits the step
                                        int ec = cand ? c : 0, em = cand ? m : 0;
#pragma unroll
                                        for (int o = 1; o < 8; o <<= 1) {
                                            const int tc = __shfl_up_sync(gmask, ec, o, 8), tm = __shfl_up_sync(gmask, em, o, 8);
                                            if (j >= o) { ec += tc; em += tm; }
                                        }
                                        ec -= cand ? c : 0; em -= cand ? m : 0;
                                        const unsigned failm = __ballot_sync(gmask, cand && (c > remc - ec || m > remm - em)) >> gsh & 0xffu;
                                        const int ff = __ffs(failm) - 1;                 // exists: the totals did not fit
                                        sc = __shfl_sync(gmask, ec, ff, 8); sm = __shfl_sync(gmask, em, ff, 8);
                                        accm &= (1u << ff) - 1;
                                        next_from = ff + 1;
                                    }
                                    if (accm) { remc -= sc; remm -= sm; }
                                }
                                cap -= __popc(accm);
                                acc_bits |= accm;
                                if (next_from >= 8 || cap <= 0) break;
                                from = next_from;
                            }
                            if (acc_bits && j == 0) atomicOr(&s_bits[sub >> 2], acc_bits << ((sub & 3) * 8));   // tracker.Remove
                        }
                    }
                }
                const int n = cap0 - cap;
                if (n > 0) {
                    added += n;
                    if (p >= 0) {                                          // first fit over the GPUs in order (mig/node.go:201-210)
                        int ex = cg;
#pragma unroll
                        for (int o = 1; o < 8; o <<= 1) { const int t = __shfl_up_sync(gmask, ex, o, 8); if (j >= o) ex += t; }
                        ex -= cg;
                        const int take = max(0, min(cg, n - ex));
                        if (take > 0) {
                            const unsigned d = (unsigned)(take * q) << ((p & 3) * 8);
                            if (p < 4) { fre.x -= d; used.x += d; } else if (p < 8) { fre.y -= d; used.y += d; }
                            else if (p < 12) { fre.z -= d; used.z += d; } else { fre.w -= d; used.w += d; }
                            um |= 1u << p;
                        }
                    }
                }
            }
            if (j == 0) st_place += added;
        }
        // ---- partitioningState[n] (planner.go:72,138): the node's geometry after the update if a pod was added, else as it was
        if (has) {
            const int x = i * G + j;
            if (added > 0 && upd) {
                pb.out_count[x] = make_uint4(used.x + fre.x, used.y + fre.y, WIDE ? used.z + fre.z : 0u, WIDE ? used.w + fre.w : 0u);
                pb.out_mask[x] = (uint16_t)(um | fm);
            } else {
                pb.out_count[x] = cnt0;
                pb.out_mask[x] = (uint16_t)gm0;
            }
        }
    }

    // ---- CTA-level reduction of what the lanes accumulated
    if (flags) atomicOr(&s_flags[0], (int)flags);
#pragma unroll
    for (int o = 8; o < 32; o <<= 1) {                 // lanes with the same j across the warp's four groups
        accD0 += __shfl_xor_sync(0xffffffffu, accD0, o); accD1 += __shfl_xor_sync(0xffffffffu, accD1, o);
    }
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
        mf.x = __vmaxu4(mf.x, __shfl_xor_sync(0xffffffffu, mf.x, o)); mf.y = __vmaxu4(mf.y, __shfl_xor_sync(0xffffffffu, mf.y, o));
        if (WIDE) { mf.z = __vmaxu4(mf.z, __shfl_xor_sync(0xffffffffu, mf.z, o)); mf.w = __vmaxu4(mf.w, __shfl_xor_sync(0xffffffffu, mf.w, o)); }
        st_place += __shfl_xor_sync(0xffffffffu, st_place, o); st_evals += __shfl_xor_sync(0xffffffffu, st_evals, o);
        st_visited += __shfl_xor_sync(0xffffffffu, st_visited, o);
    }
    if (lane < 8) {
        if (accD0) atomicAdd(&s_D[lane], (unsigned long long)accD0);
        if (WIDE && accD1) atomicAdd(&s_D[8 + lane], (unsigned long long)accD1);
    }
    if (lane < KP) { const int v = (int)byte_of(mf, lane); if (v) atomicMax(&s_maxfree[lane], v); }
    if (lane == 0) {
        atomicAdd(&s_cnt[0], (unsigned long long)st_place); atomicAdd(&s_cnt[1], (unsigned long long)st_evals);
        atomicAdd(&s_cnt[2], (unsigned long long)st_visited);
    }
}

}  // namespace

__global__ void __launch_bounds__(FUSED_THREADS, 6) k_plan_fused(DevProblem pb) {
    extern __shared__ __align__(16) unsigned s_bits[];              // one bit per padded pod position: placed by this CTA
    __shared__ uint4 s_geom[MAX_ROWS];
    __shared__ uint16_t s_gmask[MAX_ROWS];
    __shared__ int s_off[MAX_MODELS + 1];
    __shared__ FRun s_runs[FRUN_MAX];
    __shared__ unsigned long long s_prof_runs[KP + 1];
    __shared__ unsigned long long s_D[KP];
    __shared__ int s_maxfree[KP];
    __shared__ unsigned long long s_cnt[4];
    __shared__ int s_flags[1];
    __shared__ long long s_l0[KP], s_dec[KP];
    __shared__ long long s_e[FRUN_MAX];
    const int tid = threadIdx.x;
    const int n_words = pb.fp_pad_pods >> 5;

    for (int r = tid; r < pb.n_rows; r += FUSED_THREADS) { const uint4 row = pb.geom[r]; s_geom[r] = row; s_gmask[r] = nz_mask(row); }
    for (int m = tid; m <= pb.n_models; m += FUSED_THREADS) s_off[m] = pb.model_off[m];
    for (int r = tid; r < pb.fp_n_runs; r += FUSED_THREADS) s_runs[r] = pb.fp_runs[r];
    for (int w = tid; w < n_words; w += FUSED_THREADS) s_bits[w] = 0;
    if (tid <= KP) s_prof_runs[tid] = 0;
    if (tid < KP) { s_D[tid] = 0; s_maxfree[tid] = 0; }
    if (tid < 4) s_cnt[tid] = 0;
    if (tid == 0) s_flags[0] = 0;
    // the next run's copy of the small per-run state (see DevProblem::zero_next)
    if (blockIdx.x == 0) for (int x = tid; x < pb.zero_next_n8; x += FUSED_THREADS) pb.zero_next[x] = 0;
    __syncthreads();
    for (int r = tid; r < pb.fp_n_runs; r += FUSED_THREADS) atomicOr(&s_prof_runs[s_runs[r].p >= 0 ? s_runs[r].p : KP], 1ull << r);
    __syncthreads();

    if (pb.K > 8) plan_nodes<true>(pb, s_geom, s_gmask, s_off, s_runs, s_prof_runs, s_bits, s_D, s_maxfree, s_cnt, s_flags);
    else plan_nodes<false>(pb, s_geom, s_gmask, s_off, s_runs, s_prof_runs, s_bits, s_D, s_maxfree, s_cnt, s_flags);
    __syncthreads();

    // ---- this CTA's contribution to the grid-wide state
    if (tid < KP) {
        if (s_D[tid]) atomicAdd(reinterpret_cast<unsigned long long*>(&pb.tot[tid]), s_D[tid]);
        if (s_maxfree[tid]) atomicMax(&pb.maxfree[tid], s_maxfree[tid]);
    }
    if (tid < 3 && s_cnt[tid]) atomicAdd(&pb.acc[tid], s_cnt[tid]);
    if (tid == 0 && s_flags[0]) atomicOr(&pb.wave[5], s_flags[0]);
    for (int w = tid; w < n_words; w += FUSED_THREADS) {
        const unsigned b = s_bits[w];
        if (b && (b & ~__ldcg(&pb.placed[w]))) atomicOr(&pb.placed[w], b);
    }

    // ---- the CTA that finishes last verifies the hypothesis on the final state (what k_plan_sweep does for the classic path)
    if (!last_block_done(&pb.next_node[2])) return;
    if (tid < KP) { s_l0[tid] = 0; s_dec[tid] = 0; }
    __syncthreads();
    for (int r = tid; r < pb.fp_n_runs; r += FUSED_THREADS) {
        const FRun fr = s_runs[r];
        long long e = 0;
        if (fr.p >= 0) {                                              // NewSliceTracker (core/tracker.go:34-57): lacking = request - available
            const long long avail = max(0ll, __ldcg(&pb.tot[fr.p]));  // SubtractNonNegative (pkg/resource/resource.go:76-88)
            e = max(0ll, (long long)fr.q - avail);
            if (e) atomicAdd(reinterpret_cast<unsigned long long*>(&s_l0[fr.p]), (unsigned long long)(e * fr.n_real));
        }
        s_e[r] = e;
    }
    __syncthreads();
    for (int w = tid; w < n_words; w += FUSED_THREADS) {              // a 32-position word lies inside one run (runs are 64-aligned)
        const int c = __popc(__ldcg(&pb.placed[w]));
        pb.placed[w] = 0;                                             // clean for the next fused run of this problem
        if (!c) continue;
        int r = 0;
        while (r + 1 < pb.fp_n_runs && s_runs[r + 1].pos0 <= (w << 5)) ++r;
        if (s_runs[r].p >= 0 && s_e[r]) atomicAdd(reinterpret_cast<unsigned long long*>(&s_dec[s_runs[r].p]), (unsigned long long)(c * s_e[r]));
    }
    __syncthreads();
    if (tid < 32) {
        const int lane = tid;
        bool ok = true;
        long long l0 = 0, le = 0;
        if (lane < KP) {
            l0 = s_l0[lane]; le = l0 - s_dec[lane];
            const bool requested = (pb.fp_rset >> lane & 1u) != 0;
            ok = requested == (l0 > 0);
            if (requested) {                                          // saturation at the start and at the end (fast_applicable of sweep.cu)
                const long long m = max(pb.sat_m[lane], __ldcg(&pb.maxfree[lane]));
                ok = ok && le > 0 && le > (long long)(pb.G + 2) * m && l0 < (1ll << 30);
            }
        }
        int rej = __ldcg(&pb.wave[5]);
        if (!__all_sync(0xffffffffu, ok)) rej |= FREJ_HYPOTHESIS;
        const int lack_start = __popc(__ballot_sync(0xffffffffu, lane < KP && l0 > 0));
        const int lack_end = __popc(__ballot_sync(0xffffffffu, lane < KP && le > 0));
        if (lane == 0) {
            DevStats* st = pb.stats;
            const long long vis = (long long)__ldcg(&pb.acc[2]);
            st->visited_nodes = vis;
            st->candidates = vis * pb.P;
            st->geometry_evals = (long long)__ldcg(&pb.acc[1]);
            st->placements = (long long)__ldcg(&pb.acc[0]);
            st->candidate_nodes = vis;
            st->lacking_at_start = lack_start;
            st->lacking_at_end = lack_end;
            st->path = rej ? 2 : 1;
            st->pad = rej;
        }
    }
}

size_t fused_dyn_smem(const DevProblem& pb) { return (size_t)(pb.fp_pad_pods >> 5) * 4 + 16; }
constexpr size_t FUSED_DYN_MAX = 32 * 1024;
int fused_max_pad_pods() { return (int)((FUSED_DYN_MAX - 16) / 4) * 32; }

cudaError_t fused_prepare_device(int* blocks_per_sm) {
    cudaError_t e = cudaFuncSetAttribute(k_plan_fused, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)FUSED_DYN_MAX);
    if (e != cudaSuccess) return e;
    int per_sm = 4;
    if (cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, k_plan_fused, FUSED_THREADS, 4096) != cudaSuccess || per_sm < 1) per_sm = 4;
    *blocks_per_sm = per_sm;
    return cudaSuccess;
}

cudaError_t launch_fused(const DevProblem& pb, int max_blocks, cudaStream_t st) {
    int blocks = (pb.node_hi - pb.node_lo + FUSED_NODES - 1) / FUSED_NODES;
    if (blocks > max_blocks) blocks = max_blocks;
    if (blocks < 1) blocks = 1;
    k_plan_fused<<<blocks, FUSED_THREADS, fused_dyn_smem(pb), st>>>(pb);
    return cudaGetLastError();
}

}  // namespace nsp
