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
// A CTA takes tiles of `tile_nodes` nodes through three phases:
//   phase 1  8 lanes per node, lane = GPU (all lanes of a warp converged, 8-lane shuffles):
//            mig.GPU.UpdateGeometryFor (pkg/gpu/mig/gpu.go:158-212) -- each lane walks the allowed geometries of ITS GPU: SWAR
//            "row >= used" on 4 profiles per word, score = dp4a(row - used, lacking mask), strict > in table order;
//            for an MPS snapshot (template parameter) slicing.GPU.UpdateGeometryFor (pkg/gpu/slicing/gpu.go:162-232) instead,
//            thread-serial over the profiles of the lane's GPU;
//            k_prepass's duties (HasFreeCapacity, totals, largest free count, accounts against GPU maps); the node's free
//            devices per profile after the update (byte-wise sum over its GPUs) go to shared memory for phase 2
//   phase 2  ONE THREAD PER NODE: tryAddPod over all pods (planner.go:116-141) with NodeResourcesFit.  The pods are
//            (cpu, memory) int32 pairs in a padded position space, runs of identical slice requests in NewPodSorter order,
//            with minima per 8 / 64 / 512 positions; the whole structure is brought into shared memory once per CTA by an
//            asynchronous bulk copy (cp.async.bulk + mbarrier) that runs under phase 1.  A thread walks its node's runs,
//            skips what the minima rule out and tests eight pods per step in registers; there is no cross-lane
//            communication in this phase, so the warp's 32 nodes advance together with no vote / shuffle on the chain.
//            Every request on this path is for ONE device, so GPU.AddPod (mig/gpu.go:233-247) is "a free device of the
//            profile exists on the node" and the first-fit GPU choice (mig/node.go:201-210) cannot change the result:
//            used + free per GPU is invariant under AddPod.
//   phase 3  8 lanes per node again: partitioningState[n] (planner.go:138) for the nodes that added a pod after an update.
// SliceTracker.Remove (core/tracker.go:67-88) is one bit per pod POSITION (tracker keys are unique on this path), OR-ed in
// shared memory and flushed once per CTA; the final lacking vector is rebuilt from the bitmap by the last CTA.
#include <climits>
#ifdef NOSPLAN_FUSED_PROFILE
#include <cstdio>
#endif

#include "nosplan_dev.cuh"

namespace nsp {

namespace {

constexpr int FT = 640;                               // threads per CTA: one CTA per SM, 80 nodes per phase-1 step
constexpr unsigned FULL = 0xffffffffu;
constexpr unsigned HI = 0x80808080u;
constexpr int CAP_NONE = 1 << 30;                     // capacity of a run without slice request

struct __align__(16) NodeS {                          // phase 1 -> phase 2 -> phase 3, one per node of the tile
    int remc, remm;                                   // remaining cpu / memory in the pods' units (-1: nothing fits)
    int flags;                                        // 1 candidate, 2 geometry updated, 4 every pod fails the filter, 8 a pod was added
    int pad;
    unsigned cap[4];                                  // free devices per profile over the node's GPUs after the update, one byte each
};
static_assert(sizeof(NodeS) == 32, "NodeS layout");

__device__ __forceinline__ unsigned smem_u32(const void* p) { return (unsigned)__cvta_generic_to_shared(p); }
__device__ __forceinline__ void mbar_init(unsigned long long* bar, unsigned count) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(bar)), "r"(count) : "memory");
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
}
__device__ __forceinline__ void mbar_expect_tx(unsigned long long* bar, unsigned bytes) {
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(bar)), "r"(bytes) : "memory");
}
// global (this GPU's own memory) -> shared memory of this CTA, completion counted in bytes on the mbarrier; no multicast
__device__ __forceinline__ void bulk_g2s(void* dst, const void* src, unsigned bytes, unsigned long long* bar) {
    asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(smem_u32(dst)),
                 "l"(src), "r"(bytes), "r"(smem_u32(bar))
                 : "memory");
}
__device__ __forceinline__ void mbar_wait(unsigned long long* bar, unsigned parity) {
    asm volatile(
        "{\n"
        ".reg .pred P1;\n"
        "LAB_WAIT:\n"
        "mbarrier.try_wait.parity.shared::cta.b64 P1, [%0], %1;\n"
        "@P1 bra DONE;\n"
        "bra LAB_WAIT;\n"
        "DONE:\n"
        "}" ::"r"(smem_u32(bar)),
        "r"(parity)
        : "memory");
}

__device__ __forceinline__ unsigned byte_at(unsigned lo, unsigned hi, int p) {   // byte p (< 8) of the pair
    return ((p < 4 ? lo : hi) >> ((p & 3) * 8)) & 0xffu;
}

__device__ __forceinline__ unsigned get_byte(const uint4& v, int p) {
    const unsigned w = p < 8 ? (p < 4 ? v.x : v.y) : (p < 12 ? v.z : v.w);
    return (w >> ((p & 3) * 8)) & 0xffu;
}
__device__ __forceinline__ void add_byte(uint4& v, int p, unsigned d) {     // no carry: the caller checks the counts afterwards
    const unsigned b = d << ((p & 3) * 8);
    if (p < 4) v.x += b; else if (p < 8) v.y += b; else if (p < 12) v.z += b; else v.w += b;
}
__device__ __forceinline__ unsigned byte_spread(unsigned nibble) {           // bit b of the nibble -> 0xff in byte b
    return ((nibble & 1u) * 0xffu) | ((nibble >> 1 & 1u) * 0xff00u) | ((nibble >> 2 & 1u) * 0xff0000u) | ((nibble >> 3 & 1u) * 0xff000000u);
}

struct Acc {                                          // what a thread accumulates over all its tiles
    long long D0 = 0, D1 = 0;                         // sum(alloc - requested) of slice columns j and j + 8
    uint4 mf = make_uint4(0, 0, 0, 0);                // byte-wise max of the free counts seen
    unsigned flags = 0;
    long long place = 0, evals = 0;
    int visited = 0;
    int maxmem = 0;                                   // MPS: largest GPU memory seen (bounds the devices one GPU can hold)
};

// ---- phase 1 for one node: lane j of the 8-lane group = GPU j.  Executed by whole warps (converged).
template <bool WIDE, bool MPS>
__device__ __forceinline__ void node_phase1(const DevProblem& pb, int i, bool node_on, int j, int gsh, const uint4* s_geom,
                                            const uint16_t* s_gmask, const int* s_off, const unsigned (&Lw)[4], int n_req,
                                            NodeS* ns, unsigned* gs, uint4* gfree, Acc& a, unsigned long long* hdr_bar) {
    const int G = pb.G, K = pb.K, R = pb.R;
    const bool fit = pb.filter_mode == NOSPLAN_FILTER_RESOURCES_FIT;
    int ng = 0;
    uint4 used = make_uint4(0, 0, 0, 0), fre = used;
    unsigned um = 0, fm = 0;
    int model = 0;
    long long a0 = 0, r0 = 0, a1 = 0, r1 = 0;
    long long ac = 0, am = 0, ae = 0, rc = 0, rm = 0, re = 0;
    int allowed_pods = 0, npods = 0;
    if (node_on) {
        const int ng_raw = pb.node_gpu_count[i];
        const long long* al = pb.node_alloc + (size_t)i * R;
        const long long* rq = pb.node_req + (size_t)i * R;
        if (j == 0) {
            ac = al[0]; am = al[1]; ae = al[2]; rc = rq[0]; rm = rq[1]; re = rq[2];
            allowed_pods = pb.node_allowed_pods[i]; npods = pb.node_num_pods[i];
        }
        if (j < K) { a0 = al[3 + j]; r0 = rq[3 + j]; }
        if (WIDE && j + 8 < K) { a1 = al[3 + 8 + j]; r1 = rq[3 + 8 + j]; }
        if (j < G) {                                   // the row exists whatever the node's GPU count: one round trip, masked below
            const int x = i * G + j;
            used = pb.gpu_used[x]; fre = pb.gpu_free[x];
            um = pb.gpu_umask[x]; fm = pb.gpu_fmask[x];
            model = pb.gpu_model[x];                   // MPS: the same array holds the GPU's memory in GB (gpu_mem aliases it)
        }
        ng = min(max(ng_raw, 0), G);
        if (j >= ng) { used = make_uint4(0, 0, 0, 0); fre = used; um = fm = 0; model = 0; }
        if (!MPS) model = min(max(model, 0), max(pb.n_models - 1, 0));
    }
    const bool has = j < ng;
    constexpr bool mps = MPS;                         // compile-time: the MIG instantiation carries none of the MPS code
    // ---- k_prepass duties: totals, largest free count, node accounts against the GPU maps
    a.D0 += a0 - r0; a.D1 += a1 - r1;
    if (((used.x | used.y | fre.x | fre.y | (WIDE ? (used.z | used.w | fre.z | fre.w) : 0u)) & 0xe0e0e0e0u) != 0) a.flags |= FREJ_COUNTS;
    // a count without its key would make the plan's key mask depend on which GPU takes a pod: left to the classic kernels
    {
        unsigned nzu = nz_nibble(used.x) | (nz_nibble(used.y) << 4), nzf = nz_nibble(fre.x) | (nz_nibble(fre.y) << 4);
        if (WIDE) { nzu |= (nz_nibble(used.z) << 8) | (nz_nibble(used.w) << 12); nzf |= (nz_nibble(fre.z) << 8) | (nz_nibble(fre.w) << 12); }
        if ((nzu & ~um) | (nzf & ~fm)) a.flags |= FREJ_COUNTS;
    }
    a.mf.x = __vmaxu4(a.mf.x, fre.x); a.mf.y = __vmaxu4(a.mf.y, fre.y);
    if (WIDE) { a.mf.z = __vmaxu4(a.mf.z, fre.z); a.mf.w = __vmaxu4(a.mf.w, fre.w); }
    {
        unsigned fx = fre.x, fy = fre.y, ux = used.x, uy = used.y, fz = fre.z, fw = fre.w, uz = used.z, uw = used.w;
#pragma unroll
        for (int o = 1; o < 8; o <<= 1) {              // byte-wise sums over the node's GPUs (counts <= 31: no carry)
            fx += __shfl_xor_sync(FULL, fx, o); fy += __shfl_xor_sync(FULL, fy, o);
            ux += __shfl_xor_sync(FULL, ux, o); uy += __shfl_xor_sync(FULL, uy, o);
            if (WIDE) {
                fz += __shfl_xor_sync(FULL, fz, o); fw += __shfl_xor_sync(FULL, fw, o);
                uz += __shfl_xor_sync(FULL, uz, o); uw += __shfl_xor_sync(FULL, uw, o);
            }
        }
        bool bad = false;
        if (j < K) { const long long sf = byte_at(fx, fy, j), su = byte_at(ux, uy, j); bad = (a0 - r0 < sf) || (r0 > su); }
        if (WIDE && j + 8 < K) { const long long sf = byte_at(fz, fw, j), su = byte_at(uz, uw, j); bad |= (a1 - r1 < sf) || (r1 > su); }
        if (bad) a.flags |= FREJ_ACCOUNTS;
    }
    // ---- mig.GPU.UpdateGeometryFor for this lane's GPU + AllowsGeometry(current) for HasFreeCapacity
    if (hdr_bar) mbar_wait(hdr_bar, 0);               // the tables arrive by bulk copy, under the loads above
    const uint4 cnt0 = make_uint4(used.x + fre.x, used.y + fre.y, used.z + fre.z, used.w + fre.w);
    const unsigned gm0 = um | fm;
    unsigned best = 0;
    bool allowed = false;
    int nrows = 0, row0 = 0;
    if (has && !mps && pb.n_models > 0) { row0 = s_off[model]; nrows = s_off[model + 1] - row0; }
    // ---- MPS: slicing.GPU.UpdateGeometryFor (pkg/gpu/slicing/gpu.go:162-232), thread-serial over the profiles of this lane's GPU.
    //      Under the hypothesis every requested profile is missing by more than a GPU can hold, so `min(missing, spare / size)`
    //      is `spare / size`; the rest is mps_gpu_update of sweep.cu line by line: spare capacity first, then ALL originally-free
    //      keys are deleted (also what was just created on one of them), the profile is created again in what that freed, and
    //      the deleted keys are restored all-or-nothing per profile in id order.
    unsigned fm_new = fm;
    bool mps_upd = false, mps_free = false;
    if (mps && has) {
        const int mem = model;
        a.maxmem = max(a.maxmem, mem);
        auto mem_used = [&](const uint4& f) {
            int t = 0;
            for (int p = 0; p < K; ++p) t += pb.pmem[p] * (int)(get_byte(used, p) + get_byte(f, p));
            return t;
        };
        mps_free = fm != 0 || (mem - mem_used(fre) >= 1);               // HasFreeCapacity (slicing/gpu.go:131-141)
        const uint4 ofree = fre;
        const unsigned omask = fm & ((1u << K) - 1u);
        const uint4 keep = make_uint4(~byte_spread(omask & 15u), ~byte_spread(omask >> 4 & 15u), ~byte_spread(omask >> 8 & 15u), ~byte_spread(omask >> 12 & 15u));
        for (unsigned m = pb.fp_rset; m; m &= m - 1) {
            const int mp = __ffs(m) - 1, msize = pb.pmem[mp];
            int spare = mem - mem_used(fre);
            if (spare >= 1) { const int c = spare / msize; if (c > 0) { add_byte(fre, mp, (unsigned)c); fm_new |= 1u << mp; mps_upd = true; } }
            fre.x &= keep.x; fre.y &= keep.y; fre.z &= keep.z; fre.w &= keep.w;
            fm_new &= ~omask;
            spare = mem - mem_used(fre);
            if (spare >= 1) { const int c = spare / msize; if (c > 0) { add_byte(fre, mp, (unsigned)c); fm_new |= 1u << mp; mps_upd = true; spare -= c * msize; } }
            for (unsigned rest = omask; rest; rest &= rest - 1) {
                const int k = __ffs(rest) - 1;
                const int of = (int)get_byte(ofree, k), need = pb.pmem[k] * of;
                if (spare >= need) { add_byte(fre, k, (unsigned)of); fm_new |= 1u << k; spare -= need; }
            }
            // a count beyond 31 would carry in the byte-wise sums below (and beyond 255 has already wrapped): not for this kernel
            if (((fre.x | fre.y | fre.z | fre.w) & 0xe0e0e0e0u) != 0 || spare < -(1 << 24)) a.flags |= FREJ_COUNTS;
        }
    }
    uint4 fre_mps = fre;                              // MPS: the GPU after the update (kept only for a candidate node, below)
    if (mps) fre = make_uint4(cnt0.x - used.x, cnt0.y - used.y, cnt0.z - used.z, cnt0.w - used.w);
    for (int r = 0; r < nrows; ++r) {
        const uint4 row = s_geom[row0 + r];
        const unsigned tx = (row.x | HI) - used.x, ty = (row.y | HI) - used.y;   // bit 7 of a byte: row >= used
        unsigned ok = tx & ty;
        unsigned sc = __dp4a(tx & ~HI, Lw[0], 0u);
        sc = __dp4a(ty & ~HI, Lw[1], sc);
        if (WIDE) {
            const unsigned tz = (row.z | HI) - used.z, tw = (row.w | HI) - used.w;
            ok &= tz & tw;
            sc = __dp4a(tz & ~HI, Lw[2], sc);
            sc = __dp4a(tw & ~HI, Lw[3], sc);
        }
        if ((ok & HI) == HI && sc) best = max(best, (sc << 5) | (unsigned)(31 - r));   // strict > in table order
        if (fm == 0) {                                // AllowsGeometry(current) only decides HasFreeCapacity where no free key exists
            bool same = row.x == cnt0.x && row.y == cnt0.y;
            if (WIDE) same = same && row.z == cnt0.z && row.w == cnt0.w;
            allowed |= same && s_gmask[row0 + r] == gm0;
        }
    }
    const bool has_free = has && (mps ? mps_free : (fm != 0 || !allowed));
    const bool candidate = (__ballot_sync(FULL, has_free) >> gsh & 0xffu) != 0;
    int br = 0xffff;
    if (candidate && best) {
        br = row0 + 31 - (int)(best & 31u);
        const uint4 row = s_geom[br];
        fre = make_uint4(row.x - used.x, row.y - used.y, WIDE ? row.z - used.z : 0u, WIDE ? row.w - used.w : 0u);   // zero kept (:141)
    }
    bool upd_g = br != 0xffff;
    if (mps && candidate && has) {
        const bool changed = fre_mps.x != fre.x || fre_mps.y != fre.y || fre_mps.z != fre.z || fre_mps.w != fre.w || fm_new != fm;
        fre = fre_mps; fm = fm_new;
        upd_g = mps_upd;
        if (changed) br = 0xfffe;                     // phase 3 takes the row from the per-GPU copy of `fre`, not from a table row
    }
    const bool upd = (__ballot_sync(FULL, upd_g) >> gsh & 0xffu) != 0;
    // free devices per profile on the node after the update
    unsigned cx = has ? fre.x : 0u, cy = has ? fre.y : 0u, cz = (WIDE && has) ? fre.z : 0u, cw = (WIDE && has) ? fre.w : 0u;
#pragma unroll
    for (int o = 1; o < 8; o <<= 1) {
        cx += __shfl_xor_sync(FULL, cx, o); cy += __shfl_xor_sync(FULL, cy, o);
        if (WIDE) { cz += __shfl_xor_sync(FULL, cz, o); cw += __shfl_xor_sync(FULL, cw, o); }
    }
    if (candidate && has) a.evals += mps ? (long long)n_req : (long long)nrows * n_req;
    if (node_on) {
        // partitioningState[n] as it was (planner.go:72); phase 3 overwrites the rows of nodes that add a pod after an update
        if (has) {
            const int x = i * G + j;
            pb.out_count[x] = cnt0;
            pb.out_mask[x] = (uint16_t)gm0;
        }
        *gs = (unsigned)br | ((mps ? (um | fm) : um) << 16);
        if (mps) *gfree = fre;
        if (j == 0) {
            NodeS s;
            s.remc = s.remm = F_REM_MAX;
            bool all_fail = pb.filter_mode == NOSPLAN_FILTER_REJECT_ALL;
            if (fit) {
                const long long dc = ac - rc, dm = am - rm;
                const long long qc = dc >> pb.fp_cpu_shift, qm = dm >> pb.fp_mem_shift;
                if (candidate && (qc > F_REM_MAX || qm > F_REM_MAX)) a.flags |= FREJ_REM;
                s.remc = dc < 0 ? -1 : (int)min(qc, (long long)F_REM_MAX);
                s.remm = dm < 0 ? -1 : (int)min(qm, (long long)F_REM_MAX);
                if (ae - re < 0) s.remc = -1;               // no pod requests ephemeral storage: 0 <= negative fails for every pod
                all_fail = all_fail || !(npods + 1 <= allowed_pods);
            }
            s.flags = (candidate ? 1 : 0) | (upd ? 2 : 0) | (all_fail ? 4 : 0);
            s.pad = 0;
            s.cap[0] = cx; s.cap[1] = cy; s.cap[2] = cz; s.cap[3] = cw;
            *ns = s;
            if (candidate) ++a.visited;
        }
    }
}

// ---- phase 2: lane = node; the warp's 32 nodes take the runs in lockstep (the run record is warp-uniform), each lane walks
//      its own way through the run.  Returns the number of pods added on this lane's node.
// loads of the pod structure: shared-memory instructions when the whole blob is staged; otherwise generic loads (the upper
// levels may sit in shared memory while the lower ones are read from global memory through L1)
template <bool SMEM>
__device__ __forceinline__ uint2 ld_pair(const uint2* p) {
    if (SMEM) {
        uint2 v;
        asm volatile("ld.shared.v2.u32 {%0, %1}, [%2];" : "=r"(v.x), "=r"(v.y) : "r"(smem_u32(p)));
        return v;
    }
    return *p;
}
template <bool SMEM>
__device__ __forceinline__ uint4 ld_quad(const uint4* p) {
    if (SMEM) {
        uint4 v;
        asm volatile("ld.shared.v4.u32 {%0, %1, %2, %3}, [%4];" : "=r"(v.x), "=r"(v.y), "=r"(v.z), "=r"(v.w) : "r"(smem_u32(p)));
        return v;
    }
    return *p;
}

template <bool WIDE, bool FIT, bool SMEM>
__device__ __forceinline__ int warp_phase2(const NodeS& s, bool on, const FRun* s_runs,
                                           const unsigned long long* s_prof_runs, const uint2* L3, const uint2* L2, const uint2* L1,
                                           const uint2* SFX, const uint4* CUM, const uint2* PODS, unsigned* s_bits, int* s_pref) {
    const bool active = on && (s.flags & 5) == 1;
    int remc = -1, remm = -1;
    unsigned cv0 = 0, cv1 = 0, cv2 = 0, cv3 = 0;
    unsigned long long live = 0;
    if (active) {
        remc = s.remc; remm = s.remm;
        cv0 = s.cap[0]; cv1 = s.cap[1]; cv2 = s.cap[2]; cv3 = s.cap[3];
        unsigned feas = nz_nibble(cv0) | (nz_nibble(cv1) << 4);
        if (WIDE) feas |= (nz_nibble(cv2) << 8) | (nz_nibble(cv3) << 12);
        live = s_prof_runs[KP];
        for (unsigned f = feas; f; f &= f - 1) live |= s_prof_runs[__ffs(f) - 1];
    }
    // runs some node of the warp can still use, in NewPodSorter order
    unsigned long long todo = ((unsigned long long)__reduce_or_sync(FULL, (unsigned)(live >> 32)) << 32) | __reduce_or_sync(FULL, (unsigned)live);
    int added = 0;
    while (todo) {
        const int ri = __ffsll((long long)todo) - 1;
        todo &= todo - 1;
        const FRun fr = s_runs[ri];                                        // the same record for every lane
        const int p = fr.p;
        int cap = 0;
        if ((live >> ri & 1) && fr.min_c <= remc && fr.min_m <= remm) {    // else: no pod of the run can pass on this node (any more)
            if (p >= 0) {                                                  // GPU.AddPod feasibility: a free device of the profile (q == 1)
                const unsigned w = p < 8 ? (p < 4 ? cv0 : cv1) : (p < 12 ? cv2 : cv3);
                cap = (int)((w >> ((p & 3) * 8)) & 0xffu);
            } else cap = CAP_NONE;
        }
        if (!__any_sync(FULL, cap > 0)) continue;
        const int cap0 = cap;
        int cur = fr.pos0 >> 3;
        const int end = cur + fr.nblk * 8;
        if (cap > 0) {
            // the head of the run, whole groups of eight at a time: while the cumulative cost and count of the eligible pods
            // through a group fit what the node has left, every one of them is placed (each prefix of a sum that fits, fits)
            // and no pod has to be looked at.  The groups taken are recorded per run (s_pref) and turned into bits at the end.
            int k = 0;
            uint4 took = make_uint4(0, 0, 0, 0);
            while (cur + k < end) {
                const uint4 cu = ld_quad<SMEM>(CUM + cur + k);
                if ((int)cu.z > cap || (FIT && ((int)cu.x > remc || (int)cu.y > remm))) break;
                took = cu; ++k;
            }
            if (k > 0) {
                if (FIT) { remc -= (int)took.x; remm -= (int)took.y; }
                cap -= (int)took.z;
                if (took.z) atomicMax(&s_pref[ri], k);
                cur += k;
            }
        }
        for (;;) {
            // next group of eight pods this node has to look at: skip what the minima rule out
            int sub = -1;
            while (cap > 0 && cur < end) {
                const uint2 sf = ld_pair<SMEM>(SFX + cur);
                if ((int)sf.x > remc || (int)sf.y > remm) { cur = end; break; }      // nothing from here to the end of the run fits
                if ((cur & 63) == 0) {
                    const uint2 m = ld_pair<SMEM>(L3 + (cur >> 6));
                    if ((int)m.x > remc || (int)m.y > remm) { cur += 64; continue; }
                }
                if ((cur & 7) == 0) {
                    const uint2 m = ld_pair<SMEM>(L2 + (cur >> 3));
                    if ((int)m.x > remc || (int)m.y > remm) { cur += 8; continue; }
                }
                const uint2 m = ld_pair<SMEM>(L1 + cur);
                if ((int)m.x > remc || (int)m.y > remm) { ++cur; continue; }
                sub = cur++;
                break;
            }
            if (!__any_sync(FULL, sub >= 0)) break;
            if (sub >= 0) {
                // eight pods in order; a pod that does not fit now never fits later on this node: the state only gets tighter
                const uint4* pp = reinterpret_cast<const uint4*>(PODS + (size_t)sub * 8);
                const uint4 q0 = ld_quad<SMEM>(pp), q1 = ld_quad<SMEM>(pp + 1), q2 = ld_quad<SMEM>(pp + 2), q3 = ld_quad<SMEM>(pp + 3);
                unsigned bits = 0;
#define NOS_TRY(C, M, B)                                                              \
    if (cap > 0 && (int)(C) <= remc && (int)(M) <= remm) {                             \
        if (FIT) { remc -= (int)(C); remm -= (int)(M); }                               \
        --cap; bits |= 1u << (B);                                                      \
    }
                NOS_TRY(q0.x, q0.y, 0) NOS_TRY(q0.z, q0.w, 1) NOS_TRY(q1.x, q1.y, 2) NOS_TRY(q1.z, q1.w, 3)
                NOS_TRY(q2.x, q2.y, 4) NOS_TRY(q2.z, q2.w, 5) NOS_TRY(q3.x, q3.y, 6) NOS_TRY(q3.z, q3.w, 7)
#undef NOS_TRY
                if (bits) atomicOr(&s_bits[sub >> 2], bits << ((sub & 3) * 8));            // tracker.Remove
            }
        }
        const int n = cap0 - cap;                                          // devices of the run's profile taken on this node
        if (n > 0) {
            added += n;
            if (p >= 0) {
                const unsigned d = (unsigned)n << ((p & 3) * 8);
                if (p < 4) cv0 -= d; else if (p < 8) cv1 -= d; else if (p < 12) cv2 -= d; else cv3 -= d;
            }
        }
    }
    return added;
}

}  // namespace

constexpr int FUSED_XHDR = 240;                       // bytes in front of the bitmap in a rank's exchange block (NCCL transport)
constexpr int FUSED_LL_HDR = 58;                      // 32-bit words in front of the bitmap in the flag-in-word form (peer transport)

// Was the hypothesis right?  `tot`, `maxfree`, `rej`, `acc` are the grid-wide (single GPU) or job-wide (slice mode, folded)
// values; the placed bitmap is the OR of `world` bitmaps `stride` bytes apart.  Writes the statistics and the verdict
// (path 1: the plan stands, path 2: rejected, reason bits in pad).  Whole CTA; s_l0 / s_dec / s_e are shared scratch.
__device__ __forceinline__ void fused_verify(const DevProblem& pb, const FRun* runs, const long long* tot, const int* maxfree, int rej,
                                             const unsigned long long* acc, const unsigned char* bitmap, size_t stride, int world,
                                             bool clear_bitmap, long long* s_l0, long long* s_dec, long long* s_e) {
    const int tid = threadIdx.x, nthreads = blockDim.x;
    const int n_words = pb.fp_pad_pods >> 5;
    if (tid < KP) { s_l0[tid] = 0; s_dec[tid] = 0; }
    __syncthreads();
    for (int r = tid; r < pb.fp_n_runs; r += nthreads) {
        const FRun fr = runs[r];
        long long e = 0;
        if (fr.p >= 0) {                                              // NewSliceTracker (core/tracker.go:34-57): lacking = request - available
            const long long avail = max(0ll, *reinterpret_cast<const volatile long long*>(&tot[fr.p]));   // SubtractNonNegative (pkg/resource/resource.go:76-88)
            e = max(0ll, (long long)fr.q - avail);
            if (e) atomicAdd(reinterpret_cast<unsigned long long*>(&s_l0[fr.p]), (unsigned long long)(e * fr.n_real));
        }
        s_e[r] = e;
    }
    __syncthreads();
    for (int w = tid; w < n_words; w += nthreads) {                   // a 32-position word lies inside one run (runs are 64-aligned)
        unsigned v = 0;
        for (int k = 0; k < world; ++k) v |= *(reinterpret_cast<const volatile unsigned*>(bitmap + k * stride) + w);
        if (!v) continue;
        if (clear_bitmap) const_cast<unsigned*>(reinterpret_cast<const unsigned*>(bitmap))[w] = 0;   // clean for the next fused run of this problem
        const int c = __popc(v);
        int r = 0;
        while (r + 1 < pb.fp_n_runs && runs[r + 1].pos0 <= (w << 5)) ++r;
        if (runs[r].p >= 0 && s_e[r]) atomicAdd(reinterpret_cast<unsigned long long*>(&s_dec[runs[r].p]), (unsigned long long)(c * s_e[r]));
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
                int sat = pb.sat_m[lane];
                if (pb.kind != NOSPLAN_KIND_MIG)        // devices of this size the largest GPU can hold (fast_applicable of sweep.cu)
                    sat = pb.pmem[lane] > 0 ? *reinterpret_cast<const volatile int*>(&pb.gflags[1]) / pb.pmem[lane] : (1 << 30);
                const long long m = max(sat, *reinterpret_cast<const volatile int*>(&maxfree[lane]));
                ok = ok && le > 0 && le > (long long)(pb.G + 2) * m && l0 < (1ll << 30);
            }
        }
        if (!__all_sync(FULL, ok)) rej |= FREJ_HYPOTHESIS;
        const int lack_start = __popc(__ballot_sync(FULL, lane < KP && l0 > 0));
        const int lack_end = __popc(__ballot_sync(FULL, lane < KP && le > 0));
        if (lane == 0) {
            DevStats* st = pb.stats;
            const volatile unsigned long long* va = acc;
            const long long vis = (long long)va[2];
            st->visited_nodes = vis;
            st->candidates = vis * pb.P;
            st->geometry_evals = (long long)va[1];
            st->placements = (long long)va[0];
            st->candidate_nodes = vis;
            st->lacking_at_start = lack_start;
            st->lacking_at_end = lack_end;
            st->path = rej ? 2 : 1;
            st->pad = rej;
        }
    }
}

struct FusedLaunch { int tile_nodes, stage_bytes; };

template <bool WIDE, bool MPS>
__device__ __forceinline__ void fused_body(const DevProblem& pb, const FusedLaunch fl, unsigned char* s_dyn, unsigned long long (*s_wD)[KP],
                                           unsigned (*s_wmf)[4], unsigned long long (*s_wcnt)[4], int* s_flags, int* s_pref, unsigned long long* s_bar) {
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5, j = lane & 7, gsh = (lane >> 3) * 8;
    const int n_words = pb.fp_pad_pods >> 5;
    // dynamic shared memory: [placed bitmap][staged prefix of the blob: tables, runs, minima, pods][per-node state][per-GPU state]
    unsigned* s_bits = reinterpret_cast<unsigned*>(s_dyn);
    const size_t o_blob = ((size_t)n_words * 4 + 15) & ~(size_t)15;
    const unsigned char* s_blob = s_dyn + o_blob;
    NodeS* s_node = reinterpret_cast<NodeS*>(s_dyn + o_blob + (size_t)fl.stage_bytes);
    unsigned* s_gpu = reinterpret_cast<unsigned*>(s_node + fl.tile_nodes);
    uint4* s_gfree = reinterpret_cast<uint4*>(s_gpu + 8 * fl.tile_nodes);        // MPS only: every GPU's free counts after the update
    const uint4* s_geom = reinterpret_cast<const uint4*>(s_blob);
    const uint16_t* s_gmask = reinterpret_cast<const uint16_t*>(s_blob + pb.fp_off[0]);
    const int* s_off = reinterpret_cast<const int*>(s_blob + pb.fp_off[1]);
    const FRun* s_runs = reinterpret_cast<const FRun*>(s_blob + pb.fp_off[2]);
    const unsigned long long* s_prof_runs = reinterpret_cast<const unsigned long long*>(s_blob + pb.fp_off[3]);
    // levels of the pod structure: from shared memory where staged, else from global memory (L1-cached)
    auto level = [&](int k) {
        const unsigned char* base = pb.fp_off[k + 1] <= fl.stage_bytes ? s_blob : pb.fp_blob;
        return reinterpret_cast<const uint2*>(base + pb.fp_off[k]);
    };
    const uint2* L3 = level(FP_L3); const uint2* L2 = level(FP_L2); const uint2* L1 = level(FP_L1); const uint2* SFX = level(FP_SFX);
    const uint4* CUM = reinterpret_cast<const uint4*>(level(FP_CUM)); const uint2* PODS = level(FP_PODS);
    const bool smem_all = pb.fp_off[FP_END] <= fl.stage_bytes, fit = pb.filter_mode == NOSPLAN_FILTER_RESOURCES_FIT;

    const int n_req = __popc(pb.fp_rset);
    unsigned Lw[4];                                   // lacking profiles as a 0/1 byte per profile (the dp4a weights)
#pragma unroll
    for (int w = 0; w < 4; ++w) {
        unsigned v = 0;
#pragma unroll
        for (int b = 0; b < 4; ++b) v |= ((pb.fp_rset >> (w * 4 + b)) & 1u) << (8 * b);
        Lw[w] = v;
    }
    Acc a;
    const int n_nodes = pb.node_hi - pb.node_lo;
    const int n_tiles = (n_nodes + fl.tile_nodes - 1) / fl.tile_nodes;
    bool hdr_ready = false, staged_ready = fl.stage_bytes <= pb.fp_off[FP_L3];
#ifdef NOSPLAN_FUSED_PROFILE
    long long tk[8]; tk[0] = clock64();
#define TICK(k) tk[k] = clock64()
#else
#define TICK(k)
#endif
    for (int tile = blockIdx.x; tile < n_tiles; tile += gridDim.x) {
        const int base = pb.node_lo + tile * fl.tile_nodes;
        const int count = min(fl.tile_nodes, pb.node_hi - base);
        // ---- phase 1
        for (int g0 = 0; g0 < count * 8; g0 += FT) {
            const int g = g0 + tid, nl = g >> 3;
            const bool on = nl < count;
            node_phase1<WIDE, MPS>(pb, base + nl, on, j, gsh, s_geom, s_gmask, s_off, Lw, n_req, &s_node[on ? nl : 0], &s_gpu[on ? g : 0],
                              &s_gfree[on ? g : 0], a, hdr_ready ? nullptr : &s_bar[0]);
            hdr_ready = true;
        }
        __syncthreads();
        TICK(1);
        if (!staged_ready) { mbar_wait(&s_bar[1], 0); staged_ready = true; }
        TICK(2);
        // ---- phase 2: node nl on lane nl / NW of warp nl % NW -- a small tile spreads over all warps, a few lanes each, so
        //      that a warp's chain is the longest of a few nodes, not of 32
        {
            constexpr int NW = FT / 32;
            const int nl = lane * NW + warp;
            const bool on = nl < count;
            if (__any_sync(FULL, on)) {
                const NodeS& ns = s_node[on ? nl : 0];
                int added;
                if (fit) added = smem_all ? warp_phase2<WIDE, true, true>(ns, on, s_runs, s_prof_runs, L3, L2, L1, SFX, CUM, PODS, s_bits, s_pref)
                                          : warp_phase2<WIDE, true, false>(ns, on, s_runs, s_prof_runs, L3, L2, L1, SFX, CUM, PODS, s_bits, s_pref);
                else added = smem_all ? warp_phase2<WIDE, false, true>(ns, on, s_runs, s_prof_runs, L3, L2, L1, SFX, CUM, PODS, s_bits, s_pref)
                                      : warp_phase2<WIDE, false, false>(ns, on, s_runs, s_prof_runs, L3, L2, L1, SFX, CUM, PODS, s_bits, s_pref);
                if (on && added > 0) { s_node[nl].flags |= 8; a.place += added; }
            }
        }
        __syncthreads();
        TICK(3);
        // ---- phase 3: the geometry after the update, for nodes that added a pod (planner.go:138)
        for (int g = tid; g < count * 8; g += FT) {
            const int nl = g >> 3, jj = g & 7;
            if ((s_node[nl].flags & 10) != 10) continue;
            const unsigned v = s_gpu[g];
            const int br = (int)(v & 0xffffu);
            if (br == 0xffff) continue;
            const int x = (base + nl) * pb.G + jj;
            if (MPS && br == 0xfffe) {                // MPS: used + free after the update, the keys of both maps
                const uint4 u = pb.gpu_used[x], f = s_gfree[g];
                pb.out_count[x] = make_uint4(u.x + f.x, u.y + f.y, u.z + f.z, u.w + f.w);
                pb.out_mask[x] = (uint16_t)(v >> 16);
                continue;
            }
            pb.out_count[x] = s_geom[br];
            pb.out_mask[x] = (uint16_t)((v >> 16) | s_gmask[br]);
        }
        __syncthreads();
        TICK(4);
    }
    if (!hdr_ready) mbar_wait(&s_bar[0], 0);          // a CTA without tiles still has to see its copies land before it exits
    if (!staged_ready) mbar_wait(&s_bar[1], 0);

    // ---- reduction of what the threads accumulated: one shared-memory slot per warp
    if (__any_sync(FULL, a.flags != 0)) { const unsigned f = __reduce_or_sync(FULL, a.flags); if (lane == 0) atomicOr(&s_flags[0], (int)f); }
    if (MPS) { const int mm = __reduce_max_sync(FULL, a.maxmem); if (lane == 0 && mm > 0) atomicMax(&pb.gflags[1], mm); }
#pragma unroll
    for (int o = 8; o < 32; o <<= 1) {                 // lanes with the same j across the warp's four groups
        a.D0 += __shfl_xor_sync(FULL, a.D0, o); a.D1 += __shfl_xor_sync(FULL, a.D1, o);
    }
    if (lane < 8) { s_wD[warp][lane] = (unsigned long long)a.D0; s_wD[warp][8 + lane] = (unsigned long long)a.D1; }
    {   // largest free count per profile: byte-wise maximum over the warp, one redux per word half
        unsigned mx = 0, my = 0, mz = 0, mw = 0;
#pragma unroll
        for (int b = 0; b < 4; ++b) {
            mx |= __reduce_max_sync(FULL, (a.mf.x >> (8 * b)) & 0xffu) << (8 * b);
            my |= __reduce_max_sync(FULL, (a.mf.y >> (8 * b)) & 0xffu) << (8 * b);
            if (WIDE) {
                mz |= __reduce_max_sync(FULL, (a.mf.z >> (8 * b)) & 0xffu) << (8 * b);
                mw |= __reduce_max_sync(FULL, (a.mf.w >> (8 * b)) & 0xffu) << (8 * b);
            }
        }
        if (lane == 0) { s_wmf[warp][0] = mx; s_wmf[warp][1] = my; s_wmf[warp][2] = mz; s_wmf[warp][3] = mw; }
    }
    {   // placements, geometry evaluations, visited nodes: a warp's share fits 32 bits (<= 32 nodes per tile, <= 2^20 tiles)
        const unsigned pl = __reduce_add_sync(FULL, (unsigned)a.place), ev = __reduce_add_sync(FULL, (unsigned)a.evals);
        const unsigned vi = __reduce_add_sync(FULL, (unsigned)a.visited);
        if (lane == 0) { s_wcnt[warp][0] = pl; s_wcnt[warp][1] = ev; s_wcnt[warp][2] = vi; }
    }
    __syncthreads();
    // ---- this CTA's contribution to the grid-wide state
    constexpr int NW = FT / 32;
    if (tid < KP) {
        unsigned long long d = 0;
        int mfree = 0;
        for (int w = 0; w < NW; ++w) {
            d += s_wD[w][tid];
            mfree = max(mfree, (int)((s_wmf[w][tid >> 2] >> ((tid & 3) * 8)) & 0xffu));
        }
        if (d) atomicAdd(reinterpret_cast<unsigned long long*>(&pb.tot[tid]), d);
        if (mfree) atomicMax(&pb.maxfree[tid], mfree);
    } else if (tid >= 32 && tid < 35) {
        unsigned long long c = 0;
        for (int w = 0; w < NW; ++w) c += s_wcnt[w][tid - 32];
        if (c) atomicAdd(&pb.acc[tid - 32], c);
    } else if (tid == 64 && s_flags[0]) atomicOr(&pb.wave[5], s_flags[0]);
    // the heads taken whole: their eligible pods are placed pods (the cumulative record carries the group's eligibility mask)
    for (int r = 0; r < pb.fp_n_runs; ++r) {
        const int k = s_pref[r], x0 = s_runs[r].pos0 >> 3;
        for (int x = tid; x < k; x += FT) {
            const unsigned m = CUM[x0 + x].w;
            if (m) atomicOr(&s_bits[(x0 + x) >> 2], m << (((x0 + x) & 3) * 8));
        }
    }
    __syncthreads();
    for (int w = tid; w < n_words; w += FT) {        // every node places the same first pods: most words are already set by the CTAs
        const unsigned b = s_bits[w];                 // that finished earlier, and a read costs no turn in the word's atomic queue
        if (b && (b & ~__ldcg(&pb.placed[w]))) atomicOr(&pb.placed[w], b);
    }
#ifdef NOSPLAN_FUSED_PROFILE
    TICK(5);
    if (tid == 0 && (blockIdx.x == 0 || blockIdx.x == gridDim.x / 2 || blockIdx.x == gridDim.x - 1))
        printf("[fused profile] cta %d: phase1 %lld  wait-stage %lld  phase2 %lld  phase3 %lld  epilogue %lld cycles\n", blockIdx.x,
               tk[1] - tk[0], tk[2] - tk[1], tk[3] - tk[2], tk[4] - tk[3], tk[5] - tk[4]);
#endif
}

__global__ void __launch_bounds__(FT, 1) k_plan_fused(DevProblem pb, FusedLaunch fl) {
    extern __shared__ __align__(16) unsigned char s_dyn[];
    __shared__ unsigned long long s_wD[FT / 32][KP];
    __shared__ unsigned s_wmf[FT / 32][4];
    __shared__ unsigned long long s_wcnt[FT / 32][4];
    __shared__ int s_flags[1];
    __shared__ int s_pref[FRUN_MAX];                               // per run: groups of eight taken whole from its head by some node of this CTA
    __shared__ long long s_l0[KP], s_dec[KP];
    __shared__ long long s_e[FRUN_MAX];
    __shared__ __align__(8) unsigned long long s_bar[2];
    const int tid = threadIdx.x;
    const int n_words = pb.fp_pad_pods >> 5;

    // tables, runs and the pod structure (this GPU's own memory) on their way into shared memory: the header lands first
    // and is waited for in front of the geometry loop, the rest in front of the pod phase
    if (tid == 0) {
        mbar_init(&s_bar[0], 1);
        mbar_init(&s_bar[1], 1);
        unsigned char* dst = s_dyn + (((size_t)n_words * 4 + 15) & ~(size_t)15);
        const unsigned hdr = (unsigned)min(fl.stage_bytes, pb.fp_off[FP_L3]), all = (unsigned)fl.stage_bytes;
        mbar_expect_tx(&s_bar[0], hdr);
        bulk_g2s(dst, pb.fp_blob, hdr, &s_bar[0]);
        if (all > hdr) {
            mbar_expect_tx(&s_bar[1], all - hdr);
            for (unsigned o = hdr; o < all; o += 32768u) bulk_g2s(dst + o, pb.fp_blob + o, min(32768u, all - o), &s_bar[1]);
        }
    }
    for (int w = tid; w < n_words; w += FT) reinterpret_cast<unsigned*>(s_dyn)[w] = 0;
    if (tid == 0) s_flags[0] = 0;
    if (tid < FRUN_MAX) s_pref[tid] = 0;
    // the next run's copy of the small per-run state (see DevProblem::zero_next)
    if (blockIdx.x == 0) for (int x = tid; x < pb.zero_next_n8; x += FT) pb.zero_next[x] = 0;
    __syncthreads();

    if (pb.kind == NOSPLAN_KIND_MIG) {
        if (pb.K > 8) fused_body<true, false>(pb, fl, s_dyn, s_wD, s_wmf, s_wcnt, s_flags, s_pref, s_bar);
        else fused_body<false, false>(pb, fl, s_dyn, s_wD, s_wmf, s_wcnt, s_flags, s_pref, s_bar);
    } else {
        if (pb.K > 8) fused_body<true, true>(pb, fl, s_dyn, s_wD, s_wmf, s_wcnt, s_flags, s_pref, s_bar);
        else fused_body<false, true>(pb, fl, s_dyn, s_wD, s_wmf, s_wcnt, s_flags, s_pref, s_bar);
    }

    // ---- the CTA that finishes last either verifies the hypothesis on the final state (single GPU: what k_plan_sweep does for
    //      the classic path) or, in slice mode, ships this rank's part of that state to every rank in ONE block
    if (!last_block_done(&pb.next_node[2])) return;
    const FRun* s_runs = reinterpret_cast<const FRun*>(s_dyn + (((size_t)n_words * 4 + 15) & ~(size_t)15) + pb.fp_off[2]);
    if (pb.xf_on && pb.xf_box.world > 0) {
        // ---- peer-memory transport, flag-in-word: the rank's state travels as 8-byte words {value, epoch}, stored straight into
        //      every rank's mailbox over NVLink (an 8-byte store lands whole), so there is no fence and no separate flag: a
        //      consumer polls each word until it carries this Plan()'s epoch.  Words: 0..31 sum(alloc - requested) per slice
        //      column (lo, hi), 32..47 largest free count per profile, 48 reject bits, 49..56 counters (lo, hi), 58.. placed
        //      bitmap.  This CTA is producer and consumer: the sliced Plan() stays ONE kernel per rank.  Every rank's kernel
        //      reaches its stores independently of any other rank, so the wait cannot deadlock; it is bounded, and a peer
        //      that never shows up is reported.
        const int par = pb.xf_epoch & 1, W = pb.xf_box.world, D = FUSED_LL_HDR + n_words;
        const unsigned epoch = pb.xf_epoch;
        for (int i = tid; i < D; i += FT) {
            unsigned v = 0;
            if (i < 32) v = __ldcg(reinterpret_cast<const unsigned*>(pb.tot) + i);
            else if (i < 48) v = (unsigned)__ldcg(&pb.maxfree[i - 32]);
            else if (i == 48) v = (unsigned)__ldcg(&pb.wave[5]);
            else if (i < 57) v = __ldcg(reinterpret_cast<const unsigned*>(pb.acc) + (i - 49));
            else if (i >= FUSED_LL_HDR) { v = __ldcg(&pb.placed[i - FUSED_LL_HDR]); if (v) pb.placed[i - FUSED_LL_HDR] = 0; }   // clean for the next run
            for (int r = 0; r < W; ++r)
                reinterpret_cast<uint2*>(pb.xf_box.base[r] + box_ll(par, pb.xf_box.rank, pb.xf_box.xb_max))[i] = make_uint2(v, epoch);
        }
        // consume: W words per index from this rank's own mailbox (local memory)
        const unsigned char* me = pb.xf_box.base[pb.xf_box.rank];
        unsigned* s_hdr = reinterpret_cast<unsigned*>(s_dyn + (((size_t)n_words * 4 + 15) & ~(size_t)15) + pb.fp_off[FP_L3]);   // staged minima: no longer needed
        unsigned* s_or = reinterpret_cast<unsigned*>(s_dyn);                                                              // this CTA's placed bits: flushed long ago
        bool ok = true;
        for (int i = tid; i < D; i += FT) {
            // the W words of this index, one per rank: all loads in flight together, only the late ones are asked for again
            unsigned vx[PEER_MAX], pending = (1u << W) - 1u, spins = 0;
            while (pending) {
                unsigned vy[PEER_MAX];
#pragma unroll
                for (int r = 0; r < PEER_MAX; ++r)
                    if (pending >> r & 1) {
                        const volatile uint2* src = reinterpret_cast<const volatile uint2*>(me + box_ll(par, r, pb.xf_box.xb_max)) + i;
                        asm volatile("ld.volatile.global.v2.u32 {%0, %1}, [%2];" : "=r"(vx[r]), "=r"(vy[r]) : "l"(src) : "memory");
                    }
#pragma unroll
                for (int r = 0; r < PEER_MAX; ++r)
                    if ((pending >> r & 1) && vy[r] == epoch) pending &= ~(1u << r);
                if (pending) {
                    if (++spins > (1u << 22)) { ok = false; break; }
                    if (spins > 64) __nanosleep(200);
                }
            }
            unsigned acc_or = 0;
#pragma unroll
            for (int r = 0; r < PEER_MAX; ++r)
                if (r < W) {
                    const unsigned v = (pending >> r & 1) ? 0u : vx[r];
                    if (i < FUSED_LL_HDR) s_hdr[r * FUSED_LL_HDR + i] = v; else acc_or |= v;
                }
            if (i >= FUSED_LL_HDR) s_or[i - FUSED_LL_HDR] = acc_or;
        }
        if (!__syncthreads_and(ok)) {
            if (tid == 0) { pb.gflags[3] = 1; pb.stats->path = 0; pb.stats->pad = 2; }
            return;
        }
        long long* f_tot = reinterpret_cast<long long*>(&s_wD[0][0]);
        int* f_maxfree = reinterpret_cast<int*>(&s_wD[1][0]);
        unsigned long long* f_acc = &s_wcnt[0][0];
        if (tid < KP) {
            long long t = 0; int m = 0;
            for (int r = 0; r < W; ++r) {
                const unsigned* hr = s_hdr + r * FUSED_LL_HDR;
                t += (long long)(((unsigned long long)hr[2 * tid + 1] << 32) | hr[2 * tid]);
                m = max(m, (int)hr[32 + tid]);
            }
            f_tot[tid] = t; f_maxfree[tid] = m;
        }
        if (tid >= 32 && tid < 36) {
            unsigned long long a = 0;
            for (int r = 0; r < W; ++r) {
                const unsigned* hr = s_hdr + r * FUSED_LL_HDR + 49 + 2 * (tid - 32);
                a += ((unsigned long long)hr[1] << 32) | hr[0];
            }
            f_acc[tid - 32] = a;
        }
        if (tid == 64) {
            int f = 0;
            for (int r = 0; r < W; ++r) f |= (int)s_hdr[r * FUSED_LL_HDR + 48];
            s_flags[0] = f;
        }
        __syncthreads();
        fused_verify(pb, s_runs, f_tot, f_maxfree, s_flags[0], f_acc, reinterpret_cast<const unsigned char*>(s_or), 0, 1, false, s_l0, s_dec, s_e);
        return;
    }
    if (pb.xf_on) {
        // ---- the block travels by NCCL: [tot int64 x16 | maxfree int32 x16 | reject bits, pad x3 | counters u64 x4 | placed
        //      bitmap] is assembled here and k_fused_verify follows the all-gather on the stream.  Sums, maxima and ORs are
        //      order-independent, so every rank folds the W blocks into the same state and reaches the same verdict.
        unsigned char* blk = pb.xf_exchange;
        if (tid < KP) {
            reinterpret_cast<long long*>(blk)[tid] = __ldcg(&pb.tot[tid]);
            reinterpret_cast<int32_t*>(blk + 128)[tid] = __ldcg(&pb.maxfree[tid]);
        }
        if (tid < 4) {
            reinterpret_cast<int32_t*>(blk + 192)[tid] = tid == 0 ? __ldcg(&pb.wave[5]) : 0;
            reinterpret_cast<unsigned long long*>(blk + 208)[tid] = __ldcg(&pb.acc[tid]);
        }
        for (int w = tid; w < n_words + (n_words & 1); w += FT) {
            reinterpret_cast<unsigned*>(blk + FUSED_XHDR)[w] = w < n_words ? __ldcg(&pb.placed[w]) : 0u;
            if (w < n_words) pb.placed[w] = 0;                        // clean for the next fused run of this problem
        }
        return;
    }
    fused_verify(pb, s_runs, pb.tot, pb.maxfree, __ldcg(&pb.wave[5]), pb.acc, reinterpret_cast<const unsigned char*>(pb.placed), 0, 1, true,
                 s_l0, s_dec, s_e);
}

// slice mode: wait for every rank's block (peer mailbox flags, or an all-gather that is already complete), fold, verify
__global__ void __launch_bounds__(256) k_fused_verify(DevProblem pb, const unsigned char* blocks, size_t stride, int world, PeerWait wait) {
    __shared__ long long s_tot[KP];
    __shared__ int s_maxfree[KP];
    __shared__ unsigned long long s_acc[4];
    __shared__ int s_rej;
    __shared__ long long s_l0[KP], s_dec[KP];
    __shared__ long long s_e[FRUN_MAX];
    const int tid = threadIdx.x;
    cudaGridDependencySynchronize();
    if (!peer_wait(wait)) {                                            // a peer never delivered its block: nothing below would be meaningful
        if (tid == 0) { pb.gflags[3] = 1; pb.stats->path = 0; pb.stats->pad = 2; }
        return;
    }
    if (tid < KP) {
        long long t = 0; int m = 0;
        for (int r = 0; r < world; ++r) {
            t += __ldcg(reinterpret_cast<const long long*>(blocks + r * stride) + tid);
            m = max(m, __ldcg(reinterpret_cast<const int32_t*>(blocks + r * stride + 128) + tid));
        }
        s_tot[tid] = t; s_maxfree[tid] = m;
    }
    if (tid >= 32 && tid < 36) {
        unsigned long long a = 0;
        for (int r = 0; r < world; ++r) a += __ldcg(reinterpret_cast<const unsigned long long*>(blocks + r * stride + 208) + (tid - 32));
        s_acc[tid - 32] = a;
    }
    if (tid == 64) {
        int f = 0;
        for (int r = 0; r < world; ++r) f |= __ldcg(reinterpret_cast<const int32_t*>(blocks + r * stride + 192));
        s_rej = f;
    }
    __syncthreads();
    const FRun* runs = reinterpret_cast<const FRun*>(pb.fp_blob + pb.fp_off[2]);
    fused_verify(pb, runs, s_tot, s_maxfree, s_rej, s_acc, blocks + FUSED_XHDR, stride, world, false, s_l0, s_dec, s_e);
}

// ---- launch geometry: how much of the blob is staged, how many nodes a CTA takes at a time
constexpr size_t FUSED_SMEM = 200 * 1024;            // dynamic shared memory of the one CTA per SM
int fused_max_pad_pods() { return 1 << 20; }

cudaError_t fused_prepare_device() {
    return cudaFuncSetAttribute(k_plan_fused, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)FUSED_SMEM);
}

cudaError_t launch_fused(const DevProblem& pb, int sms, cudaStream_t st) {
    const int n_nodes = pb.node_hi - pb.node_lo;
    const size_t bits = (((size_t)(pb.fp_pad_pods >> 5) * 4) + 15) & ~(size_t)15;
    const size_t per_node = sizeof(NodeS) + 32 + (pb.kind != NOSPLAN_KIND_MIG ? 128 : 0);      // MPS keeps every GPU's free counts for phase 3
    auto need = [&](int stage_bytes, int tile_nodes) { return bits + (size_t)stage_bytes + (size_t)tile_nodes * per_node; };
    // one resident wave when the nodes allow it: the smallest tile (multiple of 8, at least 64 nodes) that covers them
    FusedLaunch fl;
    int tile = (n_nodes + sms - 1) / sms;
    tile = (tile + 7) & ~7;                    // 72 nodes, 139 CTAs at 10 000 nodes: measured no slower than 68 nodes on all 148 SMs
    if (tile < 64) tile = 64;
    if (tile > FT) tile = FT;
    fl.tile_nodes = tile;
    // the header (tables, runs) always; then whole levels of the pod structure, top down, while they fit
    fl.stage_bytes = pb.fp_off[FP_L3];
    for (int l = FP_L3 + 1; l <= FP_END; ++l) if (need(pb.fp_off[l], tile) <= FUSED_SMEM) fl.stage_bytes = pb.fp_off[l];
    if (need(fl.stage_bytes, tile) > FUSED_SMEM) return cudaErrorInvalidValue;
    int blocks = (n_nodes + tile - 1) / tile;
    if (blocks > sms) blocks = sms;
    if (blocks < 1) blocks = 1;
    k_plan_fused<<<blocks, FT, need(fl.stage_bytes, tile), st>>>(pb, fl);
    return cudaGetLastError();
}

// slice mode: bytes of a rank's exchange block in the flag-in-word form / as assembled for NCCL
size_t fused_exchange_bytes_ll(const DevProblem& pb) { return 8 * ((size_t)FUSED_LL_HDR + (size_t)(pb.fp_pad_pods >> 5)); }
size_t fused_exchange_bytes(const DevProblem& pb) { const size_t w = (size_t)(pb.fp_pad_pods >> 5); return FUSED_XHDR + 4 * (w + (w & 1)); }

cudaError_t launch_fused_verify(const DevProblem& pb, const void* blocks, size_t stride, int world, const PeerWait& wait, cudaStream_t st) {
    return launch_pdl(k_fused_verify, dim3(1), dim3(256), 0, st, pb, static_cast<const unsigned char*>(blocks), stride, world, wait);
}

}  // namespace nsp
