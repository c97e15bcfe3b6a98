// k_preempt_dryrun -- the whole of preemptor.SelectVictimsOnNode (pkg/scheduler/plugins/capacityscheduling/
// capacity_scheduling.go:468-675) for one preemptor over many nodes.
//
// One warp per node, lane = resource column (16 columns; quota columns double as node resource columns).  The node's
// private clone of the quota snapshot (ElasticQuotaInfos.clone, elasticquotainfo.go:37-43: one copy per NAMESPACE key) is
// the global Used minus a per-warp table in shared memory of what this node's victims have un-reserved so far, keyed by
// namespace; the aggregated used / over-quota vectors are kept in step with every removal and every reprieve.  The node's
// NodeInfo is two lane-held vectors (Allocatable, Requested; column 3 = pod counts).  Per pod a state byte lives in the
// output array: potential victim, PDB-violating, currently off the node, final victim.
//   pass 1  pods in the order given (least important first): quota predicates, removePod            (:511-596)
//   checks  no victims / Filter / max quota / aggregated min quota                                   (:598-624)
//   pass 2  potential victims most important first: PodDisruptionBudget counters                     (:850-895)
//   pass 3  reprieve the violating, then the others: addPod, Filter, quota checks, removePod again   (:636-673)
// Restated from k8s.io/kubernetes v1.25.4 (not in the reference tree): NodeInfo.RemovePod / AddPodInfo, NodeResourcesFit.
#include <climits>
#include <cstring>
#include <string>

#include "nosplan_dev.cuh"
#include "quota_ctx.cuh"

namespace nsp {

constexpr int PQR = NOSPLAN_QUOTA_MAX_RES;
constexpr unsigned P_SCALAR_BITS = 0x7ffffff0u;
constexpr int P_TOUCH = 32;                 // namespaces one node's victims may come from
constexpr int P_PDBS = 64;                  // PodDisruptionBudgets per query
constexpr unsigned P_FULL = 0xffffffffu;
constexpr uint8_t S_POTENTIAL = 1, S_VIOLATING = 2, S_OFF = 4, S_VICTIM = 8;

struct DevPreempt {
    int E, R, NS, gm, n_nodes, n_pods, pn, pprio, n_pdbs;
    unsigned fit_mask;
    const int32_t* ns_quota; const long long *qmin, *qmax, *used; const uint32_t *min_mask, *max_mask, *used_mask; const uint8_t* max_enforced;
    long long coeff[PQR], nom[PQR], pod_req[PQR], nominated[PQR], fit[PQR];
    const int32_t* node_ptr; const int32_t* ns; const int32_t* prio; const uint8_t* over; const long long* req; const uint32_t* rmask;
    const long long* node_alloc; const long long* node_req;
    const int32_t* pdb_allowed; const int32_t* pod_pdb_ptr; const int32_t* pod_pdb;
    long long* agg;          // [3][PQR] a_min, a_over, a_used of the unmodified snapshot
    unsigned* masks;         // [2] omask (min | used key sets), amin_mask
    uint8_t* state;          // [n_pods] work, then 0/1 = victim
    int32_t* status; int32_t* nviol; int32_t* overflow;
};

__device__ __forceinline__ long long p_go_int64(double x) {
    if (!(x == x) || x >= 9223372036854775808.0 || x < -9223372036854775808.0) return LLONG_MIN;
    return (long long)x;
}
__device__ __forceinline__ double p_approx(int c, long long v) { return c == 0 ? __dmul_rn((double)v, 1e-3) : (double)v; }
__device__ __forceinline__ bool p_cmp_col(int c, unsigned ymask) { return c == 0 || c == 1 || (c >= 4 && (ymask >> c & 1)); }

// aggregates over the namespace -> quota map (a composite quota counts once per namespace)
constexpr int P_AGG_THREADS = 1024, P_AGG_ROWS = P_AGG_THREADS / 16;   // 64 namespaces in flight: each step is a chain of dependent loads
__global__ void __launch_bounds__(P_AGG_THREADS) k_preempt_aggregate(DevPreempt v) {
    __shared__ long long s_min[P_AGG_ROWS][PQR], s_over[P_AGG_ROWS][PQR], s_used[P_AGG_ROWS][PQR];
    __shared__ unsigned s_om, s_am;
    const int tid = threadIdx.x, c = tid & 15, j = tid >> 4;
    if (tid == 0) { s_om = 0; s_am = 0; }
    __syncthreads();
    long long mn = 0, ov = 0, us = 0; unsigned om = 0, am = 0;
    for (int ns = j; ns < v.NS; ns += P_AGG_ROWS) {
        const int e = v.ns_quota[ns];
        if (e < 0) continue;
        if (c < v.R) {
            const long long m = v.qmin[(size_t)e * v.R + c], u = v.used[(size_t)e * v.R + c];
            mn += m; us += u; ov += max(0ll, m - u);
        }
        if (c == 0) { am |= v.min_mask[e] & P_SCALAR_BITS; om |= (v.min_mask[e] | v.used_mask[e]) & P_SCALAR_BITS; }
    }
    s_min[j][c] = mn; s_over[j][c] = ov; s_used[j][c] = us;
    if (c == 0) { if (om) atomicOr(&s_om, om); if (am) atomicOr(&s_am, am); }
    __syncthreads();
    if (tid < 3 * PQR) {                                       // one thread per (vector, column)
        const int which = tid / PQR, col = tid % PQR;
        long long (*src)[PQR] = which == 0 ? s_min : which == 1 ? s_over : s_used;
        long long a = 0;
        for (int k = 0; k < P_AGG_ROWS; ++k) a += src[k][col];
        v.agg[which * PQR + col] = a;
    }
    if (tid == 0) { v.masks[0] = s_om; v.masks[1] = s_am; }
}

}  // namespace nsp

namespace nsp {

__global__ void __launch_bounds__(256) k_preempt_dryrun(DevPreempt v) {
    __shared__ int s_tns[8][P_TOUCH];
    __shared__ long long s_trem[8][P_TOUCH][PQR];
    __shared__ int s_pdb[8][P_PDBS];
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const int c = lane;                                    // lanes 16..31 carry no column
    const bool col = c < v.R;
    int* tns = s_tns[warp];
    long long (*trem)[PQR] = s_trem[warp];
    int* pdb = s_pdb[warp];
    const int ep = (v.pn >= 0 && v.pn < v.NS) ? v.ns_quota[v.pn] : -1;
    const long long a_min = col ? v.agg[c] : 0, a_over0 = col ? v.agg[PQR + c] : 0, a_used0 = col ? v.agg[2 * PQR + c] : 0;
    const long long nom = col ? v.nom[c] : 0, preq = col ? v.pod_req[c] : 0, nominated = col ? v.nominated[c] : 0, fitq = col ? v.fit[c] : 0;
    const unsigned omask0 = v.masks[0], amin_mask = v.masks[1];
    const long long min_p = (col && ep >= 0) ? v.qmin[(size_t)ep * v.R + c] : 0, max_p = (col && ep >= 0) ? v.qmax[(size_t)ep * v.R + c] : 0;
    const unsigned mm_p = ep >= 0 ? v.min_mask[ep] : 0, xm_p = ep >= 0 ? v.max_mask[ep] : 0;
    const bool enforce_p = ep >= 0 && v.max_enforced[ep];
    const unsigned fm = v.fit_mask & P_SCALAR_BITS & ~(v.gm >= 0 ? (1u << v.gm) : 0u);
    const bool fit_zero = v.fit[0] == 0 && v.fit[1] == 0 && v.fit[2] == 0 && fm == 0;

    const int warps_total = gridDim.x * (blockDim.x >> 5);
    for (int node = blockIdx.x * (blockDim.x >> 5) + warp; node < v.n_nodes; node += warps_total) {
        const int p0 = v.node_ptr[node], p1 = v.node_ptr[node + 1];
        int n_touch = 0;
        long long a_over = a_over0, a_used = a_used0;
        unsigned omask = omask0;
        const long long alloc = col ? v.node_alloc[(size_t)node * v.R + c] : 0;
        long long requested = col ? v.node_req[(size_t)node * v.R + c] : 0;
        bool overflow = false;

        auto used_of = [&](int ns, int e) -> long long {   // lane c: Used[c] of namespace `ns` in this node's clone
            long long u = col ? v.used[(size_t)e * v.R + c] : 0;
            const unsigned hit = __ballot_sync(P_FULL, lane < n_touch && tns[lane] == ns);
            if (hit && col) u -= trem[__ffs(hit) - 1][c];
            return u;
        };
        auto guaranteed = [&](int e) -> long long {        // lane c: GetGuaranteedOverquotas(e)[c]
            if (!col) return 0;
            const bool in_min = c < 4 || (v.min_mask[e] >> c & 1);
            double pct = 0.0;
            if (in_min && a_min > 0) pct = __ddiv_rn(p_approx(c, v.qmin[(size_t)e * v.R + c]), p_approx(c, a_min));
            const bool in_res = c < 4 || (omask >> c & 1);
            return in_res ? p_go_int64(floor(__dmul_rn((double)a_over, pct))) : 0;
        };
        // the pod's request on lane c: quota request (gpu-memory filled) in rq, scalar key mask in rm; node request in rn
        auto pod_rows = [&](int i, long long& rq, long long& rn, unsigned& rm) {
            long long r = (col && c != v.gm) ? v.req[(size_t)i * v.R + c] : 0;
            long long g = col ? v.coeff[c] * r : 0;
#pragma unroll
            for (int o = 16; o; o >>= 1) g += __shfl_xor_sync(P_FULL, g, o);
            if (v.rmask) rm = v.rmask[i] & P_SCALAR_BITS;
            else rm = __ballot_sync(P_FULL, col && c >= 4 && r != 0) & P_SCALAR_BITS;
            const bool present = c < 3 || (rm >> c & 1);
            rn = (col && c != 3 && c != v.gm && present) ? r : 0;
            if (v.gm >= 0) { rm |= 1u << v.gm; if (c == v.gm) r = g; }
            rq = r;
        };
        // reserve (sign -1) / un-reserve (sign +1) pod i in the clone of its namespace; NodeInfo bookkeeping with it
        bool lte_known = false, lte_cached = false;        // usedLteWith(guaranteed + min) of the preemptor's quota: changes only with a move
        auto move_pod = [&](int i, int sign, int pns, int ev) {
            long long rq, rn; unsigned rm;
            pod_rows(i, rq, rn, rm);
            requested -= sign * rn;                        // RemovePod: Requested -= r (sign +1), AddPodInfo: += r
            if (c == 3) requested -= sign;
            if (ev < 0) return;
            lte_known = false;
            unsigned hit = __ballot_sync(P_FULL, lane < n_touch && tns[lane] == pns);
            int slot;
            if (hit) slot = __ffs(hit) - 1;
            else {
                if (n_touch >= P_TOUCH) { overflow = true; return; }
                slot = n_touch++;
                if (lane == 0) tns[slot] = pns;
                if (lane < PQR) trem[slot][lane] = 0;
                __syncwarp();
            }
            if (col && (c == 0 || c == 1 || (c >= 4 && (rm >> c & 1)))) {
                const long long mn = v.qmin[(size_t)ev * v.R + c];
                const long long old = v.used[(size_t)ev * v.R + c] - trem[slot][c];
                const long long d = sign * rq;
                trem[slot][c] += d;
                a_used -= d;
                a_over += max(0ll, mn - (old - d)) - max(0ll, mn - old);
            }
            omask |= rm;
            __syncwarp();
        };
        auto fits = [&]() -> bool {                        // NodeResourcesFit.fitsRequest
            bool bad = c == 3 && col && requested + 1 > alloc;
            if (!fit_zero && col && c != 3 && c != v.gm && (c < 3 || (fm >> c & 1))) bad = bad || fitq > alloc - requested;
            return __ballot_sync(P_FULL, bad) == 0;
        };
        auto over_max_with = [&](long long rq) -> bool {    // usedOverMaxWith of the preemptor's quota
            if (!enforce_p) return false;
            const long long up = used_of(v.pn, ep);
            return __ballot_sync(P_FULL, col && p_cmp_col(c, xm_p) && rq + up > max_p) != 0;
        };
        auto agg_over_min_with = [&](long long rq) -> bool {   // AggregatedUsedOverMinWith
            return __ballot_sync(P_FULL, col && p_cmp_col(c, amin_mask) && a_used + rq > a_min) != 0;
        };

        // ---- pass 1: potential victims, in the order given (:511-596)
        bool more_than_min = false;
        if (ep >= 0) {
            const long long up = used_of(v.pn, ep);
            more_than_min = __ballot_sync(P_FULL, col && p_cmp_col(c, mm_p) && nom + up > min_p) != 0;   // usedOverMinWith, once (:522)
        }
        // pods are taken 32 at a time: lane k fetches the scalars of pod b + k (coalesced), the warp then walks them in order
        int n_potential = 0;
        for (int b = p0; b < p1 && !overflow; b += 32) {
            const int il = b + lane;
            const bool inr = il < p1;
            const int l_ns = inr ? v.ns[il] : -1, l_prio = inr ? v.prio[il] : 0, l_over = inr ? v.over[il] : 0;
            const int l_ev = (l_ns >= 0 && l_ns < v.NS) ? v.ns_quota[l_ns] : -1;
            int l_state = 0;
            const int nb = min(32, p1 - b);
            for (int k = 0; k < nb && !overflow; ++k) {
                const int i = b + k;
                const int pns = __shfl_sync(P_FULL, l_ns, k), ev = __shfl_sync(P_FULL, l_ev, k);
                const int prio = __shfl_sync(P_FULL, l_prio, k), over = __shfl_sync(P_FULL, l_over, k);
                bool is_victim = false;
                if (ep < 0) is_victim = ev < 0 && prio < v.pprio;                                   // :582-596
                else if (ev >= 0) {
                    if (more_than_min) {
                        if (pns == v.pn) is_victim = prio < v.pprio;                                // :527-535
                        else if (over) {                                                            // :541-559
                            if (!lte_known) {
                                const unsigned ymask = omask | (mm_p & P_SCALAR_BITS);
                                const long long up = used_of(v.pn, ep);
                                const long long y = guaranteed(ep) + min_p;
                                lte_cached = __ballot_sync(P_FULL, col && p_cmp_col(c, ymask) && nom + up > y) == 0;
                                lte_known = true;
                            }
                            if (lte_cached) {
                                const unsigned vmask = omask | (v.min_mask[ev] & P_SCALAR_BITS);
                                const long long uv = used_of(pns, ev);
                                const long long yv = guaranteed(ev) + (col ? v.qmin[(size_t)ev * v.R + c] : 0);
                                is_victim = __ballot_sync(P_FULL, col && p_cmp_col(c, vmask) && uv > yv) != 0;
                            }
                        }
                    } else if (pns != v.pn && over) {                                               // :569-577
                        const long long uv = used_of(pns, ev);
                        is_victim = __ballot_sync(P_FULL, col && p_cmp_col(c, v.min_mask[ev]) &&
                                                              uv > v.qmin[(size_t)ev * v.R + c]) != 0;
                    }
                }
                if (is_victim) {
                    if (lane == k) l_state = S_POTENTIAL | S_OFF;
                    ++n_potential;
                    move_pod(i, +1, pns, ev);
                }
            }
            if (inr) v.state[il] = (uint8_t)l_state;
        }
        __syncwarp();
        // ---- the checks between the passes (:598-624)
        int status = NOSPLAN_VICTIMS_OK;
        if (overflow) status = -1;
        else if (n_potential == 0) status = NOSPLAN_VICTIMS_NONE;
        else if (!fits()) status = NOSPLAN_VICTIMS_FILTER;
        else if (ep >= 0) {
            if (over_max_with(preq)) status = NOSPLAN_VICTIMS_MAX_QUOTA;
            else if (agg_over_min_with(preq)) status = NOSPLAN_VICTIMS_MIN_QUOTA;
        }
        int n_viol = 0;
        if (status == NOSPLAN_VICTIMS_OK) {
            const int last_chunk = p0 + ((p1 - p0 - 1) & ~31);
            // ---- pass 2: PodDisruptionBudget split, most important victim first (:628-635, :850-895)
            for (int k = lane; k < v.n_pdbs; k += 32) pdb[k] = v.pdb_allowed[k];
            __syncwarp();
            for (int b = last_chunk; b >= p0; b -= 32) {
                const int il = b + lane;
                const bool mine = il < p1 && (v.state[il] & S_POTENTIAL);
                // lane k walks the budgets of pod b + k; the lanes take turns from the most important pod down
                unsigned todo = __ballot_sync(P_FULL, mine);
                while (todo) {
                    const int k = 31 - __clz(todo);
                    todo &= ~(1u << k);
                    if (lane == k) {
                        bool bad = false;
                        for (int x = v.pod_pdb_ptr[il]; x < v.pod_pdb_ptr[il + 1]; ++x)
                            if (--pdb[v.pod_pdb[x]] < 0) bad = true;
                        if (bad) v.state[il] |= S_VIOLATING;
                    }
                    __syncwarp();
                }
            }
            // ---- pass 3: reprieve the violating victims, then the others (:636-673)
            for (int round = 0; round < 2 && status == NOSPLAN_VICTIMS_OK; ++round) {
                for (int b = last_chunk; b >= p0 && status == NOSPLAN_VICTIMS_OK; b -= 32) {
                    const int il = b + lane;
                    int l_state = il < p1 ? v.state[il] : 0;
                    const int l_ns = (l_state & S_POTENTIAL) ? v.ns[il] : -1;
                    const int l_ev = (l_ns >= 0 && l_ns < v.NS) ? v.ns_quota[l_ns] : -1;
                    unsigned todo = __ballot_sync(P_FULL, (l_state & S_POTENTIAL) && ((l_state & S_VIOLATING) != 0) == (round == 0));
                    while (todo) {
                        const int k = 31 - __clz(todo);
                        todo &= ~(1u << k);
                        const int i = b + k;
                        const int pns = __shfl_sync(P_FULL, l_ns, k), ev = __shfl_sync(P_FULL, l_ev, k);
                        move_pod(i, -1, pns, ev);                                // addPod
                        bool off = false, victim = false;
                        const bool fit = fits();
                        if (!fit) { move_pod(i, +1, pns, ev); off = true; victim = true; }
                        if (ep >= 0 && (over_max_with(nom) || agg_over_min_with(nominated))) {
                            if (off) { status = NOSPLAN_VICTIMS_ERROR; break; }  // RemovePod of a pod that is not on the node
                            move_pod(i, +1, pns, ev); off = true; victim = true;
                        }
                        if (!fit && round == 0) ++n_viol;
                        if (lane == k) l_state = (l_state & ~(S_OFF | S_VICTIM)) | (off ? S_OFF : 0) | (victim ? S_VICTIM : 0);
                    }
                    if (il < p1) v.state[il] = (uint8_t)l_state;
                }
            }
        }
        // ---- results
        if (status < 0) { if (lane == 0) atomicExch(v.overflow, 1); status = NOSPLAN_VICTIMS_ERROR; }
        const bool ok = status == NOSPLAN_VICTIMS_OK;
        for (int i = p0 + lane; i < p1; i += 32) v.state[i] = (ok && (v.state[i] & S_VICTIM)) ? 1 : 0;
        if (lane == 0) { v.status[node] = status; v.nviol[node] = ok ? n_viol : 0; }
        __syncwarp();
    }
}

}  // namespace nsp

using namespace nsp;

namespace {
// what nosplan_preempt_dryrun_again needs of the last full call: the device problem and where its outputs lie
struct PreemptResident {
    bool valid = false;
    DevPreempt v;
    size_t upload_end = 0, out_end = 0, total = 0, o_state = 0, o_status = 0, o_nviol = 0, o_ovf = 0;
    int P = 0, Nn = 0, R = 0;
};
void preempt_resident_free(void* p) { delete static_cast<PreemptResident*>(p); }

// preemptor fields -> device problem, work area cleared, both kernels, outputs back to the caller
int preempt_run(cudaStream_t st, nosplan_quota_ctx* ctx, std::string* err, PreemptResident& rs, const nosplan_preempt_query* pq,
                uint8_t* victim, int32_t* node_status, int32_t* num_violating, cudaEvent_t* ev, float* ms_k, float* ms_d2h,
                int* launches, long long* d2h_bytes) {
    auto fail = [&](int code, const char* msg, const char* detail = "") { if (err) *err = std::string(msg) + detail; return code; };
#define RCU(call) do { cudaError_t e_ = (call); if (e_ != cudaSuccess) return fail(NOSPLAN_E_CUDA, #call ": ", cudaGetErrorString(e_)); } while (0)
    DevPreempt& v = rs.v;
    unsigned char* S = ctx->hs; unsigned char* A = ctx->d;
    v.pn = pq->v.preemptor_ns; v.pprio = pq->v.preemptor_priority; v.fit_mask = pq->fit_mask;
    for (int c = 0; c < PQR; ++c) {
        const bool in = c < rs.R;
        v.nom[c] = in ? pq->v.nominated_in_eq[c] : 0;
        v.pod_req[c] = in ? pq->pod_req[c] : 0;
        v.nominated[c] = in ? pq->nominated[c] : 0;
        v.fit[c] = in ? pq->fit_request[c] : 0;
    }
    RCU(cudaMemsetAsync(A + rs.upload_end, 0, rs.total - rs.upload_end, st));
    RCU(cudaEventRecord(ev[1], st));
    k_preempt_aggregate<<<1, P_AGG_THREADS, 0, st>>>(v);
    if (rs.Nn > 0) {
        // one node per warp: nodes differ widely in work (victims found or not, reprieve loops), so the block scheduler balances them
        k_preempt_dryrun<<<(rs.Nn + 7) / 8, 256, 0, st>>>(v);
    }
    RCU(cudaGetLastError());
    RCU(cudaEventRecord(ev[2], st));
    RCU(cudaMemcpyAsync(S, A + rs.upload_end, rs.out_end - rs.upload_end, cudaMemcpyDeviceToHost, st));
    RCU(cudaEventRecord(ev[3], st));
    RCU(cudaStreamSynchronize(st));
    if (ms_k) cudaEventElapsedTime(ms_k, ev[1], ev[2]);
    if (ms_d2h) cudaEventElapsedTime(ms_d2h, ev[2], ev[3]);
    if (launches) *launches = 1 + (rs.Nn > 0);
    if (d2h_bytes) *d2h_bytes = (long long)(rs.out_end - rs.upload_end);
    int32_t overflow; std::memcpy(&overflow, S + (rs.o_ovf - rs.upload_end), 4);
    if (overflow) return fail(NOSPLAN_E_LIMIT, "a node's victims come from more than 32 namespaces");
    if (rs.P) std::memcpy(victim, S + (rs.o_state - rs.upload_end), (size_t)rs.P);
    if (rs.Nn) {
        std::memcpy(node_status, S + (rs.o_status - rs.upload_end), 4 * (size_t)rs.Nn);
        std::memcpy(num_violating, S + (rs.o_nviol - rs.upload_end), 4 * (size_t)rs.Nn);
    }
    return NOSPLAN_OK;
#undef RCU
}
}  // namespace

// Another preemptor against the cluster state (quotas, nodes, pods, budgets) the last nosplan_preempt_dryrun left in HBM.
extern "C" int nosplan_preempt_dryrun_again_impl(int device, cudaStream_t st, nosplan_quota_ctx* ctx, std::string* err,
                                                 const nosplan_preempt_query* pq, uint8_t* victim, int32_t* node_status,
                                                 int32_t* num_violating, float* ms_k, float* ms_d2h, int* launches, long long* d2h_bytes) {
    auto fail = [&](int code, const char* msg) { if (err) *err = msg; return code; };
    PreemptResident* rs = ctx ? static_cast<PreemptResident*>(ctx->preempt_resident) : nullptr;
    if (!rs || !rs->valid) return fail(NOSPLAN_E_INVALID, "nosplan_preempt_dryrun_again: no cluster state resident on this handle (call nosplan_preempt_dryrun first)");
    if (!pq || !victim || !node_status || !num_violating) return fail(NOSPLAN_E_INVALID, "null argument");
    if (!pq->v.nominated_in_eq || !pq->pod_req || !pq->nominated || !pq->fit_request) return fail(NOSPLAN_E_INVALID, "nominated_in_eq, pod_req, nominated and fit_request are required");
    if (cudaSetDevice(device) != cudaSuccess) return fail(NOSPLAN_E_CUDA, "cudaSetDevice");
    cudaEvent_t* ev = nullptr;
    if (quota_ctx_events(ctx, &ev) != cudaSuccess) return fail(NOSPLAN_E_CUDA, "cudaEventCreate");
    return preempt_run(st, ctx, err, *rs, pq, victim, node_status, num_violating, ev, ms_k, ms_d2h, launches, d2h_bytes);
}

extern "C" int nosplan_preempt_dryrun_impl(int device, cudaStream_t st, nosplan_quota_ctx* ctx, std::string* err,
                                           const nosplan_quotas* qs, const nosplan_preempt_query* pq, uint8_t* victim,
                                           int32_t* node_status, int32_t* num_violating, float* ms_h2d, float* ms_k, float* ms_d2h,
                                           int* launches, long long* h2d_bytes, long long* d2h_bytes) {
    auto fail = [&](int code, const char* msg, const char* detail = "") { if (err) *err = std::string(msg) + detail; return code; };
#define PCU(call) do { cudaError_t e_ = (call); if (e_ != cudaSuccess) return fail(NOSPLAN_E_CUDA, #call ": ", cudaGetErrorString(e_)); } while (0)
    if (!qs || !pq || !victim || !node_status || !num_violating) return fail(NOSPLAN_E_INVALID, "null argument");
    const nosplan_victim_query* vq = &pq->v;
    const int E = qs->n_quotas, R = qs->n_res, NS = qs->n_namespaces, P = vq->n_pods, Nn = vq->n_nodes, B = pq->n_pdbs;
    if (R < 4 || R > PQR) return fail(NOSPLAN_E_LIMIT, "n_res must be in [4, NOSPLAN_QUOTA_MAX_RES]");
    if (E < 0 || NS < 0 || P < 0 || Nn < 0 || B < 0) return fail(NOSPLAN_E_INVALID, "negative size");
    if (qs->gpu_memory_col >= R || (qs->gpu_memory_col >= 0 && qs->gpu_memory_col < 4)) return fail(NOSPLAN_E_INVALID, "gpu_memory_col out of range");
    for (int ns = 0; ns < NS; ++ns) if (qs->ns_quota[ns] >= E) return fail(NOSPLAN_E_INVALID, "ns_quota out of range");
    if (B > P_PDBS) return fail(NOSPLAN_E_LIMIT, "n_pdbs exceeds NOSPLAN_PREEMPT_MAX_PDBS");
    if (Nn > 0 && (vq->node_ptr[0] != 0 || vq->node_ptr[Nn] != P)) return fail(NOSPLAN_E_INVALID, "node_ptr must span [0, n_pods]");
    if (!vq->nominated_in_eq || !pq->pod_req || !pq->nominated || !pq->fit_request) return fail(NOSPLAN_E_INVALID, "nominated_in_eq, pod_req, nominated and fit_request are required");
    if (Nn > 0 && (!pq->node_alloc || !pq->node_requested)) return fail(NOSPLAN_E_INVALID, "node_alloc and node_requested are required");
    if (P > 0 && !pq->pod_pdb_ptr) return fail(NOSPLAN_E_INVALID, "pod_pdb_ptr is required");
    for (int k = 0; k < Nn; ++k)
        if (vq->node_ptr[k] > vq->node_ptr[k + 1]) return fail(NOSPLAN_E_INVALID, "node_ptr must be non-decreasing");
    const int M = P > 0 ? pq->pod_pdb_ptr[P] : 0;
    if (P > 0) {
        if (pq->pod_pdb_ptr[0] != 0 || M < 0) return fail(NOSPLAN_E_INVALID, "pod_pdb_ptr must start at 0");
        for (int i = 0; i < P; ++i)
            if (pq->pod_pdb_ptr[i] > pq->pod_pdb_ptr[i + 1]) return fail(NOSPLAN_E_INVALID, "pod_pdb_ptr must be non-decreasing");
        for (int k = 0; k < M; ++k)
            if (pq->pod_pdb[k] < 0 || pq->pod_pdb[k] >= B) return fail(NOSPLAN_E_INVALID, "pod_pdb names a budget outside [0, n_pdbs)");
    }
    PCU(cudaSetDevice(device));
    if (!ctx->preempt_resident) { ctx->preempt_resident = new PreemptResident(); ctx->preempt_resident_free = preempt_resident_free; }
    PreemptResident& rs = *static_cast<PreemptResident*>(ctx->preempt_resident);
    rs.valid = false;
    size_t o = 0;
    auto al = [](size_t x) { return (x + 255) & ~(size_t)255; };
    auto put = [&](size_t bytes) { size_t at = o; o = al(o + bytes); return at; };
    const size_t ER = (size_t)E * R, NR = (size_t)Nn * R;
    const size_t o_nsq = put(4 * (size_t)NS), o_min = put(8 * ER), o_max = put(8 * ER), o_used = put(8 * ER), o_mm = put(4 * (size_t)E),
                 o_xm = put(4 * (size_t)E), o_um = put(4 * (size_t)E), o_enf = put((size_t)E),
                 o_ptr = put(4 * ((size_t)Nn + 1)), o_ns = put(4 * (size_t)P), o_prio = put(4 * (size_t)P), o_over = put((size_t)P),
                 o_req = put(8 * (size_t)P * R), o_rmask = put(4 * (size_t)P), o_alloc = put(8 * NR), o_nreq = put(8 * NR),
                 o_pdba = put(4 * (size_t)B), o_pptr = put(4 * ((size_t)P + 1)), o_ppdb = put(4 * (size_t)M);
    const size_t upload_end = o;
    const size_t o_state = put((size_t)P), o_status = put(4 * (size_t)Nn), o_nviol = put(4 * (size_t)Nn), o_ovf = put(16);
    const size_t out_end = o;
    const size_t o_agg = put(8 * 3 * PQR), o_masks = put(16);
    const size_t total = o;
    PCU(quota_ctx_ensure(ctx, total, std::max(upload_end, out_end - upload_end)));
    unsigned char* S = ctx->hs; unsigned char* A = ctx->d;
    auto cp = [&](size_t at, const void* src, size_t bytes) { if (bytes) { if (src) std::memcpy(S + at, src, bytes); else std::memset(S + at, 0, bytes); } };
    cp(o_nsq, qs->ns_quota, 4 * (size_t)NS); cp(o_min, qs->min, 8 * ER); cp(o_max, qs->max, 8 * ER); cp(o_used, qs->used_init, 8 * ER);
    cp(o_mm, qs->min_mask, 4 * (size_t)E); cp(o_xm, qs->max_mask, 4 * (size_t)E); cp(o_um, qs->used_mask, 4 * (size_t)E);
    cp(o_enf, qs->max_enforced, (size_t)E);
    cp(o_ptr, vq->node_ptr, 4 * ((size_t)Nn + 1)); cp(o_ns, vq->ns, 4 * (size_t)P); cp(o_prio, vq->priority, 4 * (size_t)P);
    cp(o_over, vq->over_quota, (size_t)P); cp(o_req, vq->request, 8 * (size_t)P * R); cp(o_rmask, vq->request_mask, 4 * (size_t)P);
    cp(o_alloc, pq->node_alloc, 8 * NR); cp(o_nreq, pq->node_requested, 8 * NR); cp(o_pdba, pq->pdb_allowed, 4 * (size_t)B);
    cp(o_pptr, pq->pod_pdb_ptr, 4 * ((size_t)P + 1)); cp(o_ppdb, pq->pod_pdb, 4 * (size_t)M);
    cudaEvent_t* ev = nullptr;
    PCU(quota_ctx_events(ctx, &ev));
    PCU(cudaEventRecord(ev[0], st));
    PCU(cudaMemcpyAsync(A, S, upload_end, cudaMemcpyHostToDevice, st));
    DevPreempt& v = rs.v; std::memset(&v, 0, sizeof v);
    v.E = E; v.R = R; v.NS = NS; v.gm = qs->gpu_memory_col; v.n_nodes = Nn; v.n_pods = P; v.pn = vq->preemptor_ns; v.pprio = vq->preemptor_priority;
    v.n_pdbs = B; v.fit_mask = pq->fit_mask;
    v.ns_quota = (const int32_t*)(A + o_nsq); v.qmin = (const long long*)(A + o_min); v.qmax = (const long long*)(A + o_max);
    v.used = (const long long*)(A + o_used); v.min_mask = (const uint32_t*)(A + o_mm); v.max_mask = (const uint32_t*)(A + o_xm);
    v.used_mask = (const uint32_t*)(A + o_um); v.max_enforced = (const uint8_t*)(A + o_enf);
    for (int c = 0; c < PQR; ++c) {
        const bool in = c < R;
        v.coeff[c] = (in && qs->gpu_memory_coeff) ? qs->gpu_memory_coeff[c] : 0;
        v.nom[c] = in ? vq->nominated_in_eq[c] : 0;
        v.pod_req[c] = in ? pq->pod_req[c] : 0;
        v.nominated[c] = in ? pq->nominated[c] : 0;
        v.fit[c] = in ? pq->fit_request[c] : 0;
    }
    v.node_ptr = (const int32_t*)(A + o_ptr); v.ns = (const int32_t*)(A + o_ns); v.prio = (const int32_t*)(A + o_prio);
    v.over = (const uint8_t*)(A + o_over); v.req = (const long long*)(A + o_req);
    v.rmask = vq->request_mask ? (const uint32_t*)(A + o_rmask) : nullptr;
    v.node_alloc = (const long long*)(A + o_alloc); v.node_req = (const long long*)(A + o_nreq);
    v.pdb_allowed = (const int32_t*)(A + o_pdba); v.pod_pdb_ptr = (const int32_t*)(A + o_pptr); v.pod_pdb = (const int32_t*)(A + o_ppdb);
    v.agg = (long long*)(A + o_agg); v.masks = (unsigned*)(A + o_masks); v.state = (uint8_t*)(A + o_state);
    v.status = (int32_t*)(A + o_status); v.nviol = (int32_t*)(A + o_nviol); v.overflow = (int32_t*)(A + o_ovf);
    rs.upload_end = upload_end; rs.out_end = out_end; rs.total = total;
    rs.o_state = o_state; rs.o_status = o_status; rs.o_nviol = o_nviol; rs.o_ovf = o_ovf;
    rs.P = P; rs.Nn = Nn; rs.R = R;
    const int rc = preempt_run(st, ctx, err, rs, pq, victim, node_status, num_violating, ev, ms_k, ms_d2h, launches, d2h_bytes);
    if (ms_h2d) cudaEventElapsedTime(ms_h2d, ev[0], ev[1]);
    if (h2d_bytes) *h2d_bytes = (long long)upload_end;
    rs.valid = rc == NOSPLAN_OK || rc == NOSPLAN_E_LIMIT;     // the uploaded state is intact either way
    return rc;
#undef PCU
}
