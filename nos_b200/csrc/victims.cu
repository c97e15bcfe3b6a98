// k_quota_victims -- first pass of preemptor.SelectVictimsOnNode (pkg/scheduler/plugins/capacityscheduling/
// capacity_scheduling.go:511-596) for one preemptor over many nodes.
//
// One warp per node, lane = resource column (16 columns), so every quota comparison (sumGreaterThan / sumLessThanEqual,
// elasticquotainfo.go:319-361) is one lane-parallel test + ballot and every GetGuaranteedOverquotas (:81-119) is one
// double division, multiplication and floor per lane.  The node's private clone of the quota snapshot
// (ElasticQuotaInfos.clone, :37-43: one copy per NAMESPACE key) is the global Used minus a small per-warp table in shared
// memory of what this node's victims have un-reserved so far, keyed by namespace; the aggregated over-quota vector
// (:145-152) is updated incrementally with each removal.  Pods of a node are visited in the order given.
#include <climits>
#include <cstring>
#include <string>

#include "nosplan_dev.cuh"
#include "quota_ctx.cuh"

namespace nsp {

constexpr int VQR = NOSPLAN_QUOTA_MAX_RES;
constexpr unsigned V_SCALAR_BITS = 0x7ffffff0u;
constexpr int V_TOUCH = 32;                 // namespaces one node's victims may come from
constexpr unsigned V_FULL = 0xffffffffu;

struct DevVictims {
    int E, R, NS, gm, n_nodes, n_pods, pn, pprio;
    const int32_t* ns_quota; const long long *qmin, *used; const uint32_t *min_mask, *used_mask;
    long long coeff[VQR], nom[VQR];
    const int32_t* node_ptr; const int32_t* ns; const int32_t* prio; const uint8_t* over; const long long* req; const uint32_t* rmask;
    long long* agg;          // [2][VQR] a_min, a_over of the unmodified snapshot
    unsigned* omask;         // [1]
    uint8_t* victim; int32_t* overflow;
};

__device__ __forceinline__ long long v_go_int64(double x) {
    if (!(x == x) || x >= 9223372036854775808.0 || x < -9223372036854775808.0) return LLONG_MIN;
    return (long long)x;
}
__device__ __forceinline__ double v_approx(int c, long long v) { return c == 0 ? __dmul_rn((double)v, 1e-3) : (double)v; }
__device__ __forceinline__ bool v_cmp_col(int c, unsigned ymask) { return c == 0 || c == 1 || (c >= 4 && (ymask >> c & 1)); }

constexpr int V_AGG_THREADS = 1024, V_AGG_ROWS = V_AGG_THREADS / 16;   // 64 namespaces in flight: each step is a chain of dependent loads
__global__ void __launch_bounds__(V_AGG_THREADS) k_victims_aggregate(DevVictims v) {
    __shared__ long long s_min[V_AGG_ROWS][VQR], s_over[V_AGG_ROWS][VQR];
    __shared__ unsigned s_om;
    const int tid = threadIdx.x, c = tid & 15, j = tid >> 4;
    if (tid == 0) s_om = 0;
    __syncthreads();
    long long mn = 0, ov = 0; unsigned om = 0;
    for (int ns = j; ns < v.NS; ns += V_AGG_ROWS) {
        const int e = v.ns_quota[ns];
        if (e < 0) continue;
        if (c < v.R) {
            const long long m = v.qmin[(size_t)e * v.R + c], u = v.used[(size_t)e * v.R + c];
            mn += m; ov += max(0ll, m - u);
        }
        if (c == 0) om |= (v.min_mask[e] | v.used_mask[e]) & V_SCALAR_BITS;
    }
    s_min[j][c] = mn; s_over[j][c] = ov;
    if (c == 0 && om) atomicOr(&s_om, om);
    __syncthreads();
    if (tid < 2 * VQR) {
        const int which = tid / VQR, col = tid % VQR;
        long long a = 0;
        for (int k = 0; k < V_AGG_ROWS; ++k) a += (which ? s_over : s_min)[k][col];
        v.agg[which * VQR + col] = a;
    }
    if (tid == 0) *v.omask = s_om;
}

__global__ void __launch_bounds__(256) k_quota_victims(DevVictims v) {
    __shared__ int s_tns[8][V_TOUCH];
    __shared__ long long s_trem[8][V_TOUCH][VQR];
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const int c = lane;                                    // lanes 16..31 carry no column
    const bool col = c < v.R;
    int* tns = s_tns[warp];
    long long (*trem)[VQR] = s_trem[warp];
    const int ep = (v.pn >= 0 && v.pn < v.NS) ? v.ns_quota[v.pn] : -1;
    const long long a_min = col ? v.agg[c] : 0, a_over0 = col ? v.agg[VQR + c] : 0, nom = col ? v.nom[c] : 0;
    const unsigned omask0 = *v.omask;
    const long long min_p = (col && ep >= 0) ? v.qmin[(size_t)ep * v.R + c] : 0;
    const unsigned mm_p = ep >= 0 ? v.min_mask[ep] : 0;

    const int warps_total = gridDim.x * (blockDim.x >> 5);
    for (int node = blockIdx.x * (blockDim.x >> 5) + warp; node < v.n_nodes; node += warps_total) {
        int n_touch = 0;
        long long a_over = a_over0;
        unsigned omask = omask0;
        auto used_of = [&](int ns, int e) -> long long {   // lane c: Used[c] of namespace `ns` in this node's clone
            long long u = col ? v.used[(size_t)e * v.R + c] : 0;
            const unsigned hit = __ballot_sync(V_FULL, lane < n_touch && tns[lane] == ns);
            if (hit && col) u -= trem[__ffs(hit) - 1][c];
            return u;
        };
        auto guaranteed = [&](int e) -> long long {        // lane c: GetGuaranteedOverquotas(e)[c]
            if (!col) return 0;
            const bool in_min = c < 4 || (v.min_mask[e] >> c & 1);
            double pct = 0.0;
            if (in_min && a_min > 0) pct = __ddiv_rn(v_approx(c, v.qmin[(size_t)e * v.R + c]), v_approx(c, a_min));
            const bool in_res = c < 4 || (omask >> c & 1);
            return in_res ? v_go_int64(floor(__dmul_rn((double)a_over, pct))) : 0;
        };
        bool more_than_min = false;
        if (ep >= 0) {
            const long long up = used_of(v.pn, ep);
            more_than_min = __ballot_sync(V_FULL, col && v_cmp_col(c, mm_p) && nom + up > min_p) != 0;   // usedOverMinWith, once (:522)
        }
        for (int i = v.node_ptr[node]; i < v.node_ptr[node + 1]; ++i) {
            const int pns = v.ns[i];
            const int ev = (pns >= 0 && pns < v.NS) ? v.ns_quota[pns] : -1;
            bool is_victim = false;
            if (ep < 0) {                                  // preemptor without quota (:582-596)
                if (lane == 0) v.victim[i] = (ev < 0 && v.prio[i] < v.pprio) ? 1 : 0;
                continue;
            }
            if (ev >= 0) {
                if (more_than_min) {
                    if (pns == v.pn) is_victim = v.prio[i] < v.pprio;                               // :527-535
                    else if (v.over[i]) {                                                           // :541-559
                        const unsigned ymask = omask | (mm_p & V_SCALAR_BITS);
                        const long long up = used_of(v.pn, ep);
                        const long long y = guaranteed(ep) + min_p;
                        const bool lte = __ballot_sync(V_FULL, col && v_cmp_col(c, ymask) && nom + up > y) == 0;
                        if (lte) {
                            const unsigned vmask = omask | (v.min_mask[ev] & V_SCALAR_BITS);
                            const long long uv = used_of(pns, ev);
                            const long long yv = guaranteed(ev) + (col ? v.qmin[(size_t)ev * v.R + c] : 0);
                            is_victim = __ballot_sync(V_FULL, col && v_cmp_col(c, vmask) && uv > yv) != 0;
                        }
                    }
                } else if (pns != v.pn && v.over[i]) {                                              // :569-577
                    const long long uv = used_of(pns, ev);
                    is_victim = __ballot_sync(V_FULL, col && v_cmp_col(c, v.min_mask[ev]) &&
                                                          uv > v.qmin[(size_t)ev * v.R + c]) != 0;
                }
            }
            if (lane == 0) v.victim[i] = is_victim ? 1 : 0;
            if (!is_victim) continue;
            // removePod: un-reserve the victim's request in this node's clone (elasticquotainfo.go:202-208)
            long long r = (col && c != v.gm) ? v.req[(size_t)i * v.R + c] : 0;
            long long g = col ? v.coeff[c] * r : 0;
#pragma unroll
            for (int o = 16; o; o >>= 1) g += __shfl_xor_sync(V_FULL, g, o);
            unsigned rm;
            if (v.rmask) rm = v.rmask[i] & V_SCALAR_BITS;
            else rm = __ballot_sync(V_FULL, col && c >= 4 && r != 0) & V_SCALAR_BITS;
            if (v.gm >= 0) { rm |= 1u << v.gm; if (c == v.gm) r = g; }
            unsigned hit = __ballot_sync(V_FULL, lane < n_touch && tns[lane] == pns);
            int slot;
            if (hit) slot = __ffs(hit) - 1;
            else {
                if (n_touch >= V_TOUCH) { if (lane == 0) atomicExch(v.overflow, 1); break; }
                slot = n_touch++;
                if (lane == 0) tns[slot] = pns;
                if (lane < VQR) trem[slot][lane] = 0;
                __syncwarp();
            }
            if (col && (c == 0 || c == 1 || (c >= 4 && (rm >> c & 1)))) {
                const long long mn = v.qmin[(size_t)ev * v.R + c];
                const long long old = v.used[(size_t)ev * v.R + c] - trem[slot][c];
                trem[slot][c] += r;
                a_over += max(0ll, mn - (old - r)) - max(0ll, mn - old);
            }
            omask |= rm;
            __syncwarp();
        }
        __syncwarp();
    }
}

}  // namespace nsp

using namespace nsp;

extern "C" int nosplan_quota_victims_impl(int device, cudaStream_t st, nosplan_quota_ctx* ctx, std::string* err,
                                          const nosplan_quotas* qs, const nosplan_victim_query* vq, uint8_t* victim,
                                          float* ms_k, int* launches) {
    auto fail = [&](int code, const char* msg, const char* detail = "") { if (err) *err = std::string(msg) + detail; return code; };
#define VCU(call) do { cudaError_t e_ = (call); if (e_ != cudaSuccess) return fail(NOSPLAN_E_CUDA, #call ": ", cudaGetErrorString(e_)); } while (0)
    if (!qs || !vq || !victim) return fail(NOSPLAN_E_INVALID, "null argument");
    const int E = qs->n_quotas, R = qs->n_res, NS = qs->n_namespaces, P = vq->n_pods, Nn = vq->n_nodes;
    if (R < 4 || R > VQR) return fail(NOSPLAN_E_LIMIT, "n_res must be in [4, NOSPLAN_QUOTA_MAX_RES]");
    if (E < 0 || NS < 0 || P < 0 || Nn < 0) return fail(NOSPLAN_E_INVALID, "negative size");
    if (qs->gpu_memory_col >= R || (qs->gpu_memory_col >= 0 && qs->gpu_memory_col < 4)) return fail(NOSPLAN_E_INVALID, "gpu_memory_col out of range");
    for (int ns = 0; ns < NS; ++ns) if (qs->ns_quota[ns] >= E) return fail(NOSPLAN_E_INVALID, "ns_quota out of range");
    if (Nn > 0 && (vq->node_ptr[0] != 0 || vq->node_ptr[Nn] != P)) return fail(NOSPLAN_E_INVALID, "node_ptr must span [0, n_pods]");
    if (!vq->nominated_in_eq) return fail(NOSPLAN_E_INVALID, "nominated_in_eq is required");
    VCU(cudaSetDevice(device));
    size_t o = 0;
    auto al = [](size_t x) { return (x + 255) & ~(size_t)255; };
    auto put = [&](size_t bytes) { size_t at = o; o = al(o + bytes); return at; };
    const size_t ER = (size_t)E * R;
    const size_t o_nsq = put(4 * (size_t)NS), o_min = put(8 * ER), o_used = put(8 * ER), o_mm = put(4 * (size_t)E), o_um = put(4 * (size_t)E),
                 o_ptr = put(4 * ((size_t)Nn + 1)), o_ns = put(4 * (size_t)P), o_prio = put(4 * (size_t)P), o_over = put((size_t)P),
                 o_req = put(8 * (size_t)P * R), o_rmask = put(4 * (size_t)P);
    const size_t upload_end = o;
    const size_t o_vic = put((size_t)P), o_ovf = put(16);
    const size_t out_end = o;
    const size_t o_agg = put(8 * 2 * VQR), o_om = put(16);
    const size_t total = o;
    VCU(quota_ctx_ensure(ctx, total, std::max(upload_end, out_end - upload_end)));
    unsigned char* S = ctx->hs; unsigned char* A = ctx->d;
    auto cp = [&](size_t at, const void* src, size_t bytes) { if (bytes) { if (src) std::memcpy(S + at, src, bytes); else std::memset(S + at, 0, bytes); } };
    cp(o_nsq, qs->ns_quota, 4 * (size_t)NS); cp(o_min, qs->min, 8 * ER); cp(o_used, qs->used_init, 8 * ER);
    cp(o_mm, qs->min_mask, 4 * (size_t)E); cp(o_um, qs->used_mask, 4 * (size_t)E);
    cp(o_ptr, vq->node_ptr, 4 * ((size_t)Nn + 1)); cp(o_ns, vq->ns, 4 * (size_t)P); cp(o_prio, vq->priority, 4 * (size_t)P);
    cp(o_over, vq->over_quota, (size_t)P); cp(o_req, vq->request, 8 * (size_t)P * R); cp(o_rmask, vq->request_mask, 4 * (size_t)P);
    cudaEvent_t* ev = nullptr;
    VCU(quota_ctx_events(ctx, &ev));
    VCU(cudaMemcpyAsync(A, S, upload_end, cudaMemcpyHostToDevice, st));
    VCU(cudaMemsetAsync(A + upload_end, 0, total - upload_end, st));
    DevVictims v; std::memset(&v, 0, sizeof v);
    v.E = E; v.R = R; v.NS = NS; v.gm = qs->gpu_memory_col; v.n_nodes = Nn; v.n_pods = P; v.pn = vq->preemptor_ns; v.pprio = vq->preemptor_priority;
    v.ns_quota = (const int32_t*)(A + o_nsq); v.qmin = (const long long*)(A + o_min); v.used = (const long long*)(A + o_used);
    v.min_mask = (const uint32_t*)(A + o_mm); v.used_mask = (const uint32_t*)(A + o_um);
    for (int c = 0; c < VQR; ++c) {
        v.coeff[c] = (c < R && qs->gpu_memory_coeff) ? qs->gpu_memory_coeff[c] : 0;
        v.nom[c] = c < R ? vq->nominated_in_eq[c] : 0;
    }
    v.node_ptr = (const int32_t*)(A + o_ptr); v.ns = (const int32_t*)(A + o_ns); v.prio = (const int32_t*)(A + o_prio);
    v.over = (const uint8_t*)(A + o_over); v.req = (const long long*)(A + o_req);
    v.rmask = vq->request_mask ? (const uint32_t*)(A + o_rmask) : nullptr;
    v.agg = (long long*)(A + o_agg); v.omask = (unsigned*)(A + o_om); v.victim = (uint8_t*)(A + o_vic); v.overflow = (int32_t*)(A + o_ovf);
    VCU(cudaEventRecord(ev[0], st));
    k_victims_aggregate<<<1, V_AGG_THREADS, 0, st>>>(v);
    if (Nn > 0) {
        int blocks = (Nn + 7) / 8;
        if (blocks > 148 * 4) blocks = 148 * 4;
        k_quota_victims<<<blocks, 256, 0, st>>>(v);
    }
    VCU(cudaGetLastError());
    VCU(cudaEventRecord(ev[1], st));
    VCU(cudaMemcpyAsync(S, A + upload_end, out_end - upload_end, cudaMemcpyDeviceToHost, st));
    VCU(cudaStreamSynchronize(st));
    if (ms_k) cudaEventElapsedTime(ms_k, ev[0], ev[1]);
    if (launches) *launches = 1 + (Nn > 0);
    int32_t overflow; std::memcpy(&overflow, S + (o_ovf - upload_end), 4);
    if (overflow) return fail(NOSPLAN_E_LIMIT, "a node's victims come from more than 32 namespaces");
    if (P) std::memcpy(victim, S + (o_vic - upload_end), (size_t)P);
    return NOSPLAN_OK;
#undef VCU
}
