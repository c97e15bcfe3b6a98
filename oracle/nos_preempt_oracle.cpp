// TEST INFRASTRUCTURE ONLY -- CPU restatement, on the packed structs of include/nosplan.h, of the whole of
// preemptor.SelectVictimsOnNode (pkg/scheduler/plugins/capacityscheduling/capacity_scheduling.go:468-675).
// Checked against oracle/nos_quota_model.py:select_victims_on_node on random cases and on the three fixtures of the reference's own
// test of this path (TestDryRunPreemption, capacity_scheduling_test.go:249-535 -> tests/golden/dryrun_preemption.json): that test
// asserts the candidate count only, which is what it pins here; the victims its `want` names are reproduced where they do not
// depend on the order of equal-priority pods.  NodeInfo / NodeResourcesFit come from k8s.io/kubernetes v1.25.4, which is not in
// the reference tree: restated.
#include <algorithm>
#include <cmath>
#include <cstdint>
#include <cstring>
#include <map>
#include <vector>

#include "../include/nosplan.h"

namespace {
constexpr int QR = NOSPLAN_QUOTA_MAX_RES;
constexpr uint32_t SCALAR_BITS = 0x7ffffff0u;

inline int64_t go_int64(double x) {
    if (!(x == x) || x >= 9223372036854775808.0 || x < -9223372036854775808.0) return INT64_MIN;
    return (int64_t)x;
}
inline double approx(int c, int64_t v) { return c == 0 ? (double)v * 1e-3 : (double)v; }
inline bool cmp_col(int c, uint32_t ymask) { return c == 0 || c == 1 || (c >= 4 && (ymask >> c & 1)); }

// one node's private view: the quota snapshot clone (per namespace key) and the NodeInfo
struct NodeView {
    const nosplan_quotas* q;
    const nosplan_preempt_query* pq;
    int R, gm;
    std::map<int, std::vector<int64_t>> removed;   // namespace -> amounts un-reserved in this node's clone
    int64_t a_min[QR], a_used[QR], a_over[QR];
    uint32_t omask, amin_mask;
    int64_t alloc[QR], requested[QR];              // NodeInfo (column 3: allowed pods / len(Pods))
    std::vector<char> on_node;

    int64_t used_of(int ns, int c) const {
        const int e = q->ns_quota[ns];
        int64_t u = q->used_init ? q->used_init[(size_t)e * R + c] : 0;
        auto it = removed.find(ns);
        if (it != removed.end()) u -= it->second[c];
        return u;
    }
    int64_t guaranteed(int e, int c) const {       // GetGuaranteedOverquotas (elasticquotainfo.go:81-119), column c
        const bool in_min = c < 4 || (q->min_mask[e] >> c & 1);
        double pct = 0.0;
        if (in_min && a_min[c] > 0) pct = approx(c, q->min[(size_t)e * R + c]) / approx(c, a_min[c]);
        const bool in_res = c < 4 || (omask >> c & 1);
        return in_res ? go_int64(std::floor((double)a_over[c] * pct)) : (int64_t)0;
    }
    // the pod's quota request (gpu-memory filled) and the scalar keys present in it
    void quota_req(int i, int64_t* req, uint32_t& rmask) const {
        const nosplan_victim_query& v = pq->v;
        int64_t g = 0;
        rmask = 0;
        for (int c = 0; c < R; ++c) {
            req[c] = c == gm ? 0 : v.request[(size_t)i * R + c];
            if (c != gm) g += q->gpu_memory_coeff[c] * req[c];
            if (c >= 4 && req[c] != 0) rmask |= 1u << c;
        }
        if (v.request_mask) rmask = v.request_mask[i] & SCALAR_BITS;
        if (gm >= 0) { req[gm] = g; rmask |= 1u << gm; }
    }
    // reserve (sign = -1) / un-reserve (sign = +1) in the clone of the pod's namespace, aggregates kept in step
    void quota_move(int i, int sign) {
        const int ns = pq->v.ns[i];
        const int e = (ns >= 0 && ns < q->n_namespaces) ? q->ns_quota[ns] : -1;
        if (e < 0) return;
        int64_t req[QR]; uint32_t rmask;
        quota_req(i, req, rmask);
        auto& rem = removed[ns];
        if (rem.empty()) rem.assign(QR, 0);
        for (int c = 0; c < R; ++c) {
            if (!(c == 0 || c == 1 || (c >= 4 && (rmask >> c & 1)))) continue;
            const int64_t mn = q->min[(size_t)e * R + c], old = used_of(ns, c), d = sign * req[c];
            rem[c] += d;
            a_used[c] -= d;
            a_over[c] += std::max<int64_t>(0, mn - (old - d)) - std::max<int64_t>(0, mn - old);
        }
        omask |= rmask;
    }
    // NodeInfo.RemovePod / AddPodInfo: Requested and len(Pods)
    void node_move(int i, int sign) {
        const nosplan_victim_query& v = pq->v;
        for (int c = 0; c < R; ++c) {
            if (c == 3 || c == gm) continue;
            const int64_t r = v.request[(size_t)i * R + c];
            const bool present = c < 3 || (v.request_mask ? (v.request_mask[i] >> c & 1) : r != 0);
            if (present) requested[c] += sign * r;
        }
        requested[3] += sign;
    }
    bool fits() const {                            // NodeResourcesFit.fitsRequest (k8s v1.25.4 plugins/noderesources/fit.go)
        const int64_t* f = pq->fit_request;
        bool ok = true;
        if (requested[3] + 1 > alloc[3]) ok = false;
        const uint32_t fm = pq->fit_mask & SCALAR_BITS & ~(gm >= 0 ? (1u << gm) : 0u);
        if (f[0] == 0 && f[1] == 0 && f[2] == 0 && fm == 0) return ok;
        for (int c = 0; c < 3; ++c) if (f[c] > alloc[c] - requested[c]) ok = false;
        for (int c = 4; c < R; ++c) if ((fm >> c & 1) && f[c] > alloc[c] - requested[c]) ok = false;
        return ok;
    }
    bool over_max_with(int ep, int pn, const int64_t* req) const {     // usedOverMaxWith (elasticquotainfo.go:214-219)
        if (!q->max_enforced[ep]) return false;
        for (int c = 0; c < R; ++c)
            if (cmp_col(c, q->max_mask[ep]) && req[c] + used_of(pn, c) > q->max[(size_t)ep * R + c]) return true;
        return false;
    }
    bool agg_over_min_with(const int64_t* req) const {                 // AggregatedUsedOverMinWith (:74-79)
        for (int c = 0; c < R; ++c)
            if (cmp_col(c, amin_mask) && a_used[c] + req[c] > a_min[c]) return true;
        return false;
    }
};
}  // namespace

extern "C" int nos_oracle_preempt_dryrun(const nosplan_quotas* q, const nosplan_preempt_query* pq, uint8_t* victim,
                                         int32_t* node_status, int32_t* num_violating) {
    if (!q || !pq || !victim || !node_status || !num_violating) return NOSPLAN_E_INVALID;
    const nosplan_victim_query& vq = pq->v;
    const int R = q->n_res, NS = q->n_namespaces, gm = q->gpu_memory_col;
    // aggregates of the snapshot (one clone per namespace key)
    int64_t a_min[QR] = {0}, a_used0[QR] = {0}, a_over0[QR] = {0};
    uint32_t omask0 = 0, amin_mask = 0;
    for (int ns = 0; ns < NS; ++ns) {
        const int e = q->ns_quota[ns];
        if (e < 0) continue;
        for (int c = 0; c < R; ++c) {
            const int64_t mn = q->min[(size_t)e * R + c], us = q->used_init ? q->used_init[(size_t)e * R + c] : 0;
            a_min[c] += mn; a_used0[c] += us;
            a_over0[c] += std::max<int64_t>(0, mn - us);
        }
        amin_mask |= q->min_mask[e] & SCALAR_BITS;
        omask0 |= (q->min_mask[e] & SCALAR_BITS) | (q->used_mask ? (q->used_mask[e] & SCALAR_BITS) : 0);
    }
    const int pn = vq.preemptor_ns;
    const int ep = (pn >= 0 && pn < NS) ? q->ns_quota[pn] : -1;
    const int64_t* nom = vq.nominated_in_eq;

    for (int node = 0; node < vq.n_nodes; ++node) {
        const int p0 = vq.node_ptr[node], p1 = vq.node_ptr[node + 1];
        NodeView nv;
        nv.q = q; nv.pq = pq; nv.R = R; nv.gm = gm;
        std::memcpy(nv.a_min, a_min, sizeof a_min); std::memcpy(nv.a_used, a_used0, sizeof a_used0);
        std::memcpy(nv.a_over, a_over0, sizeof a_over0);
        nv.omask = omask0; nv.amin_mask = amin_mask;
        for (int c = 0; c < QR; ++c) {
            nv.alloc[c] = c < R ? pq->node_alloc[(size_t)node * R + c] : 0;
            nv.requested[c] = c < R ? pq->node_requested[(size_t)node * R + c] : 0;
        }
        nv.on_node.assign(p1 - p0, 1);
        for (int i = p0; i < p1; ++i) victim[i] = 0;
        node_status[node] = NOSPLAN_VICTIMS_OK; num_violating[node] = 0;
        auto remove_pod = [&](int i) {                       // removePod closure (:492-501); false = NodeInfo.RemovePod error
            if (!nv.on_node[i - p0]) return false;
            nv.on_node[i - p0] = 0;
            nv.node_move(i, -1);
            nv.quota_move(i, +1);
            return true;
        };
        auto add_pod = [&](int i) { nv.on_node[i - p0] = 1; nv.node_move(i, +1); nv.quota_move(i, -1); };

        // ---- first pass (:511-596)
        std::vector<int> potential;
        bool more_than_min = false;
        if (ep >= 0)
            for (int c = 0; c < R; ++c)
                if (cmp_col(c, q->min_mask[ep]) && nom[c] + nv.used_of(pn, c) > q->min[(size_t)ep * R + c]) more_than_min = true;
        for (int i = p0; i < p1; ++i) {
            const int v = vq.ns[i];
            const int ev = (v >= 0 && v < NS) ? q->ns_quota[v] : -1;
            bool is_victim = false;
            if (ep < 0) is_victim = ev < 0 && vq.priority[i] < vq.preemptor_priority;
            else if (ev >= 0) {
                if (more_than_min) {
                    if (v == pn) is_victim = vq.priority[i] < vq.preemptor_priority;
                    else if (vq.over_quota[i]) {
                        const uint32_t ymask = nv.omask | (q->min_mask[ep] & SCALAR_BITS);
                        bool lte = true;
                        for (int c = 0; c < R; ++c)
                            if (cmp_col(c, ymask) && nom[c] + nv.used_of(pn, c) > nv.guaranteed(ep, c) + q->min[(size_t)ep * R + c]) lte = false;
                        if (lte) {
                            const uint32_t vmask = nv.omask | (q->min_mask[ev] & SCALAR_BITS);
                            for (int c = 0; c < R; ++c)
                                if (cmp_col(c, vmask) && nv.used_of(v, c) > nv.guaranteed(ev, c) + q->min[(size_t)ev * R + c]) is_victim = true;
                        }
                    }
                } else if (v != pn && vq.over_quota[i]) {
                    for (int c = 0; c < R; ++c)
                        if (cmp_col(c, q->min_mask[ev]) && nv.used_of(v, c) > q->min[(size_t)ev * R + c]) is_victim = true;
                }
            }
            if (is_victim) { potential.push_back(i); remove_pod(i); }
        }
        if (potential.empty()) { node_status[node] = NOSPLAN_VICTIMS_NONE; continue; }             // :598-602
        if (!nv.fits()) { node_status[node] = NOSPLAN_VICTIMS_FILTER; continue; }                   // :610-612
        if (ep >= 0) {                                                                              // :617-624
            if (nv.over_max_with(ep, pn, pq->pod_req)) { node_status[node] = NOSPLAN_VICTIMS_MAX_QUOTA; continue; }
            if (nv.agg_over_min_with(pq->pod_req)) { node_status[node] = NOSPLAN_VICTIMS_MIN_QUOTA; continue; }
        }
        // ---- most important first; PodDisruptionBudget split (:628-635, :850-895)
        std::reverse(potential.begin(), potential.end());
        std::vector<int32_t> allowed(pq->pdb_allowed, pq->pdb_allowed + pq->n_pdbs);
        std::vector<int> violating, non_violating;
        for (int i : potential) {
            bool bad = false;
            for (int k = pq->pod_pdb_ptr[i]; k < pq->pod_pdb_ptr[i + 1]; ++k)
                if (--allowed[pq->pod_pdb[k]] < 0) bad = true;
            (bad ? violating : non_violating).push_back(i);
        }
        // ---- reprieve (:636-673)
        bool error = false;
        auto reprieve = [&](int i, bool& fits) {
            add_pod(i);
            fits = nv.fits();
            if (!fits) {
                if (!remove_pod(i)) return false;
                victim[i] = 1;
            }
            if (ep >= 0 && (nv.over_max_with(ep, pn, nom) || nv.agg_over_min_with(pq->nominated))) {
                if (!remove_pod(i)) return false;            // already gone when it did not fit: RemovePod errors
                victim[i] = 1;
            }
            return true;
        };
        for (int i : violating) {
            bool fits;
            if (!reprieve(i, fits)) { error = true; break; }
            if (!fits) num_violating[node]++;
        }
        for (size_t k = 0; !error && k < non_violating.size(); ++k) {
            bool fits;
            if (!reprieve(non_violating[k], fits)) error = true;
        }
        if (error) {
            node_status[node] = NOSPLAN_VICTIMS_ERROR; num_violating[node] = 0;
            for (int i = p0; i < p1; ++i) victim[i] = 0;
        }
    }
    return NOSPLAN_OK;
}
