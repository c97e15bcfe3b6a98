// TEST INFRASTRUCTURE ONLY -- CPU restatement, on the packed structs of include/nosplan.h, of the first pass of
// preemptor.SelectVictimsOnNode (pkg/scheduler/plugins/capacityscheduling/capacity_scheduling.go:511-596): which pods of
// a node become potential victims.  Checked against oracle/nos_quota_model.py:select_potential_victims on random cases.
// The reference's own test for this path asserts nothing about the victims (capacity_scheduling_test.go:525-531 diffs
// `got` with itself), so parity here is pinned by the line-by-line restatement only.
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
}  // namespace

extern "C" int nos_oracle_quota_victims(const nosplan_quotas* q, const nosplan_victim_query* vq, uint8_t* victim) {
    if (!q || !vq || !victim) return NOSPLAN_E_INVALID;
    const int E = q->n_quotas, R = q->n_res, NS = q->n_namespaces, gm = q->gpu_memory_col;
    (void)E;
    // aggregates of the snapshot (one clone per namespace key)
    int64_t a_min[QR] = {0}, a_over0[QR] = {0};
    uint32_t omask0 = 0;
    for (int ns = 0; ns < NS; ++ns) {
        const int e = q->ns_quota[ns];
        if (e < 0) continue;
        for (int c = 0; c < R; ++c) {
            const int64_t mn = q->min[(size_t)e * R + c], us = q->used_init ? q->used_init[(size_t)e * R + c] : 0;
            a_min[c] += mn;
            a_over0[c] += std::max<int64_t>(0, mn - us);
        }
        omask0 |= (q->min_mask[e] & SCALAR_BITS) | (q->used_mask ? (q->used_mask[e] & SCALAR_BITS) : 0);
    }
    const int pn = vq->preemptor_ns;
    const int ep = (pn >= 0 && pn < NS) ? q->ns_quota[pn] : -1;
    const int64_t* nom = vq->nominated_in_eq;

    for (int node = 0; node < vq->n_nodes; ++node) {
        std::map<int, std::vector<int64_t>> removed;       // namespace -> un-reserved amounts in this node's clone
        int64_t a_over[QR];
        std::memcpy(a_over, a_over0, sizeof a_over);
        uint32_t omask = omask0;
        auto used_of = [&](int ns, int c) {
            const int e = q->ns_quota[ns];
            int64_t u = q->used_init ? q->used_init[(size_t)e * R + c] : 0;
            auto it = removed.find(ns);
            if (it != removed.end()) u -= it->second[c];
            return u;
        };
        auto guaranteed = [&](int e, int c) {              // GetGuaranteedOverquotas (elasticquotainfo.go:81-119), column c
            const bool in_min = c < 4 || (q->min_mask[e] >> c & 1);
            double pct = 0.0;
            if (in_min && a_min[c] > 0) pct = approx(c, q->min[(size_t)e * R + c]) / approx(c, a_min[c]);
            const bool in_res = c < 4 || (omask >> c & 1);
            return in_res ? go_int64(std::floor((double)a_over[c] * pct)) : (int64_t)0;
        };
        bool more_than_min = false;
        if (ep >= 0)
            for (int c = 0; c < R; ++c)
                if (cmp_col(c, q->min_mask[ep]) && nom[c] + used_of(pn, c) > q->min[(size_t)ep * R + c]) more_than_min = true;

        for (int i = vq->node_ptr[node]; i < vq->node_ptr[node + 1]; ++i) {
            victim[i] = 0;
            const int v = vq->ns[i];
            const int ev = (v >= 0 && v < NS) ? q->ns_quota[v] : -1;
            bool is_victim = false;
            if (ep < 0) {                                   // preemptor without quota (:582-596)
                if (ev >= 0) continue;
                victim[i] = vq->priority[i] < vq->preemptor_priority;
                continue;
            }
            if (ev < 0) continue;
            if (more_than_min) {
                if (v == pn) is_victim = vq->priority[i] < vq->preemptor_priority;      // :527-535
                else if (vq->over_quota[i]) {                                            // :541-559
                    const uint32_t ymask = omask | (q->min_mask[ep] & SCALAR_BITS);
                    bool lte = true;
                    for (int c = 0; c < R; ++c)
                        if (cmp_col(c, ymask) && nom[c] + used_of(pn, c) > guaranteed(ep, c) + q->min[(size_t)ep * R + c]) lte = false;
                    if (lte) {
                        const uint32_t vmask = omask | (q->min_mask[ev] & SCALAR_BITS);
                        for (int c = 0; c < R; ++c)
                            if (cmp_col(c, vmask) && used_of(v, c) > guaranteed(ev, c) + q->min[(size_t)ev * R + c]) is_victim = true;
                    }
                }
            } else if (v != pn && vq->over_quota[i]) {                                   // :569-577
                for (int c = 0; c < R; ++c)
                    if (cmp_col(c, q->min_mask[ev]) && used_of(v, c) > q->min[(size_t)ev * R + c]) is_victim = true;
            }
            if (!is_victim) continue;
            victim[i] = 1;
            // removePod -> deletePodIfPresent -> unreserveResource (elasticquotainfo.go:202-208) in this node's clone
            int64_t req[QR]; int64_t g = 0; uint32_t rmask = 0;
            for (int c = 0; c < R; ++c) {
                req[c] = c == gm ? 0 : vq->request[(size_t)i * R + c];
                if (c != gm) g += q->gpu_memory_coeff[c] * req[c];
                if (c >= 4 && req[c] != 0) rmask |= 1u << c;
            }
            if (vq->request_mask) rmask = vq->request_mask[i] & SCALAR_BITS;
            if (gm >= 0) { req[gm] = g; rmask |= 1u << gm; }
            auto& rem = removed[v];
            if (rem.empty()) rem.assign(QR, 0);
            for (int c = 0; c < R; ++c) {
                if (!(c == 0 || c == 1 || (c >= 4 && (rmask >> c & 1)))) continue;
                const int64_t mn = q->min[(size_t)ev * R + c], old = used_of(v, c);
                rem[c] += req[c];
                a_over[c] += std::max<int64_t>(0, mn - (old - req[c])) - std::max<int64_t>(0, mn - old);
            }
            omask |= rmask;
        }
    }
    return NOSPLAN_OK;
}
