// TEST INFRASTRUCTURE ONLY -- CPU restatement of the ElasticQuota accounting on the packed quota structs of
// include/nosplan.h.  Only tests/, __graft_entry__.smoke() and bench.py's CPU-baseline legs may call it.
// Pinned against oracle/nos_quota_model.py (which reproduces the vectors ported from the reference's
// elasticquotainfo_test.go / capacity_scheduling_test.go / elasticquota_test.go) on those vectors and on random cases.
//
// Follows pkg/scheduler/plugins/capacityscheduling/elasticquotainfo.go:74-361, capacity_scheduling.go:190-278,
// pkg/gpu/util/resource.go:60-86, internal/controllers/elasticquota/elasticquota.go:38-118 (paths relative to
// /root/reference).  Floating point: IEEE double, no contraction (-ffp-contract=off), same operation order as the Go code.
#include <algorithm>
#include <chrono>
#include <cmath>
#include <cstdint>
#include <cstring>
#include <vector>

#include "../include/nosplan.h"

namespace {

constexpr int QR = NOSPLAN_QUOTA_MAX_RES;
constexpr uint32_t SCALAR_BITS = 0x7ffffff0u;   // columns 4..30

// int64(f) as compiled for amd64 (CVTTSD2SQ)
inline int64_t go_int64(double x) {
    if (!(x == x) || x >= 9223372036854775808.0 || x < -9223372036854775808.0) return INT64_MIN;
    return (int64_t)x;
}

// Quantity.AsApproximateFloat64 of the quantity FromFrameworkToList builds for column c (pkg/resource/resource.go:35-50)
inline double approx(int c, int64_t v) { return c == 0 ? (double)v * 1e-3 : (double)v; }
// Quantity.Value() > 0
inline bool value_positive(int c, int64_t v) { return c == 0 ? v > 0 : v > 0; }

struct Q {
    int E, R, NS, gm;
    const nosplan_quotas* q;
    std::vector<int64_t> used;       // [E][R]
    std::vector<uint32_t> umask;     // [E]
};

void pod_request(const Q& s, const nosplan_quota_pods* p, int i, int64_t* req, uint32_t* mask) {
    const int R = s.R;
    int64_t gmem = 0;
    uint32_t m = 0;
    for (int c = 0; c < R; ++c) {
        req[c] = c == s.gm ? 0 : p->request[(size_t)i * R + c];
        if (c != s.gm) gmem += s.q->gpu_memory_coeff[c] * req[c];
        if (c >= 4 && req[c] != 0) m |= 1u << c;
    }
    if (p->request_mask) m = p->request_mask[i] & SCALAR_BITS;
    if (s.gm >= 0) { req[s.gm] = gmem; m |= 1u << s.gm; }      // the key is always set (pkg/gpu/util/resource.go:64-66)
    *mask = m;
}

// sumGreaterThan (elasticquotainfo.go:319-339): cpu and memory always, scalars only if present in y
bool sum_greater_than(int R, const int64_t* x1, const int64_t* x2, const int64_t* y, uint32_t ymask) {
    if (x1[0] + x2[0] > y[0]) return true;
    if (x1[1] + x2[1] > y[1]) return true;
    for (int c = 4; c < R; ++c)
        if ((ymask >> c & 1) && x1[c] + x2[c] > y[c]) return true;
    return false;
}

}  // namespace

static int oracle_quota(const nosplan_quotas* q, const nosplan_quota_pods* running, const nosplan_quota_pods* pending,
                        nosplan_quota_out* out, double* seconds, std::vector<int>* patch_order);

extern "C" int nos_oracle_quota(const nosplan_quotas* q, const nosplan_quota_pods* running, const nosplan_quota_pods* pending,
                                nosplan_quota_out* out, double* seconds) {
    return oracle_quota(q, running, pending, out, seconds, nullptr);
}

// patchCapacityInfoIfDifferent over every quota (elasticquota.go:44-60, :120-149): the pods whose label differs from the
// desired one, by quota row and inside a quota in the order of the reference's sort
extern "C" int nos_oracle_label_changes(const nosplan_quotas* q, const nosplan_quota_pods* running, const uint8_t* current,
                                        int32_t* pod, uint8_t* desired, int32_t* n_changes) {
    if (!q || !running || !current || !pod || !desired || !n_changes) return NOSPLAN_E_INVALID;
    std::vector<uint8_t> over((size_t)std::max(1, running->n_pods));
    nosplan_quota_out out;
    std::memset(&out, 0, sizeof out);
    out.over_quota = over.data();
    std::vector<int> order;
    const int rc = oracle_quota(q, running, nullptr, &out, nullptr, &order);
    if (rc != NOSPLAN_OK) return rc;
    int n = 0;
    for (int i : order) {
        const uint8_t want = over[i] ? NOSPLAN_LABEL_OVER_QUOTA : NOSPLAN_LABEL_IN_QUOTA;
        if (current[i] != want) { pod[n] = i; desired[n] = want; ++n; }
    }
    *n_changes = n;
    return NOSPLAN_OK;
}

static int oracle_quota(const nosplan_quotas* q, const nosplan_quota_pods* running, const nosplan_quota_pods* pending,
                        nosplan_quota_out* out, double* seconds, std::vector<int>* patch_order) {
    if (!q || !out) return NOSPLAN_E_INVALID;
    if (q->n_res < 4 || q->n_res > QR) return NOSPLAN_E_LIMIT;
    auto t0 = std::chrono::steady_clock::now();
    Q s{q->n_quotas, q->n_res, q->n_namespaces, q->gpu_memory_col, q, {}, {}};
    const int E = s.E, R = s.R;
    s.used.assign((size_t)E * R, 0);
    s.umask.assign(E, 0);
    for (int e = 0; e < E; ++e) {
        if (q->used_init) for (int c = 0; c < R; ++c) s.used[(size_t)e * R + c] = q->used_init[(size_t)e * R + c];
        s.umask[e] = q->used_mask ? (q->used_mask[e] & SCALAR_BITS) : 0;
    }
    int64_t req[QR]; uint32_t rmask;

    // ---- Q1/Q2: reserveResource of every running pod (Memory, MilliCPU and scalars only)
    const int Pr = running ? running->n_pods : 0;
    for (int i = 0; i < Pr; ++i) {
        const int ns = running->ns[i];
        const int e = (ns >= 0 && ns < s.NS) ? q->ns_quota[ns] : -1;
        if (e < 0) continue;
        pod_request(s, running, i, req, &rmask);
        int64_t* u = &s.used[(size_t)e * R];
        u[0] += req[0]; u[1] += req[1];
        for (int c = 4; c < R; ++c) if (rmask >> c & 1) u[c] += req[c];
        s.umask[e] |= rmask;
    }
    if (out->used) std::memcpy(out->used, s.used.data(), sizeof(int64_t) * (size_t)E * R);
    if (out->used_mask) std::memcpy(out->used_mask, s.umask.data(), sizeof(uint32_t) * E);

    // ---- Q3: aggregates over the namespace -> info map (a composite quota counts once per namespace)
    int64_t amin[QR] = {0}, aused[QR] = {0}, aover[QR] = {0};
    uint32_t mmask = 0, umask = 0, omask = 0;
    for (int ns = 0; ns < s.NS; ++ns) {
        const int e = q->ns_quota[ns];
        if (e < 0) continue;
        const bool has_min = (q->min_mask[e] & NOSPLAN_QUOTA_NONNIL) != 0;
        for (int c = 0; c < R; ++c) {
            const int64_t mn = has_min ? q->min[(size_t)e * R + c] : 0, us = s.used[(size_t)e * R + c];
            amin[c] += mn; aused[c] += us;
            aover[c] += std::max<int64_t>(0, mn - us);             // SubtractNonNegative then Sum
        }
        if (has_min) mmask |= q->min_mask[e] & SCALAR_BITS;
        umask |= s.umask[e];
        omask |= (has_min ? (q->min_mask[e] & SCALAR_BITS) : 0) | s.umask[e];
    }
    if (out->agg_min) std::memcpy(out->agg_min, amin, sizeof(int64_t) * R);
    if (out->agg_used) std::memcpy(out->agg_used, aused, sizeof(int64_t) * R);
    if (out->agg_overquota) std::memcpy(out->agg_overquota, aover, sizeof(int64_t) * R);
    if (out->agg_masks) { out->agg_masks[0] = mmask; out->agg_masks[1] = umask; out->agg_masks[2] = omask; }

    // ---- Q4: GetGuaranteedOverquotas for every quota
    if (out->guaranteed) {
        for (int e = 0; e < E; ++e) {
            const bool has_min = (q->min_mask[e] & NOSPLAN_QUOTA_NONNIL) != 0;
            for (int c = 0; c < R; ++c) {
                double pct = 0.0;
                const bool in_min = has_min && (c < 4 || (q->min_mask[e] >> c & 1));
                if (in_min && value_positive(c, amin[c])) pct = approx(c, q->min[(size_t)e * R + c]) / approx(c, amin[c]);
                const bool in_res = c < 4 || (omask >> c & 1);
                out->guaranteed[(size_t)e * R + c] = in_res ? go_int64(std::floor((double)aover[c] * pct)) : 0;
            }
        }
    }

    // ---- Q7: PreFilter verdict of every pending pod (partitioner context: no nominated pods)
    const int Pp = pending ? pending->n_pods : 0;
    if (out->prefilter) {
        int64_t zero[QR] = {0};
        for (int i = 0; i < Pp; ++i) {
            const int ns = pending->ns[i];
            const int e = (ns >= 0 && ns < s.NS) ? q->ns_quota[ns] : -1;
            uint8_t code = NOSPLAN_PREFILTER_SUCCESS;
            if (e >= 0) {
                pod_request(s, pending, i, req, &rmask);
                if (q->max_enforced[e] && sum_greater_than(R, req, &s.used[(size_t)e * R], &q->max[(size_t)e * R], q->max_mask[e]))
                    code = NOSPLAN_PREFILTER_OVER_MAX;
                else {
                    int64_t u[QR];
                    for (int c = 0; c < R; ++c) u[c] = aused[c] + req[c];          // used.Add(FromFrameworkToList(podRequest))
                    if (sum_greater_than(R, u, zero, amin, mmask)) code = NOSPLAN_PREFILTER_OVER_MIN;
                }
            }
            out->prefilter[i] = code;
        }
    }

    // ---- Q6: operator labelling per quota (elasticquota.go:38-118)
    if (out->over_quota || out->operator_used || out->ambiguous_order) {
        if (out->ambiguous_order) *out->ambiguous_order = 0;
        if (out->operator_used) std::memset(out->operator_used, 0, sizeof(int64_t) * (size_t)E * R);
        std::vector<std::vector<int>> by_quota(E);
        for (int i = 0; i < Pr; ++i) {
            const int ns = running->ns[i];
            const int e = (ns >= 0 && ns < s.NS) ? q->ns_quota[ns] : -1;
            if (e >= 0) by_quota[e].push_back(i);
            else if (out->over_quota) out->over_quota[i] = 0;
        }
        std::vector<int64_t> rq((size_t)Pr * R);
        for (int i = 0; i < Pr; ++i) pod_request(s, running, i, &rq[(size_t)i * R], &rmask);
        auto req_equal = [&](int a, int b) { return std::memcmp(&rq[(size_t)a * R], &rq[(size_t)b * R], sizeof(int64_t) * R) == 0; };
        auto req_lte = [&](int a, int b) {                 // quota.LessThanOrEqual on dense rows (absent key == 0 on both sides)
            for (int c = 0; c < R; ++c) if (rq[(size_t)a * R + c] > rq[(size_t)b * R + c]) return false;
            return true;
        };
        auto less = [&](int a, int b) {
            if (running->creation_ts[a] != running->creation_ts[b]) return running->creation_ts[a] < running->creation_ts[b];
            if (running->priority[a] != running->priority[b]) return running->priority[a] < running->priority[b];
            if (!req_equal(a, b)) return req_lte(a, b);
            return running->name_rank[a] < running->name_rank[b];
        };
        for (int e = 0; e < E; ++e) {
            std::vector<int>& v = by_quota[e];
            for (size_t i = 1; i < v.size(); ++i) {         // stable insertion sort (canonical for tie groups)
                int x = v[i]; size_t j = i;
                while (j > 0 && less(x, v[j - 1])) { v[j] = v[j - 1]; --j; }
                v[j] = x;
            }
            if (out->ambiguous_order)
                for (size_t i = 0; i < v.size(); ++i)
                    for (size_t j = i + 1; j < v.size() && running->creation_ts[v[j]] == running->creation_ts[v[i]] &&
                                           running->priority[v[j]] == running->priority[v[i]]; ++j)
                        if (!req_equal(v[i], v[j]) && req_lte(v[i], v[j]) == req_lte(v[j], v[i])) *out->ambiguous_order = 1;
            int64_t used[QR] = {0};
            const uint32_t mm = q->min_mask[e];
            for (int i : v) {
                bool in_quota = true;
                for (int c = 0; c < R; ++c) {
                    used[c] += rq[(size_t)i * R + c];
                    if ((mm >> c & 1) && used[c] > q->min[(size_t)e * R + c]) in_quota = false;
                }
                if (out->over_quota) out->over_quota[i] = in_quota ? 0 : 1;
                if (patch_order) patch_order->push_back(i);
            }
            if (out->operator_used)
                for (int c = 0; c < R; ++c) out->operator_used[(size_t)e * R + c] = (mm >> c & 1) ? used[c] : 0;
        }
    }
    if (seconds) *seconds = std::chrono::duration<double>(std::chrono::steady_clock::now() - t0).count();
    return NOSPLAN_OK;
}
