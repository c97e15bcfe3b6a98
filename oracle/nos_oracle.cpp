// TEST INFRASTRUCTURE ONLY -- CPU restatement of the nos partition-planning hot path on the packed
// (SoA) snapshot of include/nosplan.h.  Only tests/, __graft_entry__.smoke() and bench.py's
// cpu_baseline / --impl reference legs may link or call this file; the product (nos_b200/) never does.
//
// Two modes that must produce identical plans (asserted by tests/test_oracle_packed.py):
//   NOS_ORACLE_FAITHFUL     follows the reference object graph literally: Fork() deep-clones every node
//                           (core/snapshot.go:33-41,85-91), every gate check re-sums Allocatable and
//                           Requested over ALL nodes (core/snapshot.go:132-141), candidate nodes are
//                           pointers into the pre-fork data (core/planner.go:89) so the aliasing rules
//                           of SURVEY.md section 8c fall out of shared_ptr identity exactly as in Go.
//                           Same O(N^2 P) asymptotics as the Go planner; used for small N and as the
//                           "reference algorithm" CPU baseline.
//   NOS_ORACLE_INCREMENTAL  same results with running cluster totals and no cloning (the derived
//                           state rules are written out in plan_incremental); used at full size.
// Two more modes, NOS_ORACLE_DEVICE_MODEL and NOS_ORACLE_FUSED_MODEL, are NOT restatements of the reference: they restate, in this
// file's terms, the algorithms the CUDA path uses -- the gate replay where the cluster-wide gate is not vacuous (plan_device_model)
// and the single-launch kernel's per-node pod loop (plan_fused_model) -- so that their exactness against plan_incremental is
// checked on the CPU too (tests/test_device_model.py).  Nothing times or ships them.
// Pinning: this file is checked against oracle/nos_model.py (the dict/string model that reproduces every
// golden vector ported from the reference's tests, tests/golden/) on the golden cases and on thousands
// of randomised small snapshots.
//
// Citations are relative to /root/reference (WindowsXp-Beta/nos v0.1.2).
#include <algorithm>
#include <chrono>
#include <cstdint>
#include <cstring>
#include <functional>
#include <memory>
#include <vector>

#include "../include/nosplan.h"

#define NOS_ORACLE_FAITHFUL 0
#define NOS_ORACLE_INCREMENTAL 1
#define NOS_ORACLE_DEVICE_MODEL 2     /* model of the CUDA path's gate replay (see plan_device_model); tests only */
#define NOS_ORACLE_FUSED_MODEL 3      /* model of the single-launch kernel (see plan_fused_model); tests only */

namespace {

constexpr int KP = NOSPLAN_MAX_PROFILES;
constexpr int RMAX = 3 + NOSPLAN_MAX_PROFILES + NOSPLAN_MAX_OTHER;

struct Ctx {
    int kind, K, R;
    const int32_t* pmem;
    const int32_t* pgi;
    const int32_t* geom_off;
    const uint8_t* geom;  // [rows][16]
};

struct Gpu {
    int model, index, mem;
    uint8_t used[KP], free_[KP];
    uint16_t umask, fmask;
};

struct Node {
    int ng;
    Gpu g[NOSPLAN_MAX_GPUS];
    int64_t alloc[RMAX], req[RMAX];
    int allowed_pods, num_pods;
};

struct PodReq {            // one pod, request split by the planner's kind
    int64_t base[3];
    int n_slice;           // entries in slice columns (ascending profile id == canonical visiting order)
    int sp[KP]; int64_t sq[KP];
    int n_other;
    int oc[NOSPLAN_MAX_OTHER]; int64_t oq[NOSPLAN_MAX_OTHER];
};

// ---------------------------------------------------------------- MIG (pkg/gpu/mig/gpu.go)
inline uint16_t row_mask(const uint8_t* row) {
    uint16_t m = 0;
    for (int p = 0; p < KP; ++p) if (row[p]) m |= (uint16_t)(1u << p);
    return m;
}

// AllowsGeometry (gpu.go:215-222): cmp.Equal -> identical key set and values
bool mig_allows_current(const Ctx& c, const Gpu& g) {
    uint16_t m = g.umask | g.fmask;
    for (int r = c.geom_off[g.model]; r < c.geom_off[g.model + 1]; ++r) {
        const uint8_t* row = c.geom + 16 * r;
        if (row_mask(row) != m) continue;
        bool eq = true;
        for (int p = 0; p < KP && eq; ++p) eq = (row[p] == (uint8_t)(g.used[p] + g.free_[p]));
        if (eq) return true;
    }
    return false;
}

// CanApplyGeometry (gpu.go:97-110) for a table row (allowed by construction)
inline bool mig_can_apply(const Gpu& g, const uint8_t* row) {
    for (int p = 0; p < KP; ++p)
        if ((g.umask >> p & 1) && row[p] < g.used[p]) return false;
    return true;
}

// ApplyGeometry (gpu.go:132-152)
inline void mig_apply(Gpu& g, const uint8_t* row) {
    uint16_t m = row_mask(row);
    for (int p = 0; p < KP; ++p) g.free_[p] = (m >> p & 1) ? (uint8_t)(row[p] - g.used[p]) : 0;  // zero kept (:141)
    g.fmask = m;
}

// GPU.UpdateGeometryFor (gpu.go:158-212); canonical arg-max: lowest table index among max scorers
bool mig_gpu_update(const Ctx& c, Gpu& g, const int32_t* required, int64_t& evals) {
    int best = -1, best_score = 0;
    for (int r = c.geom_off[g.model]; r < c.geom_off[g.model + 1]; ++r) {
        const uint8_t* row = c.geom + 16 * r;
        int score = 0; bool hit = false;
        for (int p = 0; p < c.K; ++p) {
            if (required[p] <= 0) continue;
            ++evals;
            if ((int)g.free_[p] >= required[p]) continue;
            int n = std::min((int)row[p] - (int)g.used[p], required[p]);
            if (n <= 0) continue;
            if (!mig_can_apply(g, row)) continue;
            score += n; hit = true;
        }
        if (hit && score > best_score) { best_score = score; best = r; }
    }
    if (best < 0) return false;
    mig_apply(g, c.geom + 16 * best);
    return true;
}

// ---------------------------------------------------------------- MPS (pkg/gpu/slicing/gpu.go)
inline int mps_spare(const Ctx& c, const Gpu& g) {  // MemoryGB - getTotSlicesMemory (:256-265)
    int tot = 0;
    for (int p = 0; p < c.K; ++p) tot += c.pmem[p] * ((int)g.used[p] + (int)g.free_[p]);
    return g.mem - tot;
}
inline bool mps_can_create_more(const Ctx& c, const Gpu& g) { return mps_spare(c, g) >= 1; }  // :250-254
inline bool mps_create(const Ctx& c, Gpu& g, int p, int num) {                                 // :238-247
    if (mps_spare(c, g) < c.pmem[p] * num) return false;
    g.free_[p] = (uint8_t)(g.free_[p] + num);
    g.fmask |= (uint16_t)(1u << p);
    return true;
}

// GPU.UpdateGeometryFor (slicing/gpu.go:162-232); profile ids ascend by (size, name) = canonical order
bool mps_gpu_update(const Ctx& c, Gpu& g, const int32_t* required) {
    int missing[KP]; bool in_missing[KP]; bool any = false;
    for (int p = 0; p < KP; ++p) { missing[p] = 0; in_missing[p] = false; }
    for (int p = 0; p < c.K; ++p) {                                     // getMissingSlices :222-232
        if (required[p] <= 0) continue;
        int d = required[p] - (int)g.free_[p];
        if (d > 0) { missing[p] = d; in_missing[p] = true; any = true; }
    }
    if (!any) return false;
    bool updated = false;
    uint8_t ofree[KP]; uint16_t omask = g.fmask;
    std::memcpy(ofree, g.free_, KP);
    for (int mp = 0; mp < c.K; ++mp) {                                  // :173-186 ascending (size, name)
        if (!in_missing[mp]) continue;
        if (mps_can_create_more(c, g)) {                               // :188-197
            int q = missing[mp];
            for (int i = 0; i < q; ++i) {
                if (!mps_create(c, g, mp, 1)) break;
                --missing[mp]; updated = true;
            }
        }
        int q = missing[mp];                                           // :199-211
        for (int p = 0; p < KP; ++p)
            if (omask >> p & 1) { g.free_[p] = 0; g.fmask &= (uint16_t)~(1u << p); }
        for (int i = 0; i < q; ++i) {
            if (!mps_can_create_more(c, g)) break;
            if (!mps_create(c, g, mp, 1)) break;
            --missing[mp]; updated = true;
        }
        for (int p = 0; p < KP; ++p)                                   // :213-216 restore, all-or-nothing
            if (omask >> p & 1) mps_create(c, g, p, ofree[p]);
    }
    return updated;
}

inline bool mps_gpu_has_free_capacity(const Ctx& c, const Gpu& g) {    // :131-136
    return g.fmask != 0 || mps_can_create_more(c, g);
}

// ---------------------------------------------------------------- shared GPU / node ops
// GPU.AddPod (mig/gpu.go:233-247, slicing/gpu.go:142-156): partial mutation on failure is kept
inline bool gpu_add_pod(Gpu& g, const PodReq& pr) {
    for (int e = 0; e < pr.n_slice; ++e) {
        int p = pr.sp[e]; int64_t q = pr.sq[e];
        if ((int64_t)g.free_[p] < q) return false;
        g.free_[p] = (uint8_t)(g.free_[p] - q);
        g.used[p] = (uint8_t)(g.used[p] + q);
        g.fmask |= (uint16_t)(1u << p);
        g.umask |= (uint16_t)(1u << p);
    }
    return true;
}

bool node_has_free_capacity(const Ctx& c, const Node& n) {  // mig/node.go:124-139, slicing/node.go:208-215
    for (int j = 0; j < n.ng; ++j) {
        const Gpu& g = n.g[j];
        if (c.kind == NOSPLAN_KIND_MIG) {
            if (g.fmask != 0) return true;
            if (!mig_allows_current(c, g)) return true;
        } else if (mps_gpu_has_free_capacity(c, g)) return true;
    }
    return false;
}

// Node.UpdateGeometryFor (mig/node.go:145-177, slicing/node.go:120-152)
bool node_update_geometry(const Ctx& c, Node& n, const int32_t* lack, int64_t& evals) {
    bool any_lack = false;
    for (int p = 0; p < c.K; ++p) any_lack |= lack[p] > 0;
    if (n.ng == 0 || !any_lack) return false;
    int32_t required[KP];
    for (int p = 0; p < KP; ++p) required[p] = p < c.K ? lack[p] : 0;
    bool any = false;
    for (int j = 0; j < n.ng; ++j) {
        Gpu& g = n.g[j];
        bool upd;
        if (c.kind == NOSPLAN_KIND_MIG) upd = mig_gpu_update(c, g, required, evals);
        else {
            for (int p = 0; p < c.K; ++p) evals += required[p] > 0;
            upd = mps_gpu_update(c, g, required);
        }
        any = any || upd;
        for (int p = 0; p < c.K; ++p)
            if ((g.fmask >> p & 1) && required[p] > 0) required[p] = std::max(0, required[p] - (int)g.free_[p]);
    }
    // computeScalarResources (mig/node.go:179-195): slice columns := node geometry, others kept
    for (int p = 0; p < c.K; ++p) {
        int64_t t = 0;
        for (int j = 0; j < n.ng; ++j) t += (int64_t)n.g[j].used[p] + n.g[j].free_[p];
        n.alloc[3 + p] = t;
    }
    return any;
}

// Node.AddPod (mig/node.go:201-210): first GPU that takes the pod; Requested grows through the shared
// pointer, the Pods append is lost on the NodeInfo value copy
bool node_add_pod(const Ctx&, Node& n, const PodReq& pr) {
    for (int j = 0; j < n.ng; ++j) {
        if (gpu_add_pod(n.g[j], pr)) {
            for (int k = 0; k < 3; ++k) n.req[k] += pr.base[k];
            for (int e = 0; e < pr.n_slice; ++e) n.req[3 + pr.sp[e]] += pr.sq[e];
            for (int e = 0; e < pr.n_other; ++e) n.req[pr.oc[e]] += pr.oq[e];
            return true;
        }
    }
    return false;
}

// NodeResourcesFit.Filter -> fitsRequest (k8s.io/kubernetes v1.25.4 plugins/noderesources/fit.go), restated
bool resources_fit(const Node& n, const PodReq& pr) {
    bool ok = true;
    if (n.num_pods + 1 > n.allowed_pods) ok = false;
    if (pr.base[0] == 0 && pr.base[1] == 0 && pr.base[2] == 0 && pr.n_slice == 0 && pr.n_other == 0) return ok;
    for (int k = 0; k < 3; ++k)
        if (pr.base[k] > n.alloc[k] - n.req[k]) ok = false;
    for (int e = 0; e < pr.n_slice; ++e)
        if (pr.sq[e] > n.alloc[3 + pr.sp[e]] - n.req[3 + pr.sp[e]]) ok = false;
    for (int e = 0; e < pr.n_other; ++e)
        if (pr.oq[e] > n.alloc[pr.oc[e]] - n.req[pr.oc[e]]) ok = false;
    return ok;
}

// ---------------------------------------------------------------- Go 1.18 sort.SliceStable (sort/zsortfunc.go)
struct GoStable {
    std::vector<int>& d;
    std::function<bool(int, int)> less_v;  // on values
    bool lt(int i, int j) { return less_v(d[i], d[j]); }
    void swap(int i, int j) { std::swap(d[i], d[j]); }
    void insertion(int a, int b) {
        for (int i = a + 1; i < b; ++i)
            for (int j = i; j > a && lt(j, j - 1); --j) swap(j, j - 1);
    }
    void swap_range(int a, int b, int n) { for (int i = 0; i < n; ++i) swap(a + i, b + i); }
    void rotate(int a, int m, int b) {
        int i = m - a, j = b - m;
        while (i != j) {
            if (i > j) { swap_range(m - i, m, j); i -= j; }
            else { swap_range(m - i, m + j - i, i); j -= i; }
        }
        swap_range(m - i, m, i);
    }
    void sym_merge(int a, int m, int b) {
        if (m - a == 1) {
            int i = m, j = b;
            while (i < j) { int h = (int)((unsigned)(i + j) >> 1); if (lt(h, a)) i = h + 1; else j = h; }
            for (int k = a; k < i - 1; ++k) swap(k, k + 1);
            return;
        }
        if (b - m == 1) {
            int i = a, j = m;
            while (i < j) { int h = (int)((unsigned)(i + j) >> 1); if (!lt(m, h)) i = h + 1; else j = h; }
            for (int k = m; k > i; --k) swap(k, k - 1);
            return;
        }
        int mid = (int)((unsigned)(a + b) >> 1), n = mid + m, start, r;
        if (m > mid) { start = n - b; r = mid; } else { start = a; r = m; }
        int p = n - 1;
        while (start < r) { int c = (int)((unsigned)(start + r) >> 1); if (!lt(p - c, c)) start = c + 1; else r = c; }
        int end = n - start;
        if (start < m && m < end) rotate(start, m, end);
        if (a < start && start < mid) sym_merge(a, start, mid);
        if (mid < end && end < b) sym_merge(mid, end, b);
    }
    void run() {
        int n = (int)d.size(), block = 20, a = 0, b = block;
        while (b <= n) { insertion(a, b); a = b; b += block; }
        insertion(a, n);
        while (block < n) {
            a = 0; b = 2 * block;
            while (b <= n) { sym_merge(a, a + block, b); a = b; b += 2 * block; }
            int m = a + block;
            if (m < n) sym_merge(a, m, n);
            block *= 2;
        }
    }
};

// ---------------------------------------------------------------- problem set-up
struct Problem {
    Ctx c;
    int N, G, P;
    std::vector<Node> nodes;
    std::vector<PodReq> pods;
    std::vector<int32_t> prio, tkey;
    std::vector<uint8_t> pre_ok;
    int n_keys;
    int filter_mode;
    int n_models = 0;      // MIG: models in the geometry table (the device model's saturation bound reads every row)
};

int build(const nosplan_snapshot* s, const nosplan_pods* p, const uint8_t* prefilter_ok, int filter_mode, Problem& pb) {
    if (!s || !p) return NOSPLAN_E_INVALID;
    if (s->n_profiles > KP || s->n_other > NOSPLAN_MAX_OTHER || s->gpus_per_node > NOSPLAN_MAX_GPUS) return NOSPLAN_E_LIMIT;
    pb.c = Ctx{s->kind, s->n_profiles, 3 + s->n_profiles + s->n_other, s->profile_mem_gb, s->profile_gi,
               s->model_geom_offset, s->geom_counts};
    pb.N = s->n_nodes; pb.G = s->gpus_per_node; pb.P = p->n_pods; pb.filter_mode = filter_mode;
    pb.n_models = s->kind == NOSPLAN_KIND_MIG ? s->n_models : 0;
    const int R = pb.c.R, G = pb.G;
    pb.nodes.resize(pb.N);
    for (int i = 0; i < pb.N; ++i) {
        Node& n = pb.nodes[i];
        std::memset(&n, 0, sizeof(Node));
        n.ng = s->node_gpu_count[i];
        for (int j = 0; j < n.ng; ++j) {
            int x = i * G + j;
            Gpu& g = n.g[j];
            g.model = s->gpu_model[x]; g.index = s->gpu_index[x]; g.mem = s->gpu_memory_gb[x];
            std::memcpy(g.used, s->gpu_used + 16 * x, 16);
            std::memcpy(g.free_, s->gpu_free + 16 * x, 16);
            g.umask = s->gpu_used_mask[x]; g.fmask = s->gpu_free_mask[x];
        }
        for (int k = 0; k < R; ++k) { n.alloc[k] = s->node_alloc[(size_t)i * R + k]; n.req[k] = s->node_requested[(size_t)i * R + k]; }
        n.allowed_pods = s->node_allowed_pods[i]; n.num_pods = s->node_num_pods[i];
    }
    pb.pods.resize(pb.P); pb.prio.assign(p->priority, p->priority + pb.P); pb.tkey.assign(p->tracker_key, p->tracker_key + pb.P);
    pb.pre_ok.assign(pb.P, 1);
    if (prefilter_ok) pb.pre_ok.assign(prefilter_ok, prefilter_ok + pb.P);
    pb.n_keys = 0;
    for (int i = 0; i < pb.P; ++i) {
        PodReq& pr = pb.pods[i];
        std::memset(&pr, 0, sizeof(PodReq));
        for (int k = 0; k < 3; ++k) pr.base[k] = p->base[3 * (size_t)i + k];
        for (int e = p->scalar_ptr[i]; e < p->scalar_ptr[i + 1]; ++e) {
            int col = p->scalar_col[e];
            if (col < 3 || col >= R) return NOSPLAN_E_INVALID;
            if (col < 3 + pb.c.K) {
                if (pr.n_slice >= KP) return NOSPLAN_E_INVALID;
                pr.sp[pr.n_slice] = col - 3; pr.sq[pr.n_slice++] = p->scalar_qty[e];
            } else {
                if (pr.n_other >= NOSPLAN_MAX_OTHER) return NOSPLAN_E_INVALID;
                pr.oc[pr.n_other] = col; pr.oq[pr.n_other++] = p->scalar_qty[e];
            }
        }
        pb.n_keys = std::max(pb.n_keys, pb.tkey[i] + 1);
    }
    return NOSPLAN_OK;
}

// NewPodSorter (core/util.go:34-73) -> order of original pod indices
std::vector<int> sort_pods(const Problem& pb) {
    std::vector<int> order(pb.P);
    for (int i = 0; i < pb.P; ++i) order[i] = i;
    const Ctx& c = pb.c;
    auto less = [&](int a, int b) {
        if (pb.prio[a] != pb.prio[b]) return pb.prio[a] > pb.prio[b];
        const PodReq &x = pb.pods[a], &y = pb.pods[b];
        if (x.n_slice == 0) return false;
        if (y.n_slice == 0) return false;
        int pa = x.sp[0], pb_ = y.sp[0];                      // canonical "first key": lowest profile id
        if (c.pmem[pa] < c.pmem[pb_]) return true;            // mig/profile.go:84-96, slicing/profile.go:36-42
        if (c.kind == NOSPLAN_KIND_MIG && c.pgi[pa] < c.pgi[pb_]) return true;
        return false;
    };
    GoStable gs{order, less};
    gs.run();
    return order;
}

void write_partitioning(const Node& n, int i, int G, nosplan_partitioning* out) {
    for (int j = 0; j < n.ng; ++j) {
        int x = i * G + j;
        for (int p = 0; p < KP; ++p) out->gpu_count[16 * (size_t)x + p] = (uint8_t)(n.g[j].used[p] + n.g[j].free_[p]);
        out->gpu_mask[x] = n.g[j].umask | n.g[j].fmask;
    }
}

struct Tracker {  // core/tracker.go:26-88 (requestedSlices is never read by the planner and is not kept)
    int32_t lack[KP];
    std::vector<int32_t> entry;  // [n_keys][KP]
    bool empty() const { for (int p = 0; p < KP; ++p) if (lack[p] > 0) return false; return true; }
    int count() const { int c = 0; for (int p = 0; p < KP; ++p) c += lack[p] > 0; return c; }
    void remove(int key) {
        int32_t* e = &entry[(size_t)key * KP];
        for (int p = 0; p < KP; ++p) if (e[p] > 0) { lack[p] -= e[p]; e[p] = 0; }
    }
};

// lacking slices of one pod given cluster totals (core/snapshot.go:74-76,132-165)
inline bool pod_lacks(const PodReq& pr, const int64_t* totA, const int64_t* totR, int32_t* out_lack /*nullable*/) {
    bool any = false;
    for (int e = 0; e < pr.n_slice; ++e) {
        int col = 3 + pr.sp[e];
        int64_t avail = std::max<int64_t>(0, totA[col] - totR[col]);
        int64_t diff = avail - pr.sq[e];
        if (diff < 0) { any = true; if (out_lack) out_lack[pr.sp[e]] += (int32_t)(-diff); }
    }
    return any;
}

// ---------------------------------------------------------------- faithful mode (literal object graph)
using NodeP = std::shared_ptr<Node>;

int plan_faithful(Problem& pb, nosplan_partitioning* out, nosplan_stats* st) {
    const Ctx& c = pb.c; const int N = pb.N, G = pb.G, P = pb.P, R = c.R;
    std::vector<NodeP> data(N), forked; bool has_fork = false;
    for (int i = 0; i < N; ++i) data[i] = std::make_shared<Node>(pb.nodes[i]);
    auto cur = [&]() -> std::vector<NodeP>& { return has_fork ? forked : data; };
    auto totals = [&](int64_t* A, int64_t* Rq) {         // getLackingResources: re-sum over ALL nodes
        std::fill(A, A + R, 0); std::fill(Rq, Rq + R, 0);
        for (const NodeP& n : cur()) for (int k = 0; k < R; ++k) { A[k] += n->alloc[k]; Rq[k] += n->req[k]; }
    };
    for (int i = 0; i < N; ++i) write_partitioning(*data[i], i, G, out);       // GetPartitioningState (planner.go:72)
    Tracker tr; std::memset(tr.lack, 0, sizeof tr.lack); tr.entry.assign((size_t)pb.n_keys * KP, 0);
    int64_t A[RMAX], Rq[RMAX];
    for (int i = 0; i < P; ++i) {                                                // NewSliceTracker
        totals(A, Rq);
        int32_t l[KP] = {0};
        pod_lacks(pb.pods[i], A, Rq, l);
        for (int p = 0; p < KP; ++p) { tr.lack[p] += l[p]; tr.entry[(size_t)pb.tkey[i] * KP + p] += l[p]; }
    }
    st->lacking_at_start = tr.count();
    if (tr.empty()) { st->lacking_at_end = 0; return NOSPLAN_OK; }
    std::vector<int> order = sort_pods(pb);
    std::vector<std::pair<int, NodeP>> cands;                                    // GetCandidateNodes (name order == index order)
    for (int i = 0; i < N; ++i) if (node_has_free_capacity(c, *data[i])) cands.emplace_back(i, data[i]);
    st->candidate_nodes = (int64_t)cands.size();
    for (auto& [i, n] : cands) {
        if (tr.empty()) break;
        forked.resize(N);                                                        // Fork: clone everything
        for (int k = 0; k < N; ++k) forked[k] = std::make_shared<Node>(*cur()[k]);
        has_fork = true;
        ++st->visited_nodes;
        bool upd = node_update_geometry(c, *n, tr.lack, st->geometry_evals);
        if (upd) forked[i] = n;                                                  // SetNode (planner.go:109-112)
        int added = 0;
        for (int oi : order) {
            ++st->candidates;
            const PodReq& pr = pb.pods[oi];
            totals(A, Rq);
            if (pod_lacks(pr, A, Rq, nullptr)) continue;                         // planner.go:159
            Node& f = *forked[i];
            if (!pb.pre_ok[oi]) continue;
            if (pb.filter_mode == NOSPLAN_FILTER_REJECT_ALL) continue;
            if (pb.filter_mode == NOSPLAN_FILTER_RESOURCES_FIT && !resources_fit(f, pr)) continue;
            if (!node_add_pod(c, f, pr)) continue;
            write_partitioning(*n, i, G, out);                                   // planner.go:138 (from n, not the fork)
            tr.remove(pb.tkey[oi]);
            ++added; ++st->placements;
        }
        if (added == 0) { forked.clear(); has_fork = false; }                    // Revert
        else { data.swap(forked); forked.clear(); has_fork = false; }            // Commit
    }
    st->lacking_at_end = tr.count();
    return NOSPLAN_OK;
}

// ---------------------------------------------------------------- incremental mode
// State rules derived from the object graph (each candidate node is visited once, so at the start of
// its iteration the original object O_i equals the snapshot's entry D_i in value):
//   W   = O_i after UpdateGeometryFor (GPUs possibly re-partitioned, Allocatable slice columns
//         overwritten from the geometry -- mig/node.go:172-174 does this even when nothing changed)
//   F   = the forked entry the pods are added to: W itself if updated (SetNode aliases it), otherwise a
//         clone of D_i, i.e. W's GPUs (unchanged) with the snapshot-time Allocatable
//   gate totals during the iteration = totals - contribution(D_i) + contribution(F)
//   on success (Commit)            : D_i := F, committed := true
//   on failure before 1st Commit   : D_i := W   (the original object is still what `data` holds)
//   on failure after a Commit      : D_i unchanged (the original object is detached from `data`)
//   the plan entry of node i is GetPartitioning(O_i) taken at each success: W's geometry if updated
//   (including the pods' moves), else the untouched initial geometry.
int plan_incremental(Problem& pb, nosplan_partitioning* out, nosplan_stats* st) {
    const Ctx& c = pb.c; const int N = pb.N, G = pb.G, P = pb.P, R = c.R;
    std::vector<Node>& D = pb.nodes;
    for (int i = 0; i < N; ++i) write_partitioning(D[i], i, G, out);
    int64_t A[RMAX] = {0}, Rq[RMAX] = {0};
    for (int i = 0; i < N; ++i) for (int k = 0; k < R; ++k) { A[k] += D[i].alloc[k]; Rq[k] += D[i].req[k]; }
    Tracker tr; std::memset(tr.lack, 0, sizeof tr.lack); tr.entry.assign((size_t)pb.n_keys * KP, 0);
    for (int i = 0; i < P; ++i) {
        int32_t l[KP] = {0};
        pod_lacks(pb.pods[i], A, Rq, l);
        for (int p = 0; p < KP; ++p) { tr.lack[p] += l[p]; tr.entry[(size_t)pb.tkey[i] * KP + p] += l[p]; }
    }
    st->lacking_at_start = tr.count();
    if (tr.empty()) { st->lacking_at_end = 0; return NOSPLAN_OK; }
    std::vector<int> order = sort_pods(pb);
    std::vector<int> cands;
    for (int i = 0; i < N; ++i) if (node_has_free_capacity(c, D[i])) cands.push_back(i);
    st->candidate_nodes = (int64_t)cands.size();
    bool committed = false;
    for (int i : cands) {
        if (tr.empty()) break;
        ++st->visited_nodes;
        Node W = D[i];
        bool upd = node_update_geometry(c, W, tr.lack, st->geometry_evals);
        Node F = W;
        if (!upd) std::memcpy(F.alloc, D[i].alloc, sizeof F.alloc);
        int64_t gA[RMAX], gR[RMAX];
        for (int k = 0; k < R; ++k) { gA[k] = A[k] - D[i].alloc[k] + F.alloc[k]; gR[k] = Rq[k] - D[i].req[k]; }
        int added = 0;
        for (int oi : order) {
            ++st->candidates;
            const PodReq& pr = pb.pods[oi];
            int64_t tR[RMAX];
            for (int k = 0; k < R; ++k) tR[k] = gR[k] + F.req[k];
            if (pod_lacks(pr, gA, tR, nullptr)) continue;
            if (!pb.pre_ok[oi]) continue;
            if (pb.filter_mode == NOSPLAN_FILTER_REJECT_ALL) continue;
            if (pb.filter_mode == NOSPLAN_FILTER_RESOURCES_FIT && !resources_fit(F, pr)) continue;
            if (!node_add_pod(c, F, pr)) continue;
            if (upd) write_partitioning(F, i, G, out);
            tr.remove(pb.tkey[oi]);
            ++added; ++st->placements;
        }
        if (added > 0) {
            for (int k = 0; k < R; ++k) { A[k] = gA[k]; Rq[k] = gR[k] + F.req[k]; }
            D[i] = F; committed = true;
        } else if (!committed) {
            for (int k = 0; k < R; ++k) A[k] += W.alloc[k] - D[i].alloc[k];
            D[i] = W;
        }
    }
    st->lacking_at_end = tr.count();
    return NOSPLAN_OK;
}

// ---------------------------------------------------------------- model of the DEVICE algorithm (CPU tests only)
// Not a restatement of the reference: a restatement, in the oracle's own terms, of what the CUDA path does when node accounts
// disagree with the GPUs or pods ask for several profiles (nos_b200/csrc/sweep.cu: k_plan_parallel without the cluster-wide gate,
// write_gate_record, gate_scan, re-planning with the gate) -- so that the exactness argument of DESIGN.md section 4 is checked
// against plan_incremental on the CPU as well.  Every candidate node is planned from its t0 state with a SATURATED lacking vector
// and no gate, leaving (dW, placed, need, added, updated); the gate is then replayed in name order over those records with the
// running totals under the Commit / Revert rules, and a node with an attempt the gate would have refused is planned again with
// the gate on the exact totals.  Returns 1 ("not applicable": the device would fall back to the exact sweep) when the lacking
// vector is not saturated at t0 or at the end.
struct DevRec {
    int64_t dW[KP], pl[KP], need[KP];
    bool upd = false, added = false;
    int placements = 0;
    int64_t evals = 0;
    Node F;
    std::vector<int> keys;
};

static DevRec device_node(const Problem& pb, const std::vector<int>& order, int i, const int32_t* sat_lack, bool gate, const int64_t* Dtot) {
    const Ctx& c = pb.c;
    const Node& D = pb.nodes[i];
    DevRec r;
    std::memset(r.dW, 0, sizeof r.dW); std::memset(r.pl, 0, sizeof r.pl); std::memset(r.need, 0, sizeof r.need);
    Node W = D;
    r.upd = node_update_geometry(c, W, sat_lack, r.evals);
    r.F = W;
    if (!r.upd) std::memcpy(r.F.alloc, D.alloc, sizeof r.F.alloc);
    for (int p = 0; p < c.K; ++p) r.dW[p] = W.alloc[3 + p] - D.alloc[3 + p];
    for (int oi : order) {
        const PodReq& pr = pb.pods[oi];
        if (gate) {                                                       // core/snapshot.go:132-165 on the running totals
            bool lacks = false;
            for (int e = 0; e < pr.n_slice; ++e) {
                const int p = pr.sp[e];
                const int64_t avail = std::max<int64_t>(0, Dtot[p] + (r.upd ? r.dW[p] : 0) - (r.F.req[3 + p] - D.req[3 + p]));
                if (avail < pr.sq[e]) lacks = true;
            }
            if (lacks) continue;
        }
        if (!pb.pre_ok[oi]) continue;
        if (pb.filter_mode == NOSPLAN_FILTER_REJECT_ALL) continue;
        if (pb.filter_mode == NOSPLAN_FILTER_RESOURCES_FIT && !resources_fit(r.F, pr)) continue;
        if (!gate && pr.n_slice > 1)                                      // a refused attempt may leave devices moved: note what the gate would have had to grant
            for (int e = 0; e < pr.n_slice; ++e) r.need[pr.sp[e]] = std::max(r.need[pr.sp[e]], r.F.req[3 + pr.sp[e]] - D.req[3 + pr.sp[e]] + pr.sq[e]);
        if (!node_add_pod(c, r.F, pr)) continue;
        r.keys.push_back(pb.tkey[oi]);
        ++r.placements; r.added = true;
    }
    for (int p = 0; p < c.K; ++p) { r.pl[p] = r.F.req[3 + p] - D.req[3 + p]; if (!gate) r.need[p] = std::max(r.need[p], r.pl[p]); }
    return r;
}

int plan_device_model(Problem& pb, nosplan_partitioning* out, nosplan_stats* st) {
    const Ctx& c = pb.c; const int N = pb.N, G = pb.G, P = pb.P, R = c.R;
    const std::vector<Node>& D = pb.nodes;
    for (int i = 0; i < N; ++i) write_partitioning(D[i], i, G, out);
    int64_t A[RMAX] = {0}, Rq[RMAX] = {0};
    for (int i = 0; i < N; ++i) for (int k = 0; k < R; ++k) { A[k] += D[i].alloc[k]; Rq[k] += D[i].req[k]; }
    Tracker tr; std::memset(tr.lack, 0, sizeof tr.lack); tr.entry.assign((size_t)pb.n_keys * KP, 0);
    for (int i = 0; i < P; ++i) {
        int32_t l[KP] = {0};
        pod_lacks(pb.pods[i], A, Rq, l);
        for (int p = 0; p < KP; ++p) { tr.lack[p] += l[p]; tr.entry[(size_t)pb.tkey[i] * KP + p] += l[p]; }
    }
    st->lacking_at_start = tr.count();
    st->gate_replayed = -1;
    if (tr.empty()) { st->lacking_at_end = 0; return NOSPLAN_OK; }
    // the saturation bound of fast_applicable (sweep.cu): lack[p] > (G+2) * max(devices of p one GPU can hold, largest free count)
    int64_t bound[KP] = {0};
    int maxmem = 0;
    for (int i = 0; i < N; ++i) for (int j = 0; j < D[i].ng; ++j) maxmem = std::max(maxmem, D[i].g[j].mem);
    for (int p = 0; p < c.K; ++p) {
        int m = 0;
        if (c.kind == NOSPLAN_KIND_MIG) { const int rows = c.geom_off ? c.geom_off[pb.n_models] : 0; for (int r = 0; r < rows; ++r) m = std::max(m, (int)c.geom[16 * (size_t)r + p]); }
        else m = c.pmem[p] > 0 ? maxmem / c.pmem[p] : (1 << 30);
        for (int i = 0; i < N; ++i) for (int j = 0; j < D[i].ng; ++j) m = std::max(m, (int)D[i].g[j].free_[p]);
        bound[p] = (int64_t)(G + 2) * m;
    }
    auto saturated = [&](const int32_t* lack) {
        for (int p = 0; p < c.K; ++p) if (tr.lack[p] > 0 && !(lack[p] > 0 && (int64_t)lack[p] > bound[p])) return false;
        return true;
    };
    if (!saturated(tr.lack)) return 1;
    int32_t sat_lack[KP];
    for (int p = 0; p < KP; ++p) sat_lack[p] = tr.lack[p] > 0 ? (1 << 29) : 0;
    std::vector<int> order = sort_pods(pb);
    std::vector<int> cands;
    for (int i = 0; i < N; ++i) if (node_has_free_capacity(c, D[i])) cands.push_back(i);
    st->candidate_nodes = (int64_t)cands.size();
    st->visited_nodes = (int64_t)cands.size();
    st->candidates = st->visited_nodes * P;
    // ---- every node on its own, without the gate
    std::vector<DevRec> rec(cands.size());
    for (size_t k = 0; k < cands.size(); ++k) rec[k] = device_node(pb, order, cands[k], sat_lack, false, nullptr);
    // ---- the gate, replayed in name order
    int64_t Dtot[KP];
    for (int p = 0; p < KP; ++p) Dtot[p] = p < c.K ? A[3 + p] - Rq[3 + p] : 0;
    bool committed = false;
    int again = 0;
    std::vector<uint8_t> placed((size_t)pb.n_keys, 0);
    for (size_t k = 0; k < cands.size(); ++k) {
        DevRec* r = &rec[k];
        st->geometry_evals += r->evals;
        for (int key : r->keys) placed[key] = 1;                                  // keys only the gate-less run placed stay set (as on the device)
        bool refused = false;
        for (int p = 0; p < c.K; ++p) if (r->need[p] > 0 && Dtot[p] + (r->upd ? r->dW[p] : 0) - r->need[p] < 0) refused = true;
        DevRec exact;
        if (refused) { exact = device_node(pb, order, cands[k], sat_lack, true, Dtot); r = &exact; ++again; for (int key : r->keys) placed[key] = 1; }
        if (r->added) {
            for (int p = 0; p < c.K; ++p) Dtot[p] += (r->upd ? r->dW[p] : 0) - r->pl[p];
            committed = true;
            if (r->upd) write_partitioning(r->F, cands[k], G, out);
        } else if (!committed) {
            for (int p = 0; p < c.K; ++p) Dtot[p] += r->dW[p];
        }
        st->placements += r->placements;
    }
    int32_t le[KP];
    for (int p = 0; p < KP; ++p) le[p] = tr.lack[p];
    for (int key = 0; key < pb.n_keys; ++key) if (placed[key]) for (int p = 0; p < KP; ++p) le[p] -= std::max(0, tr.entry[(size_t)key * KP + p]);
    if (!saturated(le)) return 1;
    int cnt = 0;
    for (int p = 0; p < KP; ++p) cnt += le[p] > 0;
    st->lacking_at_end = cnt;
    st->gate_replayed = again;
    st->path = 1;
    return NOSPLAN_OK;
}

// ---------------------------------------------------------------- model of the single-launch kernel (CPU tests only)
// What k_plan_fused (nos_b200/csrc/fused.cu) computes, in the oracle's terms: every candidate node on its own from the t0 state,
// geometry under a saturated lacking vector, and a pod loop that keeps per node only the remaining cpu, the remaining memory
// and the free devices PER PROFILE (not per GPU) -- no cluster-wide gate, no slice-column test in NodeResourcesFit, no first-fit
// GPU choice.  DESIGN.md section 3a argues that this is the reference's plan whenever every requested profile is lacking and
// stays saturated, every request is for ONE device of at most one profile, and every node's accounts agree with its GPUs;
// tests/test_device_model.py checks the argument against plan_incremental.  Returns 1 where the kernel would reject itself or
// the host would not have chosen it.
int plan_fused_model(Problem& pb, nosplan_partitioning* out, nosplan_stats* st) {
    const Ctx& c = pb.c; const int N = pb.N, G = pb.G, P = pb.P, R = c.R;
    const std::vector<Node>& D = pb.nodes;
    const bool fit = pb.filter_mode == NOSPLAN_FILTER_RESOURCES_FIT;
    for (int i = 0; i < N; ++i) write_partitioning(D[i], i, G, out);
    // host-side eligibility (api.cu): single-profile classes, quantity 1, unique tracker keys, cpu / memory / slice requests only
    std::vector<uint8_t> key_seen((size_t)pb.n_keys, 0);
    unsigned rset = 0;
    for (int i = 0; i < P; ++i) {
        const PodReq& pr = pb.pods[i];
        if (pr.n_slice > 1 || (pr.n_slice == 1 && pr.sq[0] != 1) || pr.base[2] != 0 || (fit && pr.n_other > 0)) return 1;
        if (pr.n_slice == 0 && pr.base[0] == 0 && pr.base[1] == 0) return 1;
        if (key_seen[pb.tkey[i]]) return 1;
        key_seen[pb.tkey[i]] = 1;
        if (pr.n_slice) rset |= 1u << pr.sp[0];
    }
    if (G > 8 || rset == 0) return 1;
    int64_t A[RMAX] = {0}, Rq[RMAX] = {0};
    for (int i = 0; i < N; ++i) for (int k = 0; k < R; ++k) { A[k] += D[i].alloc[k]; Rq[k] += D[i].req[k]; }
    Tracker tr; std::memset(tr.lack, 0, sizeof tr.lack); tr.entry.assign((size_t)pb.n_keys * KP, 0);
    for (int i = 0; i < P; ++i) {
        int32_t l[KP] = {0};
        pod_lacks(pb.pods[i], A, Rq, l);
        for (int p = 0; p < KP; ++p) { tr.lack[p] += l[p]; tr.entry[(size_t)pb.tkey[i] * KP + p] += l[p]; }
    }
    st->lacking_at_start = tr.count();
    // node-side conditions (phase 1): counts that fit five bits with their keys present, accounts that agree with the GPUs
    int maxmem = 0;
    for (int i = 0; i < N; ++i) {
        const Node& n = D[i];
        for (int p = 0; p < c.K; ++p) {
            int64_t sf = 0, su = 0;
            for (int j = 0; j < n.ng; ++j) {
                const Gpu& g = n.g[j];
                if (g.used[p] > 31 || g.free_[p] > 31) return 1;
                if ((g.used[p] && !(g.umask >> p & 1)) || (g.free_[p] && !(g.fmask >> p & 1))) return 1;
                sf += g.free_[p]; su += g.used[p];
                maxmem = std::max(maxmem, g.mem);
            }
            if (n.alloc[3 + p] - n.req[3 + p] < sf || n.req[3 + p] > su) return 1;
        }
    }
    // the hypothesis at t0: exactly the requested profiles are lacking, each beyond the saturation bound
    int64_t bound[KP] = {0};
    for (int p = 0; p < c.K; ++p) {
        int m = 0;
        if (c.kind == NOSPLAN_KIND_MIG) { const int rows = c.geom_off ? c.geom_off[pb.n_models] : 0; for (int r = 0; r < rows; ++r) m = std::max(m, (int)c.geom[16 * (size_t)r + p]); }
        else m = c.pmem[p] > 0 ? maxmem / c.pmem[p] : (1 << 30);
        for (int i = 0; i < N; ++i) for (int j = 0; j < D[i].ng; ++j) m = std::max(m, (int)D[i].g[j].free_[p]);
        bound[p] = (int64_t)(G + 2) * m;
    }
    auto holds = [&](const int32_t* lack) {
        for (int p = 0; p < KP; ++p) {
            const bool requested = (rset >> p & 1u) != 0;
            if (requested != (tr.lack[p] > 0)) return false;
            if (requested && !(lack[p] > 0 && (int64_t)lack[p] > bound[p])) return false;
        }
        return true;
    };
    if (!holds(tr.lack)) return 1;
    int32_t sat_lack[KP];
    for (int p = 0; p < KP; ++p) sat_lack[p] = (rset >> p & 1u) ? (1 << 29) : 0;
    std::vector<int> order = sort_pods(pb);
    std::vector<uint8_t> placed((size_t)pb.n_keys, 0);
    for (int i = 0; i < N; ++i) {
        if (!node_has_free_capacity(c, D[i])) continue;
        ++st->visited_nodes;
        Node W = D[i];
        const bool upd = node_update_geometry(c, W, sat_lack, st->geometry_evals);
        int cap[KP] = {0};
        for (int p = 0; p < c.K; ++p) for (int j = 0; j < W.ng; ++j) { if (W.g[j].free_[p] > 31) return 1; cap[p] += W.g[j].free_[p]; }
        int64_t remc = D[i].alloc[0] - D[i].req[0], remm = D[i].alloc[1] - D[i].req[1];
        const bool all_fail = pb.filter_mode == NOSPLAN_FILTER_REJECT_ALL ||
                              (fit && (D[i].num_pods + 1 > D[i].allowed_pods || D[i].alloc[2] - D[i].req[2] < 0));
        int added = 0;
        for (int oi : order) {
            if (all_fail) break;
            const PodReq& pr = pb.pods[oi];
            if (!pb.pre_ok[oi]) continue;
            if (fit && (pr.base[0] > remc || pr.base[1] > remm)) continue;
            if (pr.n_slice) { if (cap[pr.sp[0]] == 0) continue; --cap[pr.sp[0]]; }
            remc -= pr.base[0]; remm -= pr.base[1];
            placed[pb.tkey[oi]] = 1;
            ++added; ++st->placements;
        }
        if (added && upd) write_partitioning(W, i, G, out);     // used + free per GPU and the key masks do not depend on which GPU took a pod
    }
    st->candidate_nodes = st->visited_nodes;
    st->candidates = st->visited_nodes * P;
    int32_t le[KP];
    for (int p = 0; p < KP; ++p) le[p] = tr.lack[p];
    for (int key = 0; key < pb.n_keys; ++key) if (placed[key]) for (int p = 0; p < KP; ++p) le[p] -= std::max(0, tr.entry[(size_t)key * KP + p]);
    if (!holds(le)) return 1;
    int cnt = 0;
    for (int p = 0; p < KP; ++p) cnt += le[p] > 0;
    st->lacking_at_end = cnt;
    st->path = 1; st->fused = 1;
    return NOSPLAN_OK;
}

}  // namespace

extern "C" {

int nos_oracle_plan(const nosplan_snapshot* s, const nosplan_pods* p, const uint8_t* prefilter_ok, int32_t filter_mode,
                    int32_t mode, nosplan_partitioning* out, nosplan_stats* stats, double* seconds) {
    Problem pb;
    int rc = build(s, p, prefilter_ok, filter_mode, pb);
    if (rc != NOSPLAN_OK) return rc;
    nosplan_stats st; std::memset(&st, 0, sizeof st);
    auto t0 = std::chrono::steady_clock::now();
    rc = mode == NOS_ORACLE_FAITHFUL ? plan_faithful(pb, out, &st) : mode == NOS_ORACLE_DEVICE_MODEL ? plan_device_model(pb, out, &st)
       : mode == NOS_ORACLE_FUSED_MODEL ? plan_fused_model(pb, out, &st) : plan_incremental(pb, out, &st);
    auto t1 = std::chrono::steady_clock::now();
    if (seconds) *seconds = std::chrono::duration<double>(t1 - t0).count();
    if (stats) *stats = st;
    return rc;
}

int nos_oracle_sort_pods(const nosplan_snapshot* s, const nosplan_pods* p, int32_t* order_out) {
    Problem pb;
    int rc = build(s, p, nullptr, 0, pb);
    if (rc != NOSPLAN_OK) return rc;
    std::vector<int> o = sort_pods(pb);
    for (int i = 0; i < pb.P; ++i) order_out[i] = o[i];
    return NOSPLAN_OK;
}

}  // extern "C"
