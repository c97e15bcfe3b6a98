// C-ABI of libnosplan.so (include/nosplan.h): handle, host-side preparation (NewPodSorter order, runs of
// identical slice requests, one staged H2D copy), kernel launches, D2H of the plan.
// There is no CPU planning path in this file: everything the reference's Plan() computes per node and per
// pod happens in the kernels of prepass.cu / sweep.cu.
#include <algorithm>
#include <chrono>
#include <climits>
#include <cstdlib>
#include <cstdarg>
#include <cstdio>
#include <cstring>
#include <new>
#include <string>
#include <type_traits>
#include <vector>

#include "nosplan_dev.cuh"

namespace nsp {
void launch_prepass(const DevProblem& pb, cudaStream_t st);
void launch_expand_nodes(const DevProblem& pb, const void* records, int record_bytes, int n_models, int32_t* d_flag, cudaStream_t st);
void launch_tracker_init(const DevProblem& pb, cudaStream_t st);
cudaError_t launch_sweep(const DevProblem& pb, int32_t* cand_idx, cudaStream_t st);
cudaError_t launch_parallel(const DevProblem& pb, int max_blocks, cudaStream_t st);
cudaError_t launch_waves(const DevProblem& pb, int max_blocks, cudaStream_t st);
void launch_shard_combine(const DevProblem& pb, const void* gathered, int world, size_t stride, const PeerWait& wait, cudaStream_t st);
void launch_slice_pack_totals(const DevProblem& pb, void* block, cudaStream_t st);
void launch_slice_combine_totals(const DevProblem& pb, const void* gathered, int world, const PeerWait& wait, cudaStream_t st);

cudaError_t sweep_prepare_device(int* par_blocks, int* wave_blocks);
int sweep_max_keys();
cudaError_t fused_prepare_device();
cudaError_t launch_fused(const DevProblem& pb, int sms, cudaStream_t st);
cudaError_t launch_fused_verify(const DevProblem& pb, const void* blocks, size_t stride, int world, const PeerWait& wait, cudaStream_t st);
size_t fused_exchange_bytes(const DevProblem& pb);
int fused_max_pad_pods();
}  // namespace nsp

using namespace nsp;

struct nosplan_quota_ctx;
extern "C" int nosplan_quota_recompute_impl(int device, cudaStream_t st, nosplan_quota_ctx* ctx, std::string* err,
                                            const nosplan_quotas* qs, const nosplan_quota_pods* running,
                                            const nosplan_quota_pods* pending, nosplan_quota_out* out, float* ms_h2d, float* ms_k,
                                            float* ms_d2h, int* launches, long long* h2d_bytes, long long* d2h_bytes,
                                            const nosplan_label_changes* lc, const struct nosplan_quota_reduce_hook* hook);
extern "C" void nosplan_quota_ctx_free(nosplan_quota_ctx* ctx);
extern "C" nosplan_quota_ctx* nosplan_quota_ctx_new();
extern "C" int nosplan_quota_victims_impl(int device, cudaStream_t st, nosplan_quota_ctx* ctx, std::string* err,
                                          const nosplan_quotas* qs, const nosplan_victim_query* vq, uint8_t* victim,
                                          float* ms_k, int* launches);

extern "C" int nosplan_preempt_dryrun_impl(int device, cudaStream_t st, nosplan_quota_ctx* ctx, std::string* err,
                                           const nosplan_quotas* qs, const nosplan_preempt_query* pq, uint8_t* victim,
                                           int32_t* node_status, int32_t* num_violating, float* ms_h2d, float* ms_k, float* ms_d2h,
                                           int* launches, long long* h2d_bytes, long long* d2h_bytes);

extern "C" int nosplan_quota_state_begin_impl(int device, cudaStream_t st, nosplan_quota_ctx* ctx, std::string* err,
                                              const nosplan_quotas* qs, int32_t pod_capacity);
extern "C" int nosplan_quota_state_apply_impl(int device, cudaStream_t st, nosplan_quota_ctx* ctx, nosplan_quota_ctx* scratch, std::string* err,
                                              const nosplan_quota_events* evs, nosplan_quota_out* out, int32_t* n_applied,
                                              float* ms_k, long long* h2d_bytes, long long* d2h_bytes);
extern "C" int nosplan_quota_state_prefilter_impl(int device, cudaStream_t st, nosplan_quota_ctx* ctx, nosplan_quota_ctx* scratch, std::string* err,
                                                  const nosplan_quota_pods* pending, uint8_t* verdict, float* ms_k);
extern "C" int nosplan_preempt_dryrun_again_impl(int device, cudaStream_t st, nosplan_quota_ctx* ctx, std::string* err,
                                                 const nosplan_preempt_query* pq, uint8_t* victim, int32_t* node_status,
                                                 int32_t* num_violating, float* ms_k, float* ms_d2h, int* launches, long long* d2h_bytes);

extern "C" int nosplan_parse_annotations_impl(int device, cudaStream_t st, nosplan_quota_ctx* ctx, std::string* err,
                                              const nosplan_annotations* in, nosplan_parsed_gpus* out, float* ms_h2d, float* ms_k,
                                              float* ms_d2h, long long* h2d_bytes, long long* d2h_bytes);

#include "handle.cuh"

extern "C" void nosplan_shard_ctx_free(nosplan_shard_ctx* c);

namespace {

#define fail nosplan_fail

#define CU(h, call) do { cudaError_t e_ = (call); if (e_ != cudaSuccess) \
    return fail(h, NOSPLAN_E_CUDA, "%s: %s", #call, cudaGetErrorString(e_)); } while (0)

inline size_t al(size_t x) { return (x + 255) & ~(size_t)255; }

// ---- Go 1.18 sort.SliceStable (sort/zsortfunc.go stable_func / insertionSort_func / symMerge_func / rotate_func),
// restated so that NewPodSorter (core/util.go:34-73) gives the reference's order even where its comparator is
// not a strict weak order (mig/profile.go:84-96).
template <class Less>
struct GoStable {
    int* d; int n; Less less;
    bool lt(int i, int j) { return less(d[i], d[j]); }
    void sw(int i, int j) { int t = d[i]; d[i] = d[j]; d[j] = t; }
    void insertion(int a, int b) { for (int i = a + 1; i < b; ++i) for (int j = i; j > a && lt(j, j - 1); --j) sw(j, j - 1); }
    void swap_range(int a, int b, int c) { for (int i = 0; i < c; ++i) sw(a + i, b + i); }
    void rotate(int a, int m, int b) {
        int i = m - a, j = b - m;
        while (i != j) { if (i > j) { swap_range(m - i, m, j); i -= j; } else { swap_range(m - i, m + j - i, i); j -= i; } }
        swap_range(m - i, m, i);
    }
    void sym_merge(int a, int m, int b) {
        if (m - a == 1) {
            int i = m, j = b;
            while (i < j) { int h = (int)((unsigned)(i + j) >> 1); if (lt(h, a)) i = h + 1; else j = h; }
            for (int k = a; k < i - 1; ++k) sw(k, k + 1);
            return;
        }
        if (b - m == 1) {
            int i = a, j = m;
            while (i < j) { int h = (int)((unsigned)(i + j) >> 1); if (!lt(m, h)) i = h + 1; else j = h; }
            for (int k = m; k > i; --k) sw(k, k - 1);
            return;
        }
        int mid = (int)((unsigned)(a + b) >> 1), nn = mid + m, start, r;
        if (m > mid) { start = nn - b; r = mid; } else { start = a; r = m; }
        int p = nn - 1;
        while (start < r) { int c = (int)((unsigned)(start + r) >> 1); if (!lt(p - c, c)) start = c + 1; else r = c; }
        int end = nn - start;
        if (start < m && m < end) rotate(start, m, end);
        if (a < start && start < mid) sym_merge(a, start, mid);
        if (mid < end && end < b) sym_merge(mid, end, b);
    }
    void run() {
        int block = 20, a = 0, b = block;
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


int ensure(nosplan* h, size_t dev_bytes, size_t stage_bytes, size_t out_bytes, size_t keep_bytes = 0) {
    if (dev_bytes > h->d_cap) {
        unsigned char* old = h->d_arena;
        size_t cap = dev_bytes + dev_bytes / 4 + (1 << 20);
        unsigned char* fresh = nullptr;
        if (old && !keep_bytes) { CU(h, cudaFree(old)); old = nullptr; h->d_arena = nullptr; h->d_cap = 0; }
        CU(h, cudaMalloc(&fresh, cap));
        if (old) {                                           // a resident snapshot moves with the arena
            cudaError_t e = cudaMemcpyAsync(fresh, old, keep_bytes, cudaMemcpyDeviceToDevice, h->stream);
            if (e == cudaSuccess) e = cudaStreamSynchronize(h->stream);
            if (e != cudaSuccess) { cudaFree(fresh); return fail(h, NOSPLAN_E_CUDA, "moving the resident snapshot: %s", cudaGetErrorString(e)); }
            CU(h, cudaFree(old));
        }
        h->d_arena = fresh;
        h->d_cap = cap;
    }
    if (stage_bytes > h->h_cap) {
        if (h->h_stage) { CU(h, cudaStreamSynchronize(h->stream)); CU(h, cudaFreeHost(h->h_stage)); }   // a queued copy may still read it
        h->h_stage = nullptr; h->h_cap = 0;
        size_t cap = stage_bytes + stage_bytes / 4 + (1 << 20);
        CU(h, cudaMallocHost(&h->h_stage, cap));
        h->h_cap = cap;
    }
    if (out_bytes > h->h_out_cap) {
        if (h->h_out) CU(h, cudaFreeHost(h->h_out));
        h->h_out = nullptr; h->h_out_cap = 0;
        size_t cap = out_bytes + out_bytes / 4 + 4096;
        CU(h, cudaMallocHost(&h->h_out, cap));
        h->h_out_cap = cap;
    }
    return NOSPLAN_OK;
}

}  // namespace

// NewPodSorter (core/util.go:34-73): order[k] = original index of the k-th pod.  first_prof[i] = lowest requested profile id
// of pod i (the canonical "first key" of the reference's map iteration) or -1.
static void pod_order(const nosplan_snapshot* s, const nosplan_pods* pods, const int* first_prof_ptr, std::vector<int>& order) {
    const int P = pods->n_pods, K = s->n_profiles;
    const bool mig = s->kind == NOSPLAN_KIND_MIG;
    order.resize(P);
    for (int i = 0; i < P; ++i) order[i] = i;
    const int* fp = first_prof_ptr;
        const int32_t* prio = pods->priority; const int32_t* pmem = s->profile_mem_gb; const int32_t* pgi = s->profile_gi;
        auto smaller = [=](int x, int y) {                                    // SmallerThan: mig/profile.go:84-96, slicing/profile.go:36-42
            if (pmem[x] < pmem[y]) return true;
            return mig && pgi[x] < pgi[y];
        };
        // Fast path.  Where the comparator is a strict weak order on the pods present, EVERY stable sort returns what Go's
        // sort.SliceStable returns, so a counting sort by (priority desc, profile rank) is exact.  That needs (a) SmallerThan
        // restricted to the first-requested profiles present to be asymmetric with transitive incomparability, and (b) no
        // priority level mixing slice-less pods (incomparable to everything) with pods of two different profile ranks.
        bool swo = true;
        int rank_of[KP];
        {
            bool present[KP] = {false};
            for (int i = 0; i < P; ++i) if (fp[i] >= 0) present[fp[i]] = true;
            for (int x = 0; x < K && swo; ++x) for (int y = 0; y < K && swo; ++y) {
                if (!present[x] || !present[y]) continue;
                if (smaller(x, y) && smaller(y, x)) swo = false;
                for (int z = 0; z < K && swo; ++z) {
                    if (!present[z]) continue;
                    if (smaller(x, y) && smaller(y, z) && !smaller(x, z)) swo = false;
                    const bool ixy = !smaller(x, y) && !smaller(y, x), iyz = !smaller(y, z) && !smaller(z, y), ixz = !smaller(x, z) && !smaller(z, x);
                    if (ixy && iyz && !ixz) swo = false;
                }
            }
            for (int x = 0; x < K; ++x) {
                rank_of[x] = 0;
                if (present[x]) for (int y = 0; y < K; ++y) if (present[y] && smaller(y, x)) ++rank_of[x];
            }
        }
        if (swo) {
            // one pass: priority level of every pod (distinct priorities are few: linear probe with a last-hit cache) and, per
            // level, which profile ranks occur and whether slice-less pods do
            struct Level { int32_t prio; unsigned ranks; bool slice_less; };
            std::vector<Level> levels;
            std::vector<uint16_t> lvl(P);
            size_t last = 0;
            for (int i = 0; i < P; ++i) {
                const int32_t p = prio[i];
                if (last >= levels.size() || levels[last].prio != p) {
                    size_t k = 0;
                    while (k < levels.size() && levels[k].prio != p) ++k;
                    if (k == levels.size()) {
                        if (levels.size() >= 4096) { swo = false; break; }   // pathological: leave it to the general sort
                        levels.push_back(Level{p, 0u, false});
                    }
                    last = k;
                }
                lvl[i] = (uint16_t)last;
                if (fp[i] < 0) levels[last].slice_less = true; else levels[last].ranks |= 1u << rank_of[fp[i]];
            }
            // a level mixing slice-less pods (incomparable to everything) with two different ranks breaks transitivity
            for (auto& l : levels) if (l.slice_less && (l.ranks & (l.ranks - 1))) swo = false;
            if (swo) {
                // stable counting sort by (priority desc, rank asc); levels with slice-less members have a single bucket
                std::vector<int> by_prio(levels.size());
                for (size_t k = 0; k < levels.size(); ++k) by_prio[k] = (int)k;
                std::sort(by_prio.begin(), by_prio.end(), [&](int a, int b) { return levels[a].prio > levels[b].prio; });
                const int stride = KP + 1;
                std::vector<int> base_of_level(levels.size());
                for (size_t k = 0; k < levels.size(); ++k) base_of_level[by_prio[k]] = (int)k * stride;
                std::vector<int> count(levels.size() * stride + 1, 0), bucket(P);
                for (int i = 0; i < P; ++i) {
                    const Level& l = levels[lvl[i]];
                    const int r = (l.slice_less || fp[i] < 0) ? 0 : rank_of[fp[i]];
                    bucket[i] = base_of_level[lvl[i]] + r;
                    ++count[bucket[i] + 1];
                }
                for (size_t k = 1; k < count.size(); ++k) count[k] += count[k - 1];
                for (int i = 0; i < P; ++i) order[count[bucket[i]]++] = i;
            }
        }
        if (!swo) {
            auto less = [=](int a, int b) {
                if (prio[a] != prio[b]) return prio[a] > prio[b];
                if (fp[a] < 0 || fp[b] < 0) return false;
                return smaller(fp[a], fp[b]);
            };
            GoStable<decltype(less)> gs{order.data(), P, less};
            gs.run();
        }
}

int nosplan_fail(nosplan* h, int code, const char* fmt, ...) {
    char buf[512];
    va_list ap; va_start(ap, fmt); vsnprintf(buf, sizeof buf, fmt, ap); va_end(ap);
    if (h) h->err = buf;
    return code;
}

// event-timed phases of the last call: the H2D copies (nodes, then pods), the kernels (node-only part, then the rest)
void nosplan_phase_times(nosplan* h, nosplan_stats* stats) {
    float a = 0, b = 0;
    cudaEventElapsedTime(&a, h->ev[0], h->ev[1]); cudaEventElapsedTime(&b, h->ev[9], h->ev[10]);
    stats->ms_h2d = a + b;
    if (getenv("NOSPLAN_TRACE")) {
        float c = 0, d = 0, e = 0;
        cudaEventElapsedTime(&c, h->ev[0], h->ev[3]);
        if (h->split_run) { cudaEventElapsedTime(&d, h->ev[1], h->ev[9]); cudaEventElapsedTime(&e, h->ev[10], h->ev[3]); }
        for (int k = 0; k < h->dbg_n; ++k) {
            float t = 0;
            cudaEventElapsedTime(&t, k ? h->dbg_ev[k - 1] : h->ev[0], h->dbg_ev[k]);
            fprintf(stderr, "[nosplan gpu]   copy %2d %8zu B %s %7.1f us\n", k, h->dbg_bytes[k], h->dbg_direct[k] ? "direct" : "staged", 1e3 * t);
        }
        fprintf(stderr, "[nosplan gpu] h2d nodes %.1f us, gap to pod copy %.1f us, h2d pods %.1f us, pod copy -> last kernel %.1f us, first copy -> last kernel %.1f us\n",
                1e3 * a, 1e3 * d, 1e3 * b, 1e3 * e, 1e3 * c);
    }
    if (h->split_run) {
        cudaEventElapsedTime(&a, h->ev[2], h->ev[6]); cudaEventElapsedTime(&b, h->ev[11], h->ev[3]);
        stats->ms_kernels = a + b;
    } else cudaEventElapsedTime(&stats->ms_kernels, h->ev[2], h->ev[3]);
    stats->ms_kernels += h->ms_rejected;              // a rejected single-launch attempt in front of the classic sequence
    if (h->last_fused) stats->ms_parallel = stats->ms_kernels;
    else if (h->prepass_only) stats->ms_prepass = stats->ms_kernels;
    else if (h->kernel_events) {
        cudaEventElapsedTime(&stats->ms_prepass, h->ev[2], h->ev[6]);
        cudaEventElapsedTime(&stats->ms_tracker, h->ev[h->split_run ? 11 : 6], h->ev[7]);
        cudaEventElapsedTime(&stats->ms_parallel, h->ev[7], h->ev[8]);
        cudaEventElapsedTime(&stats->ms_sweep, h->ev[8], h->ev[3]);
    }
}

extern "C" {

int nosplan_create(int cuda_device, nosplan_t** out) {
    if (!out) return NOSPLAN_E_INVALID;
    *out = nullptr;
    int n = 0;
    if (cudaGetDeviceCount(&n) != cudaSuccess || cuda_device < 0 || cuda_device >= n) return NOSPLAN_E_CUDA;
    nosplan* h = new (std::nothrow) nosplan();
    if (!h) return NOSPLAN_E_NOMEM;
    h->device = cuda_device;
    int sms = 148;
    if (cudaSetDevice(cuda_device) != cudaSuccess || cudaStreamCreateWithFlags(&h->stream, cudaStreamNonBlocking) != cudaSuccess ||
        sweep_prepare_device(&h->par_blocks, &h->wave_blocks) != cudaSuccess || fused_prepare_device() != cudaSuccess ||
        cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, cuda_device) != cudaSuccess) {
        delete h;
        return NOSPLAN_E_CUDA;
    }
    h->sms = sms;
    if (cudaMalloc(&h->d_flag, 256) != cudaSuccess || cudaMallocHost(&h->h_flag, 256) != cudaSuccess) { delete h; return NOSPLAN_E_CUDA; }
    h->h_flag[0] = -1;
    for (auto& e : h->ev)
        if (cudaEventCreate(&e) != cudaSuccess) { delete h; return NOSPLAN_E_CUDA; }
    *out = h;
    return NOSPLAN_OK;
}

void nosplan_destroy(nosplan_t* h) {
    if (!h) return;
    cudaSetDevice(h->device);
    if (h->stream) cudaStreamSynchronize(h->stream);
    for (auto& e : h->ev) if (e) cudaEventDestroy(e);
    for (auto& e : h->dbg_ev) if (e) cudaEventDestroy(e);
    if (h->d_arena) cudaFree(h->d_arena);
    if (h->d_flag) cudaFree(h->d_flag);
    if (h->h_flag) cudaFreeHost(h->h_flag);
    if (h->h_stage) cudaFreeHost(h->h_stage);
    if (h->h_out) cudaFreeHost(h->h_out);
    if (h->qctx) nosplan_quota_ctx_free(h->qctx);
    if (h->pctx) nosplan_quota_ctx_free(h->pctx);
    if (h->ectx) nosplan_quota_ctx_free(h->ectx);
    if (h->sctx) nosplan_shard_ctx_free(h->sctx);
    if (h->stream) cudaStreamDestroy(h->stream);
    delete h;
}

int nosplan_set_kernel_events(nosplan_t* h, int32_t on) {
    if (!h) return NOSPLAN_E_INVALID;
    h->kernel_events = on != 0;
    return NOSPLAN_OK;
}

const char* nosplan_last_error(const nosplan_t* h) { return h ? h->err.c_str() : "null handle"; }

void* nosplan_stream(nosplan_t* h) { return h ? (void*)h->stream : nullptr; }

int nosplan_host_pod_order(const nosplan_snapshot* s, const nosplan_pods* pods, int32_t* order_out) {
    if (!s || !pods || (!order_out && pods->n_pods > 0)) return NOSPLAN_E_INVALID;
    const int P = pods->n_pods, K = s->n_profiles;
    if (K < 0 || K > NOSPLAN_MAX_PROFILES || P < 0) return NOSPLAN_E_INVALID;
    std::vector<int> first_prof(P, -1), order;
    for (int i = 0; i < P; ++i)
        for (int e = pods->scalar_ptr[i]; e < pods->scalar_ptr[i + 1]; ++e)
            if (pods->scalar_col[e] >= 3 && pods->scalar_col[e] < 3 + K) { first_prof[i] = pods->scalar_col[e] - 3; break; }
    pod_order(s, pods, first_prof.data(), order);
    for (int i = 0; i < P; ++i) order_out[i] = order[i];
    return NOSPLAN_OK;
}

// Kernels that need only the node arrays: work-area reset + k_prepass.  nosplan_plan issues them as soon as the node DMA
// is queued, so they run while the host is still ordering the pods.
static void flip_small_state(nosplan* h) {
    {   // this run's copy of the small state; the run's first kernel clears the other copy (and k_prepass the tracker tables)
        DevProblem& pb = h->pb;
        const unsigned use = h->run_parity ^= 1u;
        const ptrdiff_t to_use = (ptrdiff_t)(use ? h->small_bytes : 0) - (ptrdiff_t)(h->bound_copy ? h->small_bytes : 0);
        auto move = [&](auto*& p) { p = reinterpret_cast<std::remove_reference_t<decltype(p)>>(reinterpret_cast<unsigned char*>(p) + to_use); };
        move(pb.tot); move(pb.lack); move(pb.stats); move(pb.maxfree); move(pb.gflags); move(pb.acc); move(pb.lack_dec);
        move(pb.next_node); move(pb.wave);
        h->bound_copy = use;
        pb.zero_next = reinterpret_cast<unsigned long long*>(h->d_work + (use ? 0 : h->small_bytes));
        pb.zero_next_n8 = (int)(h->small_bytes >> 3);
    }
}

static int run_pre(nosplan* h, bool split) {
    CU(h, cudaEventRecord(h->ev[2], h->stream));
    flip_small_state(h);
    h->fused_pending = false; h->last_fused = false; h->fused_reject = 0; h->ms_rejected = 0.f; h->prepass_only = false;
    h->placed_dirty = true;                  // the classic kernels index `placed` by tracker key
    launch_prepass(h->pb, h->stream);
    if (split || h->kernel_events) CU(h, cudaEventRecord(h->ev[6], h->stream));
    CU(h, cudaGetLastError());
    h->pre_done = split;                     // split: the pod-dependent kernels follow later, after the pod DMA
    h->split_run = split;
    return NOSPLAN_OK;
}

}  // extern "C"

// ---- the blob of the single-launch kernel (nosplan_dev.cuh): layout, header, and everything derived from the pod pairs
static void blob_layout(int n_rows, int n_models, int n_runs, int pad_pods, int* fp_off) {
    auto a16 = [](size_t x) { return (int)((x + 15) & ~(size_t)15); };
    const int n3 = (pad_pods + 511) / 512, n2 = pad_pods / 64, n1 = pad_pods / 8;
    fp_off[0] = a16(16 * (size_t)n_rows);
    fp_off[1] = fp_off[0] + a16(2 * (size_t)n_rows);
    fp_off[2] = fp_off[1] + a16(4 * ((size_t)n_models + 1));
    fp_off[3] = fp_off[2] + a16(sizeof(FRun) * (size_t)n_runs);
    fp_off[FP_L3] = fp_off[3] + a16(8 * (KP + 1));
    fp_off[FP_L2] = fp_off[FP_L3] + a16(8 * (size_t)n3);
    fp_off[FP_L1] = fp_off[FP_L2] + a16(8 * (size_t)n2);
    fp_off[FP_SFX] = fp_off[FP_L1] + a16(8 * (size_t)n1);
    fp_off[FP_CUM] = fp_off[FP_SFX] + a16(8 * (size_t)n1);
    fp_off[FP_PODS] = fp_off[FP_CUM] + a16(16 * (size_t)n1);
    fp_off[FP_END] = fp_off[FP_PODS] + a16(8 * (size_t)pad_pods);
}
// header tables; every pair of the minima and of the pods starts as F_NONE ("no eligible pod").  minima_only: the caller writes
// every pod position itself (fast_pods_fill), and blob_finish writes every per-8 section, so only the two upper levels need it
static void blob_begin(unsigned char* B, const int* fp_off, const nosplan_snapshot* s, int n_rows, int n_models, bool minima_only = false) {
    std::memset(B, 0, (size_t)fp_off[FP_L3]);
    if (n_rows) std::memcpy(B, s->geom_counts, 16 * (size_t)n_rows);
    for (int r = 0; r < n_rows; ++r) {
        uint16_t m = 0;
        for (int p = 0; p < 16; ++p) if (s->geom_counts[16 * (size_t)r + p]) m |= (uint16_t)(1u << p);
        ((uint16_t*)(B + fp_off[0]))[r] = m;
    }
    if (n_models > 0) std::memcpy(B + fp_off[1], s->model_geom_offset, 4 * ((size_t)n_models + 1));
    std::fill((int32_t*)(B + fp_off[FP_L3]), (int32_t*)(B + fp_off[minima_only ? FP_L1 : FP_END]), (int32_t)F_NONE);
}
// the pod pairs are in place: minima per 8 / 64 / 512 positions, suffix minima and cumulative sums per run, run records
static void blob_finish(unsigned char* B, const int* fp_off, FRun* fr, int nr) {
    int32_t* l3 = (int32_t*)(B + fp_off[FP_L3]); int32_t* l2 = (int32_t*)(B + fp_off[FP_L2]); int32_t* l1 = (int32_t*)(B + fp_off[FP_L1]);
    int32_t* sfx = (int32_t*)(B + fp_off[FP_SFX]); int32_t* cum = (int32_t*)(B + fp_off[FP_CUM]);
    const int32_t* fpod = (const int32_t*)(B + fp_off[FP_PODS]);
    unsigned long long* prof_runs = (unsigned long long*)(B + fp_off[3]);
    for (int r = 0; r < nr; ++r) {
        FRun& f = fr[r];
        const int x0 = f.pos0 >> 3, x1 = x0 + f.nblk * 8;
        long long sc = 0, sm = 0; int cnt = 0;
        for (int x8 = x0; x8 < x1; ++x8) {
            int32_t mc = F_NONE, mm = F_NONE; unsigned elig = 0;
            const int32_t* g = fpod + 16 * (size_t)x8;
            for (int k = 0; k < 8; ++k) {                      // no branch: an ineligible position is (F_NONE, F_NONE), neutral for the minima
                const int32_t c = g[2 * k], m = g[2 * k + 1];
                const int v = c != F_NONE;
                mc = std::min(mc, c); mm = std::min(mm, m); elig |= (unsigned)v << k; sc += v ? c : 0; sm += v ? m : 0; cnt += v;
            }
            l1[2 * (size_t)x8] = mc; l1[2 * (size_t)x8 + 1] = mm;
            int32_t* cu = cum + 4 * (size_t)x8;
            cu[0] = (int32_t)std::min<long long>(sc, F_NONE); cu[1] = (int32_t)std::min<long long>(sm, F_NONE); cu[2] = cnt; cu[3] = (int32_t)elig;
            int32_t* b2 = l2 + 2 * (size_t)(x8 >> 3); int32_t* d3 = l3 + 2 * (size_t)(x8 >> 6);
            b2[0] = std::min(b2[0], mc); b2[1] = std::min(b2[1], mm); d3[0] = std::min(d3[0], mc); d3[1] = std::min(d3[1], mm);
            f.min_c = std::min(f.min_c, mc); f.min_m = std::min(f.min_m, mm);
        }
        int32_t tc = F_NONE, tm = F_NONE;
        for (int x8 = x1 - 1; x8 >= x0; --x8) {
            tc = std::min(tc, l1[2 * (size_t)x8]); tm = std::min(tm, l1[2 * (size_t)x8 + 1]);
            sfx[2 * (size_t)x8] = tc; sfx[2 * (size_t)x8 + 1] = tm;
        }
        prof_runs[f.p >= 0 ? f.p : KP] |= 1ull << r;
    }
    std::memcpy(B + fp_off[2], fr, sizeof(FRun) * (size_t)nr);
}

// ---- the pods of the single-launch kernel in two passes over the caller's arrays, no sort (host only; nosplan_host_stage_pods
// exposes it to CPU tests).  Where NewPodSorter's comparator (core/util.go:34-73) is a strict weak order on what is present,
// its order is a bucket order -- priority descending, then the rank of the requested profile, ties in original order -- so
// pass A validates, classifies and counts per bucket, and pass B writes every pod's pair straight to its padded position.
struct FastPlan {
    int n_keys = 0, n_runs = 0, n_chunks = 0, cpu_shift = 0, mem_shift = 0, pad_pods = 0;
    unsigned rset = 0;
    FRun fr[FRUN_MAX];
    int run_of[64 * (KP + 1)];
};
// pass A.  false: not this path (the general path takes the pods, and reports what is wrong with them if anything is)
template <class CanSaturate>
static bool fast_pods_plan(const nosplan_snapshot* s, const nosplan_pods* pods, int nnz, bool mig, std::vector<uint16_t>& bk,
                           std::vector<uint8_t>& key_seen, const CanSaturate& can_saturate, FastPlan& out) {
    const int K = s->n_profiles, P = pods->n_pods;
    constexpr int LV_MAX = 64, NB = KP + 1;
    struct Level { int32_t prio; int cnt[NB]; };
    Level levels[LV_MAX];
    int n_lv = 0;
    bk.resize(P);
    key_seen.assign(P, 0);
    unsigned long long or_cpu = 0, or_mem = 0;
    long long max_cpu = 0, max_mem = 0;
    int max_key = -1;
    int32_t lv_prio[LV_MAX];
    {   // local copies of the array pointers: the byte stores below may alias anything as far as the compiler knows
        const int32_t* __restrict sptr = pods->scalar_ptr; const int32_t* __restrict scol = pods->scalar_col;
        const int64_t* __restrict sqty = pods->scalar_qty; const int64_t* __restrict base = pods->base;
        const int32_t* __restrict tkey = pods->tracker_key; const int32_t* __restrict prios = pods->priority;
        uint16_t* __restrict bkp = bk.data(); uint8_t* __restrict seen = key_seen.data();
        int cnt[LV_MAX * NB];                              // pods per (priority level, profile) bucket, filled level by level
        int a = sptr[0];
        for (int i = 0; i < P; ++i) {
            const int a1 = sptr[i + 1], n = a1 - a;
            int pr = -1;
            if (n == 1) {
                if (a < 0 || a >= nnz) return false;
                const int col = scol[a];
                if (col < 3 || col >= 3 + K || sqty[a] != 1) return false;
                pr = col - 3;
            } else if (n != 0) return false;
            a = a1;
            const long long bc = base[3 * (size_t)i], bm = base[3 * (size_t)i + 1], be = base[3 * (size_t)i + 2];
            if ((bc | bm) < 0 || be != 0 || (n == 0 && (bc | bm) == 0)) return false;
            const int k = tkey[i];
            if ((unsigned)k >= (unsigned)P || seen[k]) return false;
            seen[k] = 1;
            max_key = std::max(max_key, k);
            const int32_t prio = prios[i];
            // few distinct priorities: compare with all of them, no data-dependent branch (the priorities arrive in random order)
            int hit = -1;
            for (int q = 0; q < n_lv; ++q) hit = lv_prio[q] == prio ? q : hit;
            if (hit < 0) {
                if (n_lv >= LV_MAX) return false;
                std::memset(cnt + n_lv * NB, 0, sizeof(int) * NB);
                lv_prio[n_lv] = prio;
                hit = n_lv++;
            }
            const int b = hit * NB + pr + 1;
            cnt[b]++;
            bkp[i] = (uint16_t)b;
            or_cpu |= (unsigned long long)bc; or_mem |= (unsigned long long)bm;
            max_cpu = std::max(max_cpu, bc); max_mem = std::max(max_mem, bm);
        }
        for (int q = 0; q < n_lv; ++q) { levels[q].prio = lv_prio[q]; std::memcpy(levels[q].cnt, cnt + q * NB, sizeof(int) * NB); }
    }
    // profiles present: the comparator must be a strict weak order on them, and no two of them may share a rank
    const int32_t* pmem = s->profile_mem_gb; const int32_t* pgi = s->profile_gi;
    auto smaller = [=](int x, int y) { if (pmem[x] < pmem[y]) return true; return mig && pgi[x] < pgi[y]; };
    bool present[KP] = {false};
    for (int q = 0; q < n_lv; ++q) {
        const Level& l = levels[q];
        bool sliced = false;
        for (int pr = 0; pr < K; ++pr) if (l.cnt[pr + 1]) { present[pr] = true; sliced = true; }
        if (l.cnt[0] && sliced) return false;                         // slice-less pods are incomparable to everything
    }
    int rank_of[KP] = {0};
    unsigned ranks_used = 0;
    for (int x = 0; x < K; ++x) {
        if (!present[x]) continue;
        for (int y = 0; y < K; ++y) {
            if (!present[y]) continue;
            if (smaller(x, y) && smaller(y, x)) return false;
            if (smaller(y, x)) ++rank_of[x];
            for (int z = 0; z < K; ++z) {
                if (!present[z]) continue;
                if (smaller(x, y) && smaller(y, z) && !smaller(x, z)) return false;
                const bool ixy = !smaller(x, y) && !smaller(y, x), iyz = !smaller(y, z) && !smaller(z, y), ixz = !smaller(x, z) && !smaller(z, x);
                if (ixy && iyz && !ixz) return false;
            }
        }
        if (ranks_used >> rank_of[x] & 1u) return false;
        ranks_used |= 1u << rank_of[x];
    }
    int prof_by_rank[KP], n_present = 0;
    for (int r = 0; r < KP; ++r) for (int x = 0; x < K; ++x) if (present[x] && rank_of[x] == r) prof_by_rank[n_present++] = x;
    for (int x = 0; x < K; ++x) {
        if (!present[x]) continue;
        long long asked = 0;
        for (int q = 0; q < n_lv; ++q) asked += levels[q].cnt[x + 1];
        if (!can_saturate(x, asked)) return false;                   // the general path stages the pods for the classic kernels
    }
    // runs = non-empty buckets, priority descending, then profile rank
    int lv_order[LV_MAX];
    for (int q = 0; q < n_lv; ++q) lv_order[q] = q;
    std::sort(lv_order, lv_order + n_lv, [&](int x, int y) { return levels[x].prio > levels[y].prio; });
    int nr = 0, pos = 0, chunks = 0;
    unsigned rs = 0;
    for (int q = 0; q < n_lv; ++q) {
        const Level& l = levels[lv_order[q]];
        for (int c = -1; c < n_present; ++c) {
            const int pr = c < 0 ? -1 : prof_by_rank[c], cnt = l.cnt[pr + 1];
            if (!cnt) continue;
            if (nr >= FRUN_MAX) return false;
            FRun f; std::memset(&f, 0, sizeof f);
            f.pos0 = pos; f.nblk = (cnt + 63) >> 6; f.n_real = cnt; f.p = pr; f.q = pr >= 0 ? 1 : 0; f.min_c = f.min_m = F_NONE;
            out.run_of[lv_order[q] * NB + pr + 1] = nr;
            out.fr[nr++] = f;
            pos += f.nblk * 64; chunks += (cnt + 31) >> 5;
            if (pr >= 0) rs |= 1u << pr;
        }
    }
    const int cs = or_cpu ? __builtin_ctzll(or_cpu) : 0, ms = or_mem ? __builtin_ctzll(or_mem) : 0;
    if ((max_cpu >> cs) >= F_COST_LIMIT || (max_mem >> ms) >= F_COST_LIMIT || pos > fused_max_pad_pods() || rs == 0) return false;
    out.n_keys = max_key + 1; out.n_runs = nr; out.n_chunks = chunks; out.cpu_shift = cs; out.mem_shift = ms; out.pad_pods = pos; out.rset = rs;
    return true;
}
// pass B into a blob that blob_begin has prepared (original order == order inside a bucket), then everything derived from the pairs
static void fast_pods_fill(const nosplan_pods* pods, const uint8_t* prefilter_ok, const std::vector<uint16_t>& bk, FastPlan& fp,
                           unsigned char* B, const int* fp_off) {
    const int P = pods->n_pods, cs = fp.cpu_shift, ms = fp.mem_shift;
    int32_t* fpod = (int32_t*)(B + fp_off[FP_PODS]);
    int cur[FRUN_MAX];
    for (int r = 0; r < fp.n_runs; ++r) {                  // the padding behind every run; the pods' own positions below
        const FRun& f = fp.fr[r];
        cur[r] = f.pos0;
        std::fill(fpod + 2 * (size_t)(f.pos0 + f.n_real), fpod + 2 * (size_t)(f.pos0 + f.nblk * 64), (int32_t)F_NONE);
    }
    const int64_t* __restrict base = pods->base;
    const uint16_t* __restrict bkp = bk.data();
    for (int i = 0; i < P; ++i) {
        const int x = cur[fp.run_of[bkp[i]]]++;
        const bool ok = !prefilter_ok || prefilter_ok[i];
        fpod[2 * (size_t)x] = ok ? (int32_t)(base[3 * (size_t)i] >> cs) : (int32_t)F_NONE;
        fpod[2 * (size_t)x + 1] = ok ? (int32_t)(base[3 * (size_t)i + 1] >> ms) : (int32_t)F_NONE;
    }
    blob_finish(B, fp_off, fp.fr, fp.n_runs);
}

// Host-only: the pod structure of the single-launch kernel exactly as nosplan_plan stages it (layout in nosplan_dev.cuh), into the
// caller's buffer.  meta[0..10] = byte offsets of the sections, then runs, padded positions, cpu shift, memory shift, requested set.
// Returns the blob's size in bytes, 0 when the two-pass path does not take these pods, < 0 on error.  Needs no GPU.
extern "C" int64_t nosplan_host_stage_pods(const nosplan_snapshot* s, const nosplan_pods* pods, const uint8_t* prefilter_ok, uint8_t* blob,
                                           int64_t capacity, int32_t* meta) {
    if (!s || !pods || !meta) return NOSPLAN_E_INVALID;
    const int K = s->n_profiles, P = pods->n_pods;
    if (K < 0 || K > NOSPLAN_MAX_PROFILES || P <= 0) return NOSPLAN_E_INVALID;
    if (!pods->scalar_ptr || !pods->base || !pods->priority || !pods->tracker_key || !s->profile_mem_gb || !s->profile_gi) return NOSPLAN_E_INVALID;
    if (pods->scalar_ptr[P] < 0 || (pods->scalar_ptr[P] > 0 && (!pods->scalar_col || !pods->scalar_qty))) return NOSPLAN_E_INVALID;
    const bool mig = s->kind == NOSPLAN_KIND_MIG;
    const int n_models = mig ? s->n_models : 0;
    if (n_models < 0 || n_models > MAX_MODELS || (n_models > 0 && (!s->model_geom_offset || !s->geom_counts))) return NOSPLAN_E_INVALID;
    const int n_rows = (mig && n_models > 0) ? s->model_geom_offset[n_models] : 0;
    if (n_rows < 0 || n_rows > MAX_ROWS) return NOSPLAN_E_LIMIT;
    static thread_local std::vector<uint16_t> bk;
    static thread_local std::vector<uint8_t> key_seen;
    static const bool trace = getenv("NOSPLAN_TRACE") != nullptr;       // NOSPLAN_TRACE=1: microseconds per phase on stderr
    auto now = [] { return std::chrono::steady_clock::now(); };
    auto us = [](auto a, auto b) { return std::chrono::duration<double, std::micro>(b - a).count(); };
    FastPlan fp;
    auto always = [](int, long long) { return true; };
    const auto t0 = now();
    if (!fast_pods_plan(s, pods, pods->scalar_ptr[P], mig, bk, key_seen, always, fp)) return 0;
    const auto t1 = now();
    int fp_off[FP_NOFF];
    blob_layout(n_rows, n_models, fp.n_runs, fp.pad_pods, fp_off);
    for (int k = 0; k < FP_NOFF; ++k) meta[k] = fp_off[k];
    meta[FP_NOFF] = fp.n_runs; meta[FP_NOFF + 1] = fp.pad_pods; meta[FP_NOFF + 2] = fp.cpu_shift; meta[FP_NOFF + 3] = fp.mem_shift;
    meta[FP_NOFF + 4] = (int32_t)fp.rset;
    if (!blob || capacity < fp_off[FP_END]) return fp_off[FP_END];
    const auto t2 = now();
    blob_begin(blob, fp_off, s, n_rows, n_models, true);
    const auto t3 = now();
    fast_pods_fill(pods, prefilter_ok, bk, fp, blob, fp_off);
    if (trace) fprintf(stderr, "[nosplan_host_stage_pods] pass A %.1f us, begin %.1f us, pass B + finish %.1f us\n", us(t0, t1), us(t2, t3), us(t3, now()));
    return fp_off[FP_END];
}

// The classic kernels' view of the pods, in NewPodSorter order (h->v_order): request columns, tracker keys, flags, CSR of the
// scalar entries, the run records and per-chunk (32 pods, run-relative) minima of the requests of the PreFilter-ok pods -- the
// kernels skip chunks in which no pod can pass NodeResourcesFit.  An all-zero request passes regardless (fit.go short-circuit):
// its minimum is LLONG_MIN.  Written into the staging buffer at the offsets of h->lay.
static void stage_classic_pods(nosplan* h, const nosplan_pods* pods, int K) {
    (void)K;
    const Layout& L = h->lay;
    unsigned char* S = h->h_stage;
    const int P = pods->n_pods;
    const std::vector<int>& order = h->v_order;
    const std::vector<uint32_t>& flags = h->v_flags;
    const std::vector<RunRec>& runs = h->v_runs;
    const int n_runs = (int)runs.size();
    long long* cpu = (long long*)(S + L.pod_cpu); long long* mem = (long long*)(S + L.pod_mem); long long* eph = (long long*)(S + L.pod_eph);
    int32_t* key = (int32_t*)(S + L.pod_key); uint32_t* fl = (uint32_t*)(S + L.pod_flags);
    int32_t* sptr = (int32_t*)(S + L.pod_sptr); int32_t* scol = (int32_t*)(S + L.pod_scol); long long* sqty = (long long*)(S + L.pod_sqty);
    long long* c0 = (long long*)(S + L.cmin_cpu); long long* c1 = (long long*)(S + L.cmin_mem); long long* c2 = (long long*)(S + L.cmin_eph);
    int w = 0, chunk = -1;
    for (int r = 0; r < n_runs; ++r)
        for (int k = 0; k < runs[r].len; ++k) {
            const int si = runs[r].start + k, oi = order[si];
            if ((k & 31) == 0) { ++chunk; c0[chunk] = c1[chunk] = c2[chunk] = LLONG_MAX; }
            const long long bc = pods->base[3 * (size_t)oi], bm = pods->base[3 * (size_t)oi + 1], be = pods->base[3 * (size_t)oi + 2];
            const uint32_t f = flags[oi];
            if (f & PF_PRE_OK) {
                const bool z = (f & PF_ZERO_REQ) != 0;
                c0[chunk] = std::min(c0[chunk], z ? LLONG_MIN : bc); c1[chunk] = std::min(c1[chunk], z ? LLONG_MIN : bm);
                c2[chunk] = std::min(c2[chunk], z ? LLONG_MIN : be);
            }
            cpu[si] = bc; mem[si] = bm; eph[si] = be;
            key[si] = pods->tracker_key[oi]; fl[si] = f;
            sptr[si] = w;
            for (int e = pods->scalar_ptr[oi]; e < pods->scalar_ptr[oi + 1]; ++e) { scol[w] = pods->scalar_col[e]; sqty[w] = pods->scalar_qty[e]; ++w; }
        }
    sptr[P] = w;
    if (n_runs) std::memcpy(S + L.runs, runs.data(), sizeof(RunRec) * (size_t)n_runs);
    std::memcpy(S + L.run_chunk_off, h->v_chunk_off.data(), 4 * ((size_t)n_runs + 1));
}

// the node sections of the arena: the same offsets for every upload of a snapshot of this shape
static void layout_nodes(Layout& L, size_t& o, int n_models, int n_rows, int N, int G, int R, bool mig) {
    auto put = [&](size_t& field, size_t bytes) { field = o; o = al(o + bytes); };
    const size_t NG = (size_t)N * G;
    put(L.model_off, sizeof(int32_t) * (n_models + 1));
    put(L.geom, 16 * (size_t)n_rows);
    put(L.node_gpu_count, 4 * (size_t)N);
    // only the array of the snapshot's kind travels (model ids: MIG, memory sizes: MPS); the other pointer aliases it, its
    // values are read into registers by shared code and never used
    if (mig) { put(L.gpu_model, 4 * NG); L.gpu_mem = L.gpu_model; } else { put(L.gpu_mem, 4 * NG); L.gpu_model = L.gpu_mem; }
    put(L.gpu_used, 16 * NG); put(L.gpu_free, 16 * NG);
    put(L.gpu_umask, 2 * NG); put(L.gpu_fmask, 2 * NG);
    put(L.node_alloc, 8 * (size_t)N * R); put(L.node_req, 8 * (size_t)N * R);
    put(L.node_allowed, 4 * (size_t)N); put(L.node_npods, 4 * (size_t)N);
    L.nodes_end = o;
}

// keep: the node arrays of the last full upload are still in the arena and are planned against again (s = &h->res_snap)
static int upload_impl(nosplan_t* h, const nosplan_snapshot* s, const nosplan_pods* pods, const uint8_t* prefilter_ok,
                       int32_t filter_mode, bool keep) {
    if (!h) return NOSPLAN_E_INVALID;
    if (!s || !pods) return fail(h, NOSPLAN_E_INVALID, "null snapshot or pods");
    if (!keep) h->nodes_resident = false;
    h->uploaded = h->ran = h->pre_done = false;
    // NOSPLAN_TRACE=1: per-phase host microseconds of this call on stderr
    const bool trace = getenv("NOSPLAN_TRACE") != nullptr;
    auto now = [] { return std::chrono::steady_clock::now(); };
    auto t_prev = now();
    auto lap = [&](const char* what) {
        if (!trace) return;
        auto t = now();
        fprintf(stderr, "[nosplan_upload] %-12s %8.1f us\n", what, std::chrono::duration<double, std::micro>(t - t_prev).count());
        t_prev = t;
    };
    const int K = s->n_profiles, N = s->n_nodes, G = s->gpus_per_node, P = pods->n_pods;
    const int R = 3 + K + s->n_other;
    if (s->kind != NOSPLAN_KIND_MIG && s->kind != NOSPLAN_KIND_MPS) return fail(h, NOSPLAN_E_INVALID, "unknown kind %d", s->kind);
    if (filter_mode < 0 || filter_mode > NOSPLAN_FILTER_RESOURCES_FIT) return fail(h, NOSPLAN_E_INVALID, "unknown filter mode %d", filter_mode);
    if (K < 0 || N < 0 || P < 0 || G < 1 || s->n_other < 0) return fail(h, NOSPLAN_E_INVALID, "negative size");
    if (K > NOSPLAN_MAX_PROFILES) return fail(h, NOSPLAN_E_LIMIT, "%d profiles > NOSPLAN_MAX_PROFILES", K);
    if (G > NOSPLAN_MAX_GPUS) return fail(h, NOSPLAN_E_LIMIT, "%d GPUs per node > NOSPLAN_MAX_GPUS", G);
    if (s->n_other > NOSPLAN_MAX_OTHER) return fail(h, NOSPLAN_E_LIMIT, "%d other resources > NOSPLAN_MAX_OTHER", s->n_other);
    const bool mig = s->kind == NOSPLAN_KIND_MIG;
    const int n_models = mig ? s->n_models : 0;
    const int n_rows = (mig && n_models > 0) ? s->model_geom_offset[n_models] : 0;
    if (n_models > MAX_MODELS) return fail(h, NOSPLAN_E_LIMIT, "%d GPU models > %d", n_models, MAX_MODELS);
    if (n_rows > MAX_ROWS) return fail(h, NOSPLAN_E_LIMIT, "%d geometry rows > %d", n_rows, MAX_ROWS);
    for (int m = 0; m < n_models; ++m) {
        const int c = s->model_geom_offset[m + 1] - s->model_geom_offset[m];
        if (c < 0 || c > NOSPLAN_MAX_GEOMS) return fail(h, NOSPLAN_E_LIMIT, "model %d has %d geometries > NOSPLAN_MAX_GEOMS", m, c);
    }
    const int nnz = P > 0 ? pods->scalar_ptr[P] : 0;
    if (nnz < 0) return fail(h, NOSPLAN_E_INVALID, "negative scalar entry count");
    lap("validate");

    // ---- layout, part 1: every section whose size does not depend on the pods' contents
    Layout L; size_t o = 0;
    auto put = [&](size_t& field, size_t bytes) { field = o; o = al(o + bytes); };
    const size_t NG = (size_t)N * G;
    layout_nodes(L, o, n_models, n_rows, N, G, R, mig);
    // small per-run state, twice: run k works in copy k & 1 while its k_prepass clears the other one for run k + 1, so no
    // reset is queued in front of a run (the large tracker tables below are cleared by k_prepass for its own run)
    put(L.tot, 8 * KP); put(L.lack, 4 * KP); put(L.stats, sizeof(DevStats));
    put(L.maxfree, 4 * KP); put(L.gflags, 16); put(L.acc, 32); put(L.lack_dec, 4 * KP); put(L.next_node, 16); put(L.wave, 32);
    const size_t small_bytes = o - L.tot;
    o += small_bytes;                                        // the second copy
    // tracker tables: sized for P keys (keys <= P), reset with the rest; the single-launch plan indexes `placed` by padded pod
    // position instead (at most FRUN_MAX runs, each padded to a multiple of 64)
    const size_t pad_pods_max = (size_t)P + 64 * (size_t)FRUN_MAX;
    put(L.placed, 4 * ((pad_pods_max + 31) / 32));
    put(L.entry_lack, 4 * (size_t)P * KP);
    L.work_end = o;
    put(L.cand_flag, (size_t)N); put(L.cand_idx, 4 * (size_t)N);
    put(L.out_count, 16 * NG); put(L.out_mask, 2 * NG);
    L.out_end = o;
    // gate replay records (sweep.cu): written before they are read, never cleared
    put(L.rec_dw, 8 * (size_t)N * KP); put(L.rec_pl, 8 * (size_t)N * KP); put(L.rec_fl, 4 * (size_t)N); put(L.gate_state, 8 * (KP + 2));
    put(L.rec_chunk, 32 * (size_t)KP * ((size_t)N / 32 + 2)); put(L.rec_chunk_fl, 4 * ((size_t)N / 32 + 2));
    put(L.pod_cpu, 8 * (size_t)P); put(L.pod_mem, 8 * (size_t)P); put(L.pod_eph, 8 * (size_t)P);
    put(L.pod_key, 4 * (size_t)P); put(L.pod_flags, 4 * (size_t)P);
    put(L.pod_sptr, 4 * ((size_t)P + 1)); put(L.pod_scol, 4 * (size_t)nnz); put(L.pod_sqty, 8 * (size_t)nnz);
    const size_t part1_end = o;
    // part 2 (runs, chunk minima) is laid out once the pods are ordered; the arena is sized from bounds: runs <= P, chunks <= P
    const size_t bound = part1_end + al(sizeof(RunRec) * (size_t)P) + al(4 * ((size_t)P + 1)) + 3 * al(8 * (size_t)P) + 4096 +
                         al(8 * (pad_pods_max + pad_pods_max / 8 + pad_pods_max / 64 + pad_pods_max / 512 + 8)) + al(3 * pad_pods_max + 64) + al(sizeof(FRun) * FRUN_MAX) +
                         al(18 * (size_t)MAX_ROWS) + al(4 * (MAX_MODELS + 1)) + 4096;
    CU(h, cudaSetDevice(h->device));
    const size_t out_bytes = al(16 * NG) + al(2 * NG) + al(sizeof(DevStats));
    int rc = ensure(h, bound, bound, out_bytes, keep ? h->res_nodes_end : 0);
    if (rc != NOSPLAN_OK) return rc;
    lap("layout+alloc");

    // ---- node arrays first: their DMA runs while the host orders the pods.  If the caller keeps them in page-locked memory
    //      (cudaHostAlloc / cudaHostRegister) the copy engine reads them where they are; otherwise they pass through staging.
    unsigned char* S = h->h_stage;
    unsigned char* A = h->d_arena;
    if (!keep) {
    if (n_models > 0) std::memcpy(S + L.model_off, s->model_geom_offset, sizeof(int32_t) * (n_models + 1));
    else std::memset(S + L.model_off, 0, sizeof(int32_t));
    if (n_rows) std::memcpy(S + L.geom, s->geom_counts, 16 * (size_t)n_rows);
    }
    struct Seg { size_t off; const void* src; size_t bytes; bool direct; };
    Seg segs[11] = {
        {L.node_gpu_count, s->node_gpu_count, 4 * (size_t)N, false}, {L.gpu_model, s->gpu_model, mig ? 4 * NG : 0, false},
        {L.gpu_mem, s->gpu_memory_gb, mig ? 0 : 4 * NG, false}, {L.gpu_used, s->gpu_used, 16 * NG, false}, {L.gpu_free, s->gpu_free, 16 * NG, false},
        {L.gpu_umask, s->gpu_used_mask, 2 * NG, false}, {L.gpu_fmask, s->gpu_free_mask, 2 * NG, false},
        {L.node_alloc, s->node_alloc, 8 * (size_t)N * R, false}, {L.node_req, s->node_requested, 8 * (size_t)N * R, false},
        {L.node_allowed, s->node_allowed_pods, 4 * (size_t)N, false}, {L.node_npods, s->node_num_pods, 4 * (size_t)N, false}};
    bool any_direct = false;
    if (N && !keep) {
        for (Seg& g : segs) {
            // small arrays are cheaper to memcpy into staging than to give a DMA descriptor of their own
            cudaPointerAttributes at;
            g.direct = g.bytes >= (128u << 10) && cudaPointerGetAttributes(&at, g.src) == cudaSuccess && at.type == cudaMemoryTypeHost;
            if (g.direct) any_direct = true;
            else if (g.bytes) std::memcpy(S + g.off, g.src, g.bytes);
        }
        (void)cudaGetLastError();
    }
    CU(h, cudaEventRecord(h->ev[0], h->stream));
    h->dbg_n = 0;
    if (!keep) {
        size_t from = 0;
        auto mark = [&](size_t bytes, bool direct) {          // NOSPLAN_TRACE: one event per copy
            if (!trace || h->dbg_n >= 32) return;
            if (!h->dbg_ev[h->dbg_n]) cudaEventCreate(&h->dbg_ev[h->dbg_n]);
            cudaEventRecord(h->dbg_ev[h->dbg_n], h->stream);
            h->dbg_bytes[h->dbg_n] = bytes; h->dbg_direct[h->dbg_n] = direct; h->dbg_n++;
        };
        // cudaStreamQuery after each copy pushes it to the copy engine at once (the driver otherwise holds queued copies back for
        // a while): measured in one session, end to end 0.60 ms without, 0.55 ms with; an event record per copy and alternating
        // the copies between two streams were both measured slower.  NOSPLAN_COPY_MODE=0 switches the push off.
        const char* cm = getenv("NOSPLAN_COPY_MODE");
        const int copy_mode = cm ? atoi(cm) : 1;
        for (const Seg& g : segs) {
            if (!g.direct) continue;
            if (g.off > from) { CU(h, cudaMemcpyAsync(A + from, S + from, g.off - from, cudaMemcpyHostToDevice, h->stream)); mark(g.off - from, false); }
            CU(h, cudaMemcpyAsync(A + g.off, g.src, g.bytes, cudaMemcpyHostToDevice, h->stream));
            if (copy_mode == 1) (void)cudaStreamQuery(h->stream);
            mark(g.bytes, true);
            from = al(g.off + g.bytes);
        }
        if (L.nodes_end > from) CU(h, cudaMemcpyAsync(A + from, S + from, L.nodes_end - from, cudaMemcpyHostToDevice, h->stream));
    }
    h->direct_h2d = any_direct;
    // nothing of the caller's may be in flight when this call returns, with an error as well: outside nosplan_plan (which
    // synchronises itself) every return below waits for the copies queued above
    struct InFlight { cudaStream_t st; bool armed; ~InFlight() { if (armed) cudaStreamSynchronize(st); } } in_flight{h->stream, any_direct && !h->in_plan};
    CU(h, cudaMemsetAsync(A + L.tot, 0, L.out_end - L.tot, h->stream));       // per-run state (both copies) ... plan rows
    CU(h, cudaEventRecord(h->ev[1], h->stream));

    // ---- device problem description, node side
    DevProblem& pb = h->pb;
    std::memset(&pb, 0, sizeof pb);
    pb.kind = s->kind; pb.K = K; pb.R = R; pb.n_other = s->n_other;
    pb.node_lo = 0; pb.node_hi = N; pb.slice_mode = 0;
    pb.N = N; pb.G = G; pb.P = P; pb.filter_mode = filter_mode;
    pb.n_models = n_models; pb.n_rows = n_rows;
    for (int p = 0; p < KP; ++p) { pb.pmem[p] = p < K ? s->profile_mem_gb[p] : 0; pb.pgi[p] = p < K ? s->profile_gi[p] : 0; }
    pb.model_off = (const int32_t*)(A + L.model_off); pb.geom = (const uint4*)(A + L.geom);
    pb.node_gpu_count = (const int32_t*)(A + L.node_gpu_count);
    pb.gpu_model = (const int32_t*)(A + L.gpu_model); pb.gpu_mem = (const int32_t*)(A + L.gpu_mem);
    pb.gpu_used = (const uint4*)(A + L.gpu_used); pb.gpu_free = (const uint4*)(A + L.gpu_free);
    pb.gpu_umask = (const uint16_t*)(A + L.gpu_umask); pb.gpu_fmask = (const uint16_t*)(A + L.gpu_fmask);
    pb.node_alloc = (const long long*)(A + L.node_alloc); pb.node_req = (const long long*)(A + L.node_req);
    pb.node_allowed_pods = (const int32_t*)(A + L.node_allowed); pb.node_num_pods = (const int32_t*)(A + L.node_npods);
    pb.maxfree = (int32_t*)(A + L.maxfree); pb.gflags = (int32_t*)(A + L.gflags);
    pb.acc = (unsigned long long*)(A + L.acc);
    pb.lack_dec = (int32_t*)(A + L.lack_dec); pb.next_node = (int32_t*)(A + L.next_node); pb.whole_snapshot = 1;
    pb.wave = (int32_t*)(A + L.wave);
    pb.tot = (long long*)(A + L.tot); pb.lack = (int32_t*)(A + L.lack); pb.stats = (DevStats*)(A + L.stats);
    pb.cand_flag = (uint8_t*)(A + L.cand_flag);
    pb.out_count = (uint4*)(A + L.out_count); pb.out_mask = (uint16_t*)(A + L.out_mask);
    h->d_cand_idx = (int32_t*)(A + L.cand_idx);
    pb.rec_dw = (long long*)(A + L.rec_dw); pb.rec_pn = (int2*)(A + L.rec_pl); pb.rec_fl = (int32_t*)(A + L.rec_fl);
    pb.gate_state = (long long*)(A + L.gate_state);
    pb.rec_chunk = (long long*)(A + L.rec_chunk); pb.rec_chunk_fl = (int32_t*)(A + L.rec_chunk_fl);
    h->d_work = A + L.tot; h->work_bytes = L.work_end - L.tot; h->small_bytes = small_bytes;
    h->run_parity = 1; h->bound_copy = 0;                    // pb points at copy 0; the first run flips to it
    h->out_count_bytes = 16 * NG; h->out_mask_bytes = 2 * NG;
    // Node-side conditions of the single-launch plan (fused.cu); the pod-side ones follow below.  Where it can still apply,
    // k_prepass is not queued early: the fused kernel does that work itself once the pods are on the device.
    // (MPS as well: slicing.GPU.UpdateGeometryFor per GPU lane; NOSPLAN_NO_FUSED_MPS keeps MPS on the classic kernels)
    bool fused_ok = (mig || !getenv("NOSPLAN_NO_FUSED_MPS")) && G <= 8 && P > 0 && !getenv("NOSPLAN_NO_FUSED");
    for (int p = 0; p < K && fused_ok && !mig; ++p) if (s->profile_mem_gb[p] <= 0) fused_ok = false;
    for (int r = 0; r < n_rows && fused_ok; ++r)
        for (int p = 0; p < 16; ++p) if (s->geom_counts[16 * (size_t)r + p] > 31) { fused_ok = false; break; }
    if (h->in_plan && !fused_ok) { rc = run_pre(h, true); if (rc != NOSPLAN_OK) return rc; }
    lap("nodes queued");
    // per-node validation, while the copy engine works (k_prepass, the only kernel that can already be queued, clamps what it
    // reads; every other kernel is launched after this call has succeeded)
    for (int i = 0; i < N && !keep; ++i) {
        const int ng = s->node_gpu_count[i];
        bool bad = ng < 0 || ng > G;
        if (!bad && mig) {
            const int32_t* gm = s->gpu_model + (size_t)i * G;
            unsigned b = 0;
            for (int j = 0; j < ng; ++j) b |= (unsigned)gm[j] >= (unsigned)n_models;
            bad = b != 0;
        }
        if (bad) {
            cudaStreamSynchronize(h->stream);                    // nothing of the caller's may be in flight when we return
            const bool cnt = ng < 0 || ng > G;
            return fail(h, NOSPLAN_E_INVALID, cnt ? "node %d: gpu count out of range" : "node %d: GPU model id out of range", i);
        }
    }
    lap("validate nodes");

    // ---- results of the pod side (either path below)
    int n_keys = 0, n_runs = 0, n_chunks = 0, cpu_shift = 0, mem_shift = 0, pad_pods = 0;
    bool single_profile_classes = true, lazy = false;
    long long any_eph = 0;
    unsigned rset = 0;
    int fp_off[FP_NOFF] = {0};
    size_t up_from = 0;
    fused_ok = fused_ok && !h->force_classic;

    // ---- fast path of one nosplan_plan / nosplan_plan_resident call that can use the single-launch kernel: TWO passes over
    //      the pods in their original order.  Pass A validates, finds every pod's (priority level, requested profile) bucket
    //      and counts; where NewPodSorter's comparator is a strict weak order on what is present, every stable sort gives
    //      "buckets in (priority desc, profile rank) order, original order inside" (core/util.go:34-73, see pod_order), so pass
    //      B writes each pod's (cpu, memory) pair straight to its position of the padded space.  Anything this does not cover
    //      -- an invalid pod, several scalar entries, quantities above one, shared tracker keys, priority levels that mix
    //      slice-less and sliced pods -- leaves it to the general path below, which also reports the errors.
    // The single-launch kernel's hypothesis needs lack[p] > (G+2)*M_p for every requested profile, and lack[p] can never exceed
    // the devices of p all pods ask for together: a pod list too short for that is not worth the attempt (MIG: M_p from the
    // geometry tables; MPS: unknown here, the GPU memory sizes are not looked at on the host)
    auto can_saturate = [&](int prof, long long asked) {
        if (!mig) return true;
        int m = 0;
        for (int r = 0; r < n_rows; ++r) m = std::max(m, (int)s->geom_counts[16 * (size_t)r + prof]);
        return asked > (long long)(G + 2) * m;
    };
    bool fast_done = false;
    if (fused_ok && h->in_plan && P > 0 && !getenv("NOSPLAN_NO_FAST_PODS")) fast_done = [&]() -> bool {
        FastPlan fp;
        if (!fast_pods_plan(s, pods, nnz, mig, h->v_bucket, h->v_key_seen, can_saturate, fp)) return false;
        lap("pods: pass A");
        // layout, part 2 (the classic sections keep their room: a rejected plan stages them later), then the blob
        n_keys = fp.n_keys; n_runs = fp.n_runs; n_chunks = fp.n_chunks; cpu_shift = fp.cpu_shift; mem_shift = fp.mem_shift;
        pad_pods = fp.pad_pods; rset = fp.rset;
        single_profile_classes = true; any_eph = 0;
        o = part1_end;
        put(L.runs, sizeof(RunRec) * (size_t)n_runs);
        put(L.run_chunk_off, 4 * ((size_t)n_runs + 1));
        put(L.cmin_cpu, 8 * (size_t)n_chunks); put(L.cmin_mem, 8 * (size_t)n_chunks); put(L.cmin_eph, 8 * (size_t)n_chunks);
        blob_layout(n_rows, n_models, n_runs, pad_pods, fp_off);
        put(L.fp_blob, (size_t)fp_off[FP_END]);
        L.upload_end = o; L.total = o;
        if (L.total > bound) return false;
        h->lay = L;
        unsigned char* B = S + L.fp_blob;
        blob_begin(B, fp_off, s, n_rows, n_models, true);
        lap("pods: begin");
        fast_pods_fill(pods, prefilter_ok, h->v_bucket, fp, B, fp_off);
        lap("pods: pass B, finish");
        lazy = true; up_from = L.fp_blob;
        h->classic_staged = false;
        h->lazy_valid = true; h->lazy_pods = *pods; h->lazy_pre = prefilter_ok; h->lazy_filter = filter_mode;
        return true;
    }();
    if (fast_done) {
        CU(h, cudaEventRecord(h->ev[9], h->stream));
        CU(h, cudaMemcpyAsync(A + up_from, S + up_from, L.upload_end - up_from, cudaMemcpyHostToDevice, h->stream));
        CU(h, cudaEventRecord(h->ev[10], h->stream));
    }
    lap("pods fast");
    if (!fast_done) {
        // ---- pods, pass 1 (original order): validation, flags, first requested profile, class key of the slice request
        //      (equal keys <=> identical slice requests; CLS_MULTI = compare the entries)
        constexpr unsigned long long CLS_MULTI = ~0ull;
        std::vector<int>& first_prof = h->v_first_prof; first_prof.assign(P, -1);
        std::vector<uint32_t>& flags = h->v_flags; flags.resize(P);
        std::vector<unsigned long long>& cls = h->v_cls; cls.resize(P);
        n_keys = 0;
        std::vector<uint8_t>& key_seen = h->v_key_seen; key_seen.assign(P, 0);
        bool keys_unique = true, any_other = false, any_zero = false;
        long long max_q = 0, sum_q[KP] = {0};
        unsigned long long or_cpu = 0, or_mem = 0;
        long long max_cpu = 0, max_mem = 0;
        any_eph = 0;
        for (int i = 0; i < P; ++i) {
            const int a = pods->scalar_ptr[i], b = pods->scalar_ptr[i + 1];
            if (a > b || b > nnz) return fail(h, NOSPLAN_E_INVALID, "pod %d: bad scalar_ptr", i);
            bool other = false; int prev = 2, n_slice = 0;
            unsigned long long ck = 0;
            for (int e = a; e < b; ++e) {
                const int col = pods->scalar_col[e];
                if (col < 3 || col >= R || col <= prev) return fail(h, NOSPLAN_E_INVALID, "pod %d: scalar columns must be in [3,R) and ascending", i);
                prev = col;
                if (col < 3 + K) {
                    const long long q = pods->scalar_qty[e];
                    if (q <= 0) return fail(h, NOSPLAN_E_INVALID, "pod %d: slice request quantity must be >= 1", i);
                    // the kernels keep lacking quantities in int32 (the reference in Go's 64-bit int): refuse what could wrap
                    if (q >= (1ll << 31) || (sum_q[col - 3] += q) >= (1ll << 31))
                        return fail(h, NOSPLAN_E_LIMIT, "pod %d: requested quantities of slice column %d reach 2^31", i, col);
                    if (first_prof[i] < 0) first_prof[i] = col - 3;
                    max_q = std::max(max_q, q);
                    ck = (++n_slice == 1 && q < (1ll << 40)) ? (((unsigned long long)(col - 2)) << 40 | (unsigned long long)q) : CLS_MULTI;
                } else other = true;
            }
            cls[i] = ck;
            const long long bc = pods->base[3 * (size_t)i], bm = pods->base[3 * (size_t)i + 1], be = pods->base[3 * (size_t)i + 2];
            if ((bc | bm | be) < 0) return fail(h, NOSPLAN_E_INVALID, "pod %d: negative cpu / memory / ephemeral-storage request", i);
            uint32_t f = 0;
            if (!prefilter_ok || prefilter_ok[i]) f |= PF_PRE_OK;
            if (other) { f |= PF_HAS_OTHER; any_other = true; }
            if (a == b && (bc | bm | be) == 0) { f |= PF_ZERO_REQ; any_zero = true; }
            flags[i] = f;
            or_cpu |= (unsigned long long)bc; or_mem |= (unsigned long long)bm; any_eph |= be;
            max_cpu = std::max(max_cpu, bc); max_mem = std::max(max_mem, bm);
            const int k = pods->tracker_key[i];
            if (k < 0 || k >= P) return fail(h, NOSPLAN_E_INVALID, "pod %d: tracker_key out of range", i);
            if (key_seen[k]) keys_unique = false;
            key_seen[k] = 1;
            n_keys = std::max(n_keys, k + 1);
        }
        if (n_keys > sweep_max_keys()) return fail(h, NOSPLAN_E_LIMIT, "%d tracker keys > %d", n_keys, sweep_max_keys());
        lap("pod keys");

        // ---- NewPodSorter (core/util.go:34-73)
        std::vector<int>& order = h->v_order;
        pod_order(s, pods, first_prof.data(), order);
        lap("sort");

        // ---- pass 2a (sorted order): runs of identical slice requests; 32-pod chunks are run-relative
        std::vector<RunRec>& runs = h->v_runs; runs.clear();
        std::vector<int32_t>& run_chunk_off = h->v_chunk_off; run_chunk_off.clear();
        auto same_entries = [&](int x, int y) {
            int ex = pods->scalar_ptr[x], ey = pods->scalar_ptr[y];
            const int bx = pods->scalar_ptr[x + 1], by = pods->scalar_ptr[y + 1];
            for (;;) {
                const bool sx = ex < bx && pods->scalar_col[ex] < 3 + K, sy = ey < by && pods->scalar_col[ey] < 3 + K;
                if (!sx || !sy) return sx == sy;
                if (pods->scalar_col[ex] != pods->scalar_col[ey] || pods->scalar_qty[ex] != pods->scalar_qty[ey]) return false;
                ++ex; ++ey;
            }
        };
        single_profile_classes = true;
        n_chunks = 0;
        for (int si = 0, prev_oi = -1; si < P; ++si) {
            const int oi = order[si];
            const bool same = si > 0 && cls[oi] == cls[prev_oi] && (cls[oi] != CLS_MULTI || same_entries(prev_oi, oi));
            if (!same) {
                if (!runs.empty()) n_chunks += (runs.back().len + 31) >> 5;
                RunRec r; std::memset(&r, 0, sizeof r);
                r.start = si;
                for (int e = pods->scalar_ptr[oi]; e < pods->scalar_ptr[oi + 1] && pods->scalar_col[e] < 3 + K; ++e) {
                    const long long q = pods->scalar_qty[e];
                    r.q[pods->scalar_col[e] - 3] = (uint16_t)std::min<long long>(q, 65535);   // device counts are <= 255: anything larger can never fit
                    r.n_ent++;
                    r.p_single = (uint16_t)(pods->scalar_col[e] - 3);
                }
                if (r.n_ent > 1) single_profile_classes = false;
                runs.push_back(r);
                run_chunk_off.push_back(n_chunks);
            }
            runs.back().len++;
            prev_oi = oi;
        }
        if (!runs.empty()) n_chunks += (runs.back().len + 31) >> 5;
        n_runs = (int)runs.size();
        run_chunk_off.push_back(n_chunks);
        lap("runs");

        // ---- layout, part 2
        o = part1_end;
        put(L.runs, sizeof(RunRec) * (size_t)n_runs);
        put(L.run_chunk_off, 4 * ((size_t)n_runs + 1));
        put(L.cmin_cpu, 8 * (size_t)n_chunks); put(L.cmin_mem, 8 * (size_t)n_chunks); put(L.cmin_eph, 8 * (size_t)n_chunks);
        // ---- single-launch plan (fused.cu), pod-side conditions: single-profile classes in at most FRUN_MAX runs, one tracker
        //      entry per pod, only cpu / memory / slice requests, and requests that fit the compact int32 units
        fused_ok = fused_ok && single_profile_classes && n_runs <= FRUN_MAX && keys_unique && any_eph == 0 && !any_zero &&
                   !(any_other && filter_mode == NOSPLAN_FILTER_RESOURCES_FIT) && max_q <= 1;
        for (int x = 0; x < K && fused_ok; ++x) if (sum_q[x] > 0 && !can_saturate(x, sum_q[x])) fused_ok = false;
        if (fused_ok) {
            cpu_shift = or_cpu ? __builtin_ctzll(or_cpu) : 0; mem_shift = or_mem ? __builtin_ctzll(or_mem) : 0;
            fused_ok = (max_cpu >> cpu_shift) < F_COST_LIMIT && (max_mem >> mem_shift) < F_COST_LIMIT;
            for (const RunRec& r : runs) { pad_pods += (r.len + 63) & ~63; if (r.n_ent) rset |= 1u << r.p_single; }
            fused_ok = fused_ok && pad_pods <= fused_max_pad_pods() && rset != 0;
        }
        // the blob: header [geometry rows | key masks | model ranges | runs | runs per profile], then (cpu, memory) pairs
        // [minima per 512 | per 64 | per 8 | pods]; every section starts on a 16-byte boundary
        blob_layout(n_rows, n_models, n_runs, pad_pods, fp_off);
        if (fused_ok) put(L.fp_blob, (size_t)fp_off[FP_END]); else L.fp_blob = o;
        L.upload_end = o;
        L.total = o;
        if (L.total > bound) return fail(h, NOSPLAN_E_NOMEM, "internal: arena bound exceeded");
        h->lay = L;
        if (fused_ok) {
            // padded position space: run r occupies [pos0, pos0 + 64 * nblk); a pod that can never be placed (PreFilter verdict
            // false) and every padding slot cost F_NONE
            unsigned char* B = S + L.fp_blob;
            blob_begin(B, fp_off, s, n_rows, n_models);
            int32_t* fpod = (int32_t*)(B + fp_off[FP_PODS]);
            FRun fr[FRUN_MAX];
            int pos = 0;
            for (int r = 0; r < n_runs; ++r) {
                const RunRec& rr = runs[r];
                FRun f; std::memset(&f, 0, sizeof f);
                f.pos0 = pos; f.nblk = (rr.len + 63) >> 6; f.n_real = rr.len;
                f.p = rr.n_ent ? (int)rr.p_single : -1; f.q = rr.n_ent ? (int)rr.q[rr.p_single] : 0;
                f.min_c = f.min_m = F_NONE;
                for (int k = 0; k < rr.len; ++k) {
                    const int oi = order[rr.start + k], x = pos + k;
                    if (!(flags[oi] & PF_PRE_OK)) continue;
                    fpod[2 * (size_t)x] = (int32_t)(pods->base[3 * (size_t)oi] >> cpu_shift);
                    fpod[2 * (size_t)x + 1] = (int32_t)(pods->base[3 * (size_t)oi + 1] >> mem_shift);
                }
                fr[r] = f;
                pos += f.nblk * 64;
            }
            blob_finish(B, fp_off, fr, n_runs);
        }
        lap("blob");
        // ---- the classic kernels' pod arrays: now, unless this is one nosplan_plan / nosplan_plan_resident call planning with the
        //      single-launch kernel (then only if that plan is rejected, see nosplan_reject_to_classic)
        lazy = fused_ok && h->in_plan;
        h->classic_staged = !lazy;
        h->lazy_valid = lazy; h->lazy_pods = *pods; h->lazy_pre = prefilter_ok; h->lazy_filter = filter_mode;
        if (!lazy) stage_classic_pods(h, pods, K);
        lap("classic stage");
        CU(h, cudaEventRecord(h->ev[9], h->stream));
        up_from = lazy ? L.fp_blob : L.pod_cpu;
        CU(h, cudaMemcpyAsync(A + up_from, S + up_from, L.upload_end - up_from, cudaMemcpyHostToDevice, h->stream));
        CU(h, cudaEventRecord(h->ev[10], h->stream));

    }

    // ---- device problem description, pod side
    pb.n_keys = n_keys; pb.n_runs = n_runs;
    pb.pod_cpu = (const long long*)(A + L.pod_cpu); pb.pod_mem = (const long long*)(A + L.pod_mem); pb.pod_eph = (const long long*)(A + L.pod_eph);
    pb.pod_key = (const int32_t*)(A + L.pod_key); pb.pod_flags = (const uint32_t*)(A + L.pod_flags);
    pb.pod_sptr = (const int32_t*)(A + L.pod_sptr); pb.pod_scol = (const int32_t*)(A + L.pod_scol); pb.pod_sqty = (const long long*)(A + L.pod_sqty);
    pb.runs = (const RunRec*)(A + L.runs);
    pb.run_chunk_off = (const int32_t*)(A + L.run_chunk_off);
    pb.cmin_cpu = (const long long*)(A + L.cmin_cpu); pb.cmin_mem = (const long long*)(A + L.cmin_mem); pb.cmin_eph = (const long long*)(A + L.cmin_eph);
    pb.n_chunks = n_chunks;
    pb.fast_eligible = single_profile_classes ? 1 : 0;
    pb.no_eph = any_eph == 0 ? 1 : 0;
    for (int p = 0; p < KP; ++p) {   // MIG M_p: most devices of profile p in any allowed geometry (MPS: derived on the device from gflags[1])
        int m = 0;
        if (mig && p < K) for (int r = 0; r < n_rows; ++r) m = std::max(m, (int)s->geom_counts[16 * (size_t)r + p]);
        pb.sat_m[p] = m;
    }
    pb.placed = (unsigned*)(A + L.placed);
    pb.entry_lack = (int32_t*)(A + L.entry_lack);
    pb.fp_blob = A + L.fp_blob;
    for (int k = 0; k < FP_NOFF; ++k) pb.fp_off[k] = fp_off[k];
    pb.fp_n_runs = fused_ok ? n_runs : 0; pb.fp_pad_pods = pad_pods; pb.fp_cpu_shift = cpu_shift; pb.fp_mem_shift = mem_shift; pb.fp_rset = rset;
    h->fused_ok = fused_ok; h->fused_pending = false; h->placed_dirty = false;
    lap("pods queued");
    h->h2d_bytes = (keep ? 0 : L.nodes_end) + (L.upload_end - up_from);
    h->uploaded = true;
    if (!keep) {                                             // what a later nosplan_plan_resident / nosplan_update_nodes needs
        h->res_snap = *s;
        h->res_pmem.assign(s->profile_mem_gb, s->profile_mem_gb + K); h->res_pgi.assign(s->profile_gi, s->profile_gi + K);
        h->res_model_off.assign(n_models > 0 ? s->model_geom_offset : nullptr, n_models > 0 ? s->model_geom_offset + n_models + 1 : nullptr);
        h->res_geom.assign(n_rows ? s->geom_counts : nullptr, n_rows ? s->geom_counts + 16 * (size_t)n_rows : nullptr);
        nosplan_snapshot& r = h->res_snap;
        r.profile_mem_gb = h->res_pmem.data(); r.profile_gi = h->res_pgi.data();
        r.model_geom_offset = h->res_model_off.data(); r.geom_counts = h->res_geom.data();
        r.node_gpu_count = r.gpu_model = r.gpu_index = r.gpu_memory_gb = nullptr; r.gpu_used = r.gpu_free = nullptr;
        r.gpu_used_mask = r.gpu_free_mask = nullptr; r.node_alloc = r.node_requested = nullptr;
        r.node_allowed_pods = r.node_num_pods = nullptr;
        const size_t offs[11] = {L.node_gpu_count, L.gpu_model, L.gpu_mem, L.gpu_used, L.gpu_free, L.gpu_umask, L.gpu_fmask,
                                 L.node_alloc, L.node_req, L.node_allowed, L.node_npods};
        std::memcpy(h->res_off, offs, sizeof offs);
        h->res_nodes_end = L.nodes_end;
        h->nodes_resident = true;
    }
    if (in_flight.armed) { in_flight.armed = false; CU(h, cudaStreamSynchronize(h->stream)); }   // nothing of the caller's is read after return
    return NOSPLAN_OK;
}

extern "C" {

int nosplan_upload(nosplan_t* h, const nosplan_snapshot* s, const nosplan_pods* pods, const uint8_t* prefilter_ok,
                   int32_t filter_mode) {
    return upload_impl(h, s, pods, prefilter_ok, filter_mode, false);
}

// Kernels that need the pods: tracker tables reset, k_tracker_init, node-parallel pass, verification / sequential sweep.
static int run_post(nosplan* h) {
    if (h->split_run) CU(h, cudaEventRecord(h->ev[11], h->stream));
    launch_tracker_init(h->pb, h->stream);
    if (h->kernel_events) CU(h, cudaEventRecord(h->ev[7], h->stream));
    CU(h, cudaGetLastError());
    // single GPU: where the node-parallel pass can run at all, its last CTA verifies it and k_plan_waves (which returns at
    // once when the pass stands) holds both fallbacks; other problems go straight to k_plan_sweep
    const bool work = h->pb.N > 0 && h->pb.P > 0;
    if (work) {
        DevProblem p2 = h->pb;
        p2.verify_inline = 1;
        CU(h, launch_parallel(p2, h->par_blocks, h->stream));
        if (h->kernel_events) CU(h, cudaEventRecord(h->ev[8], h->stream));
        CU(h, launch_waves(p2, h->wave_blocks, h->stream));
    } else {
        if (h->kernel_events) CU(h, cudaEventRecord(h->ev[8], h->stream));
        CU(h, launch_sweep(h->pb, h->d_cand_idx, h->stream));
    }
    CU(h, cudaEventRecord(h->ev[3], h->stream));
    h->launches = (h->pb.N > 0) + (h->pb.P > 0) + 1 + (work ? 1 : 0);
    h->ran = true;
    return NOSPLAN_OK;
}

static int run_classic(nosplan* h);

// The whole Plan() as one kernel (fused.cu).  Its verdict is read back by nosplan_resolve.
static int run_fused(nosplan* h) {
    if (h->placed_dirty) {                                   // a classic run indexed `placed` by tracker key
        CU(h, cudaMemsetAsync(h->pb.placed, 0, 4 * (size_t)(h->pb.fp_pad_pods >> 5), h->stream));
        h->placed_dirty = false;
    }
    CU(h, cudaEventRecord(h->ev[2], h->stream));
    flip_small_state(h);
    h->pb.slice_mode = 0; h->pb.node_lo = 0; h->pb.node_hi = h->pb.N; h->pb.whole_snapshot = 1;
    CU(h, launch_fused(h->pb, h->sms, h->stream));
    CU(h, cudaEventRecord(h->ev[3], h->stream));
    h->pre_done = false; h->split_run = false;
    h->launches = 1;
    h->fused_pending = true; h->last_fused = true; h->fused_reject = 0; h->ms_rejected = 0.f;
    h->rerun = run_classic;
    h->ran = true;
    return NOSPLAN_OK;
}

static int run_classic(nosplan* h) {
    h->pb.slice_mode = 0;
    int rc = run_pre(h, false);
    if (rc != NOSPLAN_OK) return rc;
    return run_post(h);
}

int nosplan_run(nosplan_t* h) {
    if (!h) return NOSPLAN_E_INVALID;
    if (!h->uploaded) return fail(h, NOSPLAN_E_INVALID, "nosplan_run before nosplan_upload");
    CU(h, cudaSetDevice(h->device));
    return h->fused_ok ? run_fused(h) : run_classic(h);
}

// ---- node-sharded Plan() over `world` ranks (one handle per rank, the same snapshot uploaded on each) --------------
int64_t nosplan_shard_exchange_bytes(nosplan_t* h) {
    if (!h || !h->uploaded) return -1;
    return 32 + 8 * ((((int64_t)h->pb.n_keys + 31) / 32 + 1) / 2);     // 4 u64 counters + bitmap words, 8-byte granular
}

int nosplan_shard_begin(nosplan_t* h, int32_t rank, int32_t world, void* d_exchange, int64_t exchange_bytes) {
    if (!h) return NOSPLAN_E_INVALID;
    if (!h->uploaded) return fail(h, NOSPLAN_E_INVALID, "nosplan_shard_begin before nosplan_upload");
    if (world < 1 || rank < 0 || rank >= world) return fail(h, NOSPLAN_E_INVALID, "rank %d of %d", rank, world);
    const int64_t need = nosplan_shard_exchange_bytes(h);
    if (!d_exchange || exchange_bytes < need) return fail(h, NOSPLAN_E_INVALID, "exchange buffer smaller than %lld bytes", (long long)need);
    CU(h, cudaSetDevice(h->device));
    h->pb.slice_mode = 0;
    int rc = run_pre(h, false);                             // prepass replicated: cheaper than exchanging the totals
    if (rc != NOSPLAN_OK) return rc;
    CU(h, cudaMemsetAsync(d_exchange, 0, (size_t)need, h->stream));
    launch_tracker_init(h->pb, h->stream);
    if (h->kernel_events) CU(h, cudaEventRecord(h->ev[7], h->stream));
    DevProblem shard = h->pb;                               // this rank's slice of the name-ordered node list
    shard.whole_snapshot = 0;
    const int chunk = (h->pb.N + world - 1) / world;        // equal-sized slices (the last may be short) so the rows all-gather
    shard.node_lo = std::min(h->pb.N, rank * chunk);
    shard.node_hi = std::min(h->pb.N, shard.node_lo + chunk);
    shard.acc = static_cast<unsigned long long*>(d_exchange);
    shard.placed = reinterpret_cast<unsigned*>(static_cast<unsigned char*>(d_exchange) + 32);
    CU(h, cudaGetLastError());
    CU(h, launch_parallel(shard, h->par_blocks, h->stream));
    if (h->kernel_events) CU(h, cudaEventRecord(h->ev[8], h->stream));
    h->launches = (h->pb.N > 0) + (h->pb.P > 0) + (h->pb.fast_eligible && h->pb.N > 0 && h->pb.P > 0);
    return NOSPLAN_OK;
}

// ---- slice mode: every rank uploaded ONLY its own slice of the name-ordered node list -------------------------------
}  // extern "C"

int nosplan_slice_begin_ex(nosplan* h, void* d_totals) {
    if (!h) return NOSPLAN_E_INVALID;
    if (!h->uploaded) return fail(h, NOSPLAN_E_INVALID, "nosplan_slice_begin before nosplan_upload");
    if (!d_totals) return fail(h, NOSPLAN_E_INVALID, "null totals buffer");
    CU(h, cudaSetDevice(h->device));
    h->pb.slice_mode = 1;
    int rc = run_pre(h, false);                             // cluster totals of the slice
    if (rc != NOSPLAN_OK) return rc;
    launch_slice_pack_totals(h->pb, d_totals, h->stream);
    CU(h, cudaGetLastError());
    h->launches = (h->pb.N > 0) + 1;
    return NOSPLAN_OK;
}

int nosplan_slice_mid_ex(nosplan* h, int32_t rank, int32_t world, const void* d_totals_gathered, void* d_exchange, int64_t exchange_bytes,
                         const PeerWait& wait) {
    if (!h) return NOSPLAN_E_INVALID;
    if (!h->uploaded || !h->pb.slice_mode) return fail(h, NOSPLAN_E_INVALID, "nosplan_slice_mid before nosplan_slice_begin");
    if (world < 1 || rank < 0 || rank >= world) return fail(h, NOSPLAN_E_INVALID, "rank %d of %d", rank, world);
    const int64_t need = nosplan_shard_exchange_bytes(h);
    if (!d_totals_gathered || !d_exchange || exchange_bytes < need) return fail(h, NOSPLAN_E_INVALID, "exchange buffer smaller than %lld bytes", (long long)need);
    CU(h, cudaSetDevice(h->device));
    launch_slice_combine_totals(h->pb, d_totals_gathered, world, wait, h->stream);   // cluster-wide totals on every rank
    CU(h, cudaMemsetAsync(d_exchange, 0, (size_t)need, h->stream));
    launch_tracker_init(h->pb, h->stream);
    if (h->kernel_events) CU(h, cudaEventRecord(h->ev[7], h->stream));
    DevProblem shard = h->pb;                               // the whole local snapshot IS this rank's slice
    shard.whole_snapshot = 0;
    shard.acc = static_cast<unsigned long long*>(d_exchange);
    shard.placed = reinterpret_cast<unsigned*>(static_cast<unsigned char*>(d_exchange) + 32);
    CU(h, cudaGetLastError());
    CU(h, launch_parallel(shard, h->par_blocks, h->stream));
    if (h->kernel_events) CU(h, cudaEventRecord(h->ev[8], h->stream));
    h->launches += 1 + (h->pb.P > 0) + (h->pb.fast_eligible && h->pb.N > 0 && h->pb.P > 0);
    return NOSPLAN_OK;
}

int nosplan_shard_end_ex(nosplan* h, const void* d_gathered, int32_t world, size_t stride, const PeerWait& wait) {
    if (!h) return NOSPLAN_E_INVALID;
    if (!h->uploaded) return fail(h, NOSPLAN_E_INVALID, "nosplan_shard_end before nosplan_upload");
    if (!d_gathered || world < 1) return fail(h, NOSPLAN_E_INVALID, "null gathered buffer");
    CU(h, cudaSetDevice(h->device));
    launch_shard_combine(h->pb, d_gathered, world, stride, wait, h->stream);
    DevProblem verify = h->pb;
    verify.whole_snapshot = 0;                               // final lacking vector comes from the OR-ed bitmap
    CU(h, launch_sweep(verify, h->d_cand_idx, h->stream));
    if (!verify.slice_mode) CU(h, launch_waves(verify, h->wave_blocks, h->stream));
    CU(h, cudaEventRecord(h->ev[3], h->stream));
    h->launches += 2;
    h->ran = true;
    return NOSPLAN_OK;
}

// Slice mode with both exchanges fused into the planning kernels (peer-memory transport, DevProblem::xf_*).
int nosplan_run_sliced_fused(nosplan* h, const PeerBox& box, unsigned epoch, unsigned char* d_exchange, size_t xb) {
    if (!h) return NOSPLAN_E_INVALID;
    if (!h->uploaded) return fail(h, NOSPLAN_E_INVALID, "nosplan_run_sliced before nosplan_upload");
    CU(h, cudaSetDevice(h->device));
    DevProblem& pb = h->pb;
    pb.slice_mode = 1;
    pb.xf_on = 1; pb.xf_epoch = epoch; pb.xf_box = box; pb.xf_exchange = d_exchange; pb.xf_xb = xb;
    int rc = run_pre(h, false);                             // reset, k_prepass (+ push of the slice's totals)
    if (rc == NOSPLAN_OK) {
        launch_tracker_init(pb, h->stream);                 // wait + fold the totals, clear the exchange block, tracker
        if (h->kernel_events) cudaEventRecord(h->ev[7], h->stream);
        DevProblem shard = pb;                              // the whole local snapshot IS this rank's slice
        shard.whole_snapshot = 0;
        shard.acc = reinterpret_cast<unsigned long long*>(d_exchange);
        shard.placed = reinterpret_cast<unsigned*>(d_exchange + 32);
        cudaError_t e = cudaGetLastError();
        if (e == cudaSuccess) e = launch_parallel(shard, h->par_blocks, h->stream);       // node-parallel pass (+ push of the exchange block)
        if (h->kernel_events) cudaEventRecord(h->ev[8], h->stream);
        DevProblem verify = pb;
        verify.whole_snapshot = 0;
        if (e == cudaSuccess) e = launch_sweep(verify, h->d_cand_idx, h->stream);   // wait + fold the blocks, verification
        if (e == cudaSuccess) e = cudaEventRecord(h->ev[3], h->stream);
        if (e != cudaSuccess) rc = fail(h, NOSPLAN_E_CUDA, "sliced Plan(): %s", cudaGetErrorString(e));
    }
    pb.xf_on = 0;                                           // the other drivers of this handle never see the fused fields
    if (rc != NOSPLAN_OK) return rc;
    h->launches = 4;
    h->ran = true;
    return NOSPLAN_OK;
}

// Slice mode as ONE kernel per rank plus one merged exchange (fused.cu): every rank plans its slice speculatively and ships
// [totals | largest free counts | reject bits | counters | placed bitmap] once; k_fused_verify folds the W blocks and
// verifies.  box == nullptr: the block is only assembled in d_exchange and the caller moves it (an NCCL all-gather).
int nosplan_run_sliced_fused_begin(nosplan* h, const PeerBox* box, unsigned epoch, unsigned char* d_exchange, size_t xb) {
    if (!h) return NOSPLAN_E_INVALID;
    if (!h->uploaded || !h->fused_ok) return fail(h, NOSPLAN_E_INVALID, "single-launch sliced Plan() is not applicable to the uploaded problem");
    if (xb < fused_exchange_bytes(h->pb)) return fail(h, NOSPLAN_E_INVALID, "exchange block smaller than %zu bytes", fused_exchange_bytes(h->pb));
    CU(h, cudaSetDevice(h->device));
    if (h->placed_dirty) {
        CU(h, cudaMemsetAsync(h->pb.placed, 0, 4 * (size_t)(h->pb.fp_pad_pods >> 5), h->stream));
        h->placed_dirty = false;
    }
    CU(h, cudaEventRecord(h->ev[2], h->stream));
    flip_small_state(h);
    DevProblem pb = h->pb;
    pb.slice_mode = 1; pb.node_lo = 0; pb.node_hi = pb.N; pb.whole_snapshot = 0;
    pb.xf_on = 1; pb.xf_epoch = epoch; pb.xf_exchange = d_exchange; pb.xf_xb = xb;
    if (box) pb.xf_box = *box; else { std::memset(&pb.xf_box, 0, sizeof pb.xf_box); }
    CU(h, launch_fused(pb, h->sms, h->stream));
    h->pb.slice_mode = 1;
    h->pre_done = false; h->split_run = false;
    h->launches = 1;
    return NOSPLAN_OK;
}

int nosplan_run_sliced_fused_end(nosplan* h, const void* blocks, size_t stride, int world, const PeerWait& wait, int (*rerun)(nosplan*)) {
    if (blocks) {                                            // NCCL transport: the gathered blocks are folded and verified by a kernel of its own
        CU(h, launch_fused_verify(h->pb, blocks, stride, world, wait, h->stream));
        h->launches += 1;
    }
    CU(h, cudaEventRecord(h->ev[3], h->stream));
    h->fused_pending = true; h->last_fused = true; h->fused_reject = 0; h->ms_rejected = 0.f;
    h->rerun = rerun;
    h->ran = true;
    return NOSPLAN_OK;
}

extern "C" {

int nosplan_slice_begin(nosplan_t* h, void* d_totals) { return nosplan_slice_begin_ex(h, d_totals); }

int nosplan_slice_mid(nosplan_t* h, int32_t rank, int32_t world, const void* d_totals_gathered, void* d_exchange, int64_t exchange_bytes) {
    return nosplan_slice_mid_ex(h, rank, world, d_totals_gathered, d_exchange, exchange_bytes, PeerWait{nullptr, 0, 0});
}

int nosplan_shard_end(nosplan_t* h, const void* d_gathered, int32_t world) {
    return nosplan_shard_end_ex(h, d_gathered, world, h && h->uploaded ? (size_t)nosplan_shard_exchange_bytes(h) : 0, PeerWait{nullptr, 0, 0});
}

int64_t nosplan_slice_chain_bytes(nosplan_t* h) {
    if (!h || !h->uploaded) return 0;
    return (int64_t)(sizeof(ChainHead) + 4 * (((size_t)h->pb.n_keys + 31) / 32));
}

int nosplan_slice_chain(nosplan_t* h, const void* d_state_in, void* d_state_out) {
    if (!h) return NOSPLAN_E_INVALID;
    if (!h->uploaded || !h->ran || !h->pb.slice_mode) return fail(h, NOSPLAN_E_INVALID, "nosplan_slice_chain before a sliced run of this problem");
    if (!d_state_out) return fail(h, NOSPLAN_E_INVALID, "null state buffer");
    CU(h, cudaSetDevice(h->device));
    DevProblem chain = h->pb;
    chain.chain_mode = 1;
    chain.chain_in = static_cast<const unsigned char*>(d_state_in);
    chain.chain_out = static_cast<unsigned char*>(d_state_out);
    chain.xf_on = 0;
    CU(h, launch_sweep(chain, h->d_cand_idx, h->stream));
    CU(h, cudaEventRecord(h->ev[3], h->stream));
    h->launches += 1;
    return NOSPLAN_OK;
}

int nosplan_device_outputs(nosplan_t* h, void** d_gpu_count, void** d_gpu_mask) {
    if (!h) return NOSPLAN_E_INVALID;
    if (!h->uploaded) return fail(h, NOSPLAN_E_INVALID, "nosplan_device_outputs before nosplan_upload");
    { const int rc = nosplan_resolve(h); if (rc != NOSPLAN_OK) return rc; }
    if (d_gpu_count) *d_gpu_count = h->pb.out_count;
    if (d_gpu_mask) *d_gpu_mask = h->pb.out_mask;
    return NOSPLAN_OK;
}

}  // extern "C"

static int nosplan_reject_to_classic(nosplan* h, int reasons, int lacking_at_start) {
    float t = 0.f;
    cudaEventElapsedTime(&t, h->ev[2], h->ev[3]);
    if (reasons == FREJ_HYPOTHESIS && lacking_at_start == 0 && !h->pb.slice_mode) {
        // nothing is lacking at t0: Plan() returns the state as it is (planner.go:80-83) -- k_prepass alone writes exactly that
        // (GetPartitioningState), and every counter is zero; no pod has to be staged for the classic kernels
        const int rc = run_pre(h, false);
        if (rc != NOSPLAN_OK) return rc;
        CU(h, cudaMemsetAsync(h->pb.stats, 0, sizeof(DevStats), h->stream));
        CU(h, cudaEventRecord(h->ev[3], h->stream));
        h->launches += 1;
        h->fused_reject = reasons; h->ms_rejected = t; h->prepass_only = true;
        h->ran = true;
        return NOSPLAN_OK;
    }
    if (!h->classic_staged) {
        // the call that uploaded this problem staged only what the single-launch kernel reads: the pods go through the general
        // path now (the node arrays are still in the arena), and the classic sequence plans
        if (!h->lazy_valid) return fail(h, NOSPLAN_E_INVALID, "the single-launch plan was rejected and the pods of that nosplan_plan call are gone: call nosplan_plan again");
        const nosplan_pods pods = h->lazy_pods;
        h->force_classic = true; h->in_plan = true;
        int rc0 = upload_impl(h, &h->res_snap, &pods, h->lazy_pre, h->lazy_filter, true);
        h->force_classic = false; h->in_plan = false;
        if (rc0 == NOSPLAN_OK) rc0 = h->pre_done ? run_post(h) : run_classic(h);
        h->fused_reject = reasons; h->ms_rejected = t;
        h->launches += 1;
        return rc0;
    }
    const int rc = h->rerun ? h->rerun(h) : run_classic(h);
    h->fused_reject = reasons; h->ms_rejected = t;
    h->launches += 1;
    return rc;
}

int nosplan_resolve(nosplan* h) {
    if (!h || !h->fused_pending) return NOSPLAN_OK;
    CU(h, cudaSetDevice(h->device));
    DevStats* ds = reinterpret_cast<DevStats*>(h->h_out + al(h->out_count_bytes) + al(h->out_mask_bytes));
    CU(h, cudaMemcpyAsync(ds, h->pb.stats, sizeof(DevStats), cudaMemcpyDeviceToHost, h->stream));
    CU(h, cudaStreamSynchronize(h->stream));
    h->fused_pending = false;
    if (ds->path != 2) return NOSPLAN_OK;
    return nosplan_reject_to_classic(h, ds->pad, ds->lacking_at_start);
}

extern "C" {

int nosplan_download(nosplan_t* h, nosplan_partitioning* out, nosplan_stats* stats) {
    if (!h) return NOSPLAN_E_INVALID;
    if (!h->ran) return fail(h, NOSPLAN_E_INVALID, "nosplan_download before nosplan_run");
    CU(h, cudaSetDevice(h->device));
    unsigned char* O = h->h_out;
    const size_t o_mask = al(h->out_count_bytes), o_stats = o_mask + al(h->out_mask_bytes);
    const bool want = out && h->out_count_bytes;
    if (want && (!out->gpu_count || !out->gpu_mask)) return fail(h, NOSPLAN_E_INVALID, "null output buffers");
    // page-locked caller buffers receive the plan straight from the copy engine; pageable ones go through pinned staging
    bool direct = false;
    if (want) {
        cudaPointerAttributes a1, a2;
        direct = cudaPointerGetAttributes(&a1, out->gpu_count) == cudaSuccess && a1.type == cudaMemoryTypeHost &&
                 cudaPointerGetAttributes(&a2, out->gpu_mask) == cudaSuccess && a2.type == cudaMemoryTypeHost;
        (void)cudaGetLastError();
    }
    for (;;) {
        CU(h, cudaEventRecord(h->ev[4], h->stream));
        if (h->out_count_bytes) {
            CU(h, cudaMemcpyAsync(direct ? (void*)out->gpu_count : (void*)O, h->pb.out_count, h->out_count_bytes, cudaMemcpyDeviceToHost, h->stream));
            CU(h, cudaMemcpyAsync(direct ? (void*)out->gpu_mask : (void*)(O + o_mask), h->pb.out_mask, h->out_mask_bytes, cudaMemcpyDeviceToHost, h->stream));
        }
        CU(h, cudaMemcpyAsync(O + o_stats, h->pb.stats, sizeof(DevStats), cudaMemcpyDeviceToHost, h->stream));
        CU(h, cudaEventRecord(h->ev[5], h->stream));
        CU(h, cudaStreamSynchronize(h->stream));
        if (!h->fused_pending) break;
        // single-launch plan: the kernel verified its own hypothesis; rejected -> the classic sequence on the same inputs
        h->fused_pending = false;
        DevStats ds; std::memcpy(&ds, O + o_stats, sizeof ds);
        if (ds.path != 2) break;
        const int rc2 = nosplan_reject_to_classic(h, ds.pad, ds.lacking_at_start);
        if (rc2 != NOSPLAN_OK) return rc2;
    }
    if (want && !direct) {
        std::memcpy(out->gpu_count, O, h->out_count_bytes);
        std::memcpy(out->gpu_mask, O + o_mask, h->out_mask_bytes);
    }
    if (stats) {
        DevStats ds; std::memcpy(&ds, O + o_stats, sizeof ds);
        std::memset(stats, 0, sizeof *stats);
        stats->candidates = ds.candidates; stats->visited_nodes = ds.visited_nodes; stats->geometry_evals = ds.geometry_evals;
        stats->placements = ds.placements; stats->candidate_nodes = ds.candidate_nodes;
        stats->lacking_at_start = ds.lacking_at_start; stats->lacking_at_end = ds.lacking_at_end;
        nosplan_phase_times(h, stats);
        cudaEventElapsedTime(&stats->ms_d2h, h->ev[4], h->ev[5]);
        stats->path = ds.path;
        stats->waves = (h->pb.slice_mode || h->last_fused) ? 0 : ds.pad;
        stats->kernel_launches = h->launches;
        stats->fused = h->last_fused ? 1 : 0;
        stats->fused_reject = h->fused_reject;
        stats->gate_replayed = (h->pb.slice_mode || h->last_fused) ? -1 : ds.gate_again - 1;
        stats->h2d_bytes = (int64_t)h->h2d_bytes;
        stats->d2h_bytes = (int64_t)(h->out_count_bytes + h->out_mask_bytes + sizeof(DevStats));
    }
    if (h->pb.slice_mode) {
        DevStats ds; std::memcpy(&ds, O + o_stats, sizeof ds);
        if (ds.pad == 2) return fail(h, NOSPLAN_E_CUDA, "peer exchange timed out: a rank of the sliced job did not deliver its block");
        if (ds.pad) return NOSPLAN_SLICED_NOT_APPLICABLE;
    }
    return NOSPLAN_OK;
}

int nosplan_plan(nosplan_t* h, const nosplan_snapshot* snap, const nosplan_pods* pods, const uint8_t* prefilter_ok,
                 int32_t filter_mode, nosplan_partitioning* out, nosplan_stats* stats) {
    if (!h) return NOSPLAN_E_INVALID;
    h->in_plan = true;                               // download() below synchronises before this call returns
    int rc = nosplan_upload(h, snap, pods, prefilter_ok, filter_mode);
    h->in_plan = false;
    if (rc != NOSPLAN_OK) { if (h->direct_h2d) cudaStreamSynchronize(h->stream); return rc; }
    rc = h->pre_done ? run_post(h) : nosplan_run(h);       // upload() already queued the node-only kernels
    if (rc != NOSPLAN_OK) { if (h->direct_h2d) cudaStreamSynchronize(h->stream); h->lazy_valid = false; return rc; }
    rc = nosplan_download(h, out, stats);            // a rejected single-launch plan is re-planned inside this call
    h->lazy_valid = false;
    return rc;
}

int32_t nosplan_compact_record_bytes(int32_t gpus_per_node, int32_t n_other) {
    if (gpus_per_node < 1 || gpus_per_node > NOSPLAN_MAX_GPUS || n_other < 0 || n_other > NOSPLAN_MAX_OTHER) return -1;
    return (128 + 16 * n_other + 24 * gpus_per_node + 15) & ~15;            // records stay 16-byte aligned
}

// nosplan_plan with the node side as one array of records: one DMA, k_expand_nodes writes the node sections of the arena, and
// from there on the call is nosplan_plan_resident on the snapshot just expanded.
int nosplan_plan_compact(nosplan_t* h, const nosplan_snapshot_compact* c, const nosplan_pods* pods, const uint8_t* prefilter_ok,
                         int32_t filter_mode, nosplan_partitioning* out, nosplan_stats* stats) {
    if (!h) return NOSPLAN_E_INVALID;
    if (!c || !pods) return fail(h, NOSPLAN_E_INVALID, "null snapshot or pods");
    const int K = c->n_profiles, N = c->n_nodes, G = c->gpus_per_node, n_other = c->n_other;
    if (c->kind != NOSPLAN_KIND_MIG && c->kind != NOSPLAN_KIND_MPS) return fail(h, NOSPLAN_E_INVALID, "unknown kind %d", c->kind);
    if (K < 0 || N < 0 || G < 1 || n_other < 0) return fail(h, NOSPLAN_E_INVALID, "negative size");
    if (K > 8) return fail(h, NOSPLAN_E_LIMIT, "%d profiles: the compact form holds 8 (use nosplan_plan)", K);
    if (G > NOSPLAN_MAX_GPUS) return fail(h, NOSPLAN_E_LIMIT, "%d GPUs per node > NOSPLAN_MAX_GPUS", G);
    if (n_other > NOSPLAN_MAX_OTHER) return fail(h, NOSPLAN_E_LIMIT, "%d other resources > NOSPLAN_MAX_OTHER", n_other);
    const int rec = nosplan_compact_record_bytes(G, n_other);
    if (c->record_bytes != rec) return fail(h, NOSPLAN_E_INVALID, "record_bytes %d, expected %d", c->record_bytes, rec);
    if (N > 0 && !c->records) return fail(h, NOSPLAN_E_INVALID, "null records");
    const bool mig = c->kind == NOSPLAN_KIND_MIG;
    const int n_models = mig ? c->n_models : 0;
    if (n_models < 0 || n_models > MAX_MODELS || n_models > 256) return fail(h, NOSPLAN_E_LIMIT, "%d GPU models", n_models);
    const int n_rows = (mig && n_models > 0) ? c->model_geom_offset[n_models] : 0;
    if (n_rows < 0 || n_rows > MAX_ROWS) return fail(h, NOSPLAN_E_LIMIT, "%d geometry rows > %d", n_rows, MAX_ROWS);
    const int R = 3 + K + n_other;
    CU(h, cudaSetDevice(h->device));
    h->nodes_resident = false; h->uploaded = h->ran = h->pre_done = false;
    Layout L; size_t o = 0;
    layout_nodes(L, o, n_models, n_rows, N, G, R, mig);
    const size_t rec_bytes = (size_t)N * rec, o_rec = L.nodes_end;
    cudaPointerAttributes at;
    const bool direct = rec_bytes >= (128u << 10) && cudaPointerGetAttributes(&at, c->records) == cudaSuccess && at.type == cudaMemoryTypeHost;
    (void)cudaGetLastError();
    int rc = ensure(h, o_rec + al(rec_bytes) + 4096, al(L.node_gpu_count) + (direct ? 0 : al(rec_bytes)) + 4096, 0);
    if (rc != NOSPLAN_OK) return rc;
    unsigned char* S = h->h_stage; unsigned char* A = h->d_arena;
    if (n_models > 0) std::memcpy(S + L.model_off, c->model_geom_offset, sizeof(int32_t) * (n_models + 1));
    else std::memset(S + L.model_off, 0, sizeof(int32_t));
    if (n_rows) std::memcpy(S + L.geom, c->geom_counts, 16 * (size_t)n_rows);
    CU(h, cudaMemcpyAsync(A, S, L.node_gpu_count, cudaMemcpyHostToDevice, h->stream));          // the two small tables
    if (rec_bytes) {
        if (direct) CU(h, cudaMemcpyAsync(A + o_rec, c->records, rec_bytes, cudaMemcpyHostToDevice, h->stream));
        else {
            std::memcpy(S + al(L.node_gpu_count), c->records, rec_bytes);
            CU(h, cudaMemcpyAsync(A + o_rec, S + al(L.node_gpu_count), rec_bytes, cudaMemcpyHostToDevice, h->stream));
        }
        (void)cudaStreamQuery(h->stream);
    }
    DevProblem pb; std::memset(&pb, 0, sizeof pb);
    pb.kind = c->kind; pb.K = K; pb.R = R; pb.n_other = n_other; pb.N = N; pb.G = G;
    pb.node_gpu_count = (const int32_t*)(A + L.node_gpu_count);
    pb.gpu_model = (const int32_t*)(A + L.gpu_model); pb.gpu_mem = (const int32_t*)(A + L.gpu_mem);
    pb.gpu_used = (const uint4*)(A + L.gpu_used); pb.gpu_free = (const uint4*)(A + L.gpu_free);
    pb.gpu_umask = (const uint16_t*)(A + L.gpu_umask); pb.gpu_fmask = (const uint16_t*)(A + L.gpu_fmask);
    pb.node_alloc = (const long long*)(A + L.node_alloc); pb.node_req = (const long long*)(A + L.node_req);
    pb.node_allowed_pods = (const int32_t*)(A + L.node_allowed); pb.node_num_pods = (const int32_t*)(A + L.node_npods);
    CU(h, cudaMemsetAsync(h->d_flag, 0xff, 8, h->stream));                 // [0] first bad node (INT_MAX-ish = none), [1] what was wrong
    launch_expand_nodes(pb, A + o_rec, rec, n_models, h->d_flag, h->stream);
    CU(h, cudaGetLastError());
    // the checks nosplan_upload makes per node are made by k_expand_nodes (the records are strided: a host pass over them costs
    // more than the DMA saves); its verdict travels to pinned host memory behind the kernel and is read after the call's sync
    CU(h, cudaMemcpyAsync(h->h_flag, h->d_flag, 8, cudaMemcpyDeviceToHost, h->stream));
    // what nosplan_plan_resident needs: shape and small tables on the host, the node sections where upload_impl expects them
    nosplan_snapshot& r = h->res_snap;
    std::memset(&r, 0, sizeof r);
    r.kind = c->kind; r.n_profiles = K; r.n_models = n_models; r.n_nodes = N; r.gpus_per_node = G; r.n_other = n_other;
    h->res_pmem.assign(c->profile_mem_gb, c->profile_mem_gb + K); h->res_pgi.assign(c->profile_gi, c->profile_gi + K);
    h->res_model_off.assign(n_models > 0 ? c->model_geom_offset : nullptr, n_models > 0 ? c->model_geom_offset + n_models + 1 : nullptr);
    h->res_geom.assign(n_rows ? c->geom_counts : nullptr, n_rows ? c->geom_counts + 16 * (size_t)n_rows : nullptr);
    r.profile_mem_gb = h->res_pmem.data(); r.profile_gi = h->res_pgi.data();
    r.model_geom_offset = h->res_model_off.data(); r.geom_counts = h->res_geom.data();
    const size_t offs[11] = {L.node_gpu_count, L.gpu_model, L.gpu_mem, L.gpu_used, L.gpu_free, L.gpu_umask, L.gpu_fmask,
                             L.node_alloc, L.node_req, L.node_allowed, L.node_npods};
    std::memcpy(h->res_off, offs, sizeof offs);
    h->res_nodes_end = L.nodes_end;
    h->nodes_resident = true;
    h->compact_h2d = L.node_gpu_count + rec_bytes;
    h->in_plan = true;
    rc = upload_impl(h, &h->res_snap, pods, prefilter_ok, filter_mode, true);
    h->in_plan = false;
    if (rc != NOSPLAN_OK) { cudaStreamSynchronize(h->stream); h->lazy_valid = false; return rc; }
    h->h2d_bytes += h->compact_h2d;
    rc = h->pre_done ? run_post(h) : nosplan_run(h);
    if (rc != NOSPLAN_OK) { cudaStreamSynchronize(h->stream); h->lazy_valid = false; return rc; }
    rc = nosplan_download(h, out, stats);
    h->lazy_valid = false;
    if (h->h_flag[0] != -1) {                                // a record was malformed: whatever was planned is void
        h->nodes_resident = false; h->uploaded = h->ran = false;
        return fail(h, NOSPLAN_E_INVALID, h->h_flag[1] == 1 ? "node %d: gpu count out of range" : "node %d: GPU model id out of range", h->h_flag[0]);
    }
    return rc;
}

int nosplan_plan_resident(nosplan_t* h, const nosplan_pods* pods, const uint8_t* prefilter_ok, int32_t filter_mode,
                          nosplan_partitioning* out, nosplan_stats* stats) {
    if (!h) return NOSPLAN_E_INVALID;
    if (!h->nodes_resident) return fail(h, NOSPLAN_E_INVALID, "nosplan_plan_resident: no snapshot resident on this handle (call nosplan_plan or nosplan_upload first)");
    h->in_plan = true;
    int rc = upload_impl(h, &h->res_snap, pods, prefilter_ok, filter_mode, true);
    h->in_plan = false;
    if (rc != NOSPLAN_OK) return rc;
    rc = h->pre_done ? run_post(h) : nosplan_run(h);
    if (rc != NOSPLAN_OK) { h->lazy_valid = false; return rc; }
    rc = nosplan_download(h, out, stats);
    h->lazy_valid = false;
    return rc;
}

int nosplan_upload_pods(nosplan_t* h, const nosplan_pods* pods, const uint8_t* prefilter_ok, int32_t filter_mode) {
    if (!h) return NOSPLAN_E_INVALID;
    if (!h->nodes_resident) return fail(h, NOSPLAN_E_INVALID, "nosplan_upload_pods: no snapshot resident on this handle (call nosplan_plan or nosplan_upload first)");
    return upload_impl(h, &h->res_snap, pods, prefilter_ok, filter_mode, true);
}

int nosplan_update_nodes(nosplan_t* h, int32_t first, const nosplan_snapshot* rows) {
    if (!h) return NOSPLAN_E_INVALID;
    if (!h->nodes_resident) return fail(h, NOSPLAN_E_INVALID, "nosplan_update_nodes: no snapshot resident on this handle");
    if (!rows) return fail(h, NOSPLAN_E_INVALID, "null rows");
    const nosplan_snapshot& r = h->res_snap;
    const int N = r.n_nodes, G = r.gpus_per_node, R = 3 + r.n_profiles + r.n_other, n = rows->n_nodes;
    if (rows->kind != r.kind || rows->gpus_per_node != G || rows->n_profiles != r.n_profiles || rows->n_other != r.n_other)
        return fail(h, NOSPLAN_E_INVALID, "nosplan_update_nodes: rows of another shape (kind, gpus_per_node, n_profiles, n_other) than the resident snapshot");
    if (first < 0 || n < 0 || (long long)first + n > N) return fail(h, NOSPLAN_E_INVALID, "nosplan_update_nodes: nodes [%d, %d) outside the resident snapshot of %d", first, first + n, N);
    if (n == 0) return NOSPLAN_OK;
    const bool mig = r.kind == NOSPLAN_KIND_MIG;
    for (int i = 0; i < n; ++i) {                            // the checks nosplan_upload makes per node
        const int ng = rows->node_gpu_count[i];
        if (ng < 0 || ng > G) return fail(h, NOSPLAN_E_INVALID, "node %d: gpu count out of range", first + i);
        for (int j = 0; mig && j < ng; ++j)
            if ((unsigned)rows->gpu_model[(size_t)i * G + j] >= (unsigned)r.n_models) return fail(h, NOSPLAN_E_INVALID, "node %d: GPU model id out of range", first + i);
    }
    CU(h, cudaSetDevice(h->device));
    unsigned char* A = h->d_arena;
    const void* src[11] = {rows->node_gpu_count, mig ? (const void*)rows->gpu_model : nullptr, mig ? nullptr : (const void*)rows->gpu_memory_gb,
                           rows->gpu_used, rows->gpu_free, rows->gpu_used_mask, rows->gpu_free_mask, rows->node_alloc, rows->node_requested,
                           rows->node_allowed_pods, rows->node_num_pods};
    const size_t stride[11] = {4, 4 * (size_t)G, 4 * (size_t)G, 16 * (size_t)G, 16 * (size_t)G, 2 * (size_t)G, 2 * (size_t)G,
                               8 * (size_t)R, 8 * (size_t)R, 4, 4};
    for (int k = 0; k < 11; ++k) {
        if (!src[k]) continue;
        CU(h, cudaMemcpyAsync(A + h->res_off[k] + stride[k] * (size_t)first, src[k], stride[k] * (size_t)n, cudaMemcpyHostToDevice, h->stream));
    }
    CU(h, cudaStreamSynchronize(h->stream));                 // nothing of the caller's is read after return
    h->ran = false;                                          // the plan of the previous call is stale
    return NOSPLAN_OK;
}

int nosplan_quota_recompute(nosplan_t* h, const nosplan_quotas* quotas, const nosplan_quota_pods* running,
                            const nosplan_quota_pods* pending, nosplan_quota_out* out, nosplan_stats* stats) {
    if (!h) return NOSPLAN_E_INVALID;
    if (!h->qctx) h->qctx = nosplan_quota_ctx_new();
    float a = 0, b = 0, c = 0; int launches = 0; long long hb = 0, db = 0;
    const int rc = nosplan_quota_recompute_impl(h->device, h->stream, h->qctx, &h->err, quotas, running, pending, out, &a, &b, &c,
                                                &launches, &hb, &db, nullptr, nullptr);
    if (stats) {
        std::memset(stats, 0, sizeof *stats);
        stats->ms_h2d = a; stats->ms_kernels = b; stats->ms_d2h = c; stats->kernel_launches = launches;
        stats->h2d_bytes = hb; stats->d2h_bytes = db;
    }
    return rc;
}

int nosplan_quota_state_begin(nosplan_t* h, const nosplan_quotas* quotas, int32_t pod_capacity) {
    if (!h) return NOSPLAN_E_INVALID;
    if (!h->ectx) h->ectx = nosplan_quota_ctx_new();
    return nosplan_quota_state_begin_impl(h->device, h->stream, h->ectx, &h->err, quotas, pod_capacity);
}

int nosplan_quota_state_apply(nosplan_t* h, const nosplan_quota_events* events, nosplan_quota_out* out, int32_t* n_applied,
                              nosplan_stats* stats) {
    if (!h) return NOSPLAN_E_INVALID;
    if (!h->ectx) { h->err = "nosplan_quota_state_apply: no quota state resident on this handle (call nosplan_quota_state_begin first)"; return NOSPLAN_E_INVALID; }
    if (!h->qctx) h->qctx = nosplan_quota_ctx_new();
    float k = 0; long long hb = 0, db = 0;
    const int rc = nosplan_quota_state_apply_impl(h->device, h->stream, h->ectx, h->qctx, &h->err, events, out, n_applied, &k, &hb, &db);
    if (stats) {
        std::memset(stats, 0, sizeof *stats);
        stats->ms_kernels = k; stats->h2d_bytes = hb; stats->d2h_bytes = db; stats->kernel_launches = 2;
    }
    return rc;
}

int nosplan_quota_state_prefilter(nosplan_t* h, const nosplan_quota_pods* pending, uint8_t* verdict, nosplan_stats* stats) {
    if (!h) return NOSPLAN_E_INVALID;
    if (!h->ectx) { h->err = "nosplan_quota_state_prefilter: no quota state resident on this handle (call nosplan_quota_state_begin first)"; return NOSPLAN_E_INVALID; }
    if (!h->qctx) h->qctx = nosplan_quota_ctx_new();
    float k = 0;
    const int rc = nosplan_quota_state_prefilter_impl(h->device, h->stream, h->ectx, h->qctx, &h->err, pending, verdict, &k);
    if (stats) { std::memset(stats, 0, sizeof *stats); stats->ms_kernels = k; stats->kernel_launches = 2; }
    return rc;
}

int nosplan_quota_label_changes(nosplan_t* h, const nosplan_quotas* quotas, const nosplan_quota_pods* running,
                                const nosplan_label_changes* changes, nosplan_quota_out* out, nosplan_stats* stats) {
    if (!h) return NOSPLAN_E_INVALID;
    if (!changes) { h->err = "null changes"; return NOSPLAN_E_INVALID; }
    if (!h->qctx) h->qctx = nosplan_quota_ctx_new();
    float a = 0, b = 0, c = 0; int launches = 0; long long hb = 0, db = 0;
    const int rc = nosplan_quota_recompute_impl(h->device, h->stream, h->qctx, &h->err, quotas, running, nullptr, out, &a, &b, &c,
                                                &launches, &hb, &db, changes, nullptr);
    if (stats) {
        std::memset(stats, 0, sizeof *stats);
        stats->ms_h2d = a; stats->ms_kernels = b; stats->ms_d2h = c; stats->kernel_launches = launches;
        stats->h2d_bytes = hb; stats->d2h_bytes = db;
    }
    return rc;
}

int nosplan_parse_annotations(nosplan_t* h, const nosplan_annotations* in, nosplan_parsed_gpus* out, nosplan_stats* stats) {
    if (!h) return NOSPLAN_E_INVALID;
    if (!h->qctx) h->qctx = nosplan_quota_ctx_new();
    float a = 0, b = 0, c = 0; long long hb = 0, db = 0;
    const int rc = nosplan_parse_annotations_impl(h->device, h->stream, h->qctx, &h->err, in, out, &a, &b, &c, &hb, &db);
    if (stats) {
        std::memset(stats, 0, sizeof *stats);
        stats->ms_h2d = a; stats->ms_kernels = b; stats->ms_d2h = c; stats->kernel_launches = 2;
        stats->h2d_bytes = hb; stats->d2h_bytes = db;
    }
    return rc;
}

int nosplan_quota_victims(nosplan_t* h, const nosplan_quotas* quotas, const nosplan_victim_query* query, uint8_t* victim,
                          nosplan_stats* stats) {
    if (!h) return NOSPLAN_E_INVALID;
    if (!h->qctx) h->qctx = nosplan_quota_ctx_new();
    float ms = 0; int launches = 0;
    const int rc = nosplan_quota_victims_impl(h->device, h->stream, h->qctx, &h->err, quotas, query, victim, &ms, &launches);
    if (stats) {
        std::memset(stats, 0, sizeof *stats);
        stats->ms_kernels = ms; stats->kernel_launches = launches;
    }
    return rc;
}

int nosplan_preempt_dryrun(nosplan_t* h, const nosplan_quotas* quotas, const nosplan_preempt_query* query, uint8_t* victim,
                           int32_t* node_status, int32_t* num_violating, nosplan_stats* stats) {
    if (!h) return NOSPLAN_E_INVALID;
    if (!h->pctx) h->pctx = nosplan_quota_ctx_new();
    float a = 0, b = 0, c = 0; int launches = 0; long long hb = 0, db = 0;
    const int rc = nosplan_preempt_dryrun_impl(h->device, h->stream, h->pctx, &h->err, quotas, query, victim, node_status, num_violating,
                                               &a, &b, &c, &launches, &hb, &db);
    if (stats) {
        std::memset(stats, 0, sizeof *stats);
        stats->ms_h2d = a; stats->ms_kernels = b; stats->ms_d2h = c; stats->kernel_launches = launches;
        stats->h2d_bytes = hb; stats->d2h_bytes = db;
    }
    return rc;
}

int nosplan_preempt_dryrun_again(nosplan_t* h, const nosplan_preempt_query* preemptor, uint8_t* victim, int32_t* node_status,
                                 int32_t* num_violating, nosplan_stats* stats) {
    if (!h) return NOSPLAN_E_INVALID;
    if (!h->pctx) { h->err = "nosplan_preempt_dryrun_again: no cluster state resident on this handle (call nosplan_preempt_dryrun first)"; return NOSPLAN_E_INVALID; }
    float b = 0, c = 0; int launches = 0; long long db = 0;
    const int rc = nosplan_preempt_dryrun_again_impl(h->device, h->stream, h->pctx, &h->err, preemptor, victim, node_status, num_violating,
                                                     &b, &c, &launches, &db);
    if (stats) {
        std::memset(stats, 0, sizeof *stats);
        stats->ms_kernels = b; stats->ms_d2h = c; stats->kernel_launches = launches; stats->d2h_bytes = db;
    }
    return rc;
}

}  // extern "C"
