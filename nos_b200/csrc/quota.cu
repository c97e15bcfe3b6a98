// ElasticQuota accounting kernels (sm_100a) and nosplan_quota_recompute.
//
//   k_quota_used       reserveResource of every running pod into its quota row (elasticquotainfo.go:194-200, 276-292), with
//                      the gpu-memory scalar of pkg/gpu/util/resource.go:60-86 computed on the fly.  Half a warp per pod
//                      (lane = resource column, so consecutive pods are one coalesced read), block-private int64 table in
//                      shared memory, one global atomic per touched (quota, column) per block.
//   k_quota_aggregate  getAggregatedMin / Used / Overquotas over the namespace -> quota map (a composite quota counts once
//                      per namespace, :68-72, 145-174) and GetGuaranteedOverquotas of every quota (:81-119) in IEEE double
//                      with explicit round-to-nearest mul/div (no FMA contraction), floor, and amd64 int64() semantics.
//   k_quota_prefilter  the quota part of CapacityScheduling.PreFilter per pending pod (capacity_scheduling.go:190-278).
//   k_label_*          the operator's in-quota / over-quota labelling (internal/controllers/elasticquota/elasticquota.go:
//                      38-118): pods ordered by (quota, creationTimestamp, priority) with three stable LSD radix passes
//                      (CUB), ties resolved with the reference's request / name comparator, then a per-quota running sum
//                      (block scan per resource column) compared with Min.
#include <climits>
#include <cstdio>
#include <cstring>
#include <string>
#include <vector>

#include <cooperative_groups.h>
#include <cub/device/device_radix_sort.cuh>
#include <cub/device/device_select.cuh>

#include "nosplan_dev.cuh"

namespace nsp {

constexpr int QR = NOSPLAN_QUOTA_MAX_RES;
constexpr unsigned SCALAR_BITS = 0x7ffffff0u;

struct DevQuota {
    int E, R, NS, gm, Pr, Pp;
    const int32_t* ns_quota;
    const long long *qmin, *qmax, *used_init;
    const uint32_t *min_mask, *max_mask, *used_mask_in;
    const uint8_t* max_enforced;
    long long coeff[QR];
    // running pods
    const int32_t* r_ns; const long long* r_req; const uint32_t* r_mask;
    const int32_t* r_prio; const long long* r_ts; const int32_t* r_rank;
    // pending pods
    const int32_t* p_ns; const long long* p_req; const uint32_t* p_mask;
    // outputs / work
    long long* used; uint32_t* used_mask;
    long long *agg_min, *agg_used, *agg_over; uint32_t* agg_masks;
    long long* guaranteed; uint8_t* prefilter; uint8_t* over_quota; long long* operator_used; int32_t* ambiguous;
    // labelling work
    unsigned long long *key_a, *key_b; int32_t *val_a, *val_b; int32_t* seg_count; int32_t* seg_off;
    int32_t* chunk_off;        // [E+1] first 256-pod chunk of each quota's segment
    long long* chunk_sum;      // [chunks][QR] column sums of a chunk, then (in place) the running sum before the chunk
    int n_chunk_bound;
    long long fuse_ts_min, fuse_prio_min; int fuse_tbits, fuse_pbits;   // single-sort labelling (k_label_keys pass 1)
    // labelling by buckets (k_label_scatter / k_label_segments): quota e owns buckets [b_first[e], b_first[e + 1]), a pod's bucket
    // inside its quota is (creationTimestamp - min) >> b_shift[e]; per bucket: quota row, column sums, "sums published" flag
    int n_buckets; const int32_t *b_first, *b_shift, *b_quota; long long* seg_sum; int32_t *seg_flag, *seg_ticket, *seg_overflow;
    // label changes (patchCapacityInfoIfDifferent): per sorted position, pod index << 1 | over-quota and "differs from the label on the pod"
    const uint8_t* cur_label; int32_t* chg_packed; uint8_t* chg_flag;
};

__device__ __forceinline__ long long go_int64(double x) {       // int64(f) as compiled for amd64 (CVTTSD2SQ)
    if (!(x == x) || x >= 9223372036854775808.0 || x < -9223372036854775808.0) return LLONG_MIN;
    return (long long)x;
}
__device__ __forceinline__ double approx_f64(int c, long long v) {   // Quantity.AsApproximateFloat64 (cpu: milli * Pow10(-3))
    return c == 0 ? __dmul_rn((double)v, 1e-3) : (double)v;
}

// quota request of pod i, column `c` held by lane c of a 16-lane segment; returns the request mask (segment-uniform)
__device__ __forceinline__ long long pod_request16(const DevQuota& q, const long long* req, const uint32_t* mask, int i, int c,
                                                   unsigned segmask, uint32_t& rmask) {
    long long v = (c < q.R && c != q.gm) ? req[(size_t)i * q.R + c] : 0;
    long long g = (c < q.R) ? q.coeff[c] * v : 0;
#pragma unroll
    for (int o = 8; o; o >>= 1) g += __shfl_xor_sync(segmask, g, o, 16);
    uint32_t m;
    if (mask) m = mask[i] & SCALAR_BITS;
    else {
        m = (c >= 4 && v != 0) ? (1u << c) : 0;
#pragma unroll
        for (int o = 8; o; o >>= 1) m |= __shfl_xor_sync(segmask, m, o, 16);
    }
    if (q.gm >= 0) { m |= 1u << q.gm; if (c == q.gm) v = g; }
    rmask = m;
    return v;
}

__global__ void __launch_bounds__(256) k_quota_used(DevQuota q, int use_smem) {
    extern __shared__ long long s_tab[];       // [E*R] when use_smem
    const int tid = threadIdx.x, c = tid & 15, sub = (tid >> 4) & 1;
    const unsigned segmask = 0xffffu << (16 * sub);
    if (use_smem) {
        for (int x = tid; x < q.E * q.R; x += blockDim.x) s_tab[x] = 0;
        __syncthreads();
    }
    const int pods_per_block = blockDim.x / 16;
    for (int i0 = blockIdx.x * pods_per_block; i0 < q.Pr; i0 += gridDim.x * pods_per_block) {
        const int i = i0 + (tid >> 4);
        if (i < q.Pr) {                                  // whole 16-lane segment takes the same branch
            const int ns = q.r_ns[i];
            const int e = (ns >= 0 && ns < q.NS) ? q.ns_quota[ns] : -1;
            uint32_t rmask;
            const long long v = pod_request16(q, q.r_req, q.r_mask, i, c, segmask, rmask);
            if (e >= 0) {
                const bool take = c < q.R && (c == 0 || c == 1 || (c >= 4 && (rmask >> c & 1)));   // Memory, MilliCPU, scalars only
                if (take && v != 0) {
                    if (use_smem) atomicAdd((unsigned long long*)&s_tab[e * q.R + c], (unsigned long long)v);
                    else atomicAdd((unsigned long long*)&q.used[(size_t)e * q.R + c], (unsigned long long)v);
                }
                if (c == 0 && (rmask & ~q.used_mask[e])) atomicOr(&q.used_mask[e], rmask);
            }
        }
    }
    if (use_smem) {
        __syncthreads();
        for (int x = tid; x < q.E * q.R; x += blockDim.x)
            if (s_tab[x] != 0) atomicAdd((unsigned long long*)&q.used[x], (unsigned long long)s_tab[x]);
    }
}

// Event-driven accounting on a resident state: AddPodIfNotPresent / DeletePodIfPresent (elasticquotainfo.go:276-310) for a batch of
// pod events.  Half a warp per event (lane = resource column); the first lane tests and flips the pod's bit in the set of
// accounted pods (a pod occurs at most once per batch, checked on the host, so the events of a batch commute: int64 sums),
// then reserveResource / unreserveResource (:194-208) on Memory, MilliCPU and the scalars the request names.
struct DevEvents {
    int n;
    const uint8_t* op; const int32_t* pod_id; const int32_t* ns; const long long* req; const uint32_t* mask;
    unsigned* present;         // [capacity / 32] accounted pods
    int32_t* applied;          // [1] events that changed the state
};
__global__ void __launch_bounds__(256) k_quota_events(DevQuota q, DevEvents ev) {
    const int tid = threadIdx.x, c = tid & 15, sub = (tid >> 4) & 1;
    const unsigned segmask = 0xffffu << (16 * sub);
    const int i = blockIdx.x * (blockDim.x / 16) + (tid >> 4);
    if (i >= ev.n) return;                                   // whole 16-lane segment
    const int ns = ev.ns[i];
    const int e = (ns >= 0 && ns < q.NS) ? q.ns_quota[ns] : -1;
    uint32_t rmask;
    const long long v = pod_request16(q, ev.req, ev.mask, i, c, segmask, rmask);
    if (e < 0) return;                                       // namespace without quota: no ElasticQuotaInfo to account in
    const bool add = ev.op[i] == NOSPLAN_QUOTA_EVENT_ADD;
    int go = 0;
    if (c == 0) {
        const int id = ev.pod_id[i];
        const unsigned bit = 1u << (id & 31);
        const unsigned old = add ? atomicOr(&ev.present[id >> 5], bit) : atomicAnd(&ev.present[id >> 5], ~bit);
        go = add ? !(old & bit) : (old & bit) != 0;
        if (go) atomicAdd(ev.applied, 1);
    }
    go = __shfl_sync(segmask, go, 16 * sub);
    if (!go) return;
    const bool take = c < q.R && (c == 0 || c == 1 || (c >= 4 && (rmask >> c & 1)));
    if (take && v != 0) atomicAdd((unsigned long long*)&q.used[(size_t)e * q.R + c], (unsigned long long)(add ? v : -v));
    if (c == 0 && (rmask & ~q.used_mask[e])) atomicOr(&q.used_mask[e], rmask);
}

constexpr int AGG_THREADS = 1024, AGG_ROWS = AGG_THREADS / 16;   // 64 namespaces per CTA, one (namespace, column) pair per thread
// Column sums over the namespace -> quota map, many CTAs: every CTA takes 64 namespaces (all loads of a namespace's chain in
// flight at once), reduces them in shared memory and adds its 3 x 16 partial sums to the global aggregates; the CTA that
// finishes last (agg_masks[3] counts them) then derives GetGuaranteedOverquotas of every quota from the complete sums.  The
// aggregates and the counter are zeroed by launch_quota_aggregate in front of the kernel.
__global__ void __launch_bounds__(AGG_THREADS) k_quota_aggregate(DevQuota q) {
    __shared__ long long s_min[AGG_ROWS][QR], s_used[AGG_ROWS][QR], s_over[AGG_ROWS][QR];
    __shared__ unsigned s_m[3];
    __shared__ long long a_min[QR], a_over[QR];
    __shared__ int s_last;
    const int tid = threadIdx.x, c = tid & 15, j = tid >> 4;
    if (tid < 3) s_m[tid] = 0;
    __syncthreads();
    long long mn = 0, us = 0, ov = 0;
    unsigned mm = 0, um = 0, om = 0;
    for (int ns = blockIdx.x * AGG_ROWS + j; ns < q.NS; ns += gridDim.x * AGG_ROWS) {
        const int e = q.ns_quota[ns];
        if (e < 0) continue;
        const bool has_min = (q.min_mask[e] & NOSPLAN_QUOTA_NONNIL) != 0;
        if (c < q.R) {
            const long long m = has_min ? q.qmin[(size_t)e * q.R + c] : 0, u = q.used[(size_t)e * q.R + c];
            mn += m; us += u;
            ov += max(0ll, m - u);                          // SubtractNonNegative, then Sum (:145-152)
        }
        if (c == 0) {
            const unsigned mk = has_min ? (q.min_mask[e] & SCALAR_BITS) : 0;
            mm |= mk; um |= q.used_mask[e]; om |= mk | q.used_mask[e];
        }
    }
    s_min[j][c] = mn; s_used[j][c] = us; s_over[j][c] = ov;
    if (c == 0 && (mm | um | om)) { atomicOr(&s_m[0], mm); atomicOr(&s_m[1], um); atomicOr(&s_m[2], om); }
    __syncthreads();
    if (tid < q.R) {
        long long a = 0, b = 0, d = 0;
        for (int k = 0; k < AGG_ROWS; ++k) { a += s_min[k][tid]; b += s_used[k][tid]; d += s_over[k][tid]; }
        if (a) atomicAdd((unsigned long long*)&q.agg_min[tid], (unsigned long long)a);
        if (b) atomicAdd((unsigned long long*)&q.agg_used[tid], (unsigned long long)b);
        if (d) atomicAdd((unsigned long long*)&q.agg_over[tid], (unsigned long long)d);
    }
    if (tid < 3 && s_m[tid]) atomicOr(&q.agg_masks[tid], s_m[tid]);
    __threadfence();
    __syncthreads();
    if (tid == 0) s_last = atomicAdd(&q.agg_masks[3], 1u) == gridDim.x - 1;
    __syncthreads();
    if (!s_last) return;
    __threadfence();
    if (tid < QR) { a_min[tid] = tid < q.R ? __ldcg(&q.agg_min[tid]) : 0; a_over[tid] = tid < q.R ? __ldcg(&q.agg_over[tid]) : 0; }
    if (tid == 0) s_m[2] = __ldcg(&q.agg_masks[2]);
    __syncthreads();
    const unsigned omask = s_m[2];
    for (int x = tid; x < q.E * q.R; x += blockDim.x) {     // GetGuaranteedOverquotas of every quota (:81-119)
        const int e = x / q.R, cc = x - e * q.R;
        const bool has_min = (q.min_mask[e] & NOSPLAN_QUOTA_NONNIL) != 0;
        const bool in_min = has_min && (cc < 4 || (q.min_mask[e] >> cc & 1));
        double pct = 0.0;
        if (in_min && a_min[cc] > 0) pct = __ddiv_rn(approx_f64(cc, q.qmin[x]), approx_f64(cc, a_min[cc]));
        const bool in_res = cc < 4 || (omask >> cc & 1);
        q.guaranteed[x] = in_res ? go_int64(floor(__dmul_rn((double)a_over[cc], pct))) : 0;
    }
    if (tid == 0) q.agg_masks[3] = 0;
}

static cudaError_t launch_quota_aggregate(const DevQuota& q, cudaStream_t st) {
    // agg_min | agg_used | agg_over | agg_masks (+ the CTA counter in its 4th word) are consecutive sections of the arena
    const size_t bytes = (size_t)(reinterpret_cast<unsigned char*>(q.agg_masks) + 16 - reinterpret_cast<unsigned char*>(q.agg_min));
    cudaError_t e = cudaMemsetAsync(q.agg_min, 0, bytes, st);
    if (e != cudaSuccess) return e;
    int blocks = (q.NS + AGG_ROWS - 1) / AGG_ROWS;
    if (blocks < 1) blocks = 1;
    if (blocks > 148) blocks = 148;
    k_quota_aggregate<<<blocks, AGG_THREADS, 0, st>>>(q);
    return cudaGetLastError();
}

__global__ void __launch_bounds__(256) k_quota_prefilter(DevQuota q) {
    const int tid = threadIdx.x, c = tid & 15, sub = (tid >> 4) & 1;
    const unsigned segmask = 0xffffu << (16 * sub);
    const int i = blockIdx.x * (blockDim.x / 16) + (tid >> 4);
    if (i >= q.Pp) return;
    const int ns = q.p_ns[i];
    const int e = (ns >= 0 && ns < q.NS) ? q.ns_quota[ns] : -1;
    uint32_t rmask;
    const long long v = pod_request16(q, q.p_req, q.p_mask, i, c, segmask, rmask);
    uint8_t code = NOSPLAN_PREFILTER_SUCCESS;
    if (e >= 0) {
        // sumGreaterThan (:319-339): cpu and memory always, scalars only if present in y
        bool over_max = false, over_min = false;
        if (c < q.R && c != 2 && c != 3) {
            const bool in_max = c < 2 || (q.max_mask[e] >> c & 1);
            over_max = in_max && v + q.used[(size_t)e * q.R + c] > q.qmax[(size_t)e * q.R + c];
            const bool in_min = c < 2 || (q.agg_masks[0] >> c & 1);
            over_min = in_min && q.agg_used[c] + v > q.agg_min[c];
        }
        const unsigned bm = __ballot_sync(segmask, over_max) & segmask, bn = __ballot_sync(segmask, over_min) & segmask;
        if (q.max_enforced[e] && bm) code = NOSPLAN_PREFILTER_OVER_MAX;
        else if (bn) code = NOSPLAN_PREFILTER_OVER_MIN;
    }
    if (c == 0) q.prefilter[i] = code;
}

// ---------------------------------------------------------------- operator labelling
// pods per quota: block-private histogram in shared memory (a popular quota would otherwise serialise thousands of atomics on
// one word), flushed with one global atomic per quota the block saw; quota tables too large for it count with
// warp-aggregated atomics
constexpr int SEG_HIST = 1024;      // larger quota tables: zeroing and flushing the histogram per block would cost more than it saves
__device__ __forceinline__ void seg_hist_begin(const DevQuota& q, int* s_hist) {
    if (q.E <= SEG_HIST) { for (int k = threadIdx.x; k < q.E; k += blockDim.x) s_hist[k] = 0; }
    __syncthreads();
}
__device__ __forceinline__ void seg_hist_add(const DevQuota& q, int* s_hist, int e, bool valid) {
    if (q.E <= SEG_HIST) { if (valid && e >= 0) atomicAdd(&s_hist[e], 1); return; }
    const unsigned act = __ballot_sync(0xffffffffu, valid && e >= 0);
    if (valid && e >= 0) {
        const unsigned peers = __match_any_sync(act, e);
        if ((threadIdx.x & 31) == __ffs(peers) - 1) atomicAdd(&q.seg_count[e], __popc(peers));
    }
}
__device__ __forceinline__ void seg_hist_end(const DevQuota& q, int* s_hist) {
    __syncthreads();
    if (q.E <= SEG_HIST)
        for (int k = threadIdx.x; k < q.E; k += blockDim.x) { const int n = s_hist[k]; if (n) atomicAdd(&q.seg_count[k], n); }
}

__global__ void __launch_bounds__(256) k_label_keys(DevQuota q, int pass) {
    __shared__ int s_hist[SEG_HIST];
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (pass != 0) {
        seg_hist_begin(q, s_hist);
        int e = -1;
        if (i < q.Pr) {
            // all three keys in one word -- quota row | creationTimestamp - min | priority - min -- when their ranges fit 64 bits
            const int ns = q.r_ns[i];
            e = (ns >= 0 && ns < q.NS) ? q.ns_quota[ns] : -1;
            const unsigned long long ek = e < 0 ? (unsigned long long)q.E : (unsigned long long)e;
            q.key_a[i] = (ek << (q.fuse_tbits + q.fuse_pbits)) |
                         ((unsigned long long)(q.r_ts[i] - q.fuse_ts_min) << q.fuse_pbits) |
                         (unsigned long long)((long long)q.r_prio[i] - q.fuse_prio_min);
            q.val_a[i] = i;
            if (e < 0) q.over_quota[i] = 0;
        }
        seg_hist_add(q, s_hist, e, i < q.Pr);
        seg_hist_end(q, s_hist);
        return;
    }
    if (i >= q.Pr) return;
    if (pass == 0) {                                       // least significant key: priority (biased to unsigned)
        q.key_a[i] = (unsigned long long)((long long)q.r_prio[i] - (long long)INT_MIN);
        q.val_a[i] = i;
    }
}
__global__ void __launch_bounds__(256) k_label_rekey(DevQuota q, const int32_t* vals, unsigned long long* keys, int pass) {
    __shared__ int s_hist[SEG_HIST];
    const int x = blockIdx.x * blockDim.x + threadIdx.x;
    if (pass == 1) {
        if (x < q.Pr) keys[x] = (unsigned long long)q.r_ts[vals[x]] ^ 0x8000000000000000ull;     // creationTimestamp
        return;
    }
    seg_hist_begin(q, s_hist);                                                            // quota row; no quota -> sorted last
    int e = -1;
    if (x < q.Pr) {
        const int i = vals[x];
        const int ns = q.r_ns[i];
        e = (ns >= 0 && ns < q.NS) ? q.ns_quota[ns] : -1;
        keys[x] = e < 0 ? (unsigned long long)q.E : (unsigned long long)e;
        if (e < 0) q.over_quota[i] = 0;
    }
    seg_hist_add(q, s_hist, e, x < q.Pr);
    seg_hist_end(q, s_hist);
}
constexpr int LABEL_CHUNK = 256;                               // pods per chunk of the labelling scan == its block size
__global__ void k_label_offsets(DevQuota q) {                  // exclusive scan of the per-quota counts (single thread block)
    __shared__ int s_carry;
    if (threadIdx.x == 0) s_carry = 0;
    __syncthreads();
    for (int base = 0; base < q.E; base += blockDim.x) {
        const int e = base + threadIdx.x;
        const int cnt = e < q.E ? q.seg_count[e] : 0;
        __shared__ int s[256];
        s[threadIdx.x] = cnt;
        __syncthreads();
        for (int o = 1; o < (int)blockDim.x; o <<= 1) {
            const int v = threadIdx.x >= o ? s[threadIdx.x - o] : 0;
            __syncthreads();
            s[threadIdx.x] += v;
            __syncthreads();
        }
        if (e < q.E) q.seg_off[e] = s_carry + s[threadIdx.x] - cnt;
        __syncthreads();
        if (threadIdx.x == blockDim.x - 1) s_carry += s[threadIdx.x];
        __syncthreads();
    }
    if (threadIdx.x == 0) q.seg_off[q.E] = s_carry;
    __syncthreads();
    // the same again for the number of 256-pod chunks of each segment
    if (threadIdx.x == 0) s_carry = 0;
    __syncthreads();
    for (int base = 0; base < q.E; base += blockDim.x) {
        const int e = base + threadIdx.x;
        const int cnt = e < q.E ? (q.seg_count[e] + LABEL_CHUNK - 1) / LABEL_CHUNK : 0;
        __shared__ int s2[256];
        s2[threadIdx.x] = cnt;
        __syncthreads();
        for (int o = 1; o < (int)blockDim.x; o <<= 1) {
            const int v = threadIdx.x >= o ? s2[threadIdx.x - o] : 0;
            __syncthreads();
            s2[threadIdx.x] += v;
            __syncthreads();
        }
        if (e < q.E) q.chunk_off[e] = s_carry + s2[threadIdx.x] - cnt;
        __syncthreads();
        if (threadIdx.x == blockDim.x - 1) s_carry += s2[threadIdx.x];
        __syncthreads();
    }
    if (threadIdx.x == 0) q.chunk_off[q.E] = s_carry;
}

__device__ __forceinline__ long long quota_req_scalar(const DevQuota& q, int i, int c) {   // one column, thread-serial
    if (c != q.gm) return q.r_req[(size_t)i * q.R + c];
    long long g = 0;
    for (int k = 0; k < q.R; ++k) if (k != q.gm) g += q.coeff[k] * q.r_req[(size_t)i * q.R + k];
    return g;
}
// the comparator's 3rd and 4th keys (elasticquota.go:93-103) for pods that tie on (creationTimestamp, priority)
__device__ bool tie_less(const DevQuota& q, int a, int b, bool* ambiguous) {
    bool equal = true, a_le_b = true, b_le_a = true;
    for (int c = 0; c < q.R; ++c) {
        const long long x = quota_req_scalar(q, a, c), y = quota_req_scalar(q, b, c);
        equal &= x == y; a_le_b &= x <= y; b_le_a &= y <= x;
    }
    if (!equal) { if (a_le_b == b_le_a) *ambiguous = true; return a_le_b; }
    return q.r_rank[a] < q.r_rank[b];
}
// one thread per tie group (same quota, creationTimestamp, priority): stable insertion sort with the comparator
__global__ void __launch_bounds__(256) k_label_ties(DevQuota q, int32_t* order) {
    const int x = blockIdx.x * blockDim.x + threadIdx.x;
    const int n_in_quota = q.seg_off[q.E];                     // pods without a quota sort last and are not touched
    if (x >= n_in_quota) return;
    auto same = [&](int u, int v) {
        const int a = order[u], b = order[v];
        return q.ns_quota[q.r_ns[a]] == q.ns_quota[q.r_ns[b]] && q.r_ts[a] == q.r_ts[b] && q.r_prio[a] == q.r_prio[b];
    };
    if (x > 0 && same(x - 1, x)) return;                       // not a group leader
    int end = x + 1;
    while (end < n_in_quota && same(x, end)) ++end;
    if (end - x < 2) return;
    bool amb = false;
    for (int i = x + 1; i < end; ++i) {
        const int v = order[i];
        int j = i;
        while (j > x && tie_less(q, v, order[j - 1], &amb)) { order[j] = order[j - 1]; --j; }
        order[j] = v;
    }
    for (int i = x; i < end && !amb; ++i)                      // the partial order must be a chain inside the group
        for (int j = i + 1; j < end; ++j) { bool t = false; tie_less(q, order[i], order[j], &t); amb |= t; }
    if (amb) *q.ambiguous = 1;
}
// The running sum of the requests of one quota's pods in order, in-quota iff used <= min on every key of Min
// (elasticquota.go:44-60), as a chunked scan so that a quota with tens of thousands of pods does not serialise on one CTA:
//   k_label_chunks(apply = 0)  one CTA per 256-pod chunk of a quota's segment: column sums of the chunk
//   k_label_chunk_prefix       one warp per quota (lane = column): running sum before each of its chunks; final sum = status.used
//   k_label_chunks(apply = 1)  the same CTAs again: scan inside the chunk on top of that prefix, compare with Min, write the label
__device__ __forceinline__ int label_chunk_quota(const DevQuota& q, int chunk) {   // largest e with chunk_off[e] <= chunk
    int lo = 0, hi = q.E;
    while (hi - lo > 1) { const int mid = (lo + hi) >> 1; if (q.chunk_off[mid] <= chunk) lo = mid; else hi = mid; }
    return lo;
}
__global__ void __launch_bounds__(LABEL_CHUNK) k_label_chunks(DevQuota q, const int32_t* order, int apply) {
    __shared__ long long s_warp[LABEL_CHUNK / 32][QR];
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    for (int chunk = blockIdx.x; chunk < q.chunk_off[q.E]; chunk += gridDim.x) {
        const int e = label_chunk_quota(q, chunk);
        const int x = q.seg_off[e] + (chunk - q.chunk_off[e]) * LABEL_CHUNK + tid;
        const int i = x < q.seg_off[e + 1] ? order[x] : -1;
        const unsigned mm = q.min_mask[e];
        long long v[QR];
#pragma unroll
        for (int c = 0; c < QR; ++c) v[c] = (i >= 0 && c < q.R) ? quota_req_scalar(q, i, c) : 0;
#pragma unroll
        for (int c = 0; c < QR; ++c) {                          // inclusive scan across the warp, per column
            if (c >= q.R) break;
#pragma unroll
            for (int o = 1; o < 32; o <<= 1) {
                const long long t = __shfl_up_sync(0xffffffffu, v[c], o);
                if (lane >= o) v[c] += t;
            }
            if (lane == 31) s_warp[warp][c] = v[c];
        }
        __syncthreads();
        if (!apply) {
            if (tid < q.R) {
                long long t = 0;
                for (int w = 0; w < LABEL_CHUNK / 32; ++w) t += s_warp[w][tid];
                q.chunk_sum[(size_t)chunk * QR + tid] = t;
            }
        } else {
            bool in_quota = true;
#pragma unroll
            for (int c = 0; c < QR; ++c) {
                if (c >= q.R) break;
                long long pre = q.chunk_sum[(size_t)chunk * QR + c];
                for (int w = 0; w < warp; ++w) pre += s_warp[w][c];
                if ((mm >> c & 1) && v[c] + pre > q.qmin[(size_t)e * q.R + c]) in_quota = false;
            }
            if (i >= 0) {
                q.over_quota[i] = in_quota ? 0 : 1;
                if (q.cur_label) {
                    q.chg_packed[x] = (i << 1) | (in_quota ? 0 : 1);
                    q.chg_flag[x] = q.cur_label[i] != (in_quota ? NOSPLAN_LABEL_IN_QUOTA : NOSPLAN_LABEL_OVER_QUOTA);
                }
            }
        }
        __syncthreads();
    }
}
// one warp per quota, lane = chunk (32 chunks per step, every column's loads in flight together), warp scan per column
__global__ void __launch_bounds__(256) k_label_chunk_prefix(DevQuota q) {
    const int e = blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5), lane = threadIdx.x & 31;
    if (e >= q.E) return;                                      // warp-uniform
    const int c0 = q.chunk_off[e], c1 = q.chunk_off[e + 1];
    long long carry[QR];
#pragma unroll
    for (int c = 0; c < QR; ++c) carry[c] = 0;
    for (int base = c0; base < c1; base += 32) {
        const int chunk = base + lane;
        const bool in = chunk < c1;
        long long t[QR];
#pragma unroll
        for (int c = 0; c < QR; ++c) t[c] = (in && c < q.R) ? q.chunk_sum[(size_t)chunk * QR + c] : 0;
#pragma unroll
        for (int c = 0; c < QR; ++c) {
            if (c >= q.R) break;
            long long v = t[c];
#pragma unroll
            for (int o = 1; o < 32; o <<= 1) {
                const long long u = __shfl_up_sync(0xffffffffu, v, o);
                if (lane >= o) v += u;
            }
            if (in) q.chunk_sum[(size_t)chunk * QR + c] = carry[c] + v - t[c];       // running sum BEFORE the chunk
            carry[c] += __shfl_sync(0xffffffffu, v, 31);
        }
    }
    long long mine = 0;
#pragma unroll
    for (int c = 0; c < QR; ++c) if (c == lane) mine = carry[c];
    if (lane < q.R) q.operator_used[(size_t)e * q.R + lane] = (q.min_mask[e] >> lane & 1) ? mine : 0;
}

// ---- labelling without a device-wide sort.  The most significant sort key is the quota row, and everything that follows --
// ties, running sums, labels -- happens inside one quota.  The host, which counts the pods of every quota anyway, cuts a quota
// with many pods into 2^k BUCKETS by the leading bits of creationTimestamp - min (the next sort key, so buckets are in order and
// a tie group never straddles two), aiming at a few hundred pods per bucket.  Then:
//   k_label_scatter   (cooperative) counting sort by bucket: per-CTA histogram in shared memory whose atomicAdd results are the
//                     pods' ranks inside (CTA, bucket); one global atomicAdd per touched (CTA, bucket) reserves the CTA's stretch
//                     of the bucket; after the grid barrier every CTA scans the counts itself and writes its pods'
//                     (creationTimestamp - min | priority - min, pod) pairs to their slots.  Order inside a bucket: arbitrary.
//   k_label_segments  one CTA per bucket (taken in order through a ticket): column sums of the bucket published at once (they do
//                     not depend on the order), the bucket sorted in shared memory (bitonic, key then pod index: the order a
//                     stable sort of the pods in index order gives), tie groups put in the comparator's order
//                     (elasticquota.go:93-103), the sums of the quota's earlier buckets collected, running sums of the requests
//                     per column against Min (elasticquota.go:44-60), labels, status.used.
// A bucket with more than SEG_MAX pods (timestamps bunched together) raises seg_overflow: the host repeats the call with the
// device-wide sort.
constexpr int SCAT_THREADS = 256, SCAT_ITEMS = 8, SEG_NB_MAX = 8000, SEG_MAX = 2048, SEG_TARGET = 128, SEG_THREADS = 128;

__device__ __forceinline__ int label_bucket(const DevQuota& q, int i) {      // q.n_buckets: no quota
    const int ns = q.r_ns[i];
    const int e = (ns >= 0 && ns < q.NS) ? q.ns_quota[ns] : -1;
    if (e < 0) return q.n_buckets;
    const int first = q.b_first[e], n_sub = q.b_first[e + 1] - first;
    const unsigned long long t = (unsigned long long)(q.r_ts[i] - q.fuse_ts_min) >> q.b_shift[e];
    return first + (int)min((unsigned long long)(n_sub - 1), t);
}

__global__ void __launch_bounds__(SCAT_THREADS) k_label_scatter(DevQuota q, int tile) {
    extern __shared__ int s_i[];                               // [3][n_buckets + 2]
    const int NB = q.n_buckets;
    int* s_hist = s_i; int* s_base = s_i + (NB + 2); int* s_off = s_base + (NB + 2);
    __shared__ int s_wtot[SCAT_THREADS / 32];
    cooperative_groups::grid_group grid = cooperative_groups::this_grid();
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const int lo = blockIdx.x * tile, hi = min(q.Pr, lo + tile);
    for (int k = tid; k <= NB; k += SCAT_THREADS) s_hist[k] = 0;
    __syncthreads();
    int b_[SCAT_ITEMS], r_[SCAT_ITEMS];
#pragma unroll
    for (int u = 0; u < SCAT_ITEMS; ++u) {
        const int i = lo + u * SCAT_THREADS + tid;
        b_[u] = -1; r_[u] = 0;
        if (i < hi) { b_[u] = label_bucket(q, i); r_[u] = atomicAdd(&s_hist[b_[u]], 1); }
    }
    __syncthreads();
    for (int k = tid; k < NB; k += SCAT_THREADS) { const int c = s_hist[k]; s_base[k] = c ? atomicAdd(&q.seg_count[k], c) : 0; }
    grid.sync();
    // exclusive scan of the bucket counts, by every CTA for itself: contiguous bins per thread, warp scan, warp totals
    const int per = (NB + SCAT_THREADS - 1) / SCAT_THREADS;
    int mine = 0;
    for (int k = tid * per; k < min(NB, (tid + 1) * per); ++k) mine += __ldcg(&q.seg_count[k]);
    int incl = mine;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) { const int t = __shfl_up_sync(0xffffffffu, incl, o); if (lane >= o) incl += t; }
    if (lane == 31) s_wtot[warp] = incl;
    __syncthreads();
    int before = incl - mine;
    for (int w = 0; w < warp; ++w) before += s_wtot[w];
    for (int k = tid * per; k < min(NB, (tid + 1) * per); ++k) {
        s_off[k] = before;
        if (blockIdx.x == 0) q.seg_off[k] = before;
        before += __ldcg(&q.seg_count[k]);
    }
    if (blockIdx.x == 0 && tid == SCAT_THREADS - 1) q.seg_off[NB] = before;        // pods that have a quota
    __syncthreads();
#pragma unroll
    for (int u = 0; u < SCAT_ITEMS; ++u) {
        const int i = lo + u * SCAT_THREADS + tid;
        if (b_[u] < 0) continue;
        if (b_[u] == NB) { q.over_quota[i] = 0; continue; }
        const int slot = s_off[b_[u]] + s_base[b_[u]] + r_[u];
        q.key_a[slot] = ((unsigned long long)(q.r_ts[i] - q.fuse_ts_min) << q.fuse_pbits) | (unsigned long long)((long long)q.r_prio[i] - q.fuse_prio_min);
        q.val_a[slot] = i;
    }
}

__global__ void __launch_bounds__(SEG_THREADS) k_label_segments(DevQuota q) {
    __shared__ unsigned long long sk[SEG_MAX];
    __shared__ int sv[SEG_MAX];
    __shared__ long long s_warp[SEG_THREADS / 32][QR];
    __shared__ long long s_carry[QR];
    __shared__ int s_bucket;
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    if (tid == 0) s_bucket = atomicAdd(q.seg_ticket, 1);       // buckets in order: the ones a CTA waits for below have started
    __syncthreads();
    const int b = s_bucket;
    const int e = q.b_quota[b];
    const int x0 = q.seg_off[b], n = q.seg_off[b + 1] - x0;
    const int b0 = q.b_first[e];
    const bool last = b + 1 == q.b_first[e + 1];
    if (n > SEG_MAX) {                                          // timestamps bunched together: the host repeats with the device-wide sort
        if (tid == 0) { *q.seg_overflow = 1; __threadfence(); atomicExch(&q.seg_flag[b], 1); }
        return;
    }
    if (tid < QR) s_carry[tid] = 0;
    if (q.b_first[e + 1] - b0 == 1) {
        // the quota's only bucket: nobody waits for its sums
        for (int k = tid; k < n; k += SEG_THREADS) { sk[k] = q.key_a[x0 + k]; sv[k] = q.val_a[x0 + k]; }
    } else {
        // ---- this bucket's column sums, published before anything else
        long long v[QR];
#pragma unroll
        for (int c = 0; c < QR; ++c) v[c] = 0;
        for (int k = tid; k < n; k += SEG_THREADS) {
            const int i = q.val_a[x0 + k];
            sk[k] = q.key_a[x0 + k]; sv[k] = i;
#pragma unroll
            for (int c = 0; c < QR; ++c) if (c < q.R) v[c] += quota_req_scalar(q, i, c);
        }
#pragma unroll
        for (int c = 0; c < QR; ++c) {
            if (c >= q.R) break;
#pragma unroll
            for (int o = 16; o; o >>= 1) v[c] += __shfl_xor_sync(0xffffffffu, v[c], o);
            if (lane == 0) s_warp[warp][c] = v[c];
        }
        __syncthreads();
        if (tid < q.R) {
            long long t = 0;
            for (int w = 0; w < SEG_THREADS / 32; ++w) t += s_warp[w][tid];
            q.seg_sum[(size_t)b * QR + tid] = t;
        }
        __threadfence();
        __syncthreads();
        if (tid == 0) atomicExch(&q.seg_flag[b], 1);
    }
    // ---- sort.  Warps without a pod of the bucket have nothing left to do (a barrier does not wait for threads that exited).
    int np2 = 32;
    while (np2 < n) np2 <<= 1;
    const int n_warps = min(SEG_THREADS / 32, max(1, (n + 31) >> 5));
    if (warp >= n_warps) return;
    for (int k = n + tid; k < np2; k += 32 * n_warps) { sk[k] = ~0ull; sv[k] = INT_MAX; }
    __syncthreads();
    for (int k = 2; k <= np2; k <<= 1)
        for (int j = k >> 1; j > 0; j >>= 1) {
            for (int h = tid; h < (np2 >> 1); h += 32 * n_warps) {       // one compare-exchange per thread: the pair (t, t | j)
                const int t = ((h & ~(j - 1)) << 1) | (h & (j - 1)), u = t | j;
                const unsigned long long ka = sk[t], kb = sk[u];
                const int va = sv[t], vb = sv[u];
                const bool gt = ka > kb || (ka == kb && va > vb);
                if (gt == ((t & k) == 0)) { sk[t] = kb; sk[u] = ka; sv[t] = vb; sv[u] = va; }
            }
            if (np2 > 64 || j > 16) __syncthreads(); else __syncwarp();
        }
    __syncthreads();
    // ---- tie groups (same creationTimestamp and priority): one thread per group, stable insertion sort with the comparator
    for (int x = tid; x < n; x += 32 * n_warps) {
        if (x > 0 && sk[x - 1] == sk[x]) continue;              // not a group leader
        int end = x + 1;
        while (end < n && sk[end] == sk[x]) ++end;
        if (end - x < 2) continue;
        bool amb = false;
        for (int i = x + 1; i < end; ++i) {
            const int v = sv[i];
            int j = i;
            while (j > x && tie_less(q, v, sv[j - 1], &amb)) { sv[j] = sv[j - 1]; --j; }
            sv[j] = v;
        }
        for (int i = x; i < end && !amb; ++i)                   // the partial order must be a chain inside the group
            for (int j = i + 1; j < end; ++j) { bool t = false; tie_less(q, sv[i], sv[j], &t); amb |= t; }
        if (amb) *q.ambiguous = 1;
    }
    // ---- the quota's earlier buckets: their sums are the running sum this bucket starts from
    {   // one earlier bucket per thread: all the flags and sums are in flight at once
        long long acc[QR];
#pragma unroll
        for (int c = 0; c < QR; ++c) acc[c] = 0;
        for (int p = b0 + tid; p < b; p += 32 * n_warps) {
            // the bucket waited for holds a lower ticket, so it is running or done; the wait is bounded all the same (half a
            // second), and a wait that runs out makes the host repeat the call with the device-wide sort
            unsigned spins = 0;
            while (*reinterpret_cast<volatile int32_t*>(&q.seg_flag[p]) == 0) {
                if (++spins > (1u << 24)) { *q.seg_overflow = 1; break; }
                __nanosleep(32);
            }
            __threadfence();
#pragma unroll
            for (int c = 0; c < QR; ++c) if (c < q.R) acc[c] += __ldcg(&q.seg_sum[(size_t)p * QR + c]);
        }
        __syncthreads();                                        // s_warp is free again (the sort and the tie groups are done)
#pragma unroll
        for (int c = 0; c < QR; ++c) {
            if (c >= q.R) break;
#pragma unroll
            for (int o = 16; o; o >>= 1) acc[c] += __shfl_xor_sync(0xffffffffu, acc[c], o);
            if (lane == 0) s_warp[warp][c] = acc[c];
        }
        __syncthreads();
        if (tid < q.R) { long long t = 0; for (int w = 0; w < n_warps; ++w) t += s_warp[w][tid]; s_carry[tid] = t; }
        __syncthreads();
    }
    // ---- running sums in order, SEG_THREADS pods per step
    // Alongside: reserveResource of the same pods into the quota's Used (elasticquotainfo.go:194-200, 276-292; what k_quota_used
    // does with atomics when the labelling is not by buckets): Memory, MilliCPU and the scalars the request names.
    const unsigned mm = q.min_mask[e];
    long long uacc[QR], own = 0;
    unsigned macc = 0;
#pragma unroll
    for (int c = 0; c < QR; ++c) uacc[c] = 0;
    for (int base = 0; base < n; base += 32 * n_warps) {
        const int x = base + tid;
        const int i = x < n ? sv[x] : -1;
        long long v[QR];
#pragma unroll
        for (int c = 0; c < QR; ++c) v[c] = (i >= 0 && c < q.R) ? quota_req_scalar(q, i, c) : 0;
        if (i >= 0) {
            unsigned rm = 0;
            if (q.r_mask) rm = q.r_mask[i] & SCALAR_BITS;
            else {
#pragma unroll
                for (int c = 4; c < QR; ++c) if (c < q.R && c != q.gm && v[c] != 0) rm |= 1u << c;
            }
            if (q.gm >= 0) rm |= 1u << q.gm;
            macc |= rm;
#pragma unroll
            for (int c = 4; c < QR; ++c) if (rm >> c & 1) uacc[c] += v[c];
        }
#pragma unroll
        for (int c = 0; c < QR; ++c) {
            if (c >= q.R) break;
#pragma unroll
            for (int o = 1; o < 32; o <<= 1) {
                const long long t = __shfl_up_sync(0xffffffffu, v[c], o);
                if (lane >= o) v[c] += t;
            }
            if (lane == 31) s_warp[warp][c] = v[c];
        }
        __syncthreads();
        if (tid < q.R) {                                        // warp totals -> what precedes each warp; the step's total into the carry
            long long run = s_carry[tid];
            own -= run;
            for (int w = 0; w < n_warps; ++w) { const long long t = s_warp[w][tid]; s_warp[w][tid] = run; run += t; }
            s_carry[tid] = run;
            own += run;
        }
        __syncthreads();
        bool in_quota = true;
#pragma unroll
        for (int c = 0; c < QR; ++c) {
            if (c >= q.R) break;
            if ((mm >> c & 1) && v[c] + s_warp[warp][c] > q.qmin[(size_t)e * q.R + c]) in_quota = false;
        }
        if (i >= 0) {
            q.over_quota[i] = in_quota ? 0 : 1;
            if (q.cur_label) {
                q.chg_packed[x0 + x] = (i << 1) | (in_quota ? 0 : 1);
                q.chg_flag[x0 + x] = q.cur_label[i] != (in_quota ? NOSPLAN_LABEL_IN_QUOTA : NOSPLAN_LABEL_OVER_QUOTA);
            }
        }
        __syncthreads();
    }
    if (last && tid < q.R) q.operator_used[(size_t)e * q.R + tid] = (mm >> tid & 1) ? s_carry[tid] : 0;
    // Used: this bucket's share
#pragma unroll
    for (int c = 4; c < QR; ++c) {
        if (c >= q.R) break;
#pragma unroll
        for (int o = 16; o; o >>= 1) uacc[c] += __shfl_xor_sync(0xffffffffu, uacc[c], o);
        if (lane == 0) s_warp[warp][c] = uacc[c];
    }
    macc = __reduce_or_sync(0xffffffffu, macc);
    if (lane == 0) s_warp[warp][0] = (long long)macc;
    __syncthreads();
    if (tid < q.R && tid != 2 && tid != 3) {
        long long t = own;                                      // Memory, MilliCPU: every pod's
        if (tid >= 4) { t = 0; for (int w = 0; w < n_warps; ++w) t += s_warp[w][tid]; }
        if (t != 0) atomicAdd((unsigned long long*)&q.used[(size_t)e * q.R + tid], (unsigned long long)t);
    }
    if (tid == 0) {
        unsigned m = 0;
        for (int w = 0; w < n_warps; ++w) m |= (unsigned)s_warp[w][0];
        if (m & ~q.used_mask[e]) atomicOr(&q.used_mask[e], m);
    }
}

}  // namespace nsp

using namespace nsp;

// host side --------------------------------------------------------------------------------------------------
#include "quota_ctx.cuh"

extern "C" int nosplan_quota_recompute_impl(int device, cudaStream_t st, nosplan_quota_ctx* ctx, std::string* err,
                                            const nosplan_quotas* qs, const nosplan_quota_pods* running,
                                            const nosplan_quota_pods* pending, nosplan_quota_out* out, float* ms_h2d, float* ms_k,
                                            float* ms_d2h, int* launches, long long* h2d_bytes, long long* d2h_bytes,
                                            const nosplan_label_changes* lc, const nosplan_quota_reduce_hook* hook) {
    auto fail = [&](int code, const char* msg, const char* detail = "") { if (err) *err = std::string(msg) + detail; return code; };
#define QCU(call) do { cudaError_t e_ = (call); if (e_ != cudaSuccess) return fail(NOSPLAN_E_CUDA, #call ": ", cudaGetErrorString(e_)); } while (0)
    nosplan_quota_out no_out;
    if (lc && !out) { std::memset(&no_out, 0, sizeof no_out); out = &no_out; }
    if (!qs || !out) return fail(NOSPLAN_E_INVALID, "null quotas or out");
    if (lc && (!lc->n_changes || lc->capacity < 0 || (lc->capacity > 0 && (!lc->pod || !lc->desired))))
        return fail(NOSPLAN_E_INVALID, "label changes: n_changes, and pod / desired for a non-zero capacity, are required");
    const int E = qs->n_quotas, R = qs->n_res, NS = qs->n_namespaces;
    const int Pr = running ? running->n_pods : 0, Pp = pending ? pending->n_pods : 0;
    if (R < 4 || R > QR) return fail(NOSPLAN_E_LIMIT, "n_res must be in [4, NOSPLAN_QUOTA_MAX_RES]");
    if (E < 0 || NS < 0 || Pr < 0 || Pp < 0) return fail(NOSPLAN_E_INVALID, "negative size");
    if (qs->gpu_memory_col >= R || (qs->gpu_memory_col >= 0 && qs->gpu_memory_col < 4)) return fail(NOSPLAN_E_INVALID, "gpu_memory_col out of range");
    for (int ns = 0; ns < NS; ++ns) if (qs->ns_quota[ns] >= E) return fail(NOSPLAN_E_INVALID, "ns_quota out of range");
    if (lc && Pr > 0 && !lc->current) return fail(NOSPLAN_E_INVALID, "label changes: current labels of the running pods are required");
    if (lc && Pr >= (1 << 30)) return fail(NOSPLAN_E_LIMIT, "label changes: at most 2^30 - 1 running pods");
    const bool want_label = (out->over_quota || out->operator_used || out->ambiguous_order || lc) && Pr > 0 && E > 0;
    const bool want_changes = lc && want_label;
    if (want_label && (!running->priority || !running->creation_ts || !running->name_rank))
        return fail(NOSPLAN_E_INVALID, "operator labelling needs priority, creation_ts and name_rank of the running pods");
    QCU(cudaSetDevice(device));

    // ---- layout
    size_t o = 0;
    auto al = [](size_t x) { return (x + 255) & ~(size_t)255; };
    auto put = [&](size_t bytes) { size_t at = o; o = al(o + bytes); return at; };
    const size_t ER = (size_t)E * R;
    const size_t o_nsq = put(4 * (size_t)NS), o_min = put(8 * ER), o_max = put(8 * ER), o_uin = put(8 * ER),
                 o_mm = put(4 * (size_t)E), o_xm = put(4 * (size_t)E), o_um = put(4 * (size_t)E), o_enf = put((size_t)E),
                 o_rns = put(4 * (size_t)Pr), o_rreq = put(8 * (size_t)Pr * R), o_rmask = put(4 * (size_t)Pr),
                 o_rprio = put(4 * (size_t)Pr), o_rts = put(8 * (size_t)Pr), o_rrank = put(4 * (size_t)Pr),
                 o_pns = put(4 * (size_t)Pp), o_preq = put(8 * (size_t)Pp * R), o_pmask = put(4 * (size_t)Pp),
                 o_cur = put(want_changes ? (size_t)Pr : 0),
                 o_bfirst = put(4 * ((size_t)E + 2)), o_bshift = put(4 * ((size_t)E + 1)), o_bquota = put(4 * ((size_t)SEG_NB_MAX + 1));
    const size_t upload_end = o;
    // outputs (contiguous, one D2H)
    const size_t o_used = put(8 * ER), o_umask = put(4 * (size_t)E), o_amin = put(8 * QR), o_aused = put(8 * QR), o_aover = put(8 * QR),
                 o_amask = put(16), o_guar = put(8 * ER), o_pref = put((size_t)Pp), o_over = put((size_t)Pr), o_opused = put(8 * ER),
                 o_amb = put(16), o_nsel = put(16), o_ovf = put(16);
    const size_t out_end = o;
    const size_t o_sel = put(want_changes ? 4 * (size_t)Pr : 0);   // fetched by a second copy of exactly n_changes entries
    const size_t o_packed = put(want_changes ? 4 * (size_t)Pr : 0), o_flag = put(want_changes ? (size_t)Pr : 0);
    const size_t o_keya = put(8 * (size_t)Pr), o_keyb = put(8 * (size_t)Pr), o_vala = put(4 * (size_t)Pr), o_valb = put(4 * (size_t)Pr),
                 o_segc = put(4 * ((size_t)std::max(E, SEG_NB_MAX) + 2)), o_sego = put(4 * ((size_t)std::max(E, SEG_NB_MAX) + 2)),
                 o_choff = put(4 * ((size_t)E + 2)), o_segsum = put(8 * (size_t)QR * (SEG_NB_MAX + 1)), o_segflag = put(4 * ((size_t)SEG_NB_MAX + 2));
    const int n_chunk_bound = Pr / 256 + E + 1;                 // every quota's last chunk may be partial
    const size_t o_chsum = put(8 * (size_t)n_chunk_bound * QR);
    size_t cub_bytes = 0;
    if (want_label) {
        cub::DoubleBuffer<unsigned long long> dk(nullptr, nullptr); cub::DoubleBuffer<int32_t> dv(nullptr, nullptr);
        QCU(cub::DeviceRadixSort::SortPairs(nullptr, cub_bytes, dk, dv, Pr, 0, 64, st));
    }
    if (want_changes) {
        size_t sel_bytes = 0;
        QCU(cub::DeviceSelect::Flagged(nullptr, sel_bytes, (const int32_t*)nullptr, (const uint8_t*)nullptr, (int32_t*)nullptr,
                                       (int32_t*)nullptr, Pr, st));
        cub_bytes = std::max(cub_bytes, sel_bytes);
    }
    const size_t o_cub = put(cub_bytes + 256);
    const size_t total = o;
    QCU(quota_ctx_ensure(ctx, total, std::max(upload_end, out_end - upload_end + (want_changes ? 4 * (size_t)Pr + 256 : 0))));
    unsigned char* S = ctx->hs; unsigned char* A = ctx->d;
    auto cp = [&](size_t at, const void* src, size_t bytes) { if (bytes) { if (src) std::memcpy(S + at, src, bytes); else std::memset(S + at, 0, bytes); } };
    cp(o_nsq, qs->ns_quota, 4 * (size_t)NS); cp(o_min, qs->min, 8 * ER); cp(o_max, qs->max, 8 * ER); cp(o_uin, qs->used_init, 8 * ER);
    cp(o_mm, qs->min_mask, 4 * (size_t)E); cp(o_xm, qs->max_mask, 4 * (size_t)E); cp(o_um, qs->used_mask, 4 * (size_t)E);
    cp(o_enf, qs->max_enforced, (size_t)E);
    for (int e = 0; e < E; ++e) reinterpret_cast<uint32_t*>(S + o_um)[e] &= SCALAR_BITS;    // key set of the scalars only
    if (Pr) {
        cp(o_rns, running->ns, 4 * (size_t)Pr); cp(o_rreq, running->request, 8 * (size_t)Pr * R);
        cp(o_rmask, running->request_mask, 4 * (size_t)Pr); cp(o_rprio, running->priority, 4 * (size_t)Pr);
        cp(o_rts, running->creation_ts, 8 * (size_t)Pr); cp(o_rrank, running->name_rank, 4 * (size_t)Pr);
    }
    if (want_changes) cp(o_cur, lc->current, (size_t)Pr);
    if (Pp) { cp(o_pns, pending->ns, 4 * (size_t)Pp); cp(o_preq, pending->request, 8 * (size_t)Pp * R); cp(o_pmask, pending->request_mask, 4 * (size_t)Pp); }

    // ---- host analysis of the labelling keys, in front of the copies so that no kernel ever waits for it
    static thread_local bool force_global_sort = false;
    int ebits = 1; while ((1ll << ebits) <= E) ++ebits;
    long long tmin = LLONG_MAX, tmax = LLONG_MIN, pmin = LLONG_MAX, pmax = LLONG_MIN;
    bool have_keys = false, ts_ok = false, by_buckets = false;
    int tbits = 64, pbits = 32, n_buckets = 0;
    if (want_label) {
    // ranges of the two inner keys (host scan of the caller's arrays): if quota | timestamp | priority fit one 64-bit word
    // a single radix sort over exactly the bits in use replaces the three stable passes
    have_keys = running->creation_ts && running->priority;
    // ... and the number of pods in every quota: the labelling then needs no device-wide sort (k_label_scatter)
    static thread_local std::vector<int32_t> seg_n;
    const bool count_segs = have_keys && E <= SEG_NB_MAX && !force_global_sort && !getenv("NOSPLAN_QUOTA_GLOBAL_SORT");
    if (count_segs) seg_n.assign((size_t)E + 1, 0);
    int32_t* seg_np = count_segs ? seg_n.data() : nullptr;
    for (int i = 0; have_keys && i < Pr; ++i) {
        const long long t = running->creation_ts[i], pr = running->priority[i];
        tmin = std::min(tmin, t); tmax = std::max(tmax, t); pmin = std::min(pmin, pr); pmax = std::max(pmax, pr);
        if (seg_np) { const int ns = running->ns[i]; const int e = (ns >= 0 && ns < NS) ? qs->ns_quota[ns] : -1; ++seg_np[e < 0 ? E : e]; }
    }
    auto bits_of = [](unsigned long long range) { int b = 1; while (b < 64 && (range >> b)) ++b; return b; };
    ts_ok = have_keys && (unsigned long long)tmax - (unsigned long long)tmin < (1ull << 62);
    tbits = ts_ok ? bits_of((unsigned long long)(tmax - tmin)) : 64; pbits = bits_of((unsigned long long)(pmax - pmin));
    // buckets: a quota with more than 2 * SEG_TARGET pods is cut by the leading bits of its pods' timestamps
    by_buckets = count_segs && ts_ok && tbits + pbits <= 64;
    if (by_buckets) {
        int32_t* bf = (int32_t*)(S + o_bfirst); int32_t* bs = (int32_t*)(S + o_bshift); int32_t* bq = (int32_t*)(S + o_bquota);
        // about SEG_TARGET pods per bucket (small buckets: short sorts, even CTAs); more where that would be too many buckets
        for (int target = SEG_TARGET; ; target *= 2) {
            n_buckets = 0;
            bool fits = true;
            for (int e = 0; e < E && fits; ++e) {
                int lg = 0;
                while (lg < tbits && ((long long)target << lg) < seg_np[e] && seg_np[e] > 2 * target) ++lg;
                bf[e] = n_buckets; bs[e] = tbits - lg;
                if (n_buckets + (1 << lg) > SEG_NB_MAX) { fits = false; break; }
                for (int k = 0; k < (1 << lg); ++k) bq[n_buckets++] = e;
            }
            if (fits) { bf[E] = n_buckets; break; }
            if (target >= SEG_MAX / 2) { by_buckets = false; break; }
        }
    }
    }

    cudaEvent_t* ev = nullptr;
    QCU(quota_ctx_events(ctx, &ev));
    QCU(cudaEventRecord(ev[0], st));
    QCU(cudaMemcpyAsync(A, S, upload_end, cudaMemcpyHostToDevice, st));
    QCU(cudaMemsetAsync(A + upload_end, 0, total - upload_end, st));
    QCU(cudaEventRecord(ev[1], st));

    DevQuota q; std::memset(&q, 0, sizeof q);
    q.E = E; q.R = R; q.NS = NS; q.gm = qs->gpu_memory_col; q.Pr = Pr; q.Pp = Pp;
    q.ns_quota = (const int32_t*)(A + o_nsq); q.qmin = (const long long*)(A + o_min); q.qmax = (const long long*)(A + o_max);
    q.used_init = (const long long*)(A + o_uin); q.min_mask = (const uint32_t*)(A + o_mm); q.max_mask = (const uint32_t*)(A + o_xm);
    q.used_mask_in = (const uint32_t*)(A + o_um); q.max_enforced = (const uint8_t*)(A + o_enf);
    for (int c = 0; c < QR; ++c) q.coeff[c] = (c < R && qs->gpu_memory_coeff) ? qs->gpu_memory_coeff[c] : 0;
    q.r_ns = (const int32_t*)(A + o_rns); q.r_req = (const long long*)(A + o_rreq);
    q.r_mask = (running && running->request_mask) ? (const uint32_t*)(A + o_rmask) : nullptr;
    q.r_prio = (const int32_t*)(A + o_rprio); q.r_ts = (const long long*)(A + o_rts); q.r_rank = (const int32_t*)(A + o_rrank);
    q.p_ns = (const int32_t*)(A + o_pns); q.p_req = (const long long*)(A + o_preq);
    q.p_mask = (pending && pending->request_mask) ? (const uint32_t*)(A + o_pmask) : nullptr;
    q.used = (long long*)(A + o_used); q.used_mask = (uint32_t*)(A + o_umask);
    q.agg_min = (long long*)(A + o_amin); q.agg_used = (long long*)(A + o_aused); q.agg_over = (long long*)(A + o_aover);
    q.agg_masks = (uint32_t*)(A + o_amask); q.guaranteed = (long long*)(A + o_guar); q.prefilter = (uint8_t*)(A + o_pref);
    q.over_quota = (uint8_t*)(A + o_over); q.operator_used = (long long*)(A + o_opused); q.ambiguous = (int32_t*)(A + o_amb);
    q.key_a = (unsigned long long*)(A + o_keya); q.key_b = (unsigned long long*)(A + o_keyb);
    q.val_a = (int32_t*)(A + o_vala); q.val_b = (int32_t*)(A + o_valb);
    q.seg_count = (int32_t*)(A + o_segc); q.seg_off = (int32_t*)(A + o_sego);
    q.chunk_off = (int32_t*)(A + o_choff); q.chunk_sum = (long long*)(A + o_chsum); q.n_chunk_bound = n_chunk_bound;
    if (want_changes) { q.cur_label = A + o_cur; q.chg_packed = (int32_t*)(A + o_packed); q.chg_flag = A + o_flag; }
    q.b_first = (const int32_t*)(A + o_bfirst); q.b_shift = (const int32_t*)(A + o_bshift); q.b_quota = (const int32_t*)(A + o_bquota);
    q.seg_sum = (long long*)(A + o_segsum); q.seg_flag = (int32_t*)(A + o_segflag); q.seg_ticket = q.seg_flag + SEG_NB_MAX + 1;
    q.seg_overflow = (int32_t*)(A + o_ovf);

    int n_launch = 0;
    bool used_buckets = false;                                 // the labelling ran by buckets: seg_overflow must be looked at
    // Used starts from used_init (scalar key set from used_mask)
    // (pods split between ranks: used_init is counted on the first rank only, the others start from zero)
    const bool take_init = !hook || hook->first_rank;
    if (ER && take_init) QCU(cudaMemcpyAsync(q.used, q.used_init, 8 * ER, cudaMemcpyDeviceToDevice, st));
    if (E && take_init) QCU(cudaMemcpyAsync(q.used_mask, q.used_mask_in, 4 * (size_t)E, cudaMemcpyDeviceToDevice, st));
    // the labelling by buckets (k_label_scatter / k_label_segments) also accounts Used: its grid -- at most SCAT_ITEMS * SCAT_THREADS
    // pods per CTA, every CTA resident (grid barrier)
    int scat_grid = 0;
    const size_t scat_dyn = 12 * ((size_t)n_buckets + 2);
    if (by_buckets) {
        int sms = 148, per_sm = 0;
        cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, device);
        if (cudaFuncSetAttribute(k_label_scatter, cudaFuncAttributeMaxDynamicSharedMemorySize, 12 * (SEG_NB_MAX + 2)) == cudaSuccess &&
            cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, k_label_scatter, SCAT_THREADS, scat_dyn) == cudaSuccess && per_sm > 0) {
            const int need = (Pr + SCAT_ITEMS * SCAT_THREADS - 1) / (SCAT_ITEMS * SCAT_THREADS);
            const int want = std::max(need, std::min(2 * sms, (Pr + SCAT_THREADS - 1) / SCAT_THREADS));
            if (need <= sms * per_sm) scat_grid = std::min(want, sms * per_sm);
        }
    }
    if (scat_grid > 0) {
        q.fuse_ts_min = tmin; q.fuse_prio_min = pmin; q.fuse_tbits = tbits; q.fuse_pbits = pbits; q.n_buckets = n_buckets;
        int tile = (Pr + scat_grid - 1) / scat_grid;
        void* args[] = {&q, &tile};
        QCU(cudaLaunchCooperativeKernel((const void*)k_label_scatter, dim3(scat_grid), dim3(SCAT_THREADS), args, scat_dyn, st));
        k_label_segments<<<n_buckets, SEG_THREADS, 0, st>>>(q);
        n_launch += 2;
        used_buckets = true;
    } else if (Pr && E) {
        const size_t tab = 8 * ER;
        const int use_smem = tab <= 96 * 1024;
        QCU(cudaFuncSetAttribute(k_quota_used, cudaFuncAttributeMaxDynamicSharedMemorySize, 96 * 1024));
        int blocks = (Pr + 15) / 16;
        if (blocks > 148 * 4) blocks = 148 * 4;
        k_quota_used<<<blocks, 256, use_smem ? tab : 0, st>>>(q, use_smem);
        ++n_launch;
    }
    if (hook && ER) {                                          // Used rows summed, key masks OR-ed over the ranks that share the pods
        const int rc = hook->reduce(hook->ctx, q.used, ER, q.used_mask, (size_t)E, st);
        if (rc != NOSPLAN_OK) return fail(rc, "quota reduction over the ranks failed");
        n_launch += 3;
    }
    QCU(launch_quota_aggregate(q, st));
    ++n_launch;
    if (Pp) { k_quota_prefilter<<<(Pp + 15) / 16, 256, 0, st>>>(q); ++n_launch; }
    if (want_label) {
        // stable LSD radix sort: priority, then creationTimestamp, then quota row
        const int gb = (Pr + 255) / 256;
        cub::DoubleBuffer<unsigned long long> dk(q.key_a, q.key_b); cub::DoubleBuffer<int32_t> dv(q.val_a, q.val_b);
        size_t tb = cub_bytes;
        if (!used_buckets) {
        if (ts_ok && ebits + tbits + pbits <= 64) {
            q.fuse_ts_min = tmin; q.fuse_prio_min = pmin; q.fuse_tbits = tbits; q.fuse_pbits = pbits;
            k_label_keys<<<gb, 256, 0, st>>>(q, 1);
            QCU(cub::DeviceRadixSort::SortPairs(A + o_cub, tb, dk, dv, Pr, 0, ebits + tbits + pbits, st));
            n_launch -= 4;
        } else {
            k_label_keys<<<gb, 256, 0, st>>>(q, 0);
            QCU(cub::DeviceRadixSort::SortPairs(A + o_cub, tb, dk, dv, Pr, 0, 32, st));
            k_label_rekey<<<gb, 256, 0, st>>>(q, dv.Current(), dk.Current(), 1);
            QCU(cub::DeviceRadixSort::SortPairs(A + o_cub, tb, dk, dv, Pr, 0, 64, st));
            k_label_rekey<<<gb, 256, 0, st>>>(q, dv.Current(), dk.Current(), 2);
            QCU(cub::DeviceRadixSort::SortPairs(A + o_cub, tb, dk, dv, Pr, 0, ebits, st));
        }
        k_label_offsets<<<1, 256, 0, st>>>(q);
        k_label_ties<<<gb, 256, 0, st>>>(q, dv.Current());
        const int cgrid = std::min(n_chunk_bound, 148 * 8);
        k_label_chunks<<<cgrid, LABEL_CHUNK, 0, st>>>(q, dv.Current(), 0);
        k_label_chunk_prefix<<<(E + 7) / 8, 256, 0, st>>>(q);
        k_label_chunks<<<cgrid, LABEL_CHUNK, 0, st>>>(q, dv.Current(), 1);
        n_launch += 11;
        }
        if (want_changes) {                                    // sorted positions whose label differs, compacted in order
            size_t sb = cub_bytes;
            QCU(cub::DeviceSelect::Flagged(A + o_cub, sb, (const int32_t*)q.chg_packed, (const uint8_t*)q.chg_flag, (int32_t*)(A + o_sel),
                                           (int32_t*)(A + o_nsel), Pr, st));
            n_launch += 2;
        }
    }
    QCU(cudaGetLastError());
    QCU(cudaEventRecord(ev[2], st));
    // one D2H over the span of the outputs the caller asked for (all of them for nosplan_quota_recompute; for the label
    // changes alone that is the 16-byte counter, and the list follows in a second copy of exactly n_changes entries)
    size_t d_lo = out_end, d_hi = upload_end;
    auto span = [&](const void* wanted, size_t at, size_t bytes) { if (wanted && bytes) { d_lo = std::min(d_lo, at); d_hi = std::max(d_hi, at + bytes); } };
    span(out->used, o_used, 8 * ER); span(out->used_mask, o_umask, 4 * (size_t)E); span(out->agg_min, o_amin, 8 * QR);
    span(out->agg_used, o_aused, 8 * QR); span(out->agg_overquota, o_aover, 8 * QR); span(out->agg_masks, o_amask, 16);
    span(out->guaranteed, o_guar, 8 * ER); span(out->prefilter, o_pref, (size_t)Pp);
    if (want_label) { span(out->over_quota, o_over, (size_t)Pr); span(out->operator_used, o_opused, 8 * ER); span(out->ambiguous_order, o_amb, 16); }
    if (want_changes) span(lc, o_nsel, 16);
    if (used_buckets) span(out, o_ovf, 4);
    if (d_hi > d_lo) QCU(cudaMemcpyAsync(S + (d_lo - upload_end), A + d_lo, d_hi - d_lo, cudaMemcpyDeviceToHost, st));
    QCU(cudaEventRecord(ev[3], st));
    QCU(cudaStreamSynchronize(st));
    if (used_buckets) {
        int32_t ovf = 0;
        std::memcpy(&ovf, S + (o_ovf - upload_end), 4);
        if (ovf) {                                             // a bucket beyond SEG_MAX pods: once more, with the device-wide sort
            force_global_sort = true;
            const int rc = nosplan_quota_recompute_impl(device, st, ctx, err, qs, running, pending, out, ms_h2d, ms_k, ms_d2h, launches, h2d_bytes,
                                                        d2h_bytes, lc, hook);
            force_global_sort = false;
            return rc;
        }
    }
    auto back = [&](void* dst, size_t at, size_t bytes) { if (dst && bytes) std::memcpy(dst, S + (at - upload_end), bytes); };
    back(out->used, o_used, 8 * ER); back(out->used_mask, o_umask, 4 * (size_t)E);
    back(out->agg_min, o_amin, 8 * (size_t)R); back(out->agg_used, o_aused, 8 * (size_t)R); back(out->agg_overquota, o_aover, 8 * (size_t)R);
    back(out->agg_masks, o_amask, 12); back(out->guaranteed, o_guar, 8 * ER); back(out->prefilter, o_pref, (size_t)Pp);
    if (want_label) { back(out->over_quota, o_over, (size_t)Pr); back(out->operator_used, o_opused, 8 * ER); back(out->ambiguous_order, o_amb, 4); }
    else {
        if (out->over_quota && Pr) std::memset(out->over_quota, 0, (size_t)Pr);
        if (out->operator_used && ER) std::memset(out->operator_used, 0, 8 * ER);
        if (out->ambiguous_order) *out->ambiguous_order = 0;
    }
    long long d2h_extra = 0;
    int rc_changes = NOSPLAN_OK;
    if (lc) {
        int32_t n = 0;
        if (want_changes) std::memcpy(&n, S + (o_nsel - upload_end), 4);
        *lc->n_changes = n;
        const int take = std::min(n, lc->capacity);
        if (take > 0) {
            unsigned char* L = S + (out_end - upload_end);     // behind the first copy's landing area
            QCU(cudaMemcpyAsync(L, A + o_sel, 4 * (size_t)take, cudaMemcpyDeviceToHost, st));
            QCU(cudaStreamSynchronize(st));
            const int32_t* packed = (const int32_t*)L;
            for (int k = 0; k < take; ++k) {
                lc->pod[k] = packed[k] >> 1;
                lc->desired[k] = (packed[k] & 1) ? NOSPLAN_LABEL_OVER_QUOTA : NOSPLAN_LABEL_IN_QUOTA;
            }
            d2h_extra = 4 * (long long)take;
        }
        if (n > lc->capacity) rc_changes = fail(NOSPLAN_E_LIMIT, "label changes: more pods to patch than `capacity` (n_changes holds the number)");
    }
    if (ms_h2d) cudaEventElapsedTime(ms_h2d, ev[0], ev[1]);
    if (ms_k) cudaEventElapsedTime(ms_k, ev[1], ev[2]);
    if (ms_d2h) cudaEventElapsedTime(ms_d2h, ev[2], ev[3]);
    if (launches) *launches = n_launch;
    if (h2d_bytes) *h2d_bytes = (long long)upload_end;
    if (d2h_bytes) *d2h_bytes = (long long)(d_hi > d_lo ? d_hi - d_lo : 0) + d2h_extra;
    return rc_changes;
#undef QCU
}

namespace {
struct QuotaResident {
    bool valid = false;
    DevQuota q;
    int capacity = 0;
    size_t o_used = 0, o_umask = 0, o_amin = 0, o_aused = 0, o_aover = 0, o_amask = 0, o_guar = 0, o_present = 0, o_applied = 0, total = 0;
    std::vector<uint64_t> batch_seen;                        // host: pods of the batch being checked
};
void quota_resident_free(void* p) { delete static_cast<QuotaResident*>(p); }
}  // namespace

// ctx: the handle's context for the resident quota state; nothing else uses it
extern "C" int nosplan_quota_state_begin_impl(int device, cudaStream_t st, nosplan_quota_ctx* ctx, std::string* err,
                                              const nosplan_quotas* qs, int32_t pod_capacity) {
    auto fail = [&](int code, const char* msg, const char* detail = "") { if (err) *err = std::string(msg) + detail; return code; };
#define SCU(call) do { cudaError_t e_ = (call); if (e_ != cudaSuccess) return fail(NOSPLAN_E_CUDA, #call ": ", cudaGetErrorString(e_)); } while (0)
    if (!qs) return fail(NOSPLAN_E_INVALID, "null quotas");
    const int E = qs->n_quotas, R = qs->n_res, NS = qs->n_namespaces;
    if (R < 4 || R > QR) return fail(NOSPLAN_E_LIMIT, "n_res must be in [4, NOSPLAN_QUOTA_MAX_RES]");
    if (E < 0 || NS < 0 || pod_capacity < 0) return fail(NOSPLAN_E_INVALID, "negative size");
    if (qs->gpu_memory_col >= R || (qs->gpu_memory_col >= 0 && qs->gpu_memory_col < 4)) return fail(NOSPLAN_E_INVALID, "gpu_memory_col out of range");
    for (int ns = 0; ns < NS; ++ns) if (qs->ns_quota[ns] >= E) return fail(NOSPLAN_E_INVALID, "ns_quota out of range");
    SCU(cudaSetDevice(device));
    if (!ctx->quota_resident) { ctx->quota_resident = new QuotaResident(); ctx->quota_resident_free = quota_resident_free; }
    QuotaResident& rs = *static_cast<QuotaResident*>(ctx->quota_resident);
    rs.valid = false;
    size_t o = 0;
    auto al = [](size_t x) { return (x + 255) & ~(size_t)255; };
    auto put = [&](size_t bytes) { size_t at = o; o = al(o + bytes); return at; };
    const size_t ER = (size_t)E * R, words = ((size_t)pod_capacity + 31) / 32;
    const size_t o_nsq = put(4 * (size_t)NS), o_min = put(8 * ER), o_max = put(8 * ER), o_mm = put(4 * (size_t)E), o_xm = put(4 * (size_t)E),
                 o_enf = put((size_t)E);
    rs.o_used = put(8 * ER); rs.o_umask = put(4 * (size_t)E);
    const size_t upload_end = o;
    rs.o_amin = put(8 * QR); rs.o_aused = put(8 * QR); rs.o_aover = put(8 * QR); rs.o_amask = put(16); rs.o_guar = put(8 * ER);
    rs.o_present = put(4 * words); rs.o_applied = put(16);
    rs.total = o;
    SCU(quota_ctx_ensure(ctx, rs.total, std::max(upload_end, rs.o_present)));
    unsigned char* S = ctx->hs; unsigned char* A = ctx->d;
    auto cp = [&](size_t at, const void* src, size_t bytes) { if (bytes) { if (src) std::memcpy(S + at, src, bytes); else std::memset(S + at, 0, bytes); } };
    cp(o_nsq, qs->ns_quota, 4 * (size_t)NS); cp(o_min, qs->min, 8 * ER); cp(o_max, qs->max, 8 * ER);
    cp(o_mm, qs->min_mask, 4 * (size_t)E); cp(o_xm, qs->max_mask, 4 * (size_t)E); cp(o_enf, qs->max_enforced, (size_t)E);
    cp(rs.o_used, qs->used_init, 8 * ER); cp(rs.o_umask, qs->used_mask, 4 * (size_t)E);
    for (int e = 0; e < E; ++e) reinterpret_cast<uint32_t*>(S + rs.o_umask)[e] &= SCALAR_BITS;
    SCU(cudaMemcpyAsync(A, S, upload_end, cudaMemcpyHostToDevice, st));
    SCU(cudaMemsetAsync(A + upload_end, 0, rs.total - upload_end, st));
    DevQuota& q = rs.q; std::memset(&q, 0, sizeof q);
    q.E = E; q.R = R; q.NS = NS; q.gm = qs->gpu_memory_col;
    q.ns_quota = (const int32_t*)(A + o_nsq); q.qmin = (const long long*)(A + o_min); q.qmax = (const long long*)(A + o_max);
    q.min_mask = (const uint32_t*)(A + o_mm); q.max_mask = (const uint32_t*)(A + o_xm); q.max_enforced = (const uint8_t*)(A + o_enf);
    for (int c = 0; c < QR; ++c) q.coeff[c] = (c < R && qs->gpu_memory_coeff) ? qs->gpu_memory_coeff[c] : 0;
    q.used = (long long*)(A + rs.o_used); q.used_mask = (uint32_t*)(A + rs.o_umask);
    q.agg_min = (long long*)(A + rs.o_amin); q.agg_used = (long long*)(A + rs.o_aused); q.agg_over = (long long*)(A + rs.o_aover);
    q.agg_masks = (uint32_t*)(A + rs.o_amask); q.guaranteed = (long long*)(A + rs.o_guar);
    SCU(cudaStreamSynchronize(st));
    rs.capacity = pod_capacity;
    rs.batch_seen.assign(((size_t)pod_capacity + 63) / 64, 0);
    rs.valid = true;
    return NOSPLAN_OK;
#undef SCU
}

// ctx: the resident state; scratch: the handle's ordinary quota context, used to carry the events in and the outputs back
extern "C" int nosplan_quota_state_apply_impl(int device, cudaStream_t st, nosplan_quota_ctx* ctx, nosplan_quota_ctx* scratch, std::string* err,
                                              const nosplan_quota_events* evs, nosplan_quota_out* out, int32_t* n_applied,
                                              float* ms_k, long long* h2d_bytes, long long* d2h_bytes) {
    auto fail = [&](int code, const char* msg, const char* detail = "") { if (err) *err = std::string(msg) + detail; return code; };
#define SCU(call) do { cudaError_t e_ = (call); if (e_ != cudaSuccess) return fail(NOSPLAN_E_CUDA, #call ": ", cudaGetErrorString(e_)); } while (0)
    QuotaResident* rp = ctx ? static_cast<QuotaResident*>(ctx->quota_resident) : nullptr;
    if (!rp || !rp->valid) return fail(NOSPLAN_E_INVALID, "nosplan_quota_state_apply: no quota state resident on this handle (call nosplan_quota_state_begin first)");
    QuotaResident& rs = *rp;
    const int n = evs ? evs->n_events : 0;
    if (n < 0) return fail(NOSPLAN_E_INVALID, "negative event count");
    if (n > 0 && (!evs->op || !evs->pod_id || !evs->ns || !evs->request)) return fail(NOSPLAN_E_INVALID, "op, pod_id, ns and request are required");
    const int E = rs.q.E, R = rs.q.R;
    // a pod at most once per batch (the events of a batch are applied in no particular order), ids inside the capacity
    int bad = -1; const char* why = "";
    int checked = 0;
    for (; checked < n; ++checked) {
        const int id = evs->pod_id[checked];
        if (id < 0 || id >= rs.capacity) { bad = checked; why = "pod_id outside [0, pod_capacity)"; break; }
        if (evs->op[checked] > NOSPLAN_QUOTA_EVENT_DELETE) { bad = checked; why = "unknown op"; break; }
        uint64_t& w = rs.batch_seen[(size_t)id >> 6];
        const uint64_t bit = 1ull << (id & 63);
        if (w & bit) { bad = checked; why = "the same pod twice in one batch"; break; }
        w |= bit;
    }
    for (int i = 0; i < checked; ++i) rs.batch_seen[(size_t)evs->pod_id[i] >> 6] = 0;
    if (bad >= 0) { if (err) *err = std::string("nosplan_quota_state_apply: event ") + std::to_string(bad) + ": " + why; return NOSPLAN_E_INVALID; }
    SCU(cudaSetDevice(device));
    cudaEvent_t* ev = nullptr;
    SCU(quota_ctx_events(ctx, &ev));
    size_t o = 0;
    auto al = [](size_t x) { return (x + 255) & ~(size_t)255; };
    auto put = [&](size_t bytes) { size_t at = o; o = al(o + bytes); return at; };
    const size_t o_op = put((size_t)n), o_id = put(4 * (size_t)n), o_ns = put(4 * (size_t)n), o_req = put(8 * (size_t)n * R), o_mask = put(4 * (size_t)n);
    const size_t ev_bytes = o;
    const size_t ER = (size_t)E * R;
    SCU(quota_ctx_ensure(scratch, ev_bytes + 256, std::max(ev_bytes, rs.total) + 512));
    unsigned char* S = scratch->hs; unsigned char* B = scratch->d; unsigned char* A = ctx->d;
    if (n) {
        std::memcpy(S + o_op, evs->op, (size_t)n); std::memcpy(S + o_id, evs->pod_id, 4 * (size_t)n); std::memcpy(S + o_ns, evs->ns, 4 * (size_t)n);
        std::memcpy(S + o_req, evs->request, 8 * (size_t)n * R);
        if (evs->request_mask) std::memcpy(S + o_mask, evs->request_mask, 4 * (size_t)n);
        SCU(cudaMemcpyAsync(B, S, ev_bytes, cudaMemcpyHostToDevice, st));
    }
    SCU(cudaMemsetAsync(A + rs.o_applied, 0, 16, st));
    SCU(cudaEventRecord(ev[1], st));
    if (n) {
        DevEvents de;
        de.n = n; de.op = B + o_op; de.pod_id = (const int32_t*)(B + o_id); de.ns = (const int32_t*)(B + o_ns);
        de.req = (const long long*)(B + o_req); de.mask = evs->request_mask ? (const uint32_t*)(B + o_mask) : nullptr;
        de.present = (unsigned*)(A + rs.o_present); de.applied = (int32_t*)(A + rs.o_applied);
        k_quota_events<<<(n + 15) / 16, 256, 0, st>>>(rs.q, de);
    }
    const bool want_agg = out && (out->agg_min || out->agg_used || out->agg_overquota || out->agg_masks || out->guaranteed);
    if (want_agg && launch_quota_aggregate(rs.q, st) != cudaSuccess) return fail(NOSPLAN_E_CUDA, "k_quota_aggregate");
    SCU(cudaGetLastError());
    SCU(cudaEventRecord(ev[2], st));
    // outputs: the span asked for, plus the counter
    size_t d_lo = rs.o_applied, d_hi = rs.o_applied + 16;
    auto span = [&](const void* wanted, size_t at, size_t bytes) { if (wanted && bytes) { d_lo = std::min(d_lo, at); d_hi = std::max(d_hi, at + bytes); } };
    if (out) {
        span(out->used, rs.o_used, 8 * ER); span(out->used_mask, rs.o_umask, 4 * (size_t)E); span(out->agg_min, rs.o_amin, 8 * QR);
        span(out->agg_used, rs.o_aused, 8 * QR); span(out->agg_overquota, rs.o_aover, 8 * QR); span(out->agg_masks, rs.o_amask, 16);
        span(out->guaranteed, rs.o_guar, 8 * ER);
    }
    // the span may bridge the presence bitmap (it lies between the rows and the counter): fetch the counter on its own then
    const bool split = d_lo < rs.o_present;
    const size_t first_hi = split ? std::min(d_hi, rs.o_present) : d_hi;
    if (split) {
        SCU(cudaMemcpyAsync(S + d_lo, A + d_lo, first_hi - d_lo, cudaMemcpyDeviceToHost, st));
        SCU(cudaMemcpyAsync(S + first_hi + 256, A + rs.o_applied, 16, cudaMemcpyDeviceToHost, st));
    } else SCU(cudaMemcpyAsync(S + d_lo, A + d_lo, d_hi - d_lo, cudaMemcpyDeviceToHost, st));
    SCU(cudaStreamSynchronize(st));
    auto back = [&](void* dst, size_t at, size_t bytes) { if (dst && bytes) std::memcpy(dst, S + at, bytes); };
    if (out) {
        back(out->used, rs.o_used, 8 * ER); back(out->used_mask, rs.o_umask, 4 * (size_t)E); back(out->agg_min, rs.o_amin, 8 * (size_t)R);
        back(out->agg_used, rs.o_aused, 8 * (size_t)R); back(out->agg_overquota, rs.o_aover, 8 * (size_t)R); back(out->agg_masks, rs.o_amask, 12);
        back(out->guaranteed, rs.o_guar, 8 * ER);
    }
    int32_t applied = 0;
    std::memcpy(&applied, S + (split ? first_hi + 256 : rs.o_applied), 4);
    if (n_applied) *n_applied = applied;
    if (ms_k) cudaEventElapsedTime(ms_k, ev[1], ev[2]);
    if (h2d_bytes) *h2d_bytes = (long long)(n ? ev_bytes : 0);
    if (d2h_bytes) *d2h_bytes = (long long)((split ? first_hi - d_lo + 16 : d_hi - d_lo));
    return NOSPLAN_OK;
#undef SCU
}

// the quota part of PreFilter (capacity_scheduling.go:190-278) for a batch of pending pods against the resident state
extern "C" int nosplan_quota_state_prefilter_impl(int device, cudaStream_t st, nosplan_quota_ctx* ctx, nosplan_quota_ctx* scratch, std::string* err,
                                                  const nosplan_quota_pods* pending, uint8_t* verdict, float* ms_k) {
    auto fail = [&](int code, const char* msg, const char* detail = "") { if (err) *err = std::string(msg) + detail; return code; };
#define SCU(call) do { cudaError_t e_ = (call); if (e_ != cudaSuccess) return fail(NOSPLAN_E_CUDA, #call ": ", cudaGetErrorString(e_)); } while (0)
    QuotaResident* rp = ctx ? static_cast<QuotaResident*>(ctx->quota_resident) : nullptr;
    if (!rp || !rp->valid) return fail(NOSPLAN_E_INVALID, "nosplan_quota_state_prefilter: no quota state resident on this handle (call nosplan_quota_state_begin first)");
    const int Pp = pending ? pending->n_pods : 0;
    if (Pp < 0) return fail(NOSPLAN_E_INVALID, "negative size");
    if (Pp == 0) return NOSPLAN_OK;
    if (!verdict || !pending->ns || !pending->request) return fail(NOSPLAN_E_INVALID, "ns, request and verdict are required");
    const int R = rp->q.R;
    SCU(cudaSetDevice(device));
    cudaEvent_t* ev = nullptr;
    SCU(quota_ctx_events(ctx, &ev));
    auto al = [](size_t x) { return (x + 255) & ~(size_t)255; };
    const size_t o_ns = 0, o_req = al(4 * (size_t)Pp), o_mask = o_req + al(8 * (size_t)Pp * R), o_out = o_mask + al(4 * (size_t)Pp), total = o_out + al((size_t)Pp);
    SCU(quota_ctx_ensure(scratch, total, total));
    unsigned char* S = scratch->hs; unsigned char* B = scratch->d;
    std::memcpy(S + o_ns, pending->ns, 4 * (size_t)Pp); std::memcpy(S + o_req, pending->request, 8 * (size_t)Pp * R);
    if (pending->request_mask) std::memcpy(S + o_mask, pending->request_mask, 4 * (size_t)Pp);
    SCU(cudaMemcpyAsync(B, S, o_out, cudaMemcpyHostToDevice, st));
    DevQuota q = rp->q;
    q.Pp = Pp; q.p_ns = (const int32_t*)(B + o_ns); q.p_req = (const long long*)(B + o_req);
    q.p_mask = pending->request_mask ? (const uint32_t*)(B + o_mask) : nullptr;
    q.prefilter = B + o_out;
    SCU(cudaEventRecord(ev[1], st));
    if (launch_quota_aggregate(q, st) != cudaSuccess) return fail(NOSPLAN_E_CUDA, "k_quota_aggregate");          // the aggregates of the state as it is now
    k_quota_prefilter<<<(Pp + 15) / 16, 256, 0, st>>>(q);
    SCU(cudaGetLastError());
    SCU(cudaEventRecord(ev[2], st));
    SCU(cudaMemcpyAsync(S + o_out, B + o_out, (size_t)Pp, cudaMemcpyDeviceToHost, st));
    SCU(cudaStreamSynchronize(st));
    std::memcpy(verdict, S + o_out, (size_t)Pp);
    if (ms_k) cudaEventElapsedTime(ms_k, ev[1], ev[2]);
    return NOSPLAN_OK;
#undef SCU
}

extern "C" void nosplan_quota_ctx_free(nosplan_quota_ctx* ctx) {
    if (!ctx) return;
    if (ctx->d) cudaFree(ctx->d);
    if (ctx->hs) cudaFreeHost(ctx->hs);
    for (auto& e : ctx->ev) if (e) cudaEventDestroy(e);
    if (ctx->preempt_resident && ctx->preempt_resident_free) ctx->preempt_resident_free(ctx->preempt_resident);
    if (ctx->quota_resident && ctx->quota_resident_free) ctx->quota_resident_free(ctx->quota_resident);
    delete ctx;
}
extern "C" nosplan_quota_ctx* nosplan_quota_ctx_new() { return new nosplan_quota_ctx(); }
