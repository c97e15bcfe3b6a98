// Snapshot packer, device side: node status annotations -> per-GPU used / free rows of the packed snapshot.
//
// Replaces, for N nodes at once: gpu.ParseNodeAnnotations (pkg/gpu/annotation.go:87-104) with ParseSpecAnnotation /
// ParseStatusAnnotation (:29-85), StatusAnnotationList.GroupByGpuIndex (:186-196) and the GPU construction loops of
// mig.extractGPUs (pkg/gpu/mig/node.go:65-100) / slicing.extractGPUs (pkg/gpu/slicing/node.go:48-105).
//
//   k_parse_annotations   one thread per annotation: prefix test, strings.Split(key, "-") part count, strconv.Atoi of the
//                         value and of the GPU index, resource.ParseStatus (case-insensitive free / used / unknown),
//                         profile name -> id through the caller's dictionary
//   k_build_gpus          one warp per node: distinct annotated GPU indices in ascending order (the canonical resolution of
//                         the reference's map iteration) take positions 0.., GPUs without annotations are
//                         appended with index = position up to the nvidia.com/gpu.count label; counts and key-presence
//                         masks are scattered into the [N*G][16] rows
#include <climits>
#include <cstring>
#include <string>

#include "nosplan_dev.cuh"
#include "quota_ctx.cuh"

namespace nsp {

constexpr unsigned A_FULL = 0xffffffffu;
constexpr int ST_FREE = 0, ST_USED = 1, ST_UNKNOWN = 2;

struct DevAnn {
    int N, A, K, G;
    const int32_t* node_ptr; const int32_t* key_off; const char* keys; const int32_t* val_off; const char* vals;
    const int32_t* gpu_count; const int32_t* prof_off; const char* profs;
    int32_t* rec_idx;        // [A] GPU index of a valid status annotation, INT_MIN otherwise
    int32_t* rec_info;       // [A] status << 16 | profile id << 8 | quantity
    uint8_t* used; uint8_t* free_; uint16_t* umask; uint16_t* fmask; int32_t* gindex; int32_t* ngpu;
    int32_t* errors;         // [4] unknown profile names, quantities outside [0, 255], nodes with more GPUs than G, indices beyond int32
};

__device__ __forceinline__ bool a_prefix(const char* s, int n, const char* p, int pn) {
    if (n < pn) return false;
    for (int i = 0; i < pn; ++i) if (s[i] != p[i]) return false;
    return true;
}

// strconv.Atoi: optional sign, then decimal digits only; anything else (empty, blanks, underscores) is an error.
// ok = parsed and within int64; the value saturates far outside the ranges that matter here.
__device__ __forceinline__ bool a_atoi(const char* s, int n, long long& out) {
    int i = 0;
    bool neg = false;
    if (n > 0 && (s[0] == '+' || s[0] == '-')) { neg = s[0] == '-'; i = 1; }
    if (i >= n) return false;
    unsigned long long v = 0;
    for (; i < n; ++i) {
        const unsigned d = (unsigned)(s[i] - '0');
        if (d > 9) return false;
        if (v > (ULLONG_MAX - d) / 10) return false;                       // beyond uint64: out of range for Atoi too
        v = v * 10 + d;
    }
    if (neg ? v > 9223372036854775808ull : v > 9223372036854775807ull) return false;   // ErrRange
    out = neg ? -(long long)v : (long long)v;
    return true;
}

__device__ __forceinline__ char a_lower(char c) { return (c >= 'A' && c <= 'Z') ? (char)(c - 'A' + 'a') : c; }
__device__ __forceinline__ bool a_ieq(const char* s, int n, const char* w, int wn) {
    if (n != wn) return false;
    for (int i = 0; i < n; ++i) if (a_lower(s[i]) != w[i]) return false;
    return true;
}

__global__ void __launch_bounds__(256) k_parse_annotations(DevAnn d) {
    const int a = blockIdx.x * blockDim.x + threadIdx.x;
    if (a >= d.A) return;
    const char* key = d.keys + d.key_off[a];
    const int kn = d.key_off[a + 1] - d.key_off[a];
    const char* val = d.vals + d.val_off[a];
    const int vn = d.val_off[a + 1] - d.val_off[a];
    int idx = INT_MIN, info = 0;
    // a key with the spec prefix is never a status annotation (the prefixes differ), whatever ParseSpecAnnotation decides
    if (a_prefix(key, kn, "nos.nebuly.com/status-gpu", 25)) {
        int dash[5], nd = 0;
        for (int i = 0; i < kn; ++i) if (key[i] == '-') { if (nd < 5) dash[nd] = i; ++nd; }
        long long q = 0, gi = 0;
        if (nd == 4 && a_atoi(val, vn, q) && a_atoi(key + dash[1] + 1, dash[2] - dash[1] - 1, gi)) {
            const char* st = key + dash[3] + 1;
            const int sn = kn - dash[3] - 1;
            int status = -1;
            if (a_ieq(st, sn, "free", 4)) status = ST_FREE;
            else if (a_ieq(st, sn, "used", 4)) status = ST_USED;
            else if (a_ieq(st, sn, "unknown", 7)) status = ST_UNKNOWN;
            if (status >= 0) {
                if (gi > INT_MAX || gi <= INT_MIN) atomicAdd(&d.errors[3], 1);
                else {
                    idx = (int)gi;
                    int pid = 255, qty = 0;
                    if (status != ST_UNKNOWN) {                            // only used / free entries put a profile into a map
                        const char* pn = key + dash[2] + 1;
                        const int pl = dash[3] - dash[2] - 1;
                        for (int k = 0; k < d.K && pid == 255; ++k) {
                            const int o = d.prof_off[k], l = d.prof_off[k + 1] - o;
                            if (l != pl) continue;
                            bool eq = true;
                            for (int i = 0; i < l && eq; ++i) eq = d.profs[o + i] == pn[i];
                            if (eq) pid = k;
                        }
                        if (pid == 255) atomicAdd(&d.errors[0], 1);
                        if (q < 0 || q > 255) { atomicAdd(&d.errors[1], 1); q = 0; }
                        qty = (int)q;
                    }
                    info = (status << 16) | (pid << 8) | qty;
                }
            }
        }
    }
    d.rec_idx[a] = idx;
    d.rec_info[a] = info;
}

__global__ void __launch_bounds__(256) k_build_gpus(DevAnn d) {
    __shared__ int s_list[8][NOSPLAN_MAX_GPUS];
    __shared__ unsigned s_um[8][NOSPLAN_MAX_GPUS], s_fm[8][NOSPLAN_MAX_GPUS];
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const int warps_total = gridDim.x * (blockDim.x >> 5);
    for (int i = blockIdx.x * (blockDim.x >> 5) + warp; i < d.N; i += warps_total) {
        const int a0 = d.node_ptr[i], a1 = d.node_ptr[i + 1];
        // distinct annotated GPU indices, ascending
        int n_ann = 0;
        long long cur = LLONG_MIN;
        bool too_many = false;
        for (;;) {
            long long m = LLONG_MAX;
            for (int a = a0 + lane; a < a1; a += 32) {
                const int x = d.rec_idx[a];
                if (x != INT_MIN && (long long)x > cur) m = min(m, (long long)x);
            }
#pragma unroll
            for (int o = 16; o; o >>= 1) m = min(m, __shfl_xor_sync(A_FULL, m, o));
            if (m == LLONG_MAX) break;
            if (n_ann == d.G) { too_many = true; break; }
            if (lane == 0) s_list[warp][n_ann] = (int)m;
            ++n_ann; cur = m;
        }
        int total = max(n_ann, d.gpu_count[i]);                              // missing GPUs are appended (node.go:88-97)
        if (too_many || total > d.G) { if (lane == 0) atomicAdd(&d.errors[2], 1); total = min(total, d.G); }
        if (lane < NOSPLAN_MAX_GPUS) { s_um[warp][lane] = 0; s_fm[warp][lane] = 0; }
        __syncwarp();
        for (int a = a0 + lane; a < a1; a += 32) {
            const int x = d.rec_idx[a];
            if (x == INT_MIN) continue;
            const int info = d.rec_info[a], status = info >> 16, pid = (info >> 8) & 255, qty = info & 255;
            if (status == ST_UNKNOWN || pid == 255) continue;
            int pos = -1;
            for (int k = 0; k < n_ann; ++k) if (s_list[warp][k] == x) pos = k;
            if (pos < 0) continue;                                           // index beyond the G annotated GPUs kept
            const size_t row = ((size_t)i * d.G + pos) * 16 + pid;
            if (status == ST_USED) { d.used[row] = (uint8_t)qty; atomicOr(&s_um[warp][pos], 1u << pid); }
            else { d.free_[row] = (uint8_t)qty; atomicOr(&s_fm[warp][pos], 1u << pid); }
        }
        __syncwarp();
        if (lane < total) {
            const size_t x = (size_t)i * d.G + lane;
            d.umask[x] = (uint16_t)s_um[warp][lane]; d.fmask[x] = (uint16_t)s_fm[warp][lane];
            d.gindex[x] = lane < n_ann ? s_list[warp][lane] : lane;
        }
        if (lane == 0) d.ngpu[i] = total;
        __syncwarp();
    }
}

}  // namespace nsp

using namespace nsp;

extern "C" int nosplan_parse_annotations_impl(int device, cudaStream_t st, nosplan_quota_ctx* ctx, std::string* err,
                                              const nosplan_annotations* in, nosplan_parsed_gpus* out, float* ms_h2d, float* ms_k,
                                              float* ms_d2h, long long* h2d_bytes, long long* d2h_bytes) {
    auto fail = [&](int code, const char* msg, const char* detail = "") { if (err) *err = std::string(msg) + detail; return code; };
#define ACU(call) do { cudaError_t e_ = (call); if (e_ != cudaSuccess) return fail(NOSPLAN_E_CUDA, #call ": ", cudaGetErrorString(e_)); } while (0)
    if (!in || !out) return fail(NOSPLAN_E_INVALID, "null argument");
    const int N = in->n_nodes, K = in->n_profiles, G = in->gpus_per_node;
    if (N < 0 || K < 0) return fail(NOSPLAN_E_INVALID, "negative size");
    if (K > NOSPLAN_MAX_PROFILES) return fail(NOSPLAN_E_LIMIT, "more than NOSPLAN_MAX_PROFILES profile names");
    if (G < 1 || G > NOSPLAN_MAX_GPUS) return fail(NOSPLAN_E_LIMIT, "gpus_per_node must be in [1, NOSPLAN_MAX_GPUS]");
    if (N > 0 && (!in->node_ptr || !in->gpu_count)) return fail(NOSPLAN_E_INVALID, "null node arrays");
    const int A = N > 0 ? in->node_ptr[N] : 0;
    if (A < 0 || (N > 0 && in->node_ptr[0] != 0)) return fail(NOSPLAN_E_INVALID, "node_ptr must start at 0 and be non-negative");
    if (A > 0 && (!in->key_off || !in->key_bytes || !in->val_off || !in->val_bytes)) return fail(NOSPLAN_E_INVALID, "null annotation arrays");
    if (K > 0 && (!in->profile_off || !in->profile_bytes)) return fail(NOSPLAN_E_INVALID, "null profile dictionary");
    const size_t kb = A ? (size_t)in->key_off[A] : 0, vb = A ? (size_t)in->val_off[A] : 0, pb = K ? (size_t)in->profile_off[K] : 0;
    ACU(cudaSetDevice(device));
    size_t o = 0;
    auto al = [](size_t x) { return (x + 255) & ~(size_t)255; };
    auto put = [&](size_t bytes) { size_t at = o; o = al(o + bytes); return at; };
    const size_t NG = (size_t)N * G;
    const size_t o_ptr = put(4 * ((size_t)N + 1)), o_koff = put(4 * ((size_t)A + 1)), o_keys = put(kb), o_voff = put(4 * ((size_t)A + 1)),
                 o_vals = put(vb), o_cnt = put(4 * (size_t)N), o_poff = put(4 * ((size_t)K + 1)), o_prof = put(pb);
    const size_t upload_end = o;
    const size_t o_used = put(16 * NG), o_free = put(16 * NG), o_um = put(2 * NG), o_fm = put(2 * NG), o_gi = put(4 * NG), o_ng = put(4 * (size_t)N),
                 o_err = put(16);
    const size_t out_end = o;
    const size_t o_ridx = put(4 * (size_t)A), o_rinfo = put(4 * (size_t)A);
    const size_t total = o;
    ACU(quota_ctx_ensure(ctx, total, std::max(upload_end, out_end - upload_end)));
    unsigned char* S = ctx->hs; unsigned char* D = ctx->d;
    auto cp = [&](size_t at, const void* src, size_t bytes) { if (bytes) std::memcpy(S + at, src, bytes); };
    if (N > 0) cp(o_ptr, in->node_ptr, 4 * ((size_t)N + 1)); else std::memset(S + o_ptr, 0, 4);
    if (A > 0) { cp(o_koff, in->key_off, 4 * ((size_t)A + 1)); cp(o_keys, in->key_bytes, kb); cp(o_voff, in->val_off, 4 * ((size_t)A + 1)); cp(o_vals, in->val_bytes, vb); }
    cp(o_cnt, in->gpu_count, 4 * (size_t)N);
    if (K > 0) { cp(o_poff, in->profile_off, 4 * ((size_t)K + 1)); cp(o_prof, in->profile_bytes, pb); }
    cudaEvent_t* ev = nullptr;
    ACU(quota_ctx_events(ctx, &ev));
    ACU(cudaEventRecord(ev[0], st));
    ACU(cudaMemcpyAsync(D, S, upload_end, cudaMemcpyHostToDevice, st));
    ACU(cudaMemsetAsync(D + upload_end, 0, out_end - upload_end, st));
    ACU(cudaEventRecord(ev[1], st));
    DevAnn d; std::memset(&d, 0, sizeof d);
    d.N = N; d.A = A; d.K = K; d.G = G;
    d.node_ptr = (const int32_t*)(D + o_ptr); d.key_off = (const int32_t*)(D + o_koff); d.keys = (const char*)(D + o_keys);
    d.val_off = (const int32_t*)(D + o_voff); d.vals = (const char*)(D + o_vals); d.gpu_count = (const int32_t*)(D + o_cnt);
    d.prof_off = (const int32_t*)(D + o_poff); d.profs = (const char*)(D + o_prof);
    d.rec_idx = (int32_t*)(D + o_ridx); d.rec_info = (int32_t*)(D + o_rinfo);
    d.used = D + o_used; d.free_ = D + o_free; d.umask = (uint16_t*)(D + o_um); d.fmask = (uint16_t*)(D + o_fm);
    d.gindex = (int32_t*)(D + o_gi); d.ngpu = (int32_t*)(D + o_ng); d.errors = (int32_t*)(D + o_err);
    if (A > 0) k_parse_annotations<<<(A + 255) / 256, 256, 0, st>>>(d);
    if (N > 0) {
        int blocks = (N + 7) / 8;
        if (blocks > 148 * 8) blocks = 148 * 8;
        k_build_gpus<<<blocks, 256, 0, st>>>(d);
    }
    ACU(cudaGetLastError());
    ACU(cudaEventRecord(ev[2], st));
    ACU(cudaMemcpyAsync(S, D + upload_end, out_end - upload_end, cudaMemcpyDeviceToHost, st));
    ACU(cudaEventRecord(ev[3], st));
    ACU(cudaStreamSynchronize(st));
    if (ms_h2d) cudaEventElapsedTime(ms_h2d, ev[0], ev[1]);
    if (ms_k) cudaEventElapsedTime(ms_k, ev[1], ev[2]);
    if (ms_d2h) cudaEventElapsedTime(ms_d2h, ev[2], ev[3]);
    if (h2d_bytes) *h2d_bytes = (long long)upload_end;
    if (d2h_bytes) *d2h_bytes = (long long)(out_end - upload_end);
    auto back = [&](void* dst, size_t at, size_t bytes) { if (dst && bytes) std::memcpy(dst, S + (at - upload_end), bytes); };
    back(out->gpu_used, o_used, 16 * NG); back(out->gpu_free, o_free, 16 * NG);
    back(out->gpu_used_mask, o_um, 2 * NG); back(out->gpu_free_mask, o_fm, 2 * NG);
    back(out->gpu_index, o_gi, 4 * NG); back(out->node_gpu_count, o_ng, 4 * (size_t)N);
    int32_t errs[4]; std::memcpy(errs, S + (o_err - upload_end), sizeof errs);
    if (out->errors) std::memcpy(out->errors, errs, sizeof errs);
    if (errs[2]) return fail(NOSPLAN_E_LIMIT, "a node has more GPUs than gpus_per_node");
    if (errs[1]) return fail(NOSPLAN_E_LIMIT, "a device count is outside [0, 255]");
    if (errs[3]) return fail(NOSPLAN_E_LIMIT, "a GPU index does not fit 32 bits");
    if (errs[0]) return fail(NOSPLAN_E_INVALID, "an annotation names a profile that is not in the dictionary");
    return NOSPLAN_OK;
#undef ACU
}
