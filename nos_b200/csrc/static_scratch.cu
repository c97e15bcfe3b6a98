// nosplan_lacking_static -- the SliceTracker initialisation (lacking slices per profile, and how many distinct pods
// lack anything) for callers that must not allocate device memory.
//
// Who that is: a planner that shares its GPU with latency-sensitive work (an inference server on the same device,
// another controller's streams).  cudaMalloc / cudaFree are device-wide synchronisation points, so an allocation on
// the planning path stalls every other stream on the GPU; the same holds for a process that was handed a fixed
// device-memory reservation and is not allowed to grow it.  Such a caller keeps its packed pods resident in buffers
// it already owns and passes DEVICE pointers here; the only working memory the computation needs -- the per-pod-key
// table of lacking quantities that nosplan_run keeps in its arena -- lives in the fixed array below instead.
//
// What to know before using it:
//   * the table is ONE array per GPU for the whole process.  Calls that can overlap on the same GPU (two streams,
//     two threads) race on it; serialise them.  It keeps the previous call's contents, so every call first clears the
//     rows it is about to use, and clears the two outputs, on the caller's stream;
//   * at most NOSPLAN_STATIC_MAX_KEYS distinct pod keys; more is refused with NOSPLAN_E_LIMIT, nothing is launched;
//   * the pods are device data, so bad entries can only be caught on the GPU: a pod whose tracker key is outside
//     [0, n_keys) or whose scalar column is below 3 is skipped, and *d_lacking_entries comes back NEGATIVE (sign bit
//     set) to say that d_lack is incomplete;
//   * the first call on a given GPU looks up the device address of the array (cudaGetSymbolAddress, with that GPU
//     current).  Make that first call outside any stream capture;
//   * nothing here calls cudaMalloc, cudaFree, cudaMallocHost or any library routine that allocates; the work is
//     enqueued on `cuda_stream` and the call returns without synchronising: read the outputs after the stream has
//     finished.  The calling thread's current device is left set to `cuda_device`.
#include <climits>
#include <cstdio>
#include <mutex>

#include "nosplan_dev.cuh"

namespace {

constexpr int MAX_DEVICES = 64;

// [NOSPLAN_STATIC_MAX_KEYS][16] int32 = 1 MiB, held by every GPU that loads the library until the process exits.
// In HOST code this name is only a handle for the runtime, never an address a kernel may dereference.
static __device__ __align__(256) int32_t g_static_entry_lack[NOSPLAN_STATIC_MAX_KEYS * nsp::KP];

std::mutex g_setup_mutex;
int32_t* g_scratch_addr[MAX_DEVICES] = {};      // device address of g_static_entry_lack on each GPU, set once

// same arithmetic as k_tracker_init (prepass.cu): lacking = request - max(0, cluster free total), NewSliceTracker
// (internal/partitioning/core/tracker.go:34-57); `entry` is the caller-invisible table in static device memory
// The pods are DEVICE data the host cannot inspect, so the kernel itself refuses what nosplan_upload refuses on the host:
// a tracker key outside [0, n_keys) would land in rows this call did not clear (or past the table), a column below 3 is
// not a scalar column.  Such a pod is skipped and the sign bit of *bad_out (the entries counter) is raised.
__global__ void __launch_bounds__(256) k_lacking_static(int32_t* entry, int n_pods, int n_keys, int K, const int32_t* pod_key,
                                                        const int32_t* sptr, const int32_t* scol, const long long* sqty,
                                                        const long long* tot, int32_t* lack, int32_t* bad_out) {
    __shared__ int s_lack[nsp::KP];
    if (threadIdx.x < nsp::KP) s_lack[threadIdx.x] = 0;
    __syncthreads();
    const int s = blockIdx.x * blockDim.x + threadIdx.x;
    if (s < n_pods) {
        const int key = pod_key[s];
        const bool key_ok = (unsigned)key < (unsigned)n_keys;
        if (!key_ok) atomicOr(bad_out, INT_MIN);
        for (int e = sptr[s]; key_ok && e < sptr[s + 1]; ++e) {
            const int col = scol[e];
            if (col >= 3 + K) break;
            if (col < 3) { atomicOr(bad_out, INT_MIN); break; }
            const int p = col - 3;
            long long avail = tot[p];
            if (avail < 0) avail = 0;
            const long long diff = avail - sqty[e];
            if (diff < 0) {
                const int l = (int)(-diff);
                atomicAdd(&entry[(size_t)key * nsp::KP + p], l);
                atomicAdd(&s_lack[p], l);
            }
        }
    }
    __syncthreads();
    if (threadIdx.x < nsp::KP && s_lack[threadIdx.x] != 0) atomicAdd(&lack[threadIdx.x], s_lack[threadIdx.x]);
}

// number of tracker entries (distinct namespace/name) that lack at least one slice
__global__ void __launch_bounds__(256) k_count_lacking_entries(const int32_t* entry, int n_keys, int32_t* count) {
    const int k = blockIdx.x * blockDim.x + threadIdx.x;
    bool any = false;
    if (k < n_keys) {
        const int4* row = reinterpret_cast<const int4*>(entry + (size_t)k * nsp::KP);
#pragma unroll
        for (int i = 0; i < 4; ++i) { const int4 v = row[i]; any |= (v.x | v.y | v.z | v.w) != 0; }
    }
    const int n = __syncthreads_count(any);
    if (threadIdx.x == 0 && n) atomicAdd(count, n);
}

int fail(char* errbuf, int errbuf_len, int code, const char* msg, const char* detail = "") {
    if (errbuf && errbuf_len > 0) snprintf(errbuf, (size_t)errbuf_len, "%s%s", msg, detail);
    return code;
}

}  // namespace

extern "C" int nosplan_lacking_static(int cuda_device, void* cuda_stream, int32_t n_pods, int32_t n_keys, int32_t n_profiles,
                                      const int32_t* d_tracker_key, const int32_t* d_scalar_ptr, const int32_t* d_scalar_col,
                                      const int64_t* d_scalar_qty, const int64_t* d_totals, int32_t* d_lack,
                                      int32_t* d_lacking_entries, char* errbuf, int32_t errbuf_len) {
    if (n_pods < 0 || n_keys < 0 || n_profiles < 0 || n_profiles > NOSPLAN_MAX_PROFILES || cuda_device < 0 || cuda_device >= MAX_DEVICES)
        return fail(errbuf, errbuf_len, NOSPLAN_E_INVALID, "nosplan_lacking_static: size or device out of range");
    if (n_keys > NOSPLAN_STATIC_MAX_KEYS) {
        char d[96];
        snprintf(d, sizeof d, " (%d keys requested, NOSPLAN_STATIC_MAX_KEYS = %d)", n_keys, NOSPLAN_STATIC_MAX_KEYS);
        return fail(errbuf, errbuf_len, NOSPLAN_E_LIMIT,
                    "nosplan_lacking_static: more distinct pod keys than the static scratch table holds; use nosplan_plan", d);
    }
    if (!d_lack || !d_lacking_entries || !d_totals || (n_pods > 0 && (!d_tracker_key || !d_scalar_ptr)))
        return fail(errbuf, errbuf_len, NOSPLAN_E_INVALID, "nosplan_lacking_static: null device pointer");
    cudaError_t e = cudaSetDevice(cuda_device);
    if (e != cudaSuccess) return fail(errbuf, errbuf_len, NOSPLAN_E_CUDA, "cudaSetDevice: ", cudaGetErrorString(e));

    // set-up, once per GPU, with that GPU current: host code needs the DEVICE address of the array
    int32_t* scratch = nullptr;
    {
        std::lock_guard<std::mutex> lock(g_setup_mutex);
        if (!g_scratch_addr[cuda_device]) {
            void* p = nullptr;
            e = cudaGetSymbolAddress(&p, g_static_entry_lack);
            if (e != cudaSuccess || !p) return fail(errbuf, errbuf_len, NOSPLAN_E_CUDA, "cudaGetSymbolAddress: ", cudaGetErrorString(e));
            g_scratch_addr[cuda_device] = static_cast<int32_t*>(p);
        }
        scratch = g_scratch_addr[cuda_device];
    }

    cudaStream_t st = static_cast<cudaStream_t>(cuda_stream);
    // the table holds whatever the previous call left: clear the rows this call accumulates into, and the outputs
    if (n_keys > 0) {
        e = cudaMemsetAsync(scratch, 0, sizeof(int32_t) * nsp::KP * (size_t)n_keys, st);
        if (e != cudaSuccess) return fail(errbuf, errbuf_len, NOSPLAN_E_CUDA, "cudaMemsetAsync(scratch): ", cudaGetErrorString(e));
    }
    e = cudaMemsetAsync(d_lack, 0, sizeof(int32_t) * nsp::KP, st);
    if (e == cudaSuccess) e = cudaMemsetAsync(d_lacking_entries, 0, sizeof(int32_t), st);
    if (e != cudaSuccess) return fail(errbuf, errbuf_len, NOSPLAN_E_CUDA, "cudaMemsetAsync(outputs): ", cudaGetErrorString(e));
    if (n_pods > 0)
        k_lacking_static<<<(n_pods + 255) / 256, 256, 0, st>>>(scratch, n_pods, n_keys, n_profiles, d_tracker_key, d_scalar_ptr, d_scalar_col,
                                                              reinterpret_cast<const long long*>(d_scalar_qty),
                                                              reinterpret_cast<const long long*>(d_totals), d_lack, d_lacking_entries);
    if (n_keys > 0) k_count_lacking_entries<<<(n_keys + 255) / 256, 256, 0, st>>>(scratch, n_keys, d_lacking_entries);
    e = cudaGetLastError();
    if (e != cudaSuccess) return fail(errbuf, errbuf_len, NOSPLAN_E_CUDA, "kernel launch: ", cudaGetErrorString(e));
    return NOSPLAN_OK;
}
