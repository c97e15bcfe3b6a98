// nosplan_changed_nodes -- change detection for the actuator on the device.
// Actuator.Apply (internal/partitioning/core/actuator.go:39-66) compares the desired PartitioningState with the snapshot's
// current one as a whole (state/partitioning.go:40-56, NodePartitioning.Equal :33-38: GPUs as an unordered set of
// (GPUIndex, resource -> count map), zero-count keys included) and then patches EVERY node of the plan.  Here the comparison is
// made per node on the plan rows the planning kernels left in HBM, so that a caller patches only the nodes that differ:
// node i changed <=> for some GPU g < gpu_count[i]:  out_count[g] != used[g] + free[g]  or  out_mask[g] != umask[g] | fmask[g]
// (GPU g keeps its GPUIndex, so the unordered comparison is the element-wise one).  One thread per node, then one
// cub::DeviceSelect::Flagged over a counting iterator: indices ascending = name order.
#include <string>

#include <cub/device/device_select.cuh>
#include <thrust/iterator/counting_iterator.h>

#include "nosplan_dev.cuh"
#include "quota_ctx.cuh"
#include "handle.cuh"

namespace nsp {

__global__ void __launch_bounds__(256) k_changed_flags(DevProblem pb, uint8_t* flag) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= pb.N) return;
    const int ng = min(max(pb.node_gpu_count[i], 0), pb.G);
    bool changed = false;
    for (int g = 0; g < ng; ++g) {
        const size_t x = (size_t)i * pb.G + g;
        const uint4 cur = add_bytes(pb.gpu_used[x], pb.gpu_free[x]);
        const uint4 out = pb.out_count[x];
        changed |= cur.x != out.x || cur.y != out.y || cur.z != out.z || cur.w != out.w;
        changed |= pb.out_mask[x] != (uint16_t)(pb.gpu_umask[x] | pb.gpu_fmask[x]);
    }
    flag[i] = changed ? 1 : 0;
}

}  // namespace nsp

extern "C" nosplan_quota_ctx* nosplan_quota_ctx_new();

extern "C" int nosplan_changed_nodes(nosplan_t* h, int32_t* node_idx, int32_t capacity, int32_t* n_changed) {
    using namespace nsp;
    if (!h) return NOSPLAN_E_INVALID;
    if (!h->ran) return nosplan_fail(h, NOSPLAN_E_INVALID, "nosplan_changed_nodes before a Plan() on this handle");
    if (!n_changed || capacity < 0 || (capacity > 0 && !node_idx)) return nosplan_fail(h, NOSPLAN_E_INVALID, "nosplan_changed_nodes: null output");
#define DCU(call) do { cudaError_t e_ = (call); if (e_ != cudaSuccess) return nosplan_fail(h, NOSPLAN_E_CUDA, "%s: %s", #call, cudaGetErrorString(e_)); } while (0)
    DCU(cudaSetDevice(h->device));
    { const int rc = nosplan_resolve(h); if (rc != NOSPLAN_OK) return rc; }   // a single-launch plan whose verdict is still on the device
    const int N = h->pb.N;
    *n_changed = 0;
    if (N == 0) return NOSPLAN_OK;
    if (!h->qctx) h->qctx = nosplan_quota_ctx_new();
    auto al = [](size_t x) { return (x + 255) & ~(size_t)255; };
    size_t cub_bytes = 0;
    thrust::counting_iterator<int32_t> ids(0);
    DCU(cub::DeviceSelect::Flagged(nullptr, cub_bytes, ids, (const uint8_t*)nullptr, (int32_t*)nullptr, (int32_t*)nullptr, N, h->stream));
    const size_t o_flag = 0, o_idx = al((size_t)N), o_cnt = o_idx + al(4 * (size_t)N), o_cub = o_cnt + 256, total = o_cub + al(cub_bytes + 16);
    DCU(quota_ctx_ensure(h->qctx, total, 256 + 4 * (size_t)N));
    unsigned char* D = h->qctx->d; unsigned char* S = h->qctx->hs;
    k_changed_flags<<<(N + 255) / 256, 256, 0, h->stream>>>(h->pb, D + o_flag);
    DCU(cudaGetLastError());
    size_t tb = cub_bytes;
    DCU(cub::DeviceSelect::Flagged(D + o_cub, tb, ids, (const uint8_t*)(D + o_flag), (int32_t*)(D + o_idx), (int32_t*)(D + o_cnt), N, h->stream));
    DCU(cudaMemcpyAsync(S, D + o_cnt, 4, cudaMemcpyDeviceToHost, h->stream));
    DCU(cudaStreamSynchronize(h->stream));
    int32_t n = 0; std::memcpy(&n, S, 4);
    *n_changed = n;
    const int take = n < capacity ? n : capacity;
    if (take > 0) {
        DCU(cudaMemcpyAsync(S + 256, D + o_idx, 4 * (size_t)take, cudaMemcpyDeviceToHost, h->stream));
        DCU(cudaStreamSynchronize(h->stream));
        std::memcpy(node_idx, S + 256, 4 * (size_t)take);
    }
    if (n > capacity) return nosplan_fail(h, NOSPLAN_E_LIMIT, "nosplan_changed_nodes: %d nodes changed, capacity %d (n_changed holds the number)", n, capacity);
    return NOSPLAN_OK;
#undef DCU
}
