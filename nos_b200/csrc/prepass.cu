// Snapshot pre-pass and tracker initialisation kernels (grid-wide, HBM streaming).
//
//   k_prepass       one warp per node (grid-stride): reads every node row once with 128-bit loads,
//                   emits  (a) the initial partitioning state  GetPartitioningState  (core/snapshot.go:66-72,
//                              internal/partitioning/mig/partitition_calculator.go:30-46)
//                          (b) the candidate flag  HasFreeCapacity  (pkg/gpu/mig/node.go:124-139,
//                              pkg/gpu/slicing/node.go:208-215)
//                          (c) cluster totals  sum(Allocatable - Requested)  per slice resource
//                              (core/snapshot.go:132-141) reduced warp -> block -> one int64 atomic per column
//                          (d) for the node-parallel fast path: the largest free count of each profile on any GPU
//                              and whether every node's Allocatable/Requested agree with its GPUs' used/free maps
//   k_tracker_init  one thread per pod: lacking slices against the t0 totals, NewSliceTracker (core/tracker.go:34-57)
#include "nosplan_dev.cuh"

namespace nsp {

__global__ void __launch_bounds__(256) k_prepass(DevProblem pb) {
    __shared__ uint4 s_geom[MAX_ROWS];
    __shared__ uint16_t s_gmask[MAX_ROWS];
    __shared__ int s_off[MAX_MODELS + 1];
    __shared__ long long s_acc[8][KP];
    __shared__ int s_maxfree[8][KP];

    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const bool mig = pb.kind == NOSPLAN_KIND_MIG;
    if (mig) {
        for (int r = threadIdx.x; r < pb.n_rows; r += blockDim.x) {
            uint4 row = pb.geom[r];
            s_geom[r] = row;
            s_gmask[r] = nz_mask(row);
        }
        for (int m = threadIdx.x; m <= pb.n_models; m += blockDim.x) s_off[m] = pb.model_off[m];
    }
    __syncthreads();

    long long acc = 0;
    int maxfree = 0;
    bool inconsistent = false;
    const int warps_total = gridDim.x * (blockDim.x >> 5);
    for (int i = blockIdx.x * (blockDim.x >> 5) + warp; i < pb.N; i += warps_total) {
        const int ng = pb.node_gpu_count[i];
        bool has_free = false;
        if (lane < ng) {
            const int x = i * pb.G + lane;
            const uint4 used = pb.gpu_used[x], fre = pb.gpu_free[x];
            const uint16_t um = pb.gpu_umask[x], fm = pb.gpu_fmask[x];
            const uint4 cnt = add_bytes(used, fre);
            const uint16_t gm = um | fm;
            pb.out_count[x] = cnt;
            pb.out_mask[x] = gm;
            if (mig) {
                has_free = fm != 0;
                if (!has_free) {                     // AllowsGeometry(GetGeometry()) (mig/gpu.go:215-222)
                    const int m = pb.gpu_model[x];
                    bool allowed = false;
                    for (int r = s_off[m]; r < s_off[m + 1]; ++r) {
                        const uint4 row = s_geom[r];
                        allowed |= (s_gmask[r] == gm) && row.x == cnt.x && row.y == cnt.y && row.z == cnt.z && row.w == cnt.w;
                    }
                    has_free = !allowed;
                }
            } else {                                 // slicing GPU.HasFreeCapacity (slicing/gpu.go:131-136)
                int tot = 0;
#pragma unroll
                for (int p = 0; p < KP; ++p) tot += pb.pmem[p] * (int)byte_of(cnt, p);
                has_free = fm != 0 || (pb.gpu_mem[x] - tot >= 1);
            }
        }
        const unsigned any = __ballot_sync(0xffffffffu, has_free);
        if (lane == 0) pb.cand_flag[i] = any != 0;
        if (lane < pb.K) {
            const size_t o = (size_t)i * pb.R + 3 + lane;
            const long long a = pb.node_alloc[o], r = pb.node_req[o];
            acc += a - r;
            long long sum_free = 0, sum_used = 0;
            const uint8_t* fb = reinterpret_cast<const uint8_t*>(pb.gpu_free) + (size_t)i * pb.G * 16 + lane;
            const uint8_t* ub = reinterpret_cast<const uint8_t*>(pb.gpu_used) + (size_t)i * pb.G * 16 + lane;
            for (int g = 0; g < ng; ++g) {
                const int f = fb[16 * g];
                sum_free += f; sum_used += ub[16 * g];
                maxfree = max(maxfree, f);
            }
            // gate (core/snapshot.go:132-165) can only be implied by GPU.AddPod feasibility if the node accounts agree
            inconsistent |= (a - r < sum_free) || (r > sum_used);
        }
    }
    if (__any_sync(0xffffffffu, inconsistent) && lane == 0) atomicOr(&pb.gflags[0], 1);
    if (lane < KP) { s_acc[warp][lane] = acc; s_maxfree[warp][lane] = maxfree; }
    __syncthreads();
    if (warp == 0 && lane < pb.K) {
        long long t = 0;
        for (int w = 0; w < (int)(blockDim.x >> 5); ++w) t += s_acc[w][lane];
        if (t != 0) atomicAdd((unsigned long long*)&pb.tot[lane], (unsigned long long)t);
        int mf = 0;
        for (int w = 0; w < (int)(blockDim.x >> 5); ++w) mf = max(mf, s_maxfree[w][lane]);
        if (mf > 0) atomicMax(&pb.maxfree[lane], mf);
    }
}

__global__ void __launch_bounds__(256) k_tracker_init(DevProblem pb) {
    __shared__ int s_lack[KP];
    if (threadIdx.x < KP) s_lack[threadIdx.x] = 0;
    __syncthreads();
    const int s = blockIdx.x * blockDim.x + threadIdx.x;
    if (s < pb.P) {
        const int key = pb.pod_key[s];
        for (int e = pb.pod_sptr[s]; e < pb.pod_sptr[s + 1]; ++e) {
            const int col = pb.pod_scol[e];
            if (col >= 3 + pb.K) break;              // columns ascend: slice columns come first
            const int p = col - 3;
            long long avail = pb.tot[p];
            if (avail < 0) avail = 0;                 // SubtractNonNegative (pkg/resource/resource.go:76-88)
            const long long diff = avail - pb.pod_sqty[e];
            if (diff < 0) {
                const int l = (int)(-diff);
                atomicAdd(&pb.entry_lack[(size_t)key * KP + p], l);
                atomicAdd(&s_lack[p], l);
            }
        }
    }
    __syncthreads();
    if (threadIdx.x < KP && s_lack[threadIdx.x] != 0) atomicAdd(&pb.lack[threadIdx.x], s_lack[threadIdx.x]);
}

void launch_prepass(const DevProblem& pb, cudaStream_t st) {
    if (pb.N > 0) {
        int blocks = (pb.N + 7) / 8;
        if (blocks > 148 * 8) blocks = 148 * 8;
        k_prepass<<<blocks, 256, 0, st>>>(pb);
    }
}

void launch_tracker_init(const DevProblem& pb, cudaStream_t st) {
    if (pb.P > 0) k_tracker_init<<<(pb.P + 255) / 256, 256, 0, st>>>(pb);
}

}  // namespace nsp
