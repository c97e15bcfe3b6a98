// Snapshot pre-pass and tracker initialisation kernels (grid-wide, HBM streaming).
//
//   k_prepass       tiles of 32 nodes per block: phase 1 one thread per GPU row (consecutive threads read consecutive
//                   128-bit rows), phase 2 one thread per (node, profile) pair over the rows staged in shared memory,
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

constexpr int PRE_NODES = 32;                        // nodes per tile
constexpr int PRE_ROWS = PRE_NODES * NOSPLAN_MAX_GPUS; // GPU rows per tile at most

__global__ void __launch_bounds__(256) k_prepass(DevProblem pb) {
    __shared__ uint4 s_geom[MAX_ROWS];
    __shared__ uint16_t s_gmask[MAX_ROWS];
    __shared__ int s_off[MAX_MODELS + 1];
    __shared__ uint4 s_used[PRE_ROWS], s_free[PRE_ROWS];     // the tile's GPU rows, zero where the node has no such GPU
    __shared__ uint8_t s_hasfree[PRE_ROWS];
    __shared__ long long s_acc[8][KP];
    __shared__ int s_maxfree[8][KP];

    cudaTriggerProgrammaticLaunchCompletion();           // k_tracker_init may become resident; it waits for this grid itself
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const bool mig = pb.kind == NOSPLAN_KIND_MIG;
    const int G = pb.G, K = pb.K;
    if (mig) {
        for (int r = tid; r < pb.n_rows; r += blockDim.x) {
            uint4 row = pb.geom[r];
            s_geom[r] = row;
            s_gmask[r] = nz_mask(row);
        }
        for (int m = tid; m <= pb.n_models; m += blockDim.x) s_off[m] = pb.model_off[m];
    }
    __syncthreads();

    {   // resets (see DevProblem::zero_next): the next run's small state by CTA 0, this run's tracker tables by everybody
        if (blockIdx.x == 0) for (int x = tid; x < pb.zero_next_n8; x += blockDim.x) pb.zero_next[x] = 0;
        const size_t gt = (size_t)blockIdx.x * blockDim.x + tid, gn = (size_t)gridDim.x * blockDim.x;
        uint4* el = reinterpret_cast<uint4*>(pb.entry_lack);
        for (size_t x = gt; x < (size_t)pb.n_keys * (KP / 4); x += gn) el[x] = make_uint4(0, 0, 0, 0);
        for (size_t x = gt; x < (size_t)(pb.n_keys + 31) / 32; x += gn) pb.placed[x] = 0;
    }
    // this thread's profile in phase 2 never changes: (node-in-tile, profile) pairs are laid out 16 profiles per node
    const int my_p = tid & (KP - 1);
    long long acc = 0;
    int maxfree = 0, maxmem = 0;
    bool inconsistent = false;
    const int n_tiles = (pb.N + PRE_NODES - 1) / PRE_NODES;
    for (int tile = blockIdx.x; tile < n_tiles; tile += gridDim.x) {
        const int node0 = tile * PRE_NODES;
        const int rows = PRE_NODES * G;
        // the int64 account columns of phase 2 do not depend on phase 1: issue their loads now, both round trips overlap
        long long pa[PRE_NODES * KP / 256], pr_[PRE_NODES * KP / 256];
#pragma unroll
        for (int u = 0; u < PRE_NODES * KP / 256; ++u) {
            const int i = node0 + (tid + u * 256) / KP;
            const bool on = i < pb.N && my_p < K;
            pa[u] = on ? pb.node_alloc[(size_t)i * pb.R + 3 + my_p] : 0;
            pr_[u] = on ? pb.node_req[(size_t)i * pb.R + 3 + my_p] : 0;
        }
        // ---- phase 1: one thread per GPU row, consecutive threads read consecutive rows (128-bit loads)
        for (int r = tid; r < rows; r += blockDim.x) {
            const int nl = r / G, j = r - nl * G, i = node0 + nl;
            uint4 used = make_uint4(0, 0, 0, 0), fre = used;
            bool has_free = false;
            if (i < pb.N && j < min(pb.node_gpu_count[i], G)) {     // clamped: this kernel may run before the host has validated
                const int x = i * G + j;
                used = pb.gpu_used[x]; fre = pb.gpu_free[x];
                const uint16_t um = pb.gpu_umask[x], fm = pb.gpu_fmask[x];
                const uint4 cnt = add_bytes(used, fre);
                const uint16_t gm = um | fm;
                pb.out_count[x] = cnt;
                pb.out_mask[x] = gm;
                if (mig) {
                    has_free = fm != 0;
                    if (!has_free) {                     // AllowsGeometry(GetGeometry()) (mig/gpu.go:215-222)
                        const int m = min(max(pb.gpu_model[x], 0), max(pb.n_models - 1, 0));
                        bool allowed = false;
                        for (int q = s_off[m]; pb.n_models > 0 && q < s_off[m + 1]; ++q) {
                            const uint4 row = s_geom[q];
                            allowed |= (s_gmask[q] == gm) && row.x == cnt.x && row.y == cnt.y && row.z == cnt.z && row.w == cnt.w;
                        }
                        has_free = !allowed;
                    }
                } else {                                 // slicing GPU.HasFreeCapacity (slicing/gpu.go:131-136)
                    int tot = 0;
#pragma unroll
                    for (int p = 0; p < KP; ++p) tot += pb.pmem[p] * (int)byte_of(cnt, p);
                    const int gm_ = pb.gpu_mem[x];
                    maxmem = max(maxmem, gm_);
                    has_free = fm != 0 || (gm_ - tot >= 1);
                }
            }
            s_used[r] = used; s_free[r] = fre; s_hasfree[r] = has_free;
        }
        __syncthreads();
        // ---- phase 2a: HasFreeCapacity of the node = any of its GPUs
        if (tid < PRE_NODES && node0 + tid < pb.N) {
            bool any = false;
            for (int j = 0; j < G; ++j) any |= s_hasfree[tid * G + j] != 0;
            pb.cand_flag[node0 + tid] = any;
        }
        // ---- phase 2b: (node, profile) pairs: totals, largest free count, node accounts vs GPU maps
#pragma unroll
        for (int u = 0; u < PRE_NODES * KP / 256; ++u) {
            const int pr = tid + u * 256;
            const int nl = pr / KP, i = node0 + nl;          // pr % KP == my_p
            if (i >= pb.N || my_p >= K) continue;
            const long long a = pa[u], r = pr_[u];
            acc += a - r;
            int sum_free = 0, sum_used = 0;
            const uint8_t* fb = reinterpret_cast<const uint8_t*>(s_free + nl * G) + my_p;
            const uint8_t* ub = reinterpret_cast<const uint8_t*>(s_used + nl * G) + my_p;
            for (int g = 0; g < G; ++g) {
                const int f = fb[16 * g];
                sum_free += f; sum_used += ub[16 * g];
                maxfree = max(maxfree, f);
            }
            // gate (core/snapshot.go:132-165) can only be implied by GPU.AddPod feasibility if the node accounts agree
            inconsistent |= (a - r < sum_free) || (r > sum_used);
        }
        __syncthreads();
    }
    if (__any_sync(0xffffffffu, inconsistent) && lane == 0) atomicOr(&pb.gflags[0], 1);
    maxmem = __reduce_max_sync(0xffffffffu, maxmem);
    if (maxmem > 0 && lane == 0) atomicMax(&pb.gflags[1], maxmem);
    // lanes p and p+16 of a warp hold the same profile
    acc += __shfl_xor_sync(0xffffffffu, acc, 16);
    maxfree = max(maxfree, __shfl_xor_sync(0xffffffffu, maxfree, 16));
    if (lane < KP) { s_acc[warp][lane] = acc; s_maxfree[warp][lane] = maxfree; }
    __syncthreads();
    if (warp == 0 && lane < K) {
        long long t = 0;
        for (int w = 0; w < (int)(blockDim.x >> 5); ++w) t += s_acc[w][lane];
        if (t != 0) atomicAdd((unsigned long long*)&pb.tot[lane], (unsigned long long)t);
        int mf = 0;
        for (int w = 0; w < (int)(blockDim.x >> 5); ++w) mf = max(mf, s_maxfree[w][lane]);
        if (mf > 0) atomicMax(&pb.maxfree[lane], mf);
    }
    // slice mode, fused exchange 1: the CTA that finishes last pushes this rank's totals into every rank's mailbox
    if (pb.xf_on && last_block_done(&pb.next_node[1])) {
        __shared__ __align__(16) unsigned char s_blk[256];
        const int par = pb.xf_epoch & 1;
        if (tid < 64) reinterpret_cast<int32_t*>(s_blk)[tid] = 0;
        __syncthreads();
        if (tid < KP) reinterpret_cast<long long*>(s_blk)[tid] = __ldcg(&pb.tot[tid]);
        if (tid < KP) reinterpret_cast<int32_t*>(s_blk + 128)[tid] = __ldcg(&pb.maxfree[tid]);
        if (tid < 3) reinterpret_cast<int32_t*>(s_blk + 192)[tid] = __ldcg(&pb.gflags[tid]);
        __syncthreads();
        for (int x = tid; x < pb.xf_box.world * 16; x += blockDim.x) {
            const int r = x >> 4, q = x & 15;
            reinterpret_cast<uint4*>(pb.xf_box.base[r] + box_tot(par, pb.xf_box.rank))[q] = reinterpret_cast<const uint4*>(s_blk)[q];
        }
        __threadfence_system();
        __syncthreads();
        if (tid < pb.xf_box.world)
            *reinterpret_cast<volatile unsigned*>(pb.xf_box.base[tid] + box_flags(1, par) + 4 * pb.xf_box.rank) = pb.xf_epoch;
    }
}

__global__ void __launch_bounds__(256) k_tracker_init(DevProblem pb) {
    __shared__ int s_lack[KP];
    __shared__ long long s_tot[KP];
    cudaTriggerProgrammaticLaunchCompletion();
    if (threadIdx.x < KP) s_lack[threadIdx.x] = 0;
    const int s = blockIdx.x * blockDim.x + threadIdx.x;
    // the pod's own data was uploaded, not computed: fetch it while k_prepass is still finishing
    int key = 0, e0 = 0, e1 = 0, col0 = 0;
    long long qty0 = 0;
    if (s < pb.P) {
        key = pb.pod_key[s]; e0 = pb.pod_sptr[s]; e1 = pb.pod_sptr[s + 1];
        if (e0 < e1) { col0 = pb.pod_scol[e0]; qty0 = pb.pod_sqty[e0]; }
    }
    cudaGridDependencySynchronize();                     // the totals of k_prepass (and the work-area reset before it)
    if (pb.xf_on) {
        // slice mode, fused exchange 1: wait for every rank's totals in this rank's mailbox and fold them (each CTA for itself;
        // CTA 0 also publishes them for the kernels that follow), and clear this rank's exchange block for the pass to come
        const int par = pb.xf_epoch & 1, W = pb.xf_box.world;
        const unsigned char* me = pb.xf_box.base[pb.xf_box.rank];
        const bool ok = peer_wait(PeerWait{reinterpret_cast<const unsigned*>(me + box_flags(1, par)), W, pb.xf_epoch});
        const int t = threadIdx.x;
        if (t < KP) {
            long long sum = 0; int m = 0;
            for (int r = 0; r < W; ++r) {
                sum += __ldcg(reinterpret_cast<const long long*>(me + box_tot(par, r)) + t);
                m = max(m, __ldcg(reinterpret_cast<const int32_t*>(me + box_tot(par, r) + 128) + t));
            }
            s_tot[t] = sum;
            if (blockIdx.x == 0) { pb.tot[t] = sum; pb.maxfree[t] = m; }
        }
        if (blockIdx.x == 0 && t < 3) {
            int f = 0;
            for (int r = 0; r < W; ++r) {
                const int v = __ldcg(reinterpret_cast<const int32_t*>(me + box_tot(par, r) + 192) + t);
                f = t == 0 ? (f | v) : max(f, v);
            }
            pb.gflags[t] = f;
        }
        if (blockIdx.x == 0 && t == 0 && !ok) pb.gflags[3] = 1;
        unsigned long long* ex = reinterpret_cast<unsigned long long*>(pb.xf_exchange);
        for (unsigned long long i = s; i < (pb.xf_xb >> 3); i += (unsigned long long)gridDim.x * blockDim.x) ex[i] = 0;
    } else if (threadIdx.x < KP) s_tot[threadIdx.x] = pb.tot[threadIdx.x];
    __syncthreads();
    if (s < pb.P) {
        for (int e = e0; e < e1; ++e) {
            const int col = e == e0 ? col0 : pb.pod_scol[e];
            if (col >= 3 + pb.K) break;              // columns ascend: slice columns come first
            const int p = col - 3;
            long long avail = s_tot[p];
            if (avail < 0) avail = 0;                 // SubtractNonNegative (pkg/resource/resource.go:76-88)
            const long long diff = avail - (e == e0 ? qty0 : pb.pod_sqty[e]);
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

// Slice mode (every rank holds its own slice of the node list): what k_prepass reduced over the slice is exchanged as one
// 256-byte block per rank -- [tot int64 x16 | maxfree int32 x16 | gflags int32 x4] -- and folded back: sums, maxima, OR.
__global__ void k_slice_pack_totals(DevProblem pb, unsigned char* block) {
    const int t = threadIdx.x;
    if (t < KP) reinterpret_cast<long long*>(block)[t] = pb.tot[t];
    if (t < KP) reinterpret_cast<int32_t*>(block + 128)[t] = pb.maxfree[t];
    if (t < 4) reinterpret_cast<int32_t*>(block + 192)[t] = pb.gflags[t];
}

__global__ void k_slice_combine_totals(DevProblem pb, const unsigned char* gathered, int world, PeerWait wait) {
    if (!peer_wait(wait) && threadIdx.x == 0) pb.gflags[3] = 1;
    const int t = threadIdx.x;
    if (t < KP) {
        long long s = 0; int m = 0;
        for (int r = 0; r < world; ++r) {
            s += reinterpret_cast<const long long*>(gathered + (size_t)r * 256)[t];
            m = max(m, reinterpret_cast<const int32_t*>(gathered + (size_t)r * 256 + 128)[t]);
        }
        pb.tot[t] = s; pb.maxfree[t] = m;
    }
    if (t < 3) {
        int f = 0;
        for (int r = 0; r < world; ++r) {
            const int v = reinterpret_cast<const int32_t*>(gathered + (size_t)r * 256 + 192)[t];
            f = t == 0 ? (f | v) : max(f, v);
        }
        pb.gflags[t] = f;
    }
}

void launch_slice_pack_totals(const DevProblem& pb, void* block, cudaStream_t st) {
    k_slice_pack_totals<<<1, 32, 0, st>>>(pb, static_cast<unsigned char*>(block));
}
void launch_slice_combine_totals(const DevProblem& pb, const void* gathered, int world, const PeerWait& wait, cudaStream_t st) {
    k_slice_combine_totals<<<1, 32, 0, st>>>(pb, static_cast<const unsigned char*>(gathered), world, wait);
}
// nosplan_plan_compact: one record per node (layout in include/nosplan.h) -> the node sections of the arena.  One warp per node:
// lanes 0..G-1 expand the GPU slots (two 128-bit rows, masks, model / memory), all lanes the account columns.
__global__ void __launch_bounds__(256) k_expand_nodes(DevProblem pb, const unsigned char* records, int record_bytes, int n_models, int32_t* flag) {
    const int lane = threadIdx.x & 31;
    const int i = blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
    if (i >= pb.N) return;
    const unsigned char* rec = records + (size_t)i * record_bytes;
    const int K = pb.K, R = pb.R, G = pb.G, n_other = pb.n_other;
    const int ng = *reinterpret_cast<const int32_t*>(rec);
    if (ng < 0 || ng > pb.G) {                                // what nosplan_upload refuses per node; the lowest node index is reported
        if (lane == 0 && (unsigned)i < (unsigned)atomicMin(reinterpret_cast<unsigned*>(flag), (unsigned)i)) flag[1] = 1;
    }
    if (lane == 0) {
        const int4 c = *reinterpret_cast<const int4*>(rec);
        const_cast<int32_t*>(pb.node_gpu_count)[i] = min(max(c.x, 0), pb.G);
        const_cast<int32_t*>(pb.node_allowed_pods)[i] = c.y;
        const_cast<int32_t*>(pb.node_num_pods)[i] = c.z;
    }
    long long* alloc = const_cast<long long*>(pb.node_alloc) + (size_t)i * R;
    long long* req = const_cast<long long*>(pb.node_req) + (size_t)i * R;
    if (lane < 3) {
        alloc[lane] = reinterpret_cast<const long long*>(rec + 16)[lane];
        req[lane] = reinterpret_cast<const long long*>(rec + 16)[3 + lane];
    }
    if (lane < K) {
        alloc[3 + lane] = reinterpret_cast<const int32_t*>(rec + 64)[lane];
        req[3 + lane] = reinterpret_cast<const int32_t*>(rec + 64)[8 + lane];
    }
    if (lane < n_other) {
        alloc[3 + K + lane] = reinterpret_cast<const long long*>(rec + 128)[lane];
        req[3 + K + lane] = reinterpret_cast<const long long*>(rec + 128)[n_other + lane];
    }
    if (lane < G) {
        const unsigned char* g = rec + 128 + 16 * (size_t)n_other + 24 * (size_t)lane;
        const uint2 u = *reinterpret_cast<const uint2*>(g), f = *reinterpret_cast<const uint2*>(g + 8);
        const unsigned t = *reinterpret_cast<const unsigned*>(g + 16), m = *reinterpret_cast<const unsigned*>(g + 20);
        const size_t x = (size_t)i * G + lane;
        const_cast<uint4*>(pb.gpu_used)[x] = make_uint4(u.x, u.y, 0u, 0u);
        const_cast<uint4*>(pb.gpu_free)[x] = make_uint4(f.x, f.y, 0u, 0u);
        const_cast<uint16_t*>(pb.gpu_umask)[x] = (uint16_t)(t & 0xffu);
        const_cast<uint16_t*>(pb.gpu_fmask)[x] = (uint16_t)((t >> 8) & 0xffu);
        if (pb.kind == NOSPLAN_KIND_MIG) {
            const int model = (int)((t >> 16) & 0xffu);
            if (lane < ng && model >= n_models && (unsigned)i < (unsigned)atomicMin(reinterpret_cast<unsigned*>(flag), (unsigned)i)) flag[1] = 2;
            const_cast<int32_t*>(pb.gpu_model)[x] = min(model, max(n_models - 1, 0));
        }
        else const_cast<int32_t*>(pb.gpu_mem)[x] = (int32_t)(m & 0xffffu);
    }
}
void launch_expand_nodes(const DevProblem& pb, const void* records, int record_bytes, int n_models, int32_t* flag, cudaStream_t st) {
    if (pb.N > 0) k_expand_nodes<<<(pb.N + 7) / 8, 256, 0, st>>>(pb, static_cast<const unsigned char*>(records), record_bytes, n_models, flag);
}

void launch_prepass(const DevProblem& pb, cudaStream_t st) {
    if (pb.N > 0) {
        int blocks = (pb.N + PRE_NODES - 1) / PRE_NODES;      // one tile of 32 nodes per block, grid-stride beyond 8 waves
        if (blocks > 148 * 8) blocks = 148 * 8;
        k_prepass<<<blocks, 256, 0, st>>>(pb);
    } else k_prepass<<<1, 256, 0, st>>>(pb);     // no nodes: the resets (and, in fused slice mode, the push of zero totals) still happen
}

void launch_tracker_init(const DevProblem& pb, cudaStream_t st) {
    // fused slice mode: the kernel also folds the exchanged totals, so it runs even without pods
    if (pb.P > 0 || pb.xf_on) launch_pdl(k_tracker_init, dim3(pb.P > 0 ? (pb.P + 255) / 256 : 1), dim3(256), 0, st, pb);
}

}  // namespace nsp
