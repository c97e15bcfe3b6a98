// Node-sharded Plan() with the collectives issued from inside the library: one grouped NCCL call carrying the three
// all-gathers of a step (exchange block, plan count rows, plan mask rows) on the handle's stream, so a sharded Plan() costs
// one host call.  NCCL is resolved at run time (dlopen of libnccl.so.2), so single-GPU users carry no NCCL dependency
// and, inside a torch process, the already loaded NCCL is reused.
#include <dlfcn.h>
#include <nccl.h>

#include <algorithm>
#include <cstring>
#include <cstdlib>
#include <mutex>
#include <vector>

#include "handle.cuh"

extern "C" {
int64_t nosplan_shard_exchange_bytes(nosplan_t* h);
int nosplan_shard_begin(nosplan_t* h, int32_t rank, int32_t world, void* d_exchange, int64_t exchange_bytes);
int nosplan_shard_end(nosplan_t* h, const void* d_gathered, int32_t world);
int nosplan_slice_begin(nosplan_t* h, void* d_totals);
int nosplan_slice_mid(nosplan_t* h, int32_t rank, int32_t world, const void* d_totals_gathered, void* d_exchange, int64_t exchange_bytes);
}
int nosplan_run_sliced_fused(nosplan* h, const nsp::PeerBox& box, unsigned epoch, unsigned char* d_exchange, size_t xb);
namespace nsp { size_t fused_exchange_bytes(const DevProblem& pb); size_t fused_exchange_bytes_ll(const DevProblem& pb); }

namespace {

struct NcclApi {
    void* lib = nullptr;
    ncclResult_t (*GetUniqueId)(ncclUniqueId*) = nullptr;
    ncclResult_t (*CommInitRank)(ncclComm_t*, int, ncclUniqueId, int) = nullptr;
    ncclResult_t (*CommDestroy)(ncclComm_t) = nullptr;
    ncclResult_t (*AllGather)(const void*, void*, size_t, ncclDataType_t, ncclComm_t, cudaStream_t) = nullptr;
    ncclResult_t (*AllReduce)(const void*, void*, size_t, ncclDataType_t, ncclRedOp_t, ncclComm_t, cudaStream_t) = nullptr;
    ncclResult_t (*GroupStart)() = nullptr;
    ncclResult_t (*GroupEnd)() = nullptr;
    const char* (*GetErrorString)(ncclResult_t) = nullptr;
    bool ok = false;
};

NcclApi& nccl() {
    static NcclApi api;
    static std::once_flag once;
    std::call_once(once, [] {
        api.lib = dlopen("libnccl.so.2", RTLD_NOW | RTLD_GLOBAL);
        if (!api.lib) api.lib = dlopen("libnccl.so", RTLD_NOW | RTLD_GLOBAL);
        if (!api.lib) return;
        auto sym = [&](const char* n) { return dlsym(api.lib, n); };
        api.GetUniqueId = reinterpret_cast<decltype(api.GetUniqueId)>(sym("ncclGetUniqueId"));
        api.CommInitRank = reinterpret_cast<decltype(api.CommInitRank)>(sym("ncclCommInitRank"));
        api.CommDestroy = reinterpret_cast<decltype(api.CommDestroy)>(sym("ncclCommDestroy"));
        api.AllGather = reinterpret_cast<decltype(api.AllGather)>(sym("ncclAllGather"));
        api.AllReduce = reinterpret_cast<decltype(api.AllReduce)>(sym("ncclAllReduce"));
        api.GroupStart = reinterpret_cast<decltype(api.GroupStart)>(sym("ncclGroupStart"));
        api.GroupEnd = reinterpret_cast<decltype(api.GroupEnd)>(sym("ncclGroupEnd"));
        api.GetErrorString = reinterpret_cast<decltype(api.GetErrorString)>(sym("ncclGetErrorString"));
        api.ok = api.GetUniqueId && api.CommInitRank && api.CommDestroy && api.AllGather && api.GroupStart && api.GroupEnd;
    });
    return api;
}

inline size_t al(size_t x) { return (x + 255) & ~(size_t)255; }

}  // namespace

constexpr size_t PEER_XB_MAX = 64 << 10;                  // exchange blocks up to 64 KiB (524 032 tracker keys) go over peer memory

struct nosplan_shard_ctx {
    ncclComm_t comm = nullptr;
    int rank = 0, world = 1;
    // peer-memory mailboxes (slice mode): this rank's box and the IPC mappings of the others; peer_ok == false: NCCL only
    bool peer_ok = false;
    nsp::PeerBox box{};
    unsigned char* box_local = nullptr;
    unsigned epoch = 0;
    unsigned char* d = nullptr; size_t cap = 0;
    size_t o_local_x = 0, o_gather_x = 0, o_local_cnt = 0, o_local_msk = 0, o_cnt_all = 0, o_msk_all = 0;
    size_t xb = 0, chunk_rows = 0;
    bool gathered = false;
};

#define SCU(h, call) do { cudaError_t e_ = (call); if (e_ != cudaSuccess) \
    return nosplan_fail(h, NOSPLAN_E_CUDA, "%s: %s", #call, cudaGetErrorString(e_)); } while (0)
#define SNC(h, call) do { ncclResult_t r_ = (call); if (r_ != ncclSuccess) \
    return nosplan_fail(h, NOSPLAN_E_CUDA, "%s: %s", #call, nccl().GetErrorString ? nccl().GetErrorString(r_) : "nccl error"); } while (0)

extern "C" void nosplan_shard_ctx_free(nosplan_shard_ctx* c) {
    if (!c) return;
    for (int r = 0; r < c->world && r < nsp::PEER_MAX; ++r)
        if (r != c->rank && c->box.base[r]) cudaIpcCloseMemHandle(c->box.base[r]);
    if (c->box_local) cudaFree(c->box_local);
    if (c->comm && nccl().ok) nccl().CommDestroy(c->comm);
    if (c->d) cudaFree(c->d);
    delete c;
}

extern "C" int nosplan_nccl_unique_id(uint8_t* out128) {
    if (!out128) return NOSPLAN_E_INVALID;
    if (!nccl().ok) return NOSPLAN_E_CUDA;
    ncclUniqueId id;
    if (nccl().GetUniqueId(&id) != ncclSuccess) return NOSPLAN_E_CUDA;
    static_assert(sizeof(id) == 128, "ncclUniqueId size");
    std::memcpy(out128, &id, 128);
    return NOSPLAN_OK;
}

extern "C" int nosplan_shard_init(nosplan_t* h, int32_t rank, int32_t world, const uint8_t* id128) {
    if (!h) return NOSPLAN_E_INVALID;
    if (!nccl().ok) return nosplan_fail(h, NOSPLAN_E_CUDA, "libnccl.so.2 not found or incomplete");
    if (!id128 || world < 1 || rank < 0 || rank >= world) return nosplan_fail(h, NOSPLAN_E_INVALID, "rank %d of %d", rank, world);
    SCU(h, cudaSetDevice(h->device));
    if (h->sctx) { nosplan_shard_ctx_free(h->sctx); h->sctx = nullptr; }
    nosplan_shard_ctx* c = new nosplan_shard_ctx();
    c->rank = rank; c->world = world;
    ncclUniqueId id; std::memcpy(&id, id128, 128);
    ncclResult_t r = nccl().CommInitRank(&c->comm, world, id, rank);
    if (r != ncclSuccess) { delete c; return nosplan_fail(h, NOSPLAN_E_CUDA, "ncclCommInitRank: %s", nccl().GetErrorString(r)); }
    h->sctx = c;
    // Mailboxes over peer memory for the two small exchanges of slice mode.  Every step below is allowed to fail (no P2P,
    // more than PEER_MAX ranks, NOSPLAN_NO_PEER set): the ranks then agree, through one all-reduce, to stay on NCCL.
    int ok = world <= nsp::PEER_MAX && world > 1 && getenv("NOSPLAN_NO_PEER") == nullptr && nccl().AllReduce != nullptr;
    const size_t bb = nsp::box_bytes(PEER_XB_MAX);
    unsigned char* tmp = nullptr;                           // [world] IPC handles + one int
    const size_t hb = sizeof(cudaIpcMemHandle_t);
    SCU(h, cudaMalloc(&tmp, hb * (world + 1) + 256));
    cudaIpcMemHandle_t mine; std::memset(&mine, 0, sizeof mine);
    if (ok && (cudaMalloc(&c->box_local, bb) != cudaSuccess || cudaMemset(c->box_local, 0, bb) != cudaSuccess ||
               cudaIpcGetMemHandle(&mine, c->box_local) != cudaSuccess)) ok = 0;
    (void)cudaGetLastError();
    std::vector<cudaIpcMemHandle_t> all(world);
    int* d_ok = reinterpret_cast<int*>(tmp + hb * (world + 1));
    bool comm_fine = cudaMemcpyAsync(tmp + hb * world, &mine, hb, cudaMemcpyHostToDevice, h->stream) == cudaSuccess &&
                     nccl().AllGather(tmp + hb * world, tmp, hb, ncclUint8, c->comm, h->stream) == ncclSuccess &&
                     cudaMemcpyAsync(all.data(), tmp, hb * world, cudaMemcpyDeviceToHost, h->stream) == cudaSuccess &&
                     cudaStreamSynchronize(h->stream) == cudaSuccess;
    if (!comm_fine) ok = 0;
    for (int q = 0; q < world && q < nsp::PEER_MAX; ++q) c->box.base[q] = nullptr;
    if (ok) {
        for (int q = 0; q < world; ++q) {
            if (q == rank) { c->box.base[q] = c->box_local; continue; }
            void* p = nullptr;
            if (cudaIpcOpenMemHandle(&p, all[q], cudaIpcMemLazyEnablePeerAccess) != cudaSuccess) { ok = 0; (void)cudaGetLastError(); break; }
            c->box.base[q] = static_cast<unsigned char*>(p);
        }
    }
    if (comm_fine && nccl().AllReduce) {                    // unanimous or not at all
        comm_fine = cudaMemcpyAsync(d_ok, &ok, sizeof ok, cudaMemcpyHostToDevice, h->stream) == cudaSuccess &&
                    nccl().AllReduce(d_ok, d_ok, 1, ncclInt32, ncclMin, c->comm, h->stream) == ncclSuccess &&
                    cudaMemcpyAsync(&ok, d_ok, sizeof ok, cudaMemcpyDeviceToHost, h->stream) == cudaSuccess &&
                    cudaStreamSynchronize(h->stream) == cudaSuccess;
        if (!comm_fine) ok = 0;
    } else ok = 0;
    cudaFree(tmp);
    (void)cudaGetLastError();
    c->box.world = world; c->box.rank = rank; c->box.xb_max = PEER_XB_MAX;
    c->peer_ok = ok != 0;
    return NOSPLAN_OK;
}

// 1 when the two exchanges of slice mode travel over peer memory (NVLink stores + flags), 0 when they use NCCL
extern "C" int nosplan_shard_transport(nosplan_t* h) { return h && h->sctx && h->sctx->peer_ok ? 1 : 0; }

extern "C" int nosplan_run_sharded(nosplan_t* h, int32_t gather_plan) {
    if (!h) return NOSPLAN_E_INVALID;
    nosplan_shard_ctx* c = h->sctx;
    if (!c) return nosplan_fail(h, NOSPLAN_E_INVALID, "nosplan_run_sharded before nosplan_shard_init");
    if (!h->uploaded) return nosplan_fail(h, NOSPLAN_E_INVALID, "nosplan_run_sharded before nosplan_upload");
    SCU(h, cudaSetDevice(h->device));
    const int N = h->pb.N, G = h->pb.G, W = c->world;
    const size_t xb = (size_t)nosplan_shard_exchange_bytes(h);
    const size_t chunk = (size_t)((N + W - 1) / W), rows = chunk * G;
    size_t o = 0;
    auto put = [&](size_t bytes) { size_t at = o; o = al(o + bytes); return at; };
    c->xb = xb; c->chunk_rows = rows;
    c->o_local_x = put(xb); c->o_gather_x = put(xb * W);
    c->o_local_cnt = put(rows * 16); c->o_local_msk = put(rows * 2);
    c->o_cnt_all = put(rows * 16 * W); c->o_msk_all = put(rows * 2 * W);
    if (o > c->cap) {
        if (c->d) SCU(h, cudaFree(c->d));
        c->d = nullptr; c->cap = 0;
        SCU(h, cudaMalloc(&c->d, o + o / 4 + (1 << 20)));
        c->cap = o + o / 4 + (1 << 20);
    }
    int rc = nosplan_shard_begin(h, c->rank, W, c->d + c->o_local_x, (int64_t)xb);
    if (rc != NOSPLAN_OK) return rc;
    const int lo = std::min(N, c->rank * (int)chunk), hi = std::min(N, lo + (int)chunk);
    const unsigned char* send_cnt = reinterpret_cast<const unsigned char*>(h->pb.out_count) + (size_t)lo * G * 16;
    const unsigned char* send_msk = reinterpret_cast<const unsigned char*>(h->pb.out_mask) + (size_t)lo * G * 2;
    const bool do_gather = gather_plan && rows > 0;
    if (do_gather && (size_t)(hi - lo) != chunk) {              // short last slice: pad through a staging copy
        SCU(h, cudaMemsetAsync(c->d + c->o_local_cnt, 0, rows * 16, h->stream));
        SCU(h, cudaMemsetAsync(c->d + c->o_local_msk, 0, rows * 2, h->stream));
        if (hi > lo) {
            SCU(h, cudaMemcpyAsync(c->d + c->o_local_cnt, send_cnt, (size_t)(hi - lo) * G * 16, cudaMemcpyDeviceToDevice, h->stream));
            SCU(h, cudaMemcpyAsync(c->d + c->o_local_msk, send_msk, (size_t)(hi - lo) * G * 2, cudaMemcpyDeviceToDevice, h->stream));
        }
        send_cnt = c->d + c->o_local_cnt; send_msk = c->d + c->o_local_msk;
    }
    // one grouped launch: exchange block + (speculative) plan rows.  If the verification below rejects the pass, the
    // gathered rows are stale and stats.path == 0 tells the caller that every rank holds the complete plan instead.
    SNC(h, nccl().GroupStart());
    SNC(h, nccl().AllGather(c->d + c->o_local_x, c->d + c->o_gather_x, xb, ncclUint8, c->comm, h->stream));
    if (do_gather) {
        SNC(h, nccl().AllGather(send_cnt, c->d + c->o_cnt_all, rows * 16, ncclUint8, c->comm, h->stream));
        SNC(h, nccl().AllGather(send_msk, c->d + c->o_msk_all, rows * 2, ncclUint8, c->comm, h->stream));
    }
    SNC(h, nccl().GroupEnd());
    c->gathered = do_gather;
    rc = nosplan_shard_end(h, c->d + c->o_gather_x, W);
    if (rc == NOSPLAN_OK) h->launches += 1;                      // the grouped NCCL kernel
    return rc;
}

static int ensure_shard_buffer(nosplan* h, nosplan_shard_ctx* c, size_t bytes) {
    if (bytes > c->cap) {
        if (c->d) SCU(h, cudaFree(c->d));
        c->d = nullptr; c->cap = 0;
        SCU(h, cudaMalloc(&c->d, bytes + bytes / 4 + (1 << 20)));
        c->cap = bytes + bytes / 4 + (1 << 20);
    }
    return NOSPLAN_OK;
}

static int run_sliced_classic(nosplan* h);

// Slice mode.  Where the single-launch plan applies (fused.cu) a Plan() is ONE kernel per rank and ONE exchange: the last CTA
// of k_plan_fused stores the rank's block into every rank's mailbox over NVLink and raises its flag there (or the block
// travels in one NCCL all-gather), and k_fused_verify folds the W blocks and verifies the hypothesis -- every rank reaches the
// same verdict, and a rejected plan is replaced by the classic two-exchange sequence below on every rank alike.
extern "C" int nosplan_run_sliced(nosplan_t* h) {
    if (!h) return NOSPLAN_E_INVALID;
    nosplan_shard_ctx* c = h->sctx;
    if (!c) return nosplan_fail(h, NOSPLAN_E_INVALID, "nosplan_run_sliced before nosplan_shard_init");
    if (!h->uploaded) return nosplan_fail(h, NOSPLAN_E_INVALID, "nosplan_run_sliced before nosplan_upload");
    if (!h->fused_ok || h->pb.kind != NOSPLAN_KIND_MIG) return run_sliced_classic(h);      // the one-kernel exchange carries no GPU-memory maximum (MPS)
    SCU(h, cudaSetDevice(h->device));
    const int W = c->world;
    const size_t xb = nsp::fused_exchange_bytes(h->pb);
    size_t o = 0;
    auto put = [&](size_t bytes) { size_t at = o; o = al(o + bytes); return at; };
    c->xb = xb; c->chunk_rows = 0; c->gathered = false;
    c->o_local_x = put(xb); c->o_gather_x = put(xb * W);
    int rc = ensure_shard_buffer(h, c, o);
    if (rc != NOSPLAN_OK) return rc;
    if (c->peer_ok && nsp::fused_exchange_bytes_ll(h->pb) <= PEER_XB_MAX) {
        const unsigned epoch = ++c->epoch;
        rc = nosplan_run_sliced_fused_begin(h, &c->box, epoch, c->d + c->o_local_x, xb);
        if (rc != NOSPLAN_OK) return rc;
        // the kernel's last CTA pushed, waited for the other ranks' blocks, folded and verified: nothing else to launch
        return nosplan_run_sliced_fused_end(h, nullptr, 0, W, nsp::PeerWait{nullptr, 0, 0}, run_sliced_classic);
    }
    rc = nosplan_run_sliced_fused_begin(h, nullptr, 0, c->d + c->o_local_x, xb);
    if (rc != NOSPLAN_OK) return rc;
    SNC(h, nccl().AllGather(c->d + c->o_local_x, c->d + c->o_gather_x, xb, ncclUint8, c->comm, h->stream));
    rc = nosplan_run_sliced_fused_end(h, c->d + c->o_gather_x, xb, W, nsp::PeerWait{nullptr, 0, 0}, run_sliced_classic);
    if (rc == NOSPLAN_OK) h->launches += 1;                      // the NCCL kernel
    return rc;
}

// The classic form: two small exchanges per Plan() -- the 256-byte totals block between k_prepass and k_tracker_init, and the
// exchange block after the node-parallel pass.  The plan rows stay on their rank.
static int run_sliced_classic(nosplan* h) {
    nosplan_shard_ctx* c = h->sctx;
    SCU(h, cudaSetDevice(h->device));
    const int W = c->world;
    const size_t xb = (size_t)nosplan_shard_exchange_bytes(h), tb = NOSPLAN_SLICE_TOTALS_BYTES;
    size_t o = 0;
    auto put = [&](size_t bytes) { size_t at = o; o = al(o + bytes); return at; };
    const size_t o_t = put(tb), o_t_all = put(tb * W);
    c->xb = xb; c->chunk_rows = 0; c->gathered = false;
    c->o_local_x = put(xb); c->o_gather_x = put(xb * W);
    if (o > c->cap) {
        if (c->d) SCU(h, cudaFree(c->d));
        c->d = nullptr; c->cap = 0;
        SCU(h, cudaMalloc(&c->d, o + o / 4 + (1 << 20)));
        c->cap = o + o / 4 + (1 << 20);
    }
    if (c->peer_ok && xb <= PEER_XB_MAX) {
        // peer-memory transport, fused: the last CTA of each producer kernel pushes the rank's block into every mailbox and
        // the consumer kernel starts by waiting for the W flags of its own mailbox -- five stream operations per Plan(), the
        // same as on one GPU, and no collective launch on the path
        const int rc = nosplan_run_sliced_fused(h, c->box, ++c->epoch, c->d + c->o_local_x, xb);
        return rc;
    }
    int rc = nosplan_slice_begin(h, c->d + o_t);
    if (rc != NOSPLAN_OK) return rc;
    SNC(h, nccl().AllGather(c->d + o_t, c->d + o_t_all, tb, ncclUint8, c->comm, h->stream));
    rc = nosplan_slice_mid(h, c->rank, W, c->d + o_t_all, c->d + c->o_local_x, (int64_t)xb);
    if (rc != NOSPLAN_OK) return rc;
    SNC(h, nccl().AllGather(c->d + c->o_local_x, c->d + c->o_gather_x, xb, ncclUint8, c->comm, h->stream));
    rc = nosplan_shard_end(h, c->d + c->o_gather_x, W);
    if (rc == NOSPLAN_OK) h->launches += 2;                      // the two NCCL kernels
    return rc;
}

// Like nosplan_download, but takes the plan rows from the all-gathered buffers when the node-parallel pass stood.
extern "C" int nosplan_download_sharded(nosplan_t* h, nosplan_partitioning* out, nosplan_stats* stats) {
    if (!h) return NOSPLAN_E_INVALID;
    nosplan_shard_ctx* c = h->sctx;
    if (!c || !h->ran) return nosplan_fail(h, NOSPLAN_E_INVALID, "nosplan_download_sharded before nosplan_run_sharded");
    SCU(h, cudaSetDevice(h->device));
    nsp::DevStats ds;
    SCU(h, cudaMemcpyAsync(h->h_out, h->pb.stats, sizeof ds, cudaMemcpyDeviceToHost, h->stream));
    SCU(h, cudaStreamSynchronize(h->stream));
    std::memcpy(&ds, h->h_out, sizeof ds);
    const size_t nb_cnt = h->out_count_bytes, nb_msk = h->out_mask_bytes;
    if (out && nb_cnt) {
        if (!out->gpu_count || !out->gpu_mask) return nosplan_fail(h, NOSPLAN_E_INVALID, "null output buffers");
        const bool from_gather = ds.path == 1 && c->gathered;
        const void* src_cnt = from_gather ? (const void*)(c->d + c->o_cnt_all) : (const void*)h->pb.out_count;
        const void* src_msk = from_gather ? (const void*)(c->d + c->o_msk_all) : (const void*)h->pb.out_mask;
        SCU(h, cudaMemcpyAsync(h->h_out, src_cnt, nb_cnt, cudaMemcpyDeviceToHost, h->stream));
        SCU(h, cudaMemcpyAsync(h->h_out + al(nb_cnt), src_msk, nb_msk, cudaMemcpyDeviceToHost, h->stream));
        SCU(h, cudaStreamSynchronize(h->stream));
        std::memcpy(out->gpu_count, h->h_out, nb_cnt);
        std::memcpy(out->gpu_mask, h->h_out + al(nb_cnt), nb_msk);
    }
    if (stats) {
        std::memset(stats, 0, sizeof *stats);
        stats->candidates = ds.candidates; stats->visited_nodes = ds.visited_nodes; stats->geometry_evals = ds.geometry_evals;
        stats->placements = ds.placements; stats->candidate_nodes = ds.candidate_nodes;
        stats->lacking_at_start = ds.lacking_at_start; stats->lacking_at_end = ds.lacking_at_end;
        stats->path = ds.path; stats->kernel_launches = h->launches;
        nosplan_phase_times(h, stats);
        stats->h2d_bytes = (int64_t)h->h2d_bytes;
        stats->d2h_bytes = (int64_t)(nb_cnt + nb_msk + sizeof ds);
    }
    return NOSPLAN_OK;
}

// ---- quota accounting with the pods split between the ranks (collective 4) -----------------------------
#include "quota_ctx.cuh"
extern "C" int nosplan_quota_recompute_impl(int device, cudaStream_t st, nosplan_quota_ctx* ctx, std::string* err,
                                            const nosplan_quotas* qs, const nosplan_quota_pods* running,
                                            const nosplan_quota_pods* pending, nosplan_quota_out* out, float* ms_h2d, float* ms_k,
                                            float* ms_d2h, int* launches, long long* h2d_bytes, long long* d2h_bytes,
                                            const nosplan_label_changes* lc, const nosplan_quota_reduce_hook* hook);
extern "C" nosplan_quota_ctx* nosplan_quota_ctx_new();

namespace {
__global__ void k_or_fold(uint32_t* mask, const uint32_t* gathered, int n, int world) {
    const int e = blockIdx.x * blockDim.x + threadIdx.x;
    if (e >= n) return;
    uint32_t m = 0;
    for (int r = 0; r < world; ++r) m |= gathered[(size_t)r * n + e];
    mask[e] = m;
}
struct ReduceCtx { nosplan* h; };
int quota_reduce(void* vctx, long long* d_used, size_t n_used, uint32_t* d_mask, size_t E, cudaStream_t st) {
    nosplan* h = static_cast<ReduceCtx*>(vctx)->h;
    nosplan_shard_ctx* c = h->sctx;
    const size_t need = al(4 * E) + al(4 * E * (size_t)c->world);
    if (need > c->cap) {
        if (c->d && cudaFree(c->d) != cudaSuccess) return NOSPLAN_E_CUDA;
        c->d = nullptr; c->cap = 0;
        if (cudaMalloc(&c->d, need + (1 << 20)) != cudaSuccess) return NOSPLAN_E_CUDA;
        c->cap = need + (1 << 20);
    }
    // int64 sums are order-independent: the reduced rows are the single-GPU rows bit for bit.  NCCL has no bit-wise OR:
    // the key masks are gathered and folded.
    if (nccl().AllReduce(d_used, d_used, n_used, ncclInt64, ncclSum, c->comm, st) != ncclSuccess) return NOSPLAN_E_CUDA;
    uint32_t* mine = reinterpret_cast<uint32_t*>(c->d);
    uint32_t* all = reinterpret_cast<uint32_t*>(c->d + al(4 * E));
    if (cudaMemcpyAsync(mine, d_mask, 4 * E, cudaMemcpyDeviceToDevice, st) != cudaSuccess) return NOSPLAN_E_CUDA;
    if (nccl().AllGather(mine, all, 4 * E, ncclUint8, c->comm, st) != ncclSuccess) return NOSPLAN_E_CUDA;
    k_or_fold<<<(unsigned)((E + 255) / 256), 256, 0, st>>>(d_mask, all, (int)E, c->world);
    return cudaGetLastError() == cudaSuccess ? NOSPLAN_OK : NOSPLAN_E_CUDA;
}
}  // namespace

extern "C" int nosplan_quota_recompute_sharded(nosplan_t* h, const nosplan_quotas* quotas, const nosplan_quota_pods* running,
                                               const nosplan_quota_pods* pending, nosplan_quota_out* out, nosplan_stats* stats) {
    if (!h) return NOSPLAN_E_INVALID;
    if (!h->sctx) return nosplan_fail(h, NOSPLAN_E_INVALID, "nosplan_quota_recompute_sharded before nosplan_shard_init");
    if (!nccl().AllReduce) return nosplan_fail(h, NOSPLAN_E_CUDA, "ncclAllReduce not found in libnccl");
    if (!h->qctx) h->qctx = nosplan_quota_ctx_new();
    ReduceCtx rc_ctx{h};
    nosplan_quota_reduce_hook hook{quota_reduce, &rc_ctx, h->sctx->rank == 0};
    float a = 0, b = 0, c = 0; int launches = 0; long long hb = 0, db = 0;
    const int rc = nosplan_quota_recompute_impl(h->device, h->stream, h->qctx, &h->err, quotas, running, pending, out, &a, &b, &c,
                                                &launches, &hb, &db, nullptr, &hook);
    if (stats) {
        std::memset(stats, 0, sizeof *stats);
        stats->ms_h2d = a; stats->ms_kernels = b; stats->ms_d2h = c; stats->kernel_launches = launches;
        stats->h2d_bytes = hb; stats->d2h_bytes = db;
    }
    return rc;
}

