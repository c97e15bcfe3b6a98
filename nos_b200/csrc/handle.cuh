// The engine handle behind nosplan_t (include/nosplan.h).  Internal to libnosplan.
#pragma once
#include <string>
#include <vector>

#include "nosplan_dev.cuh"

struct nosplan_quota_ctx;
struct nosplan_shard_ctx;          // in-library NCCL state of the node-sharded Plan() (shard_nccl.cu)

namespace nsp {
struct Layout {            // byte offsets into the arena / staging buffer
    size_t model_off, geom, node_gpu_count, gpu_model, gpu_mem, gpu_used, gpu_free, gpu_umask, gpu_fmask,
        node_alloc, node_req, node_allowed, node_npods, nodes_end,
        tot, lack, stats, maxfree, gflags, acc, lack_dec, next_node, wave, work_end, cand_flag, cand_idx, rec_dw, rec_pl, rec_fl, rec_chunk, rec_chunk_fl, gate_state, out_count, out_mask, out_end,
        pod_cpu, pod_mem, pod_eph, pod_key, pod_flags, pod_sptr, pod_scol, pod_sqty, runs, run_chunk_off,
        cmin_cpu, cmin_mem, cmin_eph, fp_blob, upload_end, placed, entry_lack, total;
};
}  // namespace nsp

struct nosplan {
    int device = -1;
    cudaStream_t stream = nullptr;
    cudaEvent_t ev[12] = {};
    // device arena + pinned staging (grown on demand, reused across calls)
    unsigned char* d_arena = nullptr; size_t d_cap = 0;
    unsigned char* h_stage = nullptr; size_t h_cap = 0;
    unsigned char* h_out = nullptr; size_t h_out_cap = 0;
    // current problem
    bool uploaded = false, ran = false;
    bool kernel_events = true;                  // CUDA events between the kernels of a run (per-kernel times in nosplan_stats)
    bool in_plan = false, direct_h2d = false;   // H2D straight from page-locked caller memory (see nosplan_upload)
    nsp::DevProblem pb{};
    int32_t* d_cand_idx = nullptr;
    unsigned char* d_work = nullptr; size_t work_bytes = 0;      // totals, lacking vector, counters, tracker tables: reset before k_prepass
    size_t small_bytes = 0; unsigned run_parity = 1, bound_copy = 0;   // which copy of the small per-run state pb points at
    bool pre_done = false, split_run = false;                    // nosplan_plan queued the node-only kernels from inside upload()
    // host scratch reused across calls
    std::vector<int> v_first_prof, v_order;
    std::vector<uint32_t> v_flags;
    std::vector<unsigned long long> v_cls;
    std::vector<nsp::RunRec> v_runs;
    std::vector<int32_t> v_chunk_off;
    std::vector<long long> v_cmin;
    std::vector<uint8_t> v_key_seen;
    // single-launch plan (fused.cu): applicable to the uploaded problem; its verdict has not been read back yet (the host
    // runs the classic kernel sequence if the kernel rejected its own hypothesis); the classic sequence left key-indexed
    // bits in `placed`; SM count of this handle's device
    bool fused_ok = false, fused_pending = false, placed_dirty = false;
    bool last_fused = false; int fused_reject = 0; float ms_rejected = 0.f;   // what the last run was; why / how long a rejected attempt took
    int (*rerun)(nosplan*) = nullptr;
    // nosplan_plan / nosplan_plan_resident stage only what the single-launch kernel reads; the classic kernels' pod arrays are
    // staged (from the caller's arrays, still valid inside that call) only if the single-launch plan is rejected
    nsp::Layout lay{};
    bool classic_staged = true, lazy_valid = false;
    nosplan_pods lazy_pods{}; const uint8_t* lazy_pre = nullptr; int lazy_filter = 0;
    bool prepass_only = false;                  // the last run was k_prepass alone (nothing lacking after a rejected single-launch plan)
    bool force_classic = false;                 // the general pod path and the classic kernels, whatever the problem looks like
    std::vector<uint16_t> v_bucket;           // the classic form of the last run: what a rejected single-launch plan is replaced by
    int sms = 148, par_blocks = 148 * 4, wave_blocks = 148 * 4;   // SM count and resident-wave grid sizes of this handle's device
    size_t out_count_bytes = 0, out_mask_bytes = 0;
    int launches = 0;
    size_t h2d_bytes = 0, compact_h2d = 0;
    int32_t* d_flag = nullptr; int32_t* h_flag = nullptr;       // nosplan_plan_compact: first malformed record, found on the GPU
    nosplan_quota_ctx* qctx = nullptr;
    nosplan_quota_ctx* ectx = nullptr;          // the resident quota state of nosplan_quota_state_begin / apply
    nosplan_quota_ctx* pctx = nullptr;          // the preemption entry points' own scratch: holds the resident cluster state
    nosplan_shard_ctx* sctx = nullptr;
    std::string err;
    // snapshot kept resident between calls (nosplan_plan_resident / nosplan_update_nodes): shape and small tables of the last
    // full upload (node arrays of res_snap are not kept: they live in the arena), byte offsets of the node sections
    bool nodes_resident = false;
    nosplan_snapshot res_snap{};
    std::vector<int32_t> res_pmem, res_pgi, res_model_off;
    std::vector<uint8_t> res_geom;
    size_t res_off[11] = {}; size_t res_nodes_end = 0;
    // NOSPLAN_TRACE=1 only: one event per H2D copy of the node arrays
    cudaEvent_t dbg_ev[32] = {}; size_t dbg_bytes[32] = {}; bool dbg_direct[32] = {}; int dbg_n = 0;
};

int nosplan_fail(nosplan* h, int code, const char* fmt, ...);
int nosplan_resolve(nosplan* h);      // waits for a pending single-launch plan and re-plans on the classic path if it was rejected
void nosplan_phase_times(nosplan* h, nosplan_stats* stats);
int nosplan_slice_begin_ex(nosplan* h, void* d_totals);
int nosplan_slice_mid_ex(nosplan* h, int32_t rank, int32_t world, const void* d_totals_gathered, void* d_exchange, int64_t exchange_bytes,
                         const nsp::PeerWait& wait);
int nosplan_run_sliced_fused_begin(nosplan* h, const nsp::PeerBox* box, unsigned epoch, unsigned char* d_exchange, size_t xb);
int nosplan_run_sliced_fused_end(nosplan* h, const void* blocks, size_t stride, int world, const nsp::PeerWait& wait, int (*rerun)(nosplan*));
int nosplan_shard_end_ex(nosplan* h, const void* d_gathered, int32_t world, size_t stride, const nsp::PeerWait& wait);
