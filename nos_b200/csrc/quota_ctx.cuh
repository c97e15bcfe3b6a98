// Device / pinned-host scratch shared by the quota entry points of one handle (grown on demand, reused across calls).
#pragma once
#include <cuda_runtime.h>
#include <cstddef>

struct nosplan_quota_ctx {
    unsigned char* d = nullptr; size_t d_cap = 0;
    unsigned char* hs = nullptr; size_t hs_cap = 0;
    cudaEvent_t ev[4] = {};            // phase timing of the entry points, created on first use and kept with the handle
    // nosplan_preempt_dryrun leaves its cluster state in `d` for nosplan_preempt_dryrun_again (the handle gives the preemption
    // entry points a context of their own, so no other call overwrites it); owned by preempt.cu
    void* preempt_resident = nullptr;
    void (*preempt_resident_free)(void*) = nullptr;
    // nosplan_quota_state_begin keeps the quota tables, Used and the set of accounted pods in `d` for
    // nosplan_quota_state_apply (again in a context of its own); owned by quota.cu
    void* quota_resident = nullptr;
    void (*quota_resident_free)(void*) = nullptr;
};

// the context's four timing events (created once; an error path of a call has nothing to give back)
inline cudaError_t quota_ctx_events(nosplan_quota_ctx* ctx, cudaEvent_t** out) {
    for (auto& e : ctx->ev)
        if (!e) { cudaError_t r = cudaEventCreate(&e); if (r != cudaSuccess) return r; }
    *out = ctx->ev;
    return cudaSuccess;
}

// returns cudaSuccess or the failing call's error; buffers are only ever grown
inline cudaError_t quota_ctx_ensure(nosplan_quota_ctx* ctx, size_t dev_bytes, size_t host_bytes) {
    cudaError_t e;
    if (dev_bytes > ctx->d_cap) {
        if (ctx->d && (e = cudaFree(ctx->d)) != cudaSuccess) return e;
        ctx->d = nullptr; ctx->d_cap = 0;
        const size_t cap = dev_bytes + dev_bytes / 4 + (1 << 20);
        if ((e = cudaMalloc(&ctx->d, cap)) != cudaSuccess) return e;
        ctx->d_cap = cap;
    }
    if (host_bytes > ctx->hs_cap) {
        if (ctx->hs && (e = cudaFreeHost(ctx->hs)) != cudaSuccess) return e;
        ctx->hs = nullptr; ctx->hs_cap = 0;
        const size_t cap = host_bytes + host_bytes / 4 + (1 << 20);
        if ((e = cudaMallocHost(&ctx->hs, cap)) != cudaSuccess) return e;
        ctx->hs_cap = cap;
    }
    return cudaSuccess;
}

// nosplan_quota_recompute_sharded: called between "reserve this rank's pods into Used" and everything derived from Used
struct nosplan_quota_reduce_hook {
    int (*reduce)(void* ctx, long long* d_used, size_t n_used, uint32_t* d_mask, size_t n_quotas, cudaStream_t st);
    void* ctx;
    bool first_rank;
};
