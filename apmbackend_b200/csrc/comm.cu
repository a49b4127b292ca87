// Multi-GPU: keys are hash-sharded by the host, every rank owns its keys' buckets and
// baselines outright, and the only exchange is an NCCL all-gather of the compact alert
// records.  (The reference has no counterpart: its stages talk over RabbitMQ queues,
// queue.js:67-305.)
//
// NCCL is bound at run time with dlopen so that a process that already carries a NCCL
// (torch's bundled copy) shares it instead of loading a second one.
#include <dlfcn.h>

#include <algorithm>
#include <cstring>
#include <vector>

#include "ctx.h"

namespace {

typedef struct { char internal[128]; } nccl_unique_id;
typedef void* nccl_comm_t;
typedef int nccl_result_t;

struct NcclApi {
    void* handle = nullptr;
    nccl_result_t (*GetUniqueId)(nccl_unique_id*) = nullptr;
    nccl_result_t (*CommInitRank)(nccl_comm_t*, int, nccl_unique_id, int) = nullptr;
    nccl_result_t (*CommDestroy)(nccl_comm_t) = nullptr;
    nccl_result_t (*AllGather)(const void*, void*, size_t, int /*dtype*/, nccl_comm_t, cudaStream_t) = nullptr;
    const char* (*GetErrorString)(nccl_result_t) = nullptr;
    std::string error;
};

NcclApi& nccl() {
    static NcclApi api;
    if (api.handle || !api.error.empty()) return api;
    const char* names[] = {"libnccl.so.2", "libnccl.so"};
    for (const char* n : names) {
        api.handle = dlopen(n, RTLD_NOW | RTLD_GLOBAL);
        if (api.handle) break;
    }
    if (!api.handle) {
        api.error = std::string("libnccl.so.2 not found: ") + dlerror();
        return api;
    }
    api.GetUniqueId = (decltype(api.GetUniqueId))dlsym(api.handle, "ncclGetUniqueId");
    api.CommInitRank = (decltype(api.CommInitRank))dlsym(api.handle, "ncclCommInitRank");
    api.CommDestroy = (decltype(api.CommDestroy))dlsym(api.handle, "ncclCommDestroy");
    api.AllGather = (decltype(api.AllGather))dlsym(api.handle, "ncclAllGather");
    api.GetErrorString = (decltype(api.GetErrorString))dlsym(api.handle, "ncclGetErrorString");
    if (!api.GetUniqueId || !api.CommInitRank || !api.CommDestroy || !api.AllGather || !api.GetErrorString)
        api.error = "libnccl is missing an expected symbol";
    return api;
}

constexpr int NCCL_INT8 = 0;  // ncclInt8 / ncclChar

apm_status nccl_fail(apm_ctx* c, nccl_result_t r, const char* what) {
    return apm_fail(c, APM_E_COMM, std::string(what) + ": " + (nccl().GetErrorString ? nccl().GetErrorString(r) : "?"));
}

}  // namespace

void apm_comm_destroy_internal(apm_ctx* c) {
    if (c->nccl_comm && nccl().CommDestroy) nccl().CommDestroy((nccl_comm_t)c->nccl_comm);
    c->nccl_comm = nullptr;
}

extern "C" {

apm_status apm_comm_unique_id(void* out128) {
    if (!out128) return APM_E_INVALID;
    NcclApi& a = nccl();
    if (!a.error.empty()) return apm_fail(nullptr, APM_E_COMM, a.error);
    nccl_unique_id id;
    nccl_result_t r = a.GetUniqueId(&id);
    if (r) return nccl_fail(nullptr, r, "ncclGetUniqueId");
    memcpy(out128, &id, 128);
    return APM_OK;
}

apm_status apm_comm_init(apm_ctx* c, int32_t rank, int32_t nranks, const void* unique_id128) {
    if (!c || !unique_id128 || nranks < 1 || rank < 0 || rank >= nranks) return APM_E_INVALID;
    APM_ENTER(c);
    if (!c->d_alerts) return apm_fail(c, APM_E_STATE, "comm_init needs APM_F_ALERTS | APM_F_ZSCORE");
    NcclApi& a = nccl();
    if (!a.error.empty()) return apm_fail(c, APM_E_COMM, a.error);
    nccl_unique_id id;
    memcpy(&id, unique_id128, 128);
    nccl_comm_t comm = nullptr;
    nccl_result_t r = a.CommInitRank(&comm, nranks, id, rank);
    if (r) return nccl_fail(c, r, "ncclCommInitRank");
    c->nccl_comm = comm;
    c->rank = rank;
    c->nranks = nranks;
    // send area: header + this rank's alerts of all retained ticks; receive area: nranks headers, then nranks
    // record blocks of the same capacity
    const size_t G = (size_t)c->cfg.max_alerts_per_tick * c->cfg.out_ticks;
    CUDA_TRY(c, cudaMalloc((void**)&c->d_gather, ((size_t)(nranks + 1) * (G + 1)) * sizeof(apm_fs_record)));
    return APM_OK;
}

apm_status apm_gather_alerts(apm_ctx* c, apm_fs_record* out, uint64_t cap, uint64_t* n) {
    if (!c || !n) return APM_E_INVALID;
    APM_ENTER(c);
    *n = 0;
    if (!c->nccl_comm) return apm_fail(c, APM_E_STATE, "apm_comm_init has not been called");
    { apm_status fl_ = apm_flush_pending(c); if (fl_) return fl_; }
    CUDA_TRY(c, cudaStreamSynchronize(c->s_base));  // the alert blocks are written on the baseline stream
    const size_t A = c->cfg.max_alerts_per_tick, G = A * c->cfg.out_ticks;
    apm_fs_record* send_hdr = c->d_gather;                       // [1]
    apm_fs_record* send_rec = c->d_gather + 1;                   // [G]
    apm_fs_record* recv_hdr = c->d_gather + 1 + G;               // [nranks]
    apm_fs_record* recv_rec = recv_hdr + c->nranks;              // [nranks][<= G]
    // pack this rank's alerts of the retained ticks
    std::vector<unsigned int> counts(c->ticks_buffered);
    if (c->ticks_buffered) {
        CUDA_TRY(c, cudaMemcpyAsync(counts.data(), c->d_alert_count, sizeof(unsigned int) * c->ticks_buffered, cudaMemcpyDeviceToHost, c->s_main));
        CUDA_TRY(c, cudaStreamSynchronize(c->s_main));
    }
    size_t total = 0;
    for (uint32_t t = 0; t < c->ticks_buffered; ++t) {
        if (counts[t] > A) return apm_fail(c, APM_E_CAPACITY, "more alerts in one tick than max_alerts_per_tick");
        if (counts[t])
            CUDA_TRY(c, cudaMemcpyAsync(send_rec + total, c->d_alerts + (size_t)t * A, counts[t] * sizeof(apm_fs_record), cudaMemcpyDeviceToDevice, c->s_main));
        total += counts[t];
    }
    apm_fs_record hdr;
    memset(&hdr, 0, sizeof hdr);
    hdr.key = (uint32_t)total;
    hdr.lag = (uint32_t)c->rank;
    hdr.tick_ts = (int64_t)c->counters.ticks;
    CUDA_TRY(c, cudaMemcpyAsync(send_hdr, &hdr, sizeof hdr, cudaMemcpyHostToDevice, c->s_main));
    // 1) the counts, 2) the records, padded to the largest count of any rank
    nccl_result_t r = nccl().AllGather(send_hdr, recv_hdr, sizeof(apm_fs_record), NCCL_INT8, (nccl_comm_t)c->nccl_comm, c->s_main);
    if (r) return nccl_fail(c, r, "ncclAllGather (headers)");
    std::vector<apm_fs_record> hdrs(c->nranks);
    CUDA_TRY(c, cudaMemcpyAsync(hdrs.data(), recv_hdr, sizeof(apm_fs_record) * c->nranks, cudaMemcpyDeviceToHost, c->s_main));
    CUDA_TRY(c, cudaStreamSynchronize(c->s_main));
    size_t most = 0, sum = 0;
    for (int rk = 0; rk < c->nranks; ++rk) {
        if (hdrs[rk].key > G) return apm_fail(c, APM_E_COMM, "corrupt alert block header");
        most = std::max<size_t>(most, hdrs[rk].key);
        sum += hdrs[rk].key;
    }
    if (sum > cap) return apm_fail(c, APM_E_CAPACITY, "gather output buffer too small");
    if (most) {
        r = nccl().AllGather(send_rec, recv_rec, most * sizeof(apm_fs_record), NCCL_INT8, (nccl_comm_t)c->nccl_comm, c->s_main);
        if (r) return nccl_fail(c, r, "ncclAllGather (records)");
    }
    uint64_t w = 0;
    for (int rk = 0; rk < c->nranks; ++rk) {
        const uint32_t k = hdrs[rk].key;
        if (k && out)
            CUDA_TRY(c, cudaMemcpyAsync(out + w, recv_rec + (size_t)rk * most, k * sizeof(apm_fs_record), cudaMemcpyDeviceToHost, c->s_main));
        w += k;
    }
    CUDA_TRY(c, cudaStreamSynchronize(c->s_main));
    // per rank the records are (tick, lag, key)-ascending: ticks in order, each tick's block compacted in order
    c->counters.d2h_bytes += (w + c->nranks) * sizeof(apm_fs_record);
    c->counters.alerts += w;
    *n = w;
    return APM_OK;
}

}  // extern "C"
