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


// ---- device side of the gather ------------------------------------------------------------------------------
// A rank's block is {header, records[cap]}; the header record carries the count so that the receivers know how
// much of each block is valid without a host round trip.
struct GatherHeader {          // overlays the first 32 bytes of an apm_fs_record
    uint32_t count;            // records valid in this block (<= cap)
    uint32_t total;            // records the rank wanted to send
    uint32_t rank;
    uint32_t status;           // APM_GS_* bits
    uint64_t ticks;            // ticks the block covers
    uint64_t pad_;
};
static_assert(sizeof(GatherHeader) <= sizeof(apm_fs_record), "header must fit one record");
enum { APM_GS_BLOCK_FULL = 1u, APM_GS_TICK_FULL = 2u, APM_GS_OUT_FULL = 4u };

struct GatherState {
    unsigned long long cursor;  // records appended to the output since the last apm_gather_alerts
    uint32_t status;            // OR of every header's status, plus APM_GS_OUT_FULL
    uint32_t collectives;
};

__device__ __forceinline__ void copy_record(apm_fs_record* dst, const apm_fs_record* src, int part) {
    reinterpret_cast<uint4*>(dst)[part] = reinterpret_cast<const uint4*>(src)[part];
}

// One CTA per tick of [tick0, tick0 + nticks): the tick's compact alert block goes behind the blocks of the ticks
// before it (their counts are summed here: at most out_ticks values).  The last CTA writes the header.
__global__ void __launch_bounds__(256) gather_pack_kernel(const apm_fs_record* __restrict__ alerts, const unsigned int* __restrict__ counts,
                                                          uint32_t per_tick_cap, uint32_t tick0, uint32_t nticks,
                                                          apm_fs_record* __restrict__ send, uint32_t cap, uint32_t rank) {
    __shared__ uint32_t s_part[8];
    const uint32_t b = blockIdx.x, tid = threadIdx.x;
    uint32_t part = 0;
    for (uint32_t j = tid; j < b && j < nticks; j += 256) {
        const uint32_t cj = counts[tick0 + j];
        part += cj < per_tick_cap ? cj : per_tick_cap;
    }
    part = __reduce_add_sync(0xffffffffu, part);
    if ((tid & 31) == 0) s_part[tid >> 5] = part;
    __syncthreads();
    uint32_t offset = 0;
#pragma unroll
    for (int j = 0; j < 8; ++j) offset += s_part[j];
    uint32_t mine = 0, status = 0;
    if (b < nticks) {
        const uint32_t cb = counts[tick0 + b];
        mine = cb < per_tick_cap ? cb : per_tick_cap;
        if (cb > per_tick_cap) status |= APM_GS_TICK_FULL;
        const apm_fs_record* src = alerts + (size_t)(tick0 + b) * per_tick_cap;
        for (uint32_t i = tid; i < mine * 8u; i += 256) {
            const uint32_t r = i >> 3;
            if (offset + r < cap) {
                uint4 v = reinterpret_cast<const uint4*>(src + r)[i & 7u];
                if ((i & 7u) == 7u) v.w = (v.w & 0xFF00FFFFu) | ((rank & 0xFFu) << 16);  // pad_[2]: the sending rank
                reinterpret_cast<uint4*>(send + 1 + offset + r)[i & 7u] = v;
            }
        }
        if (status && tid == 0) atomicOr(reinterpret_cast<uint32_t*>(send) + 3, status);  // header.status (cleared before the launch)
    }
    if (b == gridDim.x - 1 && tid == 0) {
        GatherHeader* h = reinterpret_cast<GatherHeader*>(send);
        const uint32_t total = offset + mine;
        h->count = total < cap ? total : cap;
        h->total = total;
        h->rank = rank;
        h->ticks = nticks;
        if (total > cap) atomicOr(&h->status, (uint32_t)APM_GS_BLOCK_FULL);
    }
}

// One CTA per rank: the valid prefix of its block is appended to the output behind the prefixes of the ranks before
// it.  `out` is device memory or page-locked host memory (written over PCIe by the stores below).
__global__ void __launch_bounds__(256) gather_unpack_kernel(const apm_fs_record* __restrict__ recv, uint32_t nranks, uint32_t cap,
                                                            apm_fs_record* __restrict__ out, unsigned long long out_cap,
                                                            GatherState* __restrict__ state) {
    const uint32_t r = blockIdx.x, tid = threadIdx.x;
    const size_t stride = (size_t)cap + 1;
    unsigned long long base = state->cursor;
    for (uint32_t j = 0; j < r; ++j) base += reinterpret_cast<const GatherHeader*>(recv + j * stride)->count;
    const GatherHeader* h = reinterpret_cast<const GatherHeader*>(recv + r * stride);
    const uint32_t n = h->count <= cap ? h->count : cap;
    const apm_fs_record* src = recv + r * stride + 1;
    bool over = false;
    for (uint32_t i = tid; i < n * 8u; i += 256) {
        const uint32_t k = i >> 3;
        if (base + k < out_cap) copy_record(out + base + k, src + k, (int)(i & 7u));
        else over = true;
    }
    uint32_t st = h->status | (over ? (uint32_t)APM_GS_OUT_FULL : 0u);
    if (h->count > cap) st |= APM_GS_BLOCK_FULL;
    if (tid == 0 && st) atomicOr(&state->status, st);
    if (__syncthreads_or(over) && tid == 0) atomicOr(&state->status, (uint32_t)APM_GS_OUT_FULL);
}

// behind the unpack: the cursor moves past what all ranks appended
__global__ void gather_advance_kernel(const apm_fs_record* __restrict__ recv, uint32_t nranks, uint32_t cap, GatherState* state) {
    if (threadIdx.x != 0 || blockIdx.x != 0) return;
    unsigned long long total = 0;
    for (uint32_t j = 0; j < nranks; ++j) {
        const uint32_t cnt = reinterpret_cast<const GatherHeader*>(recv + (size_t)j * ((size_t)cap + 1))->count;
        total += cnt <= cap ? cnt : cap;
    }
    state->cursor += total;
    state->collectives += 1;
}

// pack -> all-gather -> unpack of the ticks [tick0, tick0 + nticks) on the gather stream, which has already been
// ordered behind the kernels that wrote those ticks' alert blocks.  Nothing here waits for the device.
apm_status enqueue_gather(apm_ctx* c, uint32_t tick0, uint32_t nticks, apm_fs_record* out_dev, uint64_t out_cap) {
    cudaStream_t st = c->s_comm;
    const uint32_t cap = c->gather_cap;
    CUDA_TRY(c, cudaMemsetAsync(c->d_gather_send, 0, sizeof(apm_fs_record), st));  // header: status bits are OR-ed in
    gather_pack_kernel<<<nticks ? nticks : 1, 256, 0, st>>>(c->d_alerts, c->d_alert_count, c->cfg.max_alerts_per_tick, tick0, nticks,
                                                           c->d_gather_send, cap, (uint32_t)c->rank);
    CUDA_TRY(c, cudaGetLastError());
    nccl_result_t r = nccl().AllGather(c->d_gather_send, c->d_gather_recv, ((size_t)cap + 1) * sizeof(apm_fs_record), NCCL_INT8,
                                       (nccl_comm_t)c->nccl_comm, st);
    if (r) return nccl_fail(c, r, "ncclAllGather (alert blocks)");
    GatherState* gs = (GatherState*)c->d_gather_state;
    gather_unpack_kernel<<<c->nranks, 256, 0, st>>>(c->d_gather_recv, (uint32_t)c->nranks, cap, out_dev, out_cap, gs);
    gather_advance_kernel<<<1, 32, 0, st>>>(c->d_gather_recv, (uint32_t)c->nranks, cap, gs);
    CUDA_TRY(c, cudaGetLastError());
    CUDA_TRY(c, cudaMemcpyAsync(c->h_gather_state, gs, sizeof(GatherState), cudaMemcpyDeviceToHost, st));
    c->launches += 3;
    return APM_OK;
}

// Every rank must send blocks of the same size, whatever it asked for (ranks own different numbers of keys): the
// ranks exchange their requests and all take the largest.  A collective in itself - every rank calls it.
apm_status alloc_gather(apm_ctx* c, uint32_t want) {
    uint32_t* d_req = nullptr;
    CUDA_TRY(c, cudaMalloc((void**)&d_req, sizeof(uint32_t) * 4 * ((size_t)c->nranks + 1)));
    const uint32_t mine[4] = {want, 0, 0, 0};
    CUDA_TRY(c, cudaMemcpyAsync(d_req, mine, sizeof mine, cudaMemcpyHostToDevice, c->s_comm));
    nccl_result_t r = nccl().AllGather(d_req, d_req + 4, sizeof mine, NCCL_INT8, (nccl_comm_t)c->nccl_comm, c->s_comm);
    if (r) { cudaFree(d_req); return nccl_fail(c, r, "ncclAllGather (block capacity)"); }
    std::vector<uint32_t> all((size_t)c->nranks * 4);
    CUDA_TRY(c, cudaMemcpyAsync(all.data(), d_req + 4, all.size() * sizeof(uint32_t), cudaMemcpyDeviceToHost, c->s_comm));
    CUDA_TRY(c, cudaStreamSynchronize(c->s_comm));
    cudaFree(d_req);
    uint32_t cap = want;
    for (int rk = 0; rk < c->nranks; ++rk) cap = std::max(cap, all[(size_t)rk * 4]);
    cudaFree(c->d_gather_send); cudaFree(c->d_gather_recv); cudaFree(c->d_gather_stage);
    c->d_gather_send = c->d_gather_recv = c->d_gather_stage = nullptr;
    c->gather_cap = cap;
    CUDA_TRY(c, cudaMalloc((void**)&c->d_gather_send, ((size_t)cap + 1) * sizeof(apm_fs_record)));
    CUDA_TRY(c, cudaMalloc((void**)&c->d_gather_recv, (size_t)c->nranks * ((size_t)cap + 1) * sizeof(apm_fs_record)));
    return APM_OK;
}

// device address of a caller's output buffer if it is page-locked host memory (or device memory), else nullptr
apm_fs_record* device_view(apm_fs_record* p) {
    cudaPointerAttributes at;
    if (cudaPointerGetAttributes(&at, p) != cudaSuccess) { cudaGetLastError(); return nullptr; }
    if (at.type == cudaMemoryTypeHost || at.type == cudaMemoryTypeDevice || at.type == cudaMemoryTypeManaged)
        return (apm_fs_record*)at.devicePointer;
    return nullptr;
}

}  // namespace

void apm_comm_destroy_internal(apm_ctx* c) {
    if (c->s_comm) cudaStreamSynchronize(c->s_comm);
    if (c->nccl_comm && nccl().CommDestroy) nccl().CommDestroy((nccl_comm_t)c->nccl_comm);
    c->nccl_comm = nullptr;
    cudaFree(c->d_gather_send); cudaFree(c->d_gather_recv); cudaFree(c->d_gather_stage); cudaFree(c->d_gather_state);
    cudaFreeHost(c->h_gather_state);
    c->d_gather_send = c->d_gather_recv = c->d_gather_stage = nullptr;
    c->d_gather_state = c->h_gather_state = nullptr;
    if (c->ev_comm) cudaEventDestroy(c->ev_comm);
    if (c->s_comm) cudaStreamDestroy(c->s_comm);
    c->ev_comm = nullptr;
    c->s_comm = nullptr;
}

// Bound mode (apm_bind_gather_output): the alerts of the tick whose compaction was just queued on `after` are
// all-gathered on the gather stream and appended to the bound buffer while the following ticks run.
apm_status apm_comm_tick_gather(apm_ctx* c, uint32_t alert_slot, cudaStream_t after) {
    if (!c->nccl_comm || !c->h_gather_bound) return APM_OK;
    CUDA_TRY(c, cudaEventRecord(c->ev_comm, after));
    CUDA_TRY(c, cudaStreamWaitEvent(c->s_comm, c->ev_comm, 0));
    apm_status st = enqueue_gather(c, alert_slot, 1, device_view(c->h_gather_bound), c->h_gather_bound_cap);
    if (st) return st;
    c->alerts_gather_from = alert_slot + 1;
    return APM_OK;
}

// apm_release_ticks: the gather stream may still be reading the alert blocks that are about to be reused
apm_status apm_comm_release(apm_ctx* c) {
    c->alerts_gather_from = 0;
    if (!c->s_comm) return APM_OK;
    CUDA_TRY(c, cudaStreamSynchronize(c->s_comm));
    return APM_OK;
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
    if (c->nccl_comm) return apm_fail(c, APM_E_STATE, "comm_init: the context already has a communicator");
    NcclApi& a = nccl();
    if (!a.error.empty()) return apm_fail(c, APM_E_COMM, a.error);
    {   // no first-use kernel load may happen while a collective is in flight (see apm_preload_kernels)
        apm_status pl = apm_preload_kernels(c->cfg.device);
        if (pl) return apm_fail(c, pl, apm_last_error(nullptr));
    }
    nccl_unique_id id;
    memcpy(&id, unique_id128, 128);
    nccl_comm_t comm = nullptr;
    nccl_result_t r = a.CommInitRank(&comm, nranks, id, rank);
    if (r) return nccl_fail(c, r, "ncclCommInitRank");
    c->nccl_comm = comm;
    c->rank = rank;
    c->nranks = nranks;
    CUDA_TRY(c, cudaStreamCreateWithFlags(&c->s_comm, cudaStreamNonBlocking));
    CUDA_TRY(c, cudaEventCreateWithFlags(&c->ev_comm, cudaEventDisableTiming));
    CUDA_TRY(c, cudaMalloc(&c->d_gather_state, sizeof(GatherState)));
    CUDA_TRY(c, cudaMemset(c->d_gather_state, 0, sizeof(GatherState)));
    CUDA_TRY(c, cudaMallocHost(&c->h_gather_state, sizeof(GatherState)));
    memset(c->h_gather_state, 0, sizeof(GatherState));
    // one tick always fits a block; a host that gathers many ticks at once raises it (apm_comm_set_capacity)
    return alloc_gather(c, c->cfg.max_alerts_per_tick);
}

apm_status apm_comm_set_capacity(apm_ctx* c, uint32_t records) {
    if (!c) return APM_E_INVALID;
    APM_ENTER(c);
    if (!c->nccl_comm) return apm_fail(c, APM_E_STATE, "apm_comm_init has not been called");
    if (records == 0) return apm_fail(c, APM_E_INVALID, "comm_set_capacity: at least one record");
    { apm_status fl_ = apm_flush_pending(c); if (fl_) return fl_; }
    CUDA_TRY(c, cudaStreamSynchronize(c->s_base));
    CUDA_TRY(c, cudaStreamSynchronize(c->s_comm));
    return alloc_gather(c, records);
}

apm_status apm_bind_gather_output(apm_ctx* c, apm_fs_record* pinned, uint64_t cap_records) {
    if (!c) return APM_E_INVALID;
    APM_ENTER(c);
    if (!c->nccl_comm) return apm_fail(c, APM_E_STATE, "apm_comm_init has not been called");
    { apm_status fl_ = apm_flush_pending(c); if (fl_) return fl_; }
    CUDA_TRY(c, cudaStreamSynchronize(c->s_base));
    CUDA_TRY(c, cudaStreamSynchronize(c->s_comm));
    if (pinned) {
        if (c->alerts_gather_from != c->ticks_buffered)
            return apm_fail(c, APM_E_STATE, "bind_gather_output: gather (or release) the retained ticks first");
        cudaPointerAttributes at;
        if (cudaPointerGetAttributes(&at, pinned) != cudaSuccess || at.type != cudaMemoryTypeHost) {
            cudaGetLastError();
            return apm_fail(c, APM_E_INVALID, "bind_gather_output: the buffer must be page-locked host memory (apm_host_alloc)");
        }
    }
    c->h_gather_bound = pinned;
    c->h_gather_bound_cap = pinned ? cap_records : 0;
    return APM_OK;
}

apm_status apm_gather_alerts(apm_ctx* c, apm_fs_record* out, uint64_t cap, uint64_t* n) {
    if (!c || !n) return APM_E_INVALID;
    APM_ENTER(c);
    *n = 0;
    if (!c->nccl_comm) return apm_fail(c, APM_E_STATE, "apm_comm_init has not been called");
    { apm_status fl_ = apm_flush_pending(c); if (fl_) return fl_; }  // in bound mode this queues the last tick's gather
    const bool bound = c->h_gather_bound != nullptr;
    if (bound && out != c->h_gather_bound) return apm_fail(c, APM_E_INVALID, "gather_alerts: pass the bound output buffer (apm_bind_gather_output)");
    apm_fs_record* out_dev = nullptr;
    bool staged = false;
    if (!bound) {
        // exactly one collective per call on every rank, whatever the number of pending ticks: the ranks tick in
        // step (ticks are global), and a data-dependent number of collectives could leave a rank waiting for ever
        out_dev = out ? device_view(out) : nullptr;
        uint64_t out_cap = cap;
        if (!out_dev) {  // pageable (or no) output: unpack into device memory, copy afterwards
            const size_t stage = (size_t)c->nranks * c->gather_cap;
            if (!c->d_gather_stage) CUDA_TRY(c, cudaMalloc((void**)&c->d_gather_stage, stage * sizeof(apm_fs_record)));
            out_dev = c->d_gather_stage;
            out_cap = stage;
            staged = true;
        }
        CUDA_TRY(c, cudaEventRecord(c->ev_comm, c->s_base));  // the alert blocks are written on the baseline stream
        CUDA_TRY(c, cudaStreamWaitEvent(c->s_comm, c->ev_comm, 0));
        const uint32_t from = c->alerts_gather_from < c->ticks_buffered ? c->alerts_gather_from : c->ticks_buffered;
        apm_status st = enqueue_gather(c, from, c->ticks_buffered - from, out_dev, out_cap);
        if (st) return st;
        c->alerts_gather_from = c->ticks_buffered;
    }
    CUDA_TRY(c, cudaStreamSynchronize(c->s_comm));  // the one wait of the call
    GatherState gs = *(GatherState*)c->h_gather_state;
    CUDA_TRY(c, cudaMemsetAsync(c->d_gather_state, 0, sizeof(GatherState), c->s_comm));  // the next call appends from 0
    memset(c->h_gather_state, 0, sizeof(GatherState));
    // every rank saw every header, so these errors are returned by all ranks alike
    if (gs.status & APM_GS_TICK_FULL) return apm_fail(c, APM_E_CAPACITY, "a rank had more alerts in one tick than max_alerts_per_tick");
    if (gs.status & APM_GS_BLOCK_FULL)
        return apm_fail(c, APM_E_CAPACITY, "a rank's alerts exceeded the gather block (apm_comm_set_capacity): gather more often or raise it");
    if (gs.status & APM_GS_OUT_FULL) return apm_fail(c, APM_E_CAPACITY, "gather output buffer too small");
    if (staged && gs.cursor) {
        if (gs.cursor > cap || !out) return apm_fail(c, APM_E_CAPACITY, "gather output buffer too small");
        CUDA_TRY(c, cudaMemcpyAsync(out, c->d_gather_stage, gs.cursor * sizeof(apm_fs_record), cudaMemcpyDeviceToHost, c->s_comm));
        CUDA_TRY(c, cudaStreamSynchronize(c->s_comm));
    }
    // tick-major (bound mode) or one block per rank (on demand); inside a rank's block (tick, lag, key)-ascending
    c->counters.d2h_bytes += gs.cursor * sizeof(apm_fs_record) + sizeof(GatherState);
    c->counters.alerts += gs.cursor;
    *n = gs.cursor;
    return APM_OK;
}

}  // extern "C"

// Loads this file's kernels on the current device (see apm_preload_kernels in capi.cu).
cudaError_t apm_preload_comm() {
    const void* fns[] = {(const void*)gather_advance_kernel, (const void*)gather_pack_kernel, (const void*)gather_unpack_kernel};
    cudaError_t first = cudaSuccess;
    for (const void* f : fns) {
        cudaFuncAttributes a;
        const cudaError_t e = apm_preload_one(&a, f);
        if (e != cudaSuccess && first == cudaSuccess) first = e;
    }
    return first;
}
