// K5: time-ordered transaction passthrough.
//
// Replaces the min-heap of the stats stage (binary_heap.js; stream_calc_stats.js:133 push,
// :136-155 writeHeapToQueue -> popAllLessOrEqualToScore) for hosts that do not want to keep a
// JS heap of every transaction: the device holds one word per pending transaction,
//     (endTs - base) << 24 | (arrival sequence mod 2^24),
// and at every tick hands back, in non-decreasing endTs order, the sequence numbers of all
// transactions with endTs <= edgeTs; the host maps them to the TxEntry strings it forwards to
// db_insert.  Ties are in arrival order (the reference's order among equal scores depends on the
// heap's structure and is unspecified).
#include "ctx.h"

namespace {

constexpr int PT_THREADS = 256;
constexpr int PT_ITEMS = 4;
constexpr int PT_SEQ_BITS = 24;

__device__ __forceinline__ uint32_t block_exclusive_scan(uint32_t v, uint32_t* s_warp, uint32_t* total) {
    const int lane = threadIdx.x & 31, w = threadIdx.x >> 5;
    uint32_t incl = v;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
        const uint32_t y = __shfl_up_sync(0xffffffffu, incl, o);
        if (lane >= o) incl += y;
    }
    if (lane == 31) s_warp[w] = incl;
    __syncthreads();
    uint32_t woff = 0, tot = 0;
#pragma unroll
    for (int j = 0; j < PT_THREADS / 32; ++j) {
        if (j < w) woff += s_warp[j];
        tot += s_warp[j];
    }
    __syncthreads();
    *total = tot;
    return woff + incl - v;
}

// every valid transaction of the segment joins the pending set (addData's minHeap.push, :133)
__global__ void __launch_bounds__(PT_THREADS) pt_append_kernel(const uint32_t* __restrict__ svc, const uint32_t* __restrict__ jvm,
                                                               const int64_t* __restrict__ ts, const int32_t* __restrict__ el,
                                                               uint64_t begin, uint64_t end, uint32_t num_services, uint32_t num_jvms,
                                                               uint64_t seq0, int64_t base_ms, uint64_t* __restrict__ pend,
                                                               unsigned long long* __restrict__ cursor, uint64_t cap,
                                                               unsigned long long* __restrict__ overflow) {
    __shared__ uint32_t s_warp[PT_THREADS / 32];
    __shared__ unsigned long long s_base;
    const uint64_t i0 = begin + ((uint64_t)blockIdx.x * PT_THREADS + threadIdx.x) * PT_ITEMS;
    uint64_t w[PT_ITEMS];
    uint32_t cnt = 0;
#pragma unroll
    for (int j = 0; j < PT_ITEMS; ++j) {
        const uint64_t i = i0 + j;
        w[j] = ~0ull;
        if (i < end) {
            const int64_t t = ts[i];
            if (svc[i] < num_services && jvm[i] < num_jvms && el[i] >= 0 && t >= 0) {
                int64_t d = t - base_ms;
                if (d < 0) d = 0;
                w[j] = ((uint64_t)d << PT_SEQ_BITS) | ((seq0 + (i - begin)) & ((1ull << PT_SEQ_BITS) - 1));
                cnt++;
            }
        }
    }
    uint32_t total;
    uint32_t off = block_exclusive_scan(cnt, s_warp, &total);
    if (threadIdx.x == 0) s_base = total ? atomicAdd(cursor, (unsigned long long)total) : 0ull;
    __syncthreads();
    unsigned long long pos = s_base + off;
#pragma unroll
    for (int j = 0; j < PT_ITEMS; ++j)
        if (w[j] != ~0ull) {
            if (pos < cap) pend[pos] = w[j]; else atomicAdd(overflow, 1ull);
            ++pos;
        }
}

// splits the pending words into those due now (endTs <= edgeTs) and those that stay
__global__ void __launch_bounds__(PT_THREADS) pt_partition_kernel(const uint64_t* __restrict__ pend, uint64_t n, uint64_t edge_word,
                                                                  uint64_t* __restrict__ emit, unsigned long long* __restrict__ emit_cursor,
                                                                  uint64_t* __restrict__ keep, unsigned long long* __restrict__ keep_cursor) {
    __shared__ uint32_t s_warp[PT_THREADS / 32];
    __shared__ unsigned long long s_base[2];
    const uint64_t i0 = ((uint64_t)blockIdx.x * PT_THREADS + threadIdx.x) * PT_ITEMS;
    uint64_t w[PT_ITEMS];
    uint32_t ne = 0, nk = 0;
#pragma unroll
    for (int j = 0; j < PT_ITEMS; ++j) {
        w[j] = i0 + j < n ? pend[i0 + j] : ~0ull;
        if (i0 + j < n) { if (w[j] <= edge_word) ne++; else nk++; }
    }
    uint32_t te, tk;
    const uint32_t oe = block_exclusive_scan(ne, s_warp, &te);
    const uint32_t ok = block_exclusive_scan(nk, s_warp, &tk);
    if (threadIdx.x == 0) {
        s_base[0] = te ? atomicAdd(emit_cursor, (unsigned long long)te) : 0ull;
        s_base[1] = tk ? atomicAdd(keep_cursor, (unsigned long long)tk) : 0ull;
    }
    __syncthreads();
    unsigned long long pe = s_base[0] + oe, pk = s_base[1] + ok;
#pragma unroll
    for (int j = 0; j < PT_ITEMS; ++j)
        if (i0 + j < n) {
            if (w[j] <= edge_word) emit[pe++] = w[j]; else keep[pk++] = w[j];
        }
}

}  // namespace

apm_status apm_pt_append(apm_ctx* c, const apm_tx_columns& d, uint64_t begin, uint64_t end) {
    if (!c->pt_pend[0] || end <= begin) return APM_OK;
    ProfScope prof_scope_(c, APM_PG_PASSTHROUGH);
    if (c->pt_base_ms == INT64_MIN) return apm_fail(c, APM_E_STATE, "passthrough: no tick has set the time base yet");
    const uint64_t n = end - begin, per_block = (uint64_t)PT_THREADS * PT_ITEMS;
    pt_append_kernel<<<(unsigned)((n + per_block - 1) / per_block), PT_THREADS, 0, c->s_main>>>(
        d.service_id, d.jvm_id, d.end_ts_ms, d.elapsed_ms, begin, end, c->cfg.num_services, c->cfg.num_jvms, c->pt_seq, c->pt_base_ms,
        c->pt_pend[c->pt_cur], c->pt_cursors + 0, c->pt_cap, &c->d_counters->log_overflow);
    c->launches += 1;
    CUDA_TRY(c, cudaGetLastError());
    return APM_OK;
}

// At a tick: everything with endTs <= edge_ts leaves the pending set in endTs order.
apm_status apm_pt_tick(apm_ctx* c, int64_t edge_ts, uint32_t tick_slot) {
    if (!c->pt_pend[0]) return APM_OK;
    ProfScope prof_scope_(c, APM_PG_PASSTHROUGH);
    unsigned long long cur[3];
    CUDA_TRY(c, cudaMemcpyAsync(cur, c->pt_cursors, sizeof cur, cudaMemcpyDeviceToHost, c->s_main));
    CUDA_TRY(c, cudaStreamSynchronize(c->s_main));
    const uint64_t n = cur[0];
    if (n > c->pt_cap) return apm_fail(c, APM_E_CAPACITY, "passthrough: more pending transactions than the reorder buffer holds");
    c->pt_tick_counts[tick_slot] = 0;
    if (n == 0) return APM_OK;
    int64_t d = edge_ts - c->pt_base_ms;
    if (d < 0) return APM_OK;  // nothing can be due before the base
    const uint64_t edge_word = ((uint64_t)d << PT_SEQ_BITS) | ((1ull << PT_SEQ_BITS) - 1);
    const int other = c->pt_cur ^ 1;
    CUDA_TRY(c, cudaMemsetAsync(c->pt_cursors + 1, 0, 2 * sizeof(unsigned long long), c->s_main));
    const uint64_t per_block = (uint64_t)PT_THREADS * PT_ITEMS;
    pt_partition_kernel<<<(unsigned)((n + per_block - 1) / per_block), PT_THREADS, 0, c->s_main>>>(
        c->pt_pend[c->pt_cur], n, edge_word, c->pt_emit, c->pt_cursors + 1, c->pt_pend[other], c->pt_cursors + 2);
    c->launches += 1;
    CUDA_TRY(c, cudaGetLastError());
    CUDA_TRY(c, cudaMemcpyAsync(cur, c->pt_cursors, sizeof cur, cudaMemcpyDeviceToHost, c->s_main));
    CUDA_TRY(c, cudaStreamSynchronize(c->s_main));
    const uint64_t n_emit = cur[1], n_keep = cur[2];
    // the kept words are the new pending set
    CUDA_TRY(c, cudaMemcpyAsync(c->pt_cursors + 0, &n_keep, sizeof(unsigned long long), cudaMemcpyHostToDevice, c->s_main));
    c->pt_cur = other;
    if (n_emit == 0) return APM_OK;
    if (c->pt_out_used + n_emit > c->pt_out_cap) return apm_fail(c, APM_E_WOULDBLOCK, "passthrough output full: fetch and release first");
    uint64_t* sorted = c->pt_emit;
    if (n_emit > 1) {
        const int bits = PT_SEQ_BITS + apm_bits_of((uint64_t)d);  // seq bits + the significant bits of the time field
        // the words are plain 64-bit keys: "value field" = low 32 bits, "key field" = the rest
        const int vbits = bits < 32 ? bits : 32, kbits = bits > 32 ? bits - 32 : 0;
        sorted = apm_radix_sort_words(c->s_main, c->pt_emit, c->pt_tmp, n_emit, c->sort_hist, c->sort_base, vbits, kbits, &c->launches);
        CUDA_TRY(c, cudaGetLastError());
    }
    CUDA_TRY(c, cudaMemcpyAsync(c->pt_out + c->pt_out_used, sorted, n_emit * sizeof(uint64_t), cudaMemcpyDeviceToDevice, c->s_main));
    c->pt_out_used += n_emit;
    c->pt_tick_counts[tick_slot] = (uint32_t)n_emit;
    return APM_OK;
}

extern "C" apm_status apm_fetch_passthrough(apm_ctx* c, int64_t* end_ts, uint32_t* seq, uint64_t cap, uint64_t* n, uint32_t* tick_counts,
                                            uint32_t tick_cap, uint32_t* nticks) {
    if (!c || !n) return APM_E_INVALID;
    APM_ENTER(c);
    *n = 0;
    if (nticks) *nticks = 0;
    if (!c->pt_pend[0]) return apm_fail(c, APM_E_STATE, "passthrough is off (APM_F_PASSTHROUGH)");
    if (tick_counts) {
        if (tick_cap < c->ticks_buffered) return apm_fail(c, APM_E_CAPACITY, "fetch_passthrough: tick_counts too small");
        for (uint32_t t = 0; t < c->ticks_buffered; ++t) tick_counts[t] = c->pt_tick_counts[t];
        if (nticks) *nticks = c->ticks_buffered;
    }
    if (c->pt_out_used == 0) return APM_OK;
    if (cap < c->pt_out_used) return apm_fail(c, APM_E_CAPACITY, "fetch_passthrough: output buffers too small");
    std::vector<uint64_t> words(c->pt_out_used);
    CUDA_TRY(c, cudaMemcpyAsync(words.data(), c->pt_out, words.size() * sizeof(uint64_t), cudaMemcpyDeviceToHost, c->s_main));
    CUDA_TRY(c, cudaStreamSynchronize(c->s_main));
    for (size_t i = 0; i < words.size(); ++i) {
        if (end_ts) end_ts[i] = (int64_t)(words[i] >> PT_SEQ_BITS) + c->pt_base_ms;
        if (seq) seq[i] = (uint32_t)(words[i] & ((1ull << PT_SEQ_BITS) - 1));
    }
    *n = words.size();
    c->counters.d2h_bytes += words.size() * sizeof(uint64_t);
    return APM_OK;
}

// Loads this file's kernels on the current device (see apm_preload_kernels in capi.cu).
cudaError_t apm_preload_k5_passthrough() {
    const void* fns[] = {(const void*)pt_append_kernel, (const void*)pt_partition_kernel};
    cudaError_t first = cudaSuccess;
    for (const void* f : fns) {
        cudaFuncAttributes a;
        const cudaError_t e = apm_preload_one(&a, f);
        if (e != cudaSuccess && first == cudaSuccess) first = e;
    }
    return first;
}
