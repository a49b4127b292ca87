// K0/K1: tick finder and keyed segmented reduction of a transaction segment.
//
// Replaces, for a whole batch at once, what consumeMsg does per message
// (stream_calc_stats.js:331-371): generateBucketLabel (:89-91), the
// `bucketLabel > latestBucket` test (:348) and addData (:115-134).
#include "ctx.h"

namespace {

constexpr int TF_THREADS = 256;
constexpr int TF_ITEMS = 16;
constexpr int TF_TILE = TF_THREADS * TF_ITEMS;

__device__ __forceinline__ long long warp_max_ll(long long v) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
        long long other = __shfl_xor_sync(0xffffffffu, v, o);
        v = other > v ? other : v;
    }
    return v;
}

// ---- K0a: per-tile maximum bucket label (8 B/tx read) -------------------------------------
__global__ void __launch_bounds__(TF_THREADS) tf_tile_max_kernel(const int64_t* __restrict__ ts, uint64_t n,
                                                                 int64_t bucket_ms, double inv_bucket,
                                                                 long long* __restrict__ tile_max) {
    __shared__ long long s_max[TF_THREADS / 32];
    const uint64_t base = (uint64_t)blockIdx.x * TF_TILE;
    long long m = -1;
#pragma unroll
    for (int i = 0; i < TF_ITEMS; ++i) {
        uint64_t idx = base + (uint64_t)i * TF_THREADS + threadIdx.x;
        if (idx < n) {
            int64_t t = __ldg(ts + idx);
            long long lab = t >= 0 ? apm_label(t, bucket_ms, inv_bucket) : -1;
            m = lab > m ? lab : m;
        }
    }
    m = warp_max_ll(m);
    if ((threadIdx.x & 31) == 0) s_max[threadIdx.x >> 5] = m;
    __syncthreads();
    if (threadIdx.x < 32) {
        long long v = threadIdx.x < TF_THREADS / 32 ? s_max[threadIdx.x] : -1;
        v = warp_max_ll(v);
        if (threadIdx.x == 0) tile_max[blockIdx.x] = v;
    }
}

// ---- K0b: exclusive running maximum over the tiles, carry-in = latestBucket ----------------
__global__ void __launch_bounds__(1024) tf_scan_kernel(const long long* __restrict__ tile_max, uint64_t ntiles,
                                                       long long latest, long long* __restrict__ tile_prefix) {
    __shared__ long long s_warp[32];
    const uint64_t per = (ntiles + 1023) / 1024;
    const uint64_t lo = (uint64_t)threadIdx.x * per;
    const uint64_t hi = lo + per < ntiles ? lo + per : ntiles;
    long long m = -1;
    for (uint64_t i = lo; i < hi; ++i) m = tile_max[i] > m ? tile_max[i] : m;
    // inclusive scan (max) across the block
    long long v = m;
    const int lane = threadIdx.x & 31, w = threadIdx.x >> 5;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
        long long other = __shfl_up_sync(0xffffffffu, v, o);
        if (lane >= o && other > v) v = other;
    }
    if (lane == 31) s_warp[w] = v;
    __syncthreads();
    if (w == 0) {
        long long x = s_warp[lane];
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) {
            long long other = __shfl_up_sync(0xffffffffu, x, o);
            if (lane >= o && other > x) x = other;
        }
        s_warp[lane] = x;
    }
    __syncthreads();
    long long excl = __shfl_up_sync(0xffffffffu, v, 1);
    if (lane == 0) excl = -1;
    if (w > 0 && s_warp[w - 1] > excl) excl = s_warp[w - 1];
    long long run = excl > latest ? excl : latest;
    for (uint64_t i = lo; i < hi; ++i) {
        tile_prefix[i] = run;
        run = tile_max[i] > run ? tile_max[i] : run;
    }
}

// ---- K0c: emit the positions where the label exceeds everything before it ------------------
__global__ void __launch_bounds__(TF_THREADS) tf_emit_kernel(const int64_t* __restrict__ ts, uint64_t n,
                                                             int64_t bucket_ms, double inv_bucket,
                                                             const long long* __restrict__ tile_max,
                                                             const long long* __restrict__ tile_prefix,
                                                             TickMark* __restrict__ ticks, unsigned int* __restrict__ nticks,
                                                             unsigned int cap) {
    const long long carry = tile_prefix[blockIdx.x];
    if (tile_max[blockIdx.x] <= carry) return;  // no tick inside this tile
    __shared__ long long s_warp[TF_THREADS / 32];
    const uint64_t base = (uint64_t)blockIdx.x * TF_TILE + (uint64_t)threadIdx.x * TF_ITEMS;
    long long lab[TF_ITEMS];
    long long m = -1;
#pragma unroll
    for (int i = 0; i < TF_ITEMS; ++i) {
        uint64_t idx = base + i;
        long long l = -1;
        if (idx < n) {
            int64_t t = ts[idx];
            if (t >= 0) l = apm_label(t, bucket_ms, inv_bucket);
        }
        lab[i] = l;
        m = l > m ? l : m;
    }
    long long v = m;
    const int lane = threadIdx.x & 31, w = threadIdx.x >> 5;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
        long long other = __shfl_up_sync(0xffffffffu, v, o);
        if (lane >= o && other > v) v = other;
    }
    if (lane == 31) s_warp[w] = v;
    __syncthreads();
    long long excl = __shfl_up_sync(0xffffffffu, v, 1);
    if (lane == 0) excl = -1;
    for (int j = 0; j < w; ++j) excl = s_warp[j] > excl ? s_warp[j] : excl;
    long long run = excl > carry ? excl : carry;
#pragma unroll
    for (int i = 0; i < TF_ITEMS; ++i) {
        if (lab[i] > run) {
            run = lab[i];
            unsigned int k = atomicAdd(nticks, 1u);
            if (k < cap) {
                ticks[k].pos = base + i;
                ticks[k].label = lab[i];
            }
        }
    }
}

// ---- K1: keyed segmented reduction ------------------------------------------------------------
// v1: one transaction per thread, accumulators updated with L2 reductions.
struct ReduceParams {
    const uint32_t* svc;
    const uint32_t* jvm;
    const int64_t* ts;
    const int32_t* el;
    uint64_t begin, end;
    int64_t bucket_ms;
    double inv_bucket;
    int64_t latest;
    int keep, S;
    uint32_t num_services, num_jvms, K;
    Acc* acc;
    uint8_t* seen;
    SlotMeta* slot_meta;
    uint64_t* const* logs;
    uint64_t log_cap;
    DevCounters* counters;
    int with_log;
};

__global__ void __launch_bounds__(256) reduce_kernel_v1(const ReduceParams p) {
    const uint64_t stride = (uint64_t)gridDim.x * blockDim.x;
    unsigned long long n_ing = 0, n_drop = 0, n_inv = 0, n_fut = 0, n_ovf = 0;
    for (uint64_t i = p.begin + (uint64_t)blockIdx.x * blockDim.x + threadIdx.x; i < p.end; i += stride) {
        const uint32_t svc = __ldg(p.svc + i), jvm = __ldg(p.jvm + i);
        const int64_t t = __ldg(p.ts + i);
        const int32_t el = __ldg(p.el + i);
        if (svc >= p.num_services || jvm >= p.num_jvms || el < 0 || t < 0) { n_inv++; continue; }
        const int64_t lab = apm_label(t, p.bucket_ms, p.inv_bucket);
        if (lab > p.latest) { n_fut++; continue; }
        if (lab < p.latest - p.keep) { n_drop++; continue; }
        const int slot = (int)(lab % p.S);
        const uint32_t key = jvm * p.num_services + svc;
        p.seen[key] = 1;
        Acc* a = p.acc + (size_t)slot * p.K + key;
        const uint32_t v = (uint32_t)el;
        atomicAdd(&a->count, 1u);
        atomicAdd((unsigned long long*)&a->sum, (unsigned long long)v);
        const unsigned long long sq = (unsigned long long)v * v;
        const unsigned long long old = atomicAdd((unsigned long long*)&a->sumsq_lo, sq);
        if (old + sq < old) atomicAdd(&a->sumsq_hi, 1u);
        atomicMin(&a->vmin, v);
        atomicMax(&a->vmax, v);
        n_ing++;
        if (p.with_log) {
            SlotMeta* sm = p.slot_meta + slot;
            const unsigned long long pos = atomicAdd(&sm->count, 1ull);
            if (pos < p.log_cap) p.logs[slot][pos] = ((uint64_t)key << 32) | v;
            else n_ovf++;
            atomicMax(&sm->max_value, v);
        }
    }
    // block-level counter aggregation
    __shared__ unsigned long long s_c[5];
    if (threadIdx.x < 5) s_c[threadIdx.x] = 0;
    __syncthreads();
    if (n_ing) atomicAdd(&s_c[0], n_ing);
    if (n_drop) atomicAdd(&s_c[1], n_drop);
    if (n_inv) atomicAdd(&s_c[2], n_inv);
    if (n_fut) atomicAdd(&s_c[3], n_fut);
    if (n_ovf) atomicAdd(&s_c[4], n_ovf);
    __syncthreads();
    if (threadIdx.x == 0) {
        if (s_c[0]) atomicAdd(&p.counters->ingested, s_c[0]);
        if (s_c[1]) atomicAdd(&p.counters->dropped, s_c[1]);
        if (s_c[2]) atomicAdd(&p.counters->invalid, s_c[2]);
        if (s_c[3]) atomicAdd(&p.counters->future, s_c[3]);
        if (s_c[4]) atomicAdd(&p.counters->log_overflow, s_c[4]);
    }
}

}  // namespace

apm_status apm_launch_find_ticks(apm_ctx* c, const int64_t* d_ts, uint64_t n, unsigned int* n_ticks_out) {
    ProfScope prof_scope_(c, APM_PG_TICKFIND);
    *n_ticks_out = 0;
    if (n == 0) return APM_OK;
    const uint64_t ntiles = (n + TF_TILE - 1) / TF_TILE;
    if (ntiles > c->tf_tiles_cap) return apm_fail(c, APM_E_CAPACITY, "chunk larger than max_batch");
    CUDA_TRY(c, cudaMemsetAsync(c->d_nticks, 0, sizeof(unsigned int), c->s_main));
    tf_tile_max_kernel<<<(unsigned)ntiles, TF_THREADS, 0, c->s_main>>>(d_ts, n, c->cfg.bucket_ms, c->inv_bucket,
                                                                      c->tf_tile_max);
    tf_scan_kernel<<<1, 1024, 0, c->s_main>>>(c->tf_tile_max, ntiles, c->latest, c->tf_tile_prefix);
    tf_emit_kernel<<<(unsigned)ntiles, TF_THREADS, 0, c->s_main>>>(d_ts, n, c->cfg.bucket_ms, c->inv_bucket,
                                                                  c->tf_tile_max, c->tf_tile_prefix, c->d_ticks,
                                                                  c->d_nticks, c->max_ticks_per_chunk);
    c->launches += 3;
    CUDA_TRY(c, cudaGetLastError());
    CUDA_TRY(c, cudaMemcpyAsync(c->h_nticks, c->d_nticks, sizeof(unsigned int), cudaMemcpyDeviceToHost, c->s_main));
    CUDA_TRY(c, cudaStreamSynchronize(c->s_main));
    unsigned int nt = *c->h_nticks;
    if (nt > c->max_ticks_per_chunk) return apm_fail(c, APM_E_CAPACITY, "too many ticks in one chunk");
    if (nt) {
        CUDA_TRY(c, cudaMemcpyAsync(c->h_ticks, c->d_ticks, nt * sizeof(TickMark), cudaMemcpyDeviceToHost, c->s_main));
        CUDA_TRY(c, cudaStreamSynchronize(c->s_main));
        c->counters.d2h_bytes += nt * sizeof(TickMark);
    }
    c->counters.d2h_bytes += sizeof(unsigned int);
    *n_ticks_out = nt;
    return APM_OK;
}

apm_status apm_launch_reduce(apm_ctx* c, const apm_tx_columns& d, uint64_t begin, uint64_t end) {
    ProfScope prof_scope_(c, APM_PG_REDUCE);
    if (end <= begin) return APM_OK;
    ReduceParams p;
    p.svc = d.service_id; p.jvm = d.jvm_id; p.ts = d.end_ts_ms; p.el = d.elapsed_ms;
    p.begin = begin; p.end = end;
    p.bucket_ms = c->cfg.bucket_ms; p.inv_bucket = c->inv_bucket; p.latest = c->latest;
    p.keep = c->keep; p.S = c->S;
    p.num_services = c->cfg.num_services; p.num_jvms = c->cfg.num_jvms; p.K = c->K;
    p.acc = c->acc; p.seen = c->seen; p.slot_meta = c->slot_meta; p.logs = c->d_log_ptrs; p.log_cap = c->log_cap;
    p.counters = c->d_counters;
    p.with_log = (c->cfg.flags & APM_F_PERCENTILES) && c->cfg.num_percentiles > 0;
    const uint64_t n = end - begin;
    uint64_t blocks = (n + 255) / 256;
    const uint64_t max_blocks = (uint64_t)APM_NUM_SMS * 8;
    if (blocks > max_blocks) blocks = max_blocks;
    reduce_kernel_v1<<<(unsigned)blocks, 256, 0, c->s_main>>>(p);
    c->launches += 1;
    CUDA_TRY(c, cudaGetLastError());
    return APM_OK;
}
