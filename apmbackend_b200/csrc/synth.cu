// libapm_synth.so: synthetic stream generator (device + host twins) and the few raw device
// memory helpers the ctypes tests and bench.py need.  Bench/test support, not product API.
#include <cuda_runtime.h>
#include <stdint.h>
#include <string.h>

#include "../../include/apm_b200.h"
#include "synth_common.h"

namespace {
__global__ void __launch_bounds__(256) synth_kernel(const synth_params p, uint64_t count, uint32_t* svc, uint32_t* jvm,
                                                    int64_t* ts, int32_t* el) {
    const uint64_t stride = (uint64_t)gridDim.x * blockDim.x;
    for (uint64_t j = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x; j < count; j += stride)
        syn_tx(p, p.first + j, svc + j, jvm + j, ts + j, el + j);
}

// ---- one rank's shard of a global stream (bench support for the key-sharded multi-GPU job) -------------------------
// Every rank walks the same global candidate sequence and keeps the transactions whose key it owns, in arrival
// order, with the key replaced by its dense id on the owner.  Two passes over tiles of SH_TILE candidates:
// A counts the owned transactions and the largest bucket label per tile; the host turns those into a prefix sum and a
// running maximum; B regenerates the tile, writes the owned transactions at their final positions and emits a tick
// mark (position in the SHARD, label) wherever the GLOBAL running maximum of the label grows.
constexpr int SHT_THREADS = 256, SHT_ITEMS = 16, SHT_TILE = SHT_THREADS * SHT_ITEMS;

struct shard_params {
    synth_params p;
    const uint32_t* owner;     // [num_services * num_jvms] global key -> rank
    const uint32_t* local_id;  // global key -> dense id on its owner
    uint32_t rank;
    uint32_t pad_;
    int64_t bucket_ms;
};

__global__ void __launch_bounds__(SHT_THREADS) shard_count_kernel(const shard_params sp, uint64_t count, uint32_t* __restrict__ tile_count,
                                                                  long long* __restrict__ tile_max) {
    __shared__ uint32_t s_cnt[SHT_THREADS / 32];
    __shared__ long long s_max[SHT_THREADS / 32];
    const uint64_t base = (uint64_t)blockIdx.x * SHT_TILE + (uint64_t)threadIdx.x * SHT_ITEMS;
    uint32_t cnt = 0;
    long long mx = -1;
    for (int i = 0; i < SHT_ITEMS; ++i) {
        const uint64_t j = base + i;
        if (j >= count) break;
        uint32_t svc, jvm; int64_t ts; int32_t el;
        syn_tx(sp.p, sp.p.first + j, &svc, &jvm, &ts, &el);
        const long long lab = ts / sp.bucket_ms;
        mx = lab > mx ? lab : mx;
        cnt += sp.owner[(uint64_t)jvm * sp.p.num_services + svc] == sp.rank ? 1u : 0u;
    }
    cnt = __reduce_add_sync(0xffffffffu, cnt);
    for (int o = 16; o > 0; o >>= 1) { const long long y = __shfl_xor_sync(0xffffffffu, mx, o); mx = y > mx ? y : mx; }
    if ((threadIdx.x & 31) == 0) { s_cnt[threadIdx.x >> 5] = cnt; s_max[threadIdx.x >> 5] = mx; }
    __syncthreads();
    if (threadIdx.x == 0) {
        uint32_t c = 0; long long m = -1;
        for (int w = 0; w < SHT_THREADS / 32; ++w) { c += s_cnt[w]; m = s_max[w] > m ? s_max[w] : m; }
        tile_count[blockIdx.x] = c;
        tile_max[blockIdx.x] = m;
    }
}

struct shard_mark { unsigned long long pos; long long label; };

__global__ void __launch_bounds__(SHT_THREADS) shard_emit_kernel(const shard_params sp, uint64_t count, const unsigned long long* __restrict__ tile_prefix,
                                                                 const long long* __restrict__ tile_carry, const long long* __restrict__ tile_max,
                                                                 uint32_t* svc_out, uint32_t* jvm_out, int64_t* ts_out, int32_t* el_out,
                                                                 shard_mark* __restrict__ marks, unsigned int* __restrict__ nmarks, unsigned int marks_cap) {
    __shared__ uint32_t s_cnt[SHT_THREADS / 32];
    __shared__ long long s_max[SHT_THREADS / 32];
    const int lane = threadIdx.x & 31, w = threadIdx.x >> 5;
    const uint64_t base = (uint64_t)blockIdx.x * SHT_TILE + (uint64_t)threadIdx.x * SHT_ITEMS;
    uint32_t svc[SHT_ITEMS], jvm[SHT_ITEMS]; int64_t ts[SHT_ITEMS]; int32_t el[SHT_ITEMS];
    uint32_t mine = 0;  // bit i: candidate i is owned
    uint32_t cnt = 0;
    long long mx = -1;
#pragma unroll
    for (int i = 0; i < SHT_ITEMS; ++i) {
        const uint64_t j = base + i;
        if (j < count) {
            syn_tx(sp.p, sp.p.first + j, &svc[i], &jvm[i], &ts[i], &el[i]);
            const long long lab = ts[i] / sp.bucket_ms;
            mx = lab > mx ? lab : mx;
            if (sp.owner[(uint64_t)jvm[i] * sp.p.num_services + svc[i]] == sp.rank) { mine |= 1u << i; ++cnt; }
        }
    }
    // exclusive scans over the CTA: owned transactions before this thread, largest label before this thread
    uint32_t incl = cnt;
    long long imax = mx;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
        const uint32_t y = __shfl_up_sync(0xffffffffu, incl, o);
        const long long ym = __shfl_up_sync(0xffffffffu, imax, o);
        if (lane >= o) { incl += y; imax = ym > imax ? ym : imax; }
    }
    if (lane == 31) { s_cnt[w] = incl; s_max[w] = imax; }
    __syncthreads();
    uint32_t before = incl - cnt;
    long long run = __shfl_up_sync(0xffffffffu, imax, 1);
    if (lane == 0) run = -1;
    for (int k = 0; k < w; ++k) { before += s_cnt[k]; run = s_max[k] > run ? s_max[k] : run; }
    const long long carry = tile_carry[blockIdx.x];
    run = carry > run ? carry : run;
    unsigned long long pos = tile_prefix[blockIdx.x] + before;
    const bool has_tick = tile_max[blockIdx.x] > carry;
#pragma unroll
    for (int i = 0; i < SHT_ITEMS; ++i) {
        const uint64_t j = base + i;
        if (j >= count) break;
        if (has_tick) {
            const long long lab = ts[i] / sp.bucket_ms;
            if (lab > run) {
                run = lab;
                const unsigned int k = atomicAdd(nmarks, 1u);
                if (k < marks_cap) { marks[k].pos = pos; marks[k].label = lab; }
            }
        }
        if (mine & (1u << i)) {
            svc_out[pos] = sp.local_id[(uint64_t)jvm[i] * sp.p.num_services + svc[i]];
            jvm_out[pos] = 0;
            ts_out[pos] = ts[i];
            el_out[pos] = el[i];
            ++pos;
        }
    }
}

// ring pre-fill: a plausible history for every (series, key) column - the level follows the
// service's duration scale (mean ~ 1.13 x base, p95 ~ 2.28 x, p99 ~ 3.2 x for lognormal(0, 0.5)),
// plus a small deterministic noise, all in tenths like the values toFixed(1) produces
__host__ __device__ __forceinline__ double ring_fill_value(uint64_t i, uint64_t ncol, uint64_t seed, const uint32_t* base,
                                                           uint32_t num_services, uint32_t K) {
    const uint64_t col = i % ncol;
    const uint32_t series = (uint32_t)(col / K), key = (uint32_t)(col % K);
    const uint32_t mult = series == 0 ? 113u : (series == 1 ? 228u : 320u);
    const uint64_t hn = syn_mix(seed ^ (i * 0xD1B54A32D192ED03ull) ^ 0x5555);
    const uint32_t tenths = base[key % num_services] * mult / 10u + (uint32_t)(hn % 41u);
    return (double)tenths / 10.0;
}

__global__ void __launch_bounds__(256) ring_fill_kernel(double* out, uint64_t rows, uint64_t ncol, uint64_t seed,
                                                        const uint32_t* base, uint32_t num_services, uint32_t K) {
    const uint64_t total = rows * ncol, stride = (uint64_t)gridDim.x * blockDim.x;
    for (uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += stride)
        out[i] = ring_fill_value(i, ncol, seed, base, num_services, K);
}
}  // namespace

extern "C" {

int synth_device_count(void) {
    int n = 0;
    if (cudaGetDeviceCount(&n) != cudaSuccess) return 0;
    return n;
}
int synth_set_device(int d) { return (int)cudaSetDevice(d); }
int synth_malloc(void** p, uint64_t bytes) { return (int)cudaMalloc(p, bytes); }
int synth_free(void* p) { return (int)cudaFree(p); }
int synth_malloc_host(void** p, uint64_t bytes) { return (int)cudaMallocHost(p, bytes); }
int synth_free_host(void* p) { return (int)cudaFreeHost(p); }
int synth_h2d(void* d, const void* h, uint64_t bytes) { return (int)cudaMemcpy(d, h, bytes, cudaMemcpyHostToDevice); }
int synth_d2h(void* h, const void* d, uint64_t bytes) { return (int)cudaMemcpy(h, d, bytes, cudaMemcpyDeviceToHost); }
int synth_sync(void) { return (int)cudaDeviceSynchronize(); }
int synth_mem_info(uint64_t* free_bytes, uint64_t* total_bytes) {
    size_t f = 0, t = 0;
    const cudaError_t e = cudaMemGetInfo(&f, &t);
    *free_bytes = f; *total_bytes = t;
    return (int)e;
}
const char* synth_error_string(int e) { return cudaGetErrorString((cudaError_t)e); }

// tables in p must be device pointers
int synth_generate_device(const synth_params* p, uint64_t count, apm_tx_columns* cols) {
    synth_kernel<<<148 * 8, 256>>>(*p, count, cols->service_id, cols->jvm_id, cols->end_ts_ms, cols->elapsed_ms);
    cudaError_t e = cudaGetLastError();
    if (e != cudaSuccess) return (int)e;
    return (int)cudaDeviceSynchronize();
}

// tables in p must be host pointers
void synth_generate_host(const synth_params* p, uint64_t count, apm_tx_columns* cols) {
    for (uint64_t j = 0; j < count; ++j)
        syn_tx(*p, p->first + j, cols->service_id + j, cols->jvm_id + j, cols->end_ts_ms + j, cols->elapsed_ms + j);
}

// base: device pointer to the stream's duration-scale table
int synth_ring_fill_device(double* out, uint64_t rows, uint64_t ncol, uint64_t seed, const uint32_t* base, uint32_t num_services,
                           uint32_t K) {
    ring_fill_kernel<<<148 * 8, 256>>>(out, rows, ncol, seed, base, num_services, K);
    cudaError_t e = cudaGetLastError();
    if (e != cudaSuccess) return (int)e;
    return (int)cudaDeviceSynchronize();
}

// the same values for a selection of columns, packed [rows][ncols_sel]; base: host pointer
void synth_ring_fill_host(double* out, uint64_t rows, uint64_t ncol, uint64_t seed, const uint32_t* base, uint32_t num_services,
                          uint32_t K, const uint64_t* cols, uint64_t ncols_sel) {
    for (uint64_t r = 0; r < rows; ++r)
        for (uint64_t c = 0; c < ncols_sel; ++c)
            out[r * ncols_sel + c] = ring_fill_value(r * ncol + cols[c], ncol, seed, base, num_services, K);
}

// pass A of the shard generator: per tile of 4096 candidates the owned transactions and the largest label (device arrays)
int synth_shard_tiles(uint64_t count) { return (int)((count + SHT_TILE - 1) / SHT_TILE); }
int synth_shard_count_device(const synth_params* p, uint64_t count, const uint32_t* owner, uint32_t rank, int64_t bucket_ms,
                             uint32_t* tile_count, long long* tile_max) {
    shard_params sp; sp.p = *p; sp.owner = owner; sp.local_id = nullptr; sp.rank = rank; sp.pad_ = 0; sp.bucket_ms = bucket_ms;
    shard_count_kernel<<<(unsigned)synth_shard_tiles(count), SHT_THREADS>>>(sp, count, tile_count, tile_max);
    cudaError_t e = cudaGetLastError();
    if (e != cudaSuccess) return (int)e;
    return (int)cudaDeviceSynchronize();
}
// pass B: tile_prefix = owned transactions before each tile (plus what earlier calls wrote), tile_carry = running
// maximum of the label before each tile (starting from latestBucket).  Marks come back unordered.
int synth_shard_emit_device(const synth_params* p, uint64_t count, const uint32_t* owner, const uint32_t* local_id, uint32_t rank,
                            int64_t bucket_ms, const unsigned long long* tile_prefix, const long long* tile_carry,
                            const long long* tile_max, apm_tx_columns* cols, void* marks, unsigned int* nmarks, unsigned int marks_cap) {
    shard_params sp; sp.p = *p; sp.owner = owner; sp.local_id = local_id; sp.rank = rank; sp.pad_ = 0; sp.bucket_ms = bucket_ms;
    shard_emit_kernel<<<(unsigned)synth_shard_tiles(count), SHT_THREADS>>>(sp, count, tile_prefix, tile_carry, tile_max, cols->service_id,
                                                                        cols->jvm_id, cols->end_ts_ms, cols->elapsed_ms,
                                                                        (shard_mark*)marks, nmarks, marks_cap);
    cudaError_t e = cudaGetLastError();
    if (e != cudaSuccess) return (int)e;
    return (int)cudaDeviceSynchronize();
}

// L2 flush helper for the bench: overwrite a buffer larger than L2
int synth_memset(void* p, int v, uint64_t bytes) { return (int)cudaMemset(p, v, bytes); }

}  // extern "C"
