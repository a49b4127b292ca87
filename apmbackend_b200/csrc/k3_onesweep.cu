// K3a': single-pass-per-digit LSD radix sort ("onesweep": chained scan with decoupled look-back).
//
// Same job as k3_sort.cu - order one bucket's value log by (key, elapsed) so every key's
// durations form one ascending run (what binaryConcat builds, util_methods.js:84-106) - with
// less memory traffic and fewer launches: one histogram kernel reads the log once for all
// digits, then each digit is ONE kernel that reads the words once and writes them once.  The
// per-tile digit offsets that the three-kernel version gets from a global scan come from a
// look-back over the predecessor tiles' published counts instead.
//
// Only the significant bits are sorted: the logical word is the low `vbits` of the value
// field with the key field packed directly above it, cut into equal digits of 8 bits (up to 11 on request).
#include <cstdlib>
#include <type_traits>

#include "ctx.h"

namespace {

constexpr int OS_THREADS = 256;
constexpr int OS_WARPS = OS_THREADS / 32;
constexpr int OS_ITEMS = 16;
constexpr int OS_TILE = OS_THREADS * OS_ITEMS;  // 4096 words
constexpr int OS_MAX_BITS = 11;
constexpr int OS_MAX_PASSES = 6;
constexpr uint32_t OS_FLAG_AGG = 1u << 30, OS_FLAG_PREFIX = 1u << 31, OS_COUNT_MASK = (1u << 30) - 1u;

struct OsPlan {
    int npass;
    int vbits;                  // significant bits of the value field
    int shift[OS_MAX_PASSES];   // logical bit position of each digit
    int bits[OS_MAX_PASSES];
};

__device__ __forceinline__ uint64_t os_logical(uint64_t x, int vbits) {
    return vbits >= 32 ? x : ((x & ((1ull << vbits) - 1ull)) | ((x >> 32) << vbits));
}
__device__ __forceinline__ uint32_t os_digit(uint64_t x, int vbits, int shift, int bits) {
    return (uint32_t)(os_logical(x, vbits) >> shift) & ((1u << bits) - 1u);
}

// lanes of the warp that hold the same digit
__device__ __forceinline__ uint32_t os_match(uint32_t d, int bits) {
    uint32_t peers = 0xffffffffu;
    for (int b = 0; b < bits; ++b) {
        const uint32_t bal = __ballot_sync(0xffffffffu, (d >> b) & 1u);
        peers &= ((d >> b) & 1u) ? bal : ~bal;
    }
    return peers;
}

// ---- all digits' global histograms in one read ------------------------------------------------
__global__ void __launch_bounds__(OS_THREADS) os_hist_kernel(const uint64_t* __restrict__ src, uint64_t n, const OsPlan plan,
                                                             uint32_t* __restrict__ ghist /* [npass][2048] */) {
    extern __shared__ uint32_t s_hist[];  // [npass][2048]
    const int nb = plan.npass << OS_MAX_BITS;
    for (int i = threadIdx.x; i < nb; i += OS_THREADS) s_hist[i] = 0;
    __syncthreads();
    // four independent loads in flight per thread: the loop is bound by the load latency, not by the atomics
    const uint64_t stride = (uint64_t)gridDim.x * OS_THREADS;
    uint64_t i = (uint64_t)blockIdx.x * OS_THREADS + threadIdx.x;
    for (; i + 3 * stride < n; i += 4 * stride) {
        uint64_t x[4];
#pragma unroll
        for (int u = 0; u < 4; ++u) x[u] = __ldg(src + i + (uint64_t)u * stride);
#pragma unroll
        for (int u = 0; u < 4; ++u)
            for (int p = 0; p < plan.npass; ++p)
                atomicAdd(&s_hist[(p << OS_MAX_BITS) + os_digit(x[u], plan.vbits, plan.shift[p], plan.bits[p])], 1u);
    }
    for (; i < n; i += stride) {
        const uint64_t x = __ldg(src + i);
        for (int p = 0; p < plan.npass; ++p)
            atomicAdd(&s_hist[(p << OS_MAX_BITS) + os_digit(x, plan.vbits, plan.shift[p], plan.bits[p])], 1u);
    }
    __syncthreads();
    for (int i = threadIdx.x; i < nb; i += OS_THREADS)
        if (s_hist[i]) atomicAdd(&ghist[i], s_hist[i]);
}

// exclusive scan of each digit's bins -> first output position of every bin
__global__ void __launch_bounds__(1024) os_scan_kernel(uint32_t* __restrict__ ghist) {
    __shared__ uint32_t s_warp[32];
    uint32_t* h = ghist + ((size_t)blockIdx.x << OS_MAX_BITS);
    const int lane = threadIdx.x & 31, w = threadIdx.x >> 5;
    const uint32_t a = h[2 * threadIdx.x], b = h[2 * threadIdx.x + 1];
    uint32_t v = a + b;
    for (int o = 1; o < 32; o <<= 1) {
        const uint32_t y = __shfl_up_sync(0xffffffffu, v, o);
        if (lane >= o) v += y;
    }
    if (lane == 31) s_warp[w] = v;
    __syncthreads();
    if (w == 0) {
        uint32_t x = s_warp[lane];
        for (int o = 1; o < 32; o <<= 1) {
            const uint32_t y = __shfl_up_sync(0xffffffffu, x, o);
            if (lane >= o) x += y;
        }
        s_warp[lane] = x;
    }
    __syncthreads();
    const uint32_t excl = v - (a + b) + (w ? s_warp[w - 1] : 0u);
    h[2 * threadIdx.x] = excl;
    h[2 * threadIdx.x + 1] = excl + a;
}

// ---- one digit: rank, look back, scatter ----------------------------------------------------------
// NARROW (the logical word fits 32 bits - e.g. 10 k keys x durations below 131 s = 31 bits): the words travel
// between the passes as packed u32 `key << vbits | value`; the first pass packs the log's u64 words while it
// reads them and the last one unpacks while it writes, so the middle passes move half the bytes, hold half
// the registers and shared memory, and a digit is one shift and one mask.
template <bool NARROW>
struct OsSmemLayout {
    // [items x 4096][wcnt u16 x 8 x bins][tile_cnt u32 x bins][gbase u32 x bins][misc]
    static constexpr size_t item_bytes = NARROW ? 4 : 8;
    static size_t bytes(int bins) { return (size_t)OS_TILE * item_bytes + (size_t)OS_WARPS * bins * 2 + (size_t)bins * 4 * 2 + 64; }
};

// BITS: digit width known at compile time (the planner's default, 8), 0 = taken from the argument.  The ballots of
// the ranking loop and the per-bin loops unroll only then; with run-time bits the kernel executes twice the instructions.
template <int BITS>
__device__ __forceinline__ uint32_t os_match_t(uint32_t d, int bits_rt) {
    if (BITS == 0) return os_match(d, bits_rt);
    uint32_t peers = 0xffffffffu;
#pragma unroll
    for (int b = 0; b < (BITS ? BITS : 1); ++b) {
        const bool bit = (d >> b) & 1u;
        const uint32_t bal = __ballot_sync(0xffffffffu, bit);
        peers &= bit ? bal : ~bal;
    }
    return peers;
}

template <bool NARROW, int BITS>
__global__ void __launch_bounds__(OS_THREADS, 3)
os_pass_kernel(const void* __restrict__ src_v, void* __restrict__ dst_v, uint32_t n, int vbits, int shift, int bits_rt, int first, int last,
               const uint32_t* __restrict__ gbase, volatile uint32_t* status /* [tiles][bins] */, uint32_t* ticket) {
    using T = typename std::conditional<NARROW, uint32_t, uint64_t>::type;
    extern __shared__ __align__(16) unsigned char os_raw[];
    const int bits = BITS ? BITS : bits_rt;
    const int bins = 1 << bits;
    const uint32_t dmask = (uint32_t)bins - 1u;
    constexpr size_t IB = OsSmemLayout<NARROW>::item_bytes;
    T* s_items = reinterpret_cast<T*>(os_raw);
    uint16_t* s_wcnt = reinterpret_cast<uint16_t*>(os_raw + (size_t)OS_TILE * IB);          // [warp][bin]
    uint32_t* s_cnt = reinterpret_cast<uint32_t*>(os_raw + (size_t)OS_TILE * IB + (size_t)OS_WARPS * bins * 2);  // tile count, then local start
    uint32_t* s_gbase = s_cnt + bins;
    uint32_t* s_misc = s_gbase + bins;  // [0] tile id, [1..8] scan scratch

    const uint32_t tid = threadIdx.x, lane = tid & 31, w = tid >> 5;
    if (tid == 0) s_misc[0] = atomicAdd(ticket, 1u);  // tiles start in index order: a predecessor is always running or done
    for (uint32_t i = tid; i < (uint32_t)(OS_WARPS * bins / 2); i += OS_THREADS) reinterpret_cast<uint32_t*>(s_wcnt)[i] = 0;
    __syncthreads();
    const uint32_t tile = s_misc[0];
    const uint32_t tile_base = tile * OS_TILE;  // n < 2^30
    const uint32_t in_tile = tile_base >= n ? 0u : (n - tile_base < (uint32_t)OS_TILE ? n - tile_base : (uint32_t)OS_TILE);
    const uint32_t lbase = w * (32 * OS_ITEMS) + lane;  // this thread's first word inside the tile
    const bool in64 = !NARROW || first, out64 = !NARROW || last;

    // Words past the end are all ones: top bin of every digit, behind every real word of the (last) tile, so they are
    // ranked like real words - no validity test in the ranking loop - and simply not written out at the end.
    T item[OS_ITEMS];
#pragma unroll
    for (int i = 0; i < OS_ITEMS; ++i) {
        const uint32_t li = lbase + i * 32;
        if (in64) {
            const uint64_t x = li < in_tile ? __ldcs(reinterpret_cast<const uint64_t*>(src_v) + tile_base + li) : ~0ull;
            item[i] = NARROW ? (T)os_logical(x, vbits) : (T)x;
        } else {
            item[i] = li < in_tile ? (T)__ldcs(reinterpret_cast<const uint32_t*>(src_v) + tile_base + li) : (T)~0u;
        }
    }
    uint16_t* my_wcnt = s_wcnt + (size_t)w * bins;
    const uint32_t lanes_below = (1u << lane) - 1u;
    uint16_t rank[OS_ITEMS];  // position inside the warp's share of the bin
#pragma unroll
    for (int i = 0; i < OS_ITEMS; ++i) {
        const uint32_t d = NARROW ? (((uint32_t)item[i] >> shift) & dmask) : os_digit((uint64_t)item[i], vbits, shift, bits);
        // eight ballots: __match_any_sync gives the same mask in one instruction but was measured slower here (142 vs 122 us per sort)
        const uint32_t peers = os_match_t<BITS>(d, bits);
        const uint32_t before = my_wcnt[d];
        __syncwarp();
        if ((peers & lanes_below) == 0) my_wcnt[d] = (uint16_t)(before + __popc(peers));  // the lowest lane of the group
        __syncwarp();
        rank[i] = (uint16_t)(before + __popc(peers & lanes_below));
    }
    __syncthreads();

    // per bin: offsets of the warps inside the bin and the tile's count, published at once so that
    // successor tiles can already look back through this tile
    for (uint32_t d = tid; d < (uint32_t)bins; d += OS_THREADS) {
        uint32_t run = 0;
#pragma unroll
        for (int j = 0; j < OS_WARPS; ++j) {
            const uint32_t c = s_wcnt[(size_t)j * bins + d];
            s_wcnt[(size_t)j * bins + d] = (uint16_t)run;
            run += c;
        }
        s_cnt[d] = run;
        status[(size_t)tile * bins + d] = run | (tile == 0 ? OS_FLAG_PREFIX : OS_FLAG_AGG);
    }
    // decoupled look-back: add the predecessors' counts until one of them carries an inclusive prefix
    for (uint32_t d = tid; d < (uint32_t)bins; d += OS_THREADS) {
        uint32_t excl = 0;
#ifndef OS_MICRO_NO_LOOKBACK  // scripts/micro_sort.cu: the cost of the chain
        if (tile != 0) {
            for (int32_t t = (int32_t)tile - 1; t >= 0; --t) {
                volatile uint32_t* prev = status + (size_t)t * bins + d;
                uint32_t s;
                while ((s = *prev) == 0) { /* the predecessor holds a ticket, so it is resident and will publish */ }
                excl += s & OS_COUNT_MASK;
                if (s & OS_FLAG_PREFIX) break;
            }
            status[(size_t)tile * bins + d] = (excl + s_cnt[d]) | OS_FLAG_PREFIX;
        }
#endif
        s_gbase[d] = gbase[d] + excl;
    }
    __syncthreads();

    // exclusive scan of the tile counts over the bins -> first local position of each bin
    {
        constexpr int PER_CT = BITS ? ((1 << BITS) / OS_THREADS > 0 ? (1 << BITS) / OS_THREADS : 1) : 0;
        const int per = BITS ? PER_CT : (bins / OS_THREADS > 0 ? bins / OS_THREADS : 1);  // bins >= 256 is guaranteed by the planner
        const int d0 = tid * per;
        uint32_t local[8];
        uint32_t sum = 0;
#pragma unroll
        for (int j = 0; j < 8; ++j)
            if (j < per) { local[j] = s_cnt[d0 + j]; sum += local[j]; }
        uint32_t v = sum;
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) {
            const uint32_t y = __shfl_up_sync(0xffffffffu, v, o);
            if (lane >= (uint32_t)o) v += y;
        }
        if (lane == 31) s_misc[1 + w] = v;
        __syncthreads();
        uint32_t woff = 0;
#pragma unroll
        for (int j = 0; j < OS_WARPS; ++j)
            if ((uint32_t)j < w) woff += s_misc[1 + j];
        uint32_t run = woff + v - sum;
#pragma unroll
        for (int j = 0; j < 8; ++j)
            if (j < per) { s_cnt[d0 + j] = run; run += local[j]; }
    }
    __syncthreads();

#pragma unroll
    for (int i = 0; i < OS_ITEMS; ++i) {
        const uint32_t d = NARROW ? (((uint32_t)item[i] >> shift) & dmask) : os_digit((uint64_t)item[i], vbits, shift, bits);
        s_items[s_cnt[d] + my_wcnt[d] + rank[i]] = item[i];
    }
    __syncthreads();
    const uint32_t vmask = vbits >= 32 ? 0xffffffffu : ((1u << vbits) - 1u);
    for (uint32_t j = tid; j < in_tile; j += OS_THREADS) {  // the padding sits behind in_tile
        const T x = s_items[j];
        const uint32_t d = NARROW ? (((uint32_t)x >> shift) & dmask) : os_digit((uint64_t)x, vbits, shift, bits);
        const uint32_t pos = s_gbase[d] + (j - s_cnt[d]);
        if (!NARROW) reinterpret_cast<uint64_t*>(dst_v)[pos] = (uint64_t)x;
        else if (out64) reinterpret_cast<uint64_t*>(dst_v)[pos] = ((uint64_t)((uint32_t)x >> vbits) << 32) | ((uint32_t)x & vmask);
        else reinterpret_cast<uint32_t*>(dst_v)[pos] = (uint32_t)x;
    }
}

}  // namespace

size_t apm_onesweep_scratch_bytes(uint64_t max_n) {
    const uint64_t tiles = (max_n + OS_TILE - 1) / OS_TILE;
    // [tickets: 64 B][ghist: 6 x 2048 u32][status: 6 x tiles x 2048 u32]
    return 64 + (size_t)OS_MAX_PASSES * 2048 * 4 + (size_t)OS_MAX_PASSES * tiles * 2048 * 4;
}

// Sorts n words between a and b on `st`; returns the buffer holding the result, or nullptr when
// the scratch is too small for this n (the caller falls back to the three-kernel sort).
uint64_t* apm_onesweep_sort(cudaStream_t st, uint64_t* a, uint64_t* b, uint64_t n, unsigned char* scratch, size_t scratch_bytes,
                            int vbits, int kbits, uint64_t* launches, cudaError_t* err) {
    *err = cudaSuccess;
    int total = vbits + kbits;
    if (total < 1) total = 1;
    OsPlan plan;
    // Widest digit the planner uses (APM_SORT_BITS, 8..11).  A 4096-word tile has two words per 11-bit bin, so
    // the per-bin work of a pass (warp offsets, publish, look-back, scan, 8 KB of status per tile) is as large as
    // its per-word work: four 8-bit passes sort a 31-bit word faster than three 11-bit ones (measured 139 vs
    // 150 us for 1.67 M words; 37-bit words: 5 x 8 bits 97 us, 4 x 10 bits 135 us).
    static int digit_bits = 0;
    if (!digit_bits) {
        const char* e = getenv("APM_SORT_BITS");
        digit_bits = e ? atoi(e) : 8;
        if (digit_bits < 8 || digit_bits > OS_MAX_BITS) digit_bits = 8;
    }
    plan.npass = (total + digit_bits - 1) / digit_bits;
    if (plan.npass > OS_MAX_PASSES) return nullptr;
    int per = (total + plan.npass - 1) / plan.npass;
    if (per < 8) per = 8;  // at least 256 bins: one per thread in the per-bin phases
    plan.vbits = vbits;
    for (int p = 0; p < plan.npass; ++p) { plan.shift[p] = p * per; plan.bits[p] = per; }
    const int bins = 1 << per;
    const uint64_t tiles = (n + OS_TILE - 1) / OS_TILE;
    const size_t status_bytes = (size_t)plan.npass * tiles * bins * 4;
    const size_t head = 64 + (size_t)OS_MAX_PASSES * 2048 * 4;
    if (head + status_bytes > scratch_bytes || n >= (1ull << 30)) return nullptr;
    uint32_t* tickets = reinterpret_cast<uint32_t*>(scratch);
    uint32_t* ghist = reinterpret_cast<uint32_t*>(scratch + 64);
    uint32_t* status = reinterpret_cast<uint32_t*>(scratch + head);
    // 32 significant bits or fewer: the passes work on packed u32 words (APM_SORT_NARROW=0 keeps the u64 form)
    static int narrow_ok = -1;
    if (narrow_ok < 0) {
        const char* e = getenv("APM_SORT_NARROW");
        narrow_ok = e ? atoi(e) != 0 : 1;
    }
    const bool narrow = narrow_ok && vbits + kbits <= 32 && vbits < 32;
    static unsigned long long attr_done = 0;  // per device ordinal
    if (apm_first_on_device(&attr_done)) {
        cudaError_t ea = cudaFuncSetAttribute(os_pass_kernel<false, 0>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)OsSmemLayout<false>::bytes(2048));
        if (ea == cudaSuccess) ea = cudaFuncSetAttribute(os_pass_kernel<true, 0>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)OsSmemLayout<true>::bytes(2048));
        if (ea == cudaSuccess) ea = cudaFuncSetAttribute(os_pass_kernel<false, 8>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)OsSmemLayout<false>::bytes(256));
        if (ea == cudaSuccess) ea = cudaFuncSetAttribute(os_pass_kernel<true, 8>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)OsSmemLayout<true>::bytes(256));
        if (ea == cudaSuccess) ea = cudaFuncSetAttribute(os_hist_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, OS_MAX_PASSES * 2048 * 4);
        if (ea != cudaSuccess) { attr_done = 0; *err = ea; return nullptr; }
    }
    *err = cudaMemsetAsync(scratch, 0, head + status_bytes, st);
    if (*err != cudaSuccess) return nullptr;
    unsigned hg = (unsigned)((n + OS_THREADS * 16 - 1) / (OS_THREADS * 16));
    const unsigned hg_cap = (unsigned)apm_sm_count() * 4;
    if (hg > hg_cap) hg = hg_cap;
    os_hist_kernel<<<hg, OS_THREADS, (size_t)plan.npass * 2048 * 4, st>>>(a, n, plan, ghist);
    os_scan_kernel<<<plan.npass, 1024, 0, st>>>(ghist);
    uint64_t* src = a;
    uint64_t* dst = b;
    for (int p = 0; p < plan.npass; ++p) {
        const int first = p == 0, last = p == plan.npass - 1;
        auto kernel = narrow ? (per == 8 ? os_pass_kernel<true, 8> : os_pass_kernel<true, 0>)
                             : (per == 8 ? os_pass_kernel<false, 8> : os_pass_kernel<false, 0>);
        const size_t smem = narrow ? OsSmemLayout<true>::bytes(bins) : OsSmemLayout<false>::bytes(bins);
        kernel<<<(unsigned)tiles, OS_THREADS, smem, st>>>(src, dst, (uint32_t)n, plan.vbits, plan.shift[p], plan.bits[p], first, last,
                                                          ghist + ((size_t)p << OS_MAX_BITS), status + (size_t)p * tiles * bins,
                                                          tickets + p);
        uint64_t* t = src; src = dst; dst = t;
    }
    if (launches) *launches += 2 + plan.npass;
    *err = cudaGetLastError();
    return src;
}

// Loads this file's kernels on the current device (see apm_preload_kernels in capi.cu).
cudaError_t apm_preload_k3_onesweep() {
    const void* fns[] = {(const void*)os_hist_kernel, (const void*)os_scan_kernel, (const void*)os_pass_kernel<false, 0>, (const void*)os_pass_kernel<true, 0>, (const void*)os_pass_kernel<false, 8>, (const void*)os_pass_kernel<true, 8>};
    cudaError_t first = cudaSuccess;
    for (const void* f : fns) {
        cudaFuncAttributes a;
        const cudaError_t e = apm_preload_one(&a, f);
        if (e != cudaSuccess && first == cudaSuccess) first = e;
    }
    return first;
}
