// Stand-alone percentile operator: the Array.prototype pair binaryConcat + calcPercentile
// (util_methods.js:102-106, 112-142) for one array of durations that already sits in device
// memory - sort ascending (duplicates kept), then apply the reference's index rule.
#include <mutex>
#include <string>

#include "ctx.h"

namespace {

__global__ void __launch_bounds__(256) widen_kernel(const int32_t* __restrict__ v, uint64_t* __restrict__ w, uint32_t n) {
    const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i < n) w[i] = (uint64_t)(uint32_t)v[i];  // key field 0: only the value bits are sorted
}

// max over the values, for the number of radix passes
__global__ void __launch_bounds__(256) max_kernel(const int32_t* __restrict__ v, uint32_t n, uint32_t* __restrict__ out) {
    uint32_t m = 0;
    for (uint32_t i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += gridDim.x * blockDim.x) {
        const uint32_t x = (uint32_t)v[i];
        m = x > m ? x : m;
    }
    m = __reduce_max_sync(0xffffffffu, m);
    if ((threadIdx.x & 31) == 0 && m) atomicMax(out, m);
}

struct PctParams {
    int npct;
    double pfrac[APM_MAX_PCT];
};

// narrow the sorted words back to int32 and evaluate calcPercentile for each requested percentile
__global__ void __launch_bounds__(256) narrow_pct_kernel(const uint64_t* __restrict__ w, uint32_t n, int32_t* __restrict__ sorted,
                                                         const PctParams p, double* __restrict__ pct_out) {
    const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i < n && sorted) sorted[i] = (int32_t)(uint32_t)w[i];
    if (blockIdx.x == 0 && (int)threadIdx.x < p.npct) {
        const double nan = __longlong_as_double(0x7ff8000000000000ll);
        double r = nan;
        if (n > 0) {
            const double f = p.pfrac[threadIdx.x];
            const double index = __dsub_rn(__dmul_rn(f, (double)n), 1.0);  // (percentile / 100.0) * length - 1.0
            uint32_t i0;
            bool pair = false;
            if (f == 0.0) i0 = 0;
            else if (f == 1.0) i0 = n - 1;
            else if (n == 1 || index == trunc(index)) i0 = index < 0 ? 0u : (uint32_t)index;
            else {
                const double ci = ceil(index);
                i0 = ci < 0 ? 0u : (uint32_t)ci;
                pair = i0 != n - 1;
            }
            const double v0 = (double)(uint32_t)w[i0];
            r = pair ? __dmul_rn(__dadd_rn(v0, (double)(uint32_t)w[i0 + 1]), 0.5) : v0;
        }
        pct_out[threadIdx.x] = r;
    }
}

// Scratch layout for n values (all offsets 256-byte aligned):
//   words A [n] u64 | words B [n] u64 | hist [256 * tiles] u32 | totals [256] u32 | max u32 | pct [4] f64
struct ScratchPlan {
    size_t off_a, off_b, off_hist, off_totals, off_max, off_pct, bytes;
};

size_t align256(size_t x) { return (x + 255) & ~(size_t)255; }

ScratchPlan plan_for(uint64_t n) {
    ScratchPlan s;
    const uint64_t m = n ? n : 1;
    s.off_a = 0;
    s.off_b = align256(s.off_a + m * 8);
    s.off_hist = align256(s.off_b + m * 8);
    s.off_totals = align256(s.off_hist + 256 * apm_radix_sort_tiles(m) * 4);
    s.off_max = align256(s.off_totals + 256 * 4);
    s.off_pct = align256(s.off_max + 4);
    s.bytes = align256(s.off_pct + APM_MAX_PCT * 8);
    return s;
}

apm_status run_percentiles(unsigned char* scratch, const int32_t* d_values, uint32_t n, const double* pcts, int32_t npct,
                           double* out_pct, int32_t* d_sorted) {
    const ScratchPlan sp = plan_for(n);
    uint64_t* a = (uint64_t*)(scratch + sp.off_a);
    uint64_t* b = (uint64_t*)(scratch + sp.off_b);
    uint32_t* hist = (uint32_t*)(scratch + sp.off_hist);
    uint32_t* totals = (uint32_t*)(scratch + sp.off_totals);
    uint32_t* dmax = (uint32_t*)(scratch + sp.off_max);
    double* dpct = (double*)(scratch + sp.off_pct);
    PctParams pp;
    pp.npct = npct;
    for (int i = 0; i < APM_MAX_PCT; ++i) pp.pfrac[i] = i < npct ? pcts[i] / 100.0 : 0.0;
    cudaStream_t st = 0;
    const uint64_t* sorted_words = a;
    if (n > 0) {
        // the scratch keeps the previous call's contents: the one word that is accumulated into is cleared
        CUDA_TRY(nullptr, cudaMemsetAsync(dmax, 0, sizeof(uint32_t), st));
        const unsigned g = (n + 255) / 256;
        max_kernel<<<g < (unsigned)apm_sm_count() * 8 ? g : (unsigned)apm_sm_count() * 8, 256, 0, st>>>(d_values, n, dmax);
        widen_kernel<<<g, 256, 0, st>>>(d_values, a, n);
        uint32_t hmax = 0;
        CUDA_TRY(nullptr, cudaMemcpyAsync(&hmax, dmax, sizeof hmax, cudaMemcpyDeviceToHost, st));
        CUDA_TRY(nullptr, cudaStreamSynchronize(st));
        if (hmax > 0x7FFFFFFFu) return apm_fail(nullptr, APM_E_INVALID, "percentiles: negative duration in the input");
        sorted_words = apm_radix_sort_words(st, a, b, n, hist, totals, apm_bits_of(hmax), 0, nullptr);
    }
    narrow_pct_kernel<<<n ? (n + 255) / 256 : 1, 256, 0, st>>>(sorted_words, n, d_sorted, pp, dpct);
    CUDA_TRY(nullptr, cudaGetLastError());
    if (npct > 0) CUDA_TRY(nullptr, cudaMemcpyAsync(out_pct, dpct, npct * sizeof(double), cudaMemcpyDeviceToHost, st));
    CUDA_TRY(nullptr, cudaStreamSynchronize(st));
    return APM_OK;
}

apm_status check_args(const int32_t* d_values, uint32_t n, const double* pcts, int32_t npct, double* out_pct) {
    if ((!d_values && n) || npct < 0 || npct > APM_MAX_PCT || (npct && (!pcts || !out_pct)))
        return apm_fail(nullptr, APM_E_INVALID, "percentiles: bad argument");
    for (int i = 0; i < npct; ++i)
        if (!(pcts[i] >= 0.0 && pcts[i] <= 100.0)) return apm_fail(nullptr, APM_E_INVALID, "percentiles: percentile out of [0,100]");
    return APM_OK;
}

// ---------------------------------------------------------------------------------------------
// Scratch for apm_sorted_percentiles_noalloc: a fixed array inside the library image, so the
// call never has to ask the CUDA allocator for memory.  Every GPU that loads the library
// carries these bytes until the process exits.  In host code `g_noalloc_scratch` is only a
// handle for the symbol - it is NOT a device address - so the address each device gave the
// array is looked up once with cudaGetSymbolAddress and kept here.
// ---------------------------------------------------------------------------------------------
constexpr uint32_t kNoAllocMaxValues = APM_NOALLOC_MAX_VALUES;
constexpr size_t kNoAllocScratchBytes = 2 * ((size_t)kNoAllocMaxValues * 8) + 256 * ((kNoAllocMaxValues + 4095) / 4096) * 4 +
                                        256 * 4 + 4 + APM_MAX_PCT * 8 + 6 * 256 /* alignment slack */;
static __device__ __align__(256) unsigned char g_noalloc_scratch[kNoAllocScratchBytes];

constexpr int kMaxDevices = 64;
unsigned char* g_noalloc_ptr[kMaxDevices];  // per device ordinal, nullptr until that device's set-up ran
std::mutex g_noalloc_mutex;

}  // namespace

extern "C" {

apm_status apm_sorted_percentiles(int32_t device, const int32_t* d_values, uint32_t n, const double* pcts, int32_t npct,
                                  double* out_pct, int32_t* d_sorted);

apm_status apm_sorted_percentiles_noalloc(int32_t device, const int32_t* d_values, uint32_t n, const double* pcts,
                                          int32_t npct, double* out_pct, int32_t* d_sorted) {
    apm_status st = check_args(d_values, n, pcts, npct, out_pct);
    if (st) return st;
    if (n > kNoAllocMaxValues)
        return apm_fail(nullptr, APM_E_CAPACITY,
                        "apm_sorted_percentiles_noalloc: n exceeds APM_NOALLOC_MAX_VALUES (" + std::to_string(kNoAllocMaxValues) +
                            "); use apm_sorted_percentiles for larger arrays");
    if (device < 0 || device >= kMaxDevices) return apm_fail(nullptr, APM_E_INVALID, "apm_sorted_percentiles_noalloc: device ordinal");
    ApmDeviceGuard guard(device);
    unsigned char* scratch;
    {
        std::lock_guard<std::mutex> lock(g_noalloc_mutex);
        if (!g_noalloc_ptr[device]) {
            // first call on this device: resolve the array's device address (this device is current)
            void* p = nullptr;
            CUDA_TRY(nullptr, cudaGetSymbolAddress(&p, g_noalloc_scratch));
            g_noalloc_ptr[device] = (unsigned char*)p;
        }
        scratch = g_noalloc_ptr[device];
    }
    if (plan_for(n).bytes > kNoAllocScratchBytes) return apm_fail(nullptr, APM_E_CAPACITY, "apm_sorted_percentiles_noalloc: scratch plan exceeds the static array");
    return run_percentiles(scratch, d_values, n, pcts, npct, out_pct, d_sorted);
}

// the same operator for an array in host memory (what a JavaScript caller has): copies in, sorts, copies out
apm_status apm_sorted_percentiles_host(int32_t device, const int32_t* h_values, uint32_t n, const double* pcts, int32_t npct,
                                       double* out_pct, int32_t* h_sorted) {
    apm_status st = check_args(h_values, n, pcts, npct, out_pct);
    if (st) return st;
    ApmDeviceGuard guard(device);
    int32_t* d = nullptr;
    if (n) {
        CUDA_TRY(nullptr, cudaMalloc((void**)&d, (size_t)n * sizeof(int32_t)));
        cudaError_t e = cudaMemcpy(d, h_values, (size_t)n * sizeof(int32_t), cudaMemcpyHostToDevice);
        if (e != cudaSuccess) { cudaFree(d); return apm_fail_cuda(nullptr, e, "cudaMemcpy", __FILE__, __LINE__); }
    }
    st = apm_sorted_percentiles(device, d, n, pcts, npct, out_pct, h_sorted ? d : nullptr);
    if (!st && n && h_sorted) {
        cudaError_t e = cudaMemcpy(h_sorted, d, (size_t)n * sizeof(int32_t), cudaMemcpyDeviceToHost);
        if (e != cudaSuccess) st = apm_fail_cuda(nullptr, e, "cudaMemcpy", __FILE__, __LINE__);
    }
    cudaFree(d);
    return st;
}

apm_status apm_sorted_percentiles(int32_t device, const int32_t* d_values, uint32_t n, const double* pcts, int32_t npct,
                                  double* out_pct, int32_t* d_sorted) {
    apm_status st = check_args(d_values, n, pcts, npct, out_pct);
    if (st) return st;
    ApmDeviceGuard guard(device);
    unsigned char* scratch = nullptr;
    CUDA_TRY(nullptr, cudaMalloc((void**)&scratch, plan_for(n).bytes));
    st = run_percentiles(scratch, d_values, n, pcts, npct, out_pct, d_sorted);
    cudaFree(scratch);
    return st;
}

}  // extern "C"

// Loads this file's kernels on the current device (see apm_preload_kernels in capi.cu).
cudaError_t apm_preload_k3_percentile_op() {
    const void* fns[] = {(const void*)max_kernel, (const void*)narrow_pct_kernel, (const void*)widen_kernel};
    cudaError_t first = cudaSuccess;
    for (const void* f : fns) {
        cudaFuncAttributes a;
        const cudaError_t e = apm_preload_one(&a, f);
        if (e != cudaSuccess && first == cudaSuccess) first = e;
    }
    return first;
}
