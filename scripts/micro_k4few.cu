// What one 96-slot turn of baseline_few_kernel's adders costs, without the loader: the chunk data sits in shared
// memory, two warps take the chunks in turn.  Variants of the hand-over are timed against the bare chain.
// nvcc -gencode arch=compute_100a,code=sm_100a -O3 -fmad=false -o /tmp/micro_k4few scripts/micro_k4few.cu && /tmp/micro_k4few
#include <cstdio>
#include <cuda_runtime.h>
#include <stdint.h>

constexpr int ROWS = 96;

struct alignas(16) Mail { double sum; uint32_t cnt; uint32_t seq; };

__device__ __forceinline__ uint32_t smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }

// MODE 0: bar.sync / bar.arrive + mailbox     MODE 1: polled mailbox with sequence number
// MODE 2: one warp, LDS then chain, no hand-over   MODE 3: chain only (registers loaded once)
// MODE 4: like 0, but the mailbox carries the sum alone (8 bytes)
template <int MODE>
__global__ void __launch_bounds__(64) turns(const double* __restrict__ x, double* __restrict__ out, long long* __restrict__ cycles, int nchunks) {
    __shared__ double buf[1][ROWS][32];
    __shared__ Mail mail[2][32];
    const int lane = threadIdx.x & 31, w = threadIdx.x >> 5;
    for (int i = threadIdx.x; i < ROWS * 32; i += 64) (&buf[0][0][0])[i] = x[i];
    if (threadIdx.x < 64) { Mail m; m.sum = 0; m.cnt = 0; m.seq = 0; mail[w][lane] = m; }
    __syncthreads();
    if ((MODE == 2 || MODE == 3) && w == 1) return;
    double sum = 0.0;
    uint32_t cnt = 0;
    double v[ROWS];
    if (MODE == 3) {
#pragma unroll
        for (int r = 0; r < ROWS; ++r) v[r] = buf[0][r][lane];
    }
    const long long t0 = clock64();
    const int step = (MODE == 2 || MODE == 3) ? 1 : 2;
    for (int c = (MODE == 2 || MODE == 3) ? 0 : w; c < nchunks; c += step) {
        if (MODE != 3) {
#pragma unroll
            for (int r = 0; r < ROWS; ++r) v[r] = buf[0][r][lane];
        }
        if (MODE == 0 || MODE == 1 || MODE == 4) {
            uint32_t hi_max = 0;
#pragma unroll
            for (int r = 0; r < ROWS; ++r) {
                const uint32_t hi = (uint32_t)__double2hiint(v[r]) & 0x7fffffffu;
                hi_max = hi > hi_max ? hi : hi_max;
            }
            if (!__all_sync(0xffffffffu, hi_max < 0x7ff00000u)) cnt += 1000;
        }
        if (c > 0) {
            if (MODE == 0) {
                asm volatile("bar.sync %0, 64;" ::"r"(1u + (w ^ 1u)) : "memory");
                unsigned long long a, b;
                asm volatile("ld.volatile.shared.v2.b64 {%0, %1}, [%2];" : "=l"(a), "=l"(b) : "r"(smem_u32(&mail[w ^ 1][lane])) : "memory");
                sum = __longlong_as_double((long long)a);
                cnt = (uint32_t)b;
            } else if (MODE == 4) {
                asm volatile("bar.sync %0, 64;" ::"r"(1u + (w ^ 1u)) : "memory");
                unsigned long long a;
                asm volatile("ld.volatile.shared.b64 %0, [%1];" : "=l"(a) : "r"(smem_u32(&mail[w ^ 1][lane])) : "memory");
                sum = __longlong_as_double((long long)a);
            } else if (MODE == 1) {
                unsigned long long a, b;
                do {
                    asm volatile("ld.volatile.shared.v2.b64 {%0, %1}, [%2];" : "=l"(a), "=l"(b) : "r"(smem_u32(&mail[w ^ 1][lane])) : "memory");
                } while (!__all_sync(0xffffffffu, (uint32_t)(b >> 32) == (uint32_t)c));
                sum = __longlong_as_double((long long)a);
                cnt = (uint32_t)b;
            }
        }
#pragma unroll
        for (int r = 0; r < ROWS; ++r) sum = __dadd_rn(sum, v[r]);
        cnt += ROWS;
        if (c + 1 < nchunks) {
            if (MODE == 0) {
                asm volatile("st.volatile.shared.v2.b64 [%0], {%1, %2};" ::"r"(smem_u32(&mail[w][lane])), "l"((unsigned long long)__double_as_longlong(sum)), "l"((unsigned long long)cnt) : "memory");
                asm volatile("bar.arrive %0, 64;" ::"r"(1u + w) : "memory");
            } else if (MODE == 4) {
                asm volatile("st.volatile.shared.b64 [%0], %1;" ::"r"(smem_u32(&mail[w][lane])), "l"((unsigned long long)__double_as_longlong(sum)) : "memory");
                asm volatile("bar.arrive %0, 64;" ::"r"(1u + w) : "memory");
            } else if (MODE == 1) {
                asm volatile("st.volatile.shared.v2.b64 [%0], {%1, %2};" ::"r"(smem_u32(&mail[w][lane])), "l"((unsigned long long)__double_as_longlong(sum)),
                             "l"(((unsigned long long)(c + 1) << 32) | cnt) : "memory");
            }
        }
    }
    const long long t1 = clock64();
    out[threadIdx.x] = sum + cnt;
    if (lane == 0) cycles[w] = t1 - t0;
}

template <int MODE>
void run(const char* name, const double* x, double* o, long long* c, int nchunks) {
    for (int rep = 0; rep < 2; ++rep) {
        cudaMemset(c, 0, 16);
        turns<MODE><<<1, 64>>>(x, o, c, nchunks);
        long long h[2];
        cudaMemcpy(h, c, 16, cudaMemcpyDeviceToHost);
        const long long m = h[0] > h[1] ? h[0] : h[1];
        if (rep) printf("%-44s %8.1f cycles per 96-slot chunk = %.2f per slot  (%s)\n", name, (double)m / nchunks, (double)m / nchunks / ROWS,
                        cudaGetErrorString(cudaGetLastError()));
    }
}

int main() {
    double *x, *o;
    long long* c;
    cudaMalloc(&x, 2 * ROWS * 32 * 8);
    cudaMalloc(&o, 64 * 8);
    cudaMalloc(&c, 16);
    double h[2 * ROWS * 32];
    for (int i = 0; i < 2 * ROWS * 32; ++i) h[i] = 100.0 + (i % 977) * 0.1;
    cudaMemcpy(x, h, sizeof(h), cudaMemcpyHostToDevice);
    const int n = 400;
    run<3>("chain only (registers loaded once)", x, o, c, n);
    run<2>("one warp: LDS then chain", x, o, c, n);
    run<1>("two warps, polled mailbox", x, o, c, n);
    run<0>("two warps, named barrier + 16-byte mailbox", x, o, c, n);
    run<4>("two warps, named barrier + 8-byte mailbox", x, o, c, n);
    return 0;
}
