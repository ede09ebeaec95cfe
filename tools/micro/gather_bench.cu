// Microbenchmark: random 32-byte gathers from an L2-resident 16 MB region,
// (a) per-lane LDG.256, (b) per-lane cp.async.bulk of 32 bytes into shared memory.
#include <cuda_runtime.h>
#include <cstdio>
#include <cstdint>
#include "../../igloo_b200/csrc/device_utils.cuh"
using namespace igl;

constexpr int kThreads = 256, kRows = 2;

__global__ void __launch_bounds__(kThreads) gather_ldg(const ulonglong4* tab, uint64_t mask, long long iters, unsigned long long* sink) {
    uint64_t x = splitmix64(blockIdx.x * 1315423911ull + threadIdx.x);
    unsigned long long acc = 0;
    for (long long it = 0; it < iters; ++it) {
        ulonglong4 q[kRows];
#pragma unroll
        for (int r = 0; r < kRows; ++r) { x = x * 6364136223846793005ull + 1442695040888963407ull; q[r] = ld_nc_u64x4(tab + ((x >> 20) & mask)); }
#pragma unroll
        for (int r = 0; r < kRows; ++r) acc += q[r].x + q[r].w;
    }
    if (acc == 0x1234567) *sink = acc;
}

__global__ void __launch_bounds__(kThreads) gather_tma(const ulonglong4* tab, uint64_t mask, long long iters, unsigned long long* sink) {
    __shared__ __align__(128) ulonglong4 buf[2][kThreads * kRows];
    __shared__ uint64_t bar[2];
    if (threadIdx.x == 0) { mbar_init(&bar[0], kThreads); mbar_init(&bar[1], kThreads); mbar_fence_init(); }
    __syncthreads();
    uint64_t x = splitmix64(blockIdx.x * 1315423911ull + threadIdx.x);
    unsigned long long acc = 0;
    auto issue = [&](int s) {
        mbar_arrive_expect_tx(&bar[s], 32 * kRows);
#pragma unroll
        for (int r = 0; r < kRows; ++r) {
            x = x * 6364136223846793005ull + 1442695040888963407ull;
            bulk_g2s(&buf[s][r * kThreads + threadIdx.x], tab + ((x >> 20) & mask), 32, &bar[s]);
        }
    };
    issue(0);
    for (long long it = 0; it < iters; ++it) {
        const int s = it & 1;
        if (it + 1 < iters) issue(s ^ 1);
        mbar_wait(&bar[s], (uint32_t)(it >> 1) & 1u);
#pragma unroll
        for (int r = 0; r < kRows; ++r) { const ulonglong4 q = buf[s][r * kThreads + threadIdx.x]; acc += q.x + q.w; }
        __syncthreads();   // buffer s is free again for iteration it + 2
    }
    if (acc == 0x1234567) *sink = acc;
}

int main() {
    const size_t bytes = 16u << 20;
    ulonglong4* tab; unsigned long long* sink;
    cudaMalloc(&tab, bytes); cudaMemset(tab, 1, bytes); cudaMalloc(&sink, 8);
    const uint64_t mask = bytes / 32 - 1;
    int sms = 0; cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, 0);
    cudaEvent_t e0, e1; cudaEventCreate(&e0); cudaEventCreate(&e1);
    for (int ctas = 2; ctas <= 6; ctas += 2) {
        const long long iters = 4000;
        for (int mode = 0; mode < 2; ++mode) {
            for (int rep = 0; rep < 2; ++rep) {
                cudaEventRecord(e0);
                if (mode == 0) gather_ldg<<<sms * ctas, kThreads>>>(tab, mask, iters, sink);
                else gather_tma<<<sms * ctas, kThreads>>>(tab, mask, iters, sink);
                cudaEventRecord(e1); cudaEventSynchronize(e1);
            }
            float ms; cudaEventElapsedTime(&ms, e0, e1);
            const double n = (double)sms * ctas * kThreads * kRows * iters;
            printf("%s ctas/SM=%d  %.1f G gathers/s  (%.2f cyc/gather/SM @1.9GHz)  err=%s\n", mode ? "TMA32" : "LDG256", ctas, n / ms / 1e6,
                   1.9e9 / (n / (ms / 1e3) / sms), cudaGetErrorString(cudaGetLastError()));
        }
    }
    return 0;
}
