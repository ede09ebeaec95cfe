// atomic_bench.cu — what one insert into an L2-resident table region can cost on B200.
// N inserts walk a 2^S-slot table of 16-byte slots region by region (16 MB regions, random slot inside the
// region: the access pattern of join_build_unique_kernel on a radix-partitioned build side).
//   nvcc -O3 -gencode arch=compute_100a,code=sm_100a -o atomic_bench atomic_bench.cu && ./atomic_bench
#include <cstdint>
#include <cstdio>
#include <cuda_runtime.h>

struct Slot { unsigned long long key, val; };
__device__ __forceinline__ uint64_t mix(uint64_t k) { k ^= k >> 32; k *= 0xD6E8FEB86659FD93ull; k ^= k >> 32; k *= 0xD6E8FEB86659FD93ull; k ^= k >> 32; return k; }

template <int MODE>
__global__ void bench(Slot* t, long long n, int region_slots_log2, int rows_per_region_log2, unsigned long long* sink) {
    const long long stride = (long long)gridDim.x * blockDim.x;
    unsigned long long acc = 0;
    for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += stride) {
        const uint64_t h = mix((uint64_t)i);
        const uint64_t slot = ((uint64_t)(i >> rows_per_region_log2) << region_slots_log2) + ((h >> 20) & ((1ull << region_slots_log2) - 1) & ~1ull);
        const unsigned long long k = h | 1;
        if (MODE == 0) { unsigned long long p = atomicCAS(&t[slot].key, 0ull, k); if (p == 0) t[slot].val = i; acc += p; }
        if (MODE == 1) { acc += atomicCAS(&t[slot].key, 0ull, k); }
        if (MODE == 2) { acc += atomicExch(&t[slot].key, k); }
        if (MODE == 3) { acc += atomicCAS((unsigned int*)&t[slot].key, 0u, (unsigned int)k); }
        if (MODE == 4) { *reinterpret_cast<ulonglong2*>(&t[slot]) = make_ulonglong2(k, (unsigned long long)i); }
        if (MODE == 5) { atomicAdd((unsigned int*)&t[slot].key, 1u); }            // no return value: RED
        if (MODE == 6) { acc += t[slot].key; }                                      // plain 8-byte load
        if (MODE == 7) { unsigned long long p = atomicCAS(&t[slot].key, 0ull, k); acc += p; t[slot].val = i; }   // store without waiting for the CAS
    }
    if (acc == 0x1234567) *sink = acc;
}

int main() {
    const int S = 26;                       // 2^26 slots = 1 GiB
    const long long n = 1ll << 25;          // load 0.5
    Slot* t; unsigned long long* sink;
    cudaMalloc(&t, sizeof(Slot) << S); cudaMalloc(&sink, 8);
    cudaEvent_t a, b; cudaEventCreate(&a); cudaEventCreate(&b);
    const char* names[] = {"cas64 + val store (dependent)", "cas64", "exch64", "cas32", "st.v2.u64 (16 B)", "red.add32", "ld64", "cas64 + val store (independent)"};
    for (int region_log2 = 20; region_log2 <= 26; region_log2 += 6) {   // 16 MB regions / the whole table at random
        printf("region = 2^%d slots (%lld MB)\n", region_log2, (16ll << region_log2) >> 20);
        for (int mode = 0; mode < 8; ++mode) {
            float best = 1e9f;
            for (int rep = 0; rep < 3; ++rep) {
                cudaMemset(t, 0, sizeof(Slot) << S);
                cudaEventRecord(a);
                const int grid = 148 * 8, th = 256, rl = region_log2, rr = region_log2 - 1;
                switch (mode) {
                    case 0: bench<0><<<grid, th>>>(t, n, rl, rr, sink); break;
                    case 1: bench<1><<<grid, th>>>(t, n, rl, rr, sink); break;
                    case 2: bench<2><<<grid, th>>>(t, n, rl, rr, sink); break;
                    case 3: bench<3><<<grid, th>>>(t, n, rl, rr, sink); break;
                    case 4: bench<4><<<grid, th>>>(t, n, rl, rr, sink); break;
                    case 5: bench<5><<<grid, th>>>(t, n, rl, rr, sink); break;
                    case 6: bench<6><<<grid, th>>>(t, n, rl, rr, sink); break;
                    case 7: bench<7><<<grid, th>>>(t, n, rl, rr, sink); break;
                }
                cudaEventRecord(b); cudaEventSynchronize(b);
                float ms; cudaEventElapsedTime(&ms, a, b);
                if (ms < best) best = ms;
            }
            printf("  %-36s %7.3f ms  %6.1f G ops/s\n", names[mode], best, n / best * 1e-6);
        }
    }
    if (cudaGetLastError() != cudaSuccess) { printf("CUDA error\n"); return 1; }
    return 0;
}
