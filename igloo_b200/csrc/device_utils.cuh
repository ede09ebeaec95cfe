// device_utils.cuh — sm_100a device helpers: mbarrier, 1-D bulk async copies
// (TMA engine, SASS UBLKCP), named barriers, streaming loads/stores, hashing and
// the single-pass ordered scan ("decoupled look-back") used by every kernel that
// compacts rows.
#pragma once

#include <cuda_runtime.h>
#include <stdint.h>

namespace igl {

__device__ __forceinline__ uint32_t smem_u32(const void* p) {
    return static_cast<uint32_t>(__cvta_generic_to_shared(p));
}

// ---- mbarrier ---------------------------------------------------------------
__device__ __forceinline__ void mbar_init(uint64_t* bar, uint32_t count) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(bar)), "r"(count) : "memory");
}
__device__ __forceinline__ void mbar_fence_init() {
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
}
__device__ __forceinline__ void mbar_arrive(uint64_t* bar) {
    asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(smem_u32(bar)) : "memory");
}
__device__ __forceinline__ void mbar_arrive_expect_tx(uint64_t* bar, uint32_t bytes) {
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(bar)), "r"(bytes) : "memory");
}
__device__ __forceinline__ bool mbar_try_wait(uint64_t* bar, uint32_t parity) {
    uint32_t ok;
    asm volatile(
        "{\n\t.reg .pred p;\n\t"
        "mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\t"
        "selp.u32 %0, 1, 0, p;\n\t}"
        : "=r"(ok)
        : "r"(smem_u32(bar)), "r"(parity)
        : "memory");
    return ok != 0;
}
__device__ __forceinline__ void mbar_wait(uint64_t* bar, uint32_t parity) {
    while (!mbar_try_wait(bar, parity)) {
    }
}

// ---- bulk async copy global -> shared (TMA engine, no tensor map needed) -----
// dst, src 16-byte aligned; bytes a multiple of 16.
__device__ __forceinline__ void bulk_g2s(void* smem_dst, const void* gmem_src, uint32_t bytes, uint64_t* bar) {
    asm volatile(
        "cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(
            smem_u32(smem_dst)),
        "l"(gmem_src), "r"(bytes), "r"(smem_u32(bar))
        : "memory");
}

// ---- named barriers (producer/consumer split inside one CTA) -----------------
__device__ __forceinline__ void bar_sync(int id, int nthreads) {
    asm volatile("bar.sync %0, %1;" ::"r"(id), "r"(nthreads) : "memory");
}
__device__ __forceinline__ void bar_arrive(int id, int nthreads) {
    asm volatile("bar.arrive %0, %1;" ::"r"(id), "r"(nthreads) : "memory");
}

// ---- streaming global access --------------------------------------------------
__device__ __forceinline__ uint64_t ld_volatile_u64(const uint64_t* p) {
    uint64_t v;
    asm volatile("ld.volatile.global.u64 %0, [%1];" : "=l"(v) : "l"(p) : "memory");
    return v;
}
__device__ __forceinline__ void st_volatile_u64(uint64_t* p, uint64_t v) {
    asm volatile("st.volatile.global.u64 [%0], %1;" ::"l"(p), "l"(v) : "memory");
}
__device__ __forceinline__ void st_cs_u32(void* p, uint32_t v) {
    asm volatile("st.global.cs.u32 [%0], %1;" ::"l"(p), "r"(v) : "memory");
}
__device__ __forceinline__ void st_cs_u64(void* p, uint64_t v) {
    asm volatile("st.global.cs.u64 [%0], %1;" ::"l"(p), "l"(v) : "memory");
}
__device__ __forceinline__ ulonglong2 ld_nc_u64x2(const void* p) {
    ulonglong2 v;
    asm volatile("ld.global.nc.L1::no_allocate.v2.u64 {%0, %1}, [%2];" : "=l"(v.x), "=l"(v.y) : "l"(p));
    return v;
}

// One 32-byte sector (a two-slot bucket of the join table) in a single request.
__device__ __forceinline__ ulonglong4 ld_nc_u64x4(const void* p) {
    ulonglong4 v;
    asm volatile("ld.global.nc.L1::no_allocate.v4.u64 {%0, %1, %2, %3}, [%4];"
                 : "=l"(v.x), "=l"(v.y), "=l"(v.z), "=l"(v.w) : "l"(p));
    return v;
}

// ---- hashing ------------------------------------------------------------------
__host__ __device__ __forceinline__ uint64_t splitmix64(uint64_t x) {
    x += 0x9E3779B97F4A7C15ull;
    x = (x ^ (x >> 30)) * 0xBF58476D1CE4E5B9ull;
    x = (x ^ (x >> 27)) * 0x94D049BB133111EBull;
    return x ^ (x >> 31);
}
// Join / partition hash: one multiply-xorshift round pair (cheaper than the full
// splitmix, still avalanches the high bits the table index is taken from).
__host__ __device__ __forceinline__ uint64_t mix64(uint64_t k) {
    k ^= k >> 32;
    k *= 0xD6E8FEB86659FD93ull;
    k ^= k >> 32;
    k *= 0xD6E8FEB86659FD93ull;
    k ^= k >> 32;
    return k;
}

// ---- ordered single-pass scan: tile status words -----------------------------
// status[t] = flag << 62 | value ; flag 1 = this tile's own count is known,
// flag 2 = the inclusive prefix up to and including this tile is known.
constexpr uint64_t kFlagAggregate = 1ull << 62;
constexpr uint64_t kFlagPrefix = 2ull << 62;
constexpr uint64_t kValueMask = (1ull << 62) - 1;

// Publishes a tile's own count (one lane).  Tile 0's count is already a prefix.
__device__ __forceinline__ void lookback_publish(uint64_t* status, int64_t tile, uint64_t count) {
    st_volatile_u64(status + tile, (tile == 0 ? kFlagPrefix : kFlagAggregate) | count);
}

// Called by one full warp after lookback_publish.  Walks back over the
// predecessors 128 at a time until it meets a known prefix, publishes this
// tile's inclusive prefix and returns the exclusive one (same value in every
// lane).  Word j of lane l is predecessor idx - (32*j + l): every load
// instruction reads 32 consecutive words, i.e. 8 sector requests instead of 32 —
// hundreds of control warps poll the same few lines of L2, and that traffic, not
// latency, is what bounds the scan.  A failed poll backs off with nanosleep.
__device__ __forceinline__ uint64_t lookback_exclusive(uint64_t* status, int64_t tile, uint64_t count, int lane) {
    if (tile == 0) return 0;
    uint64_t excl = 0;
    int64_t idx = tile - 1;   // nearest predecessor not yet accounted for
    while (true) {
        uint64_t w[4];
        bool pending;
        unsigned ns = 32;
        while (true) {
            pending = false;
#pragma unroll
            for (int j = 0; j < 4; ++j) {
                const int64_t i = idx - (32 * j + lane);
                w[j] = (i >= 0) ? ld_volatile_u64(status + i) : kFlagPrefix;   // before tile 0: prefix 0
                pending |= (w[j] >> 62) == 0;
            }
            if (!__any_sync(0xffffffffu, pending)) break;
            __nanosleep(ns);
            if (ns < 256) ns <<= 1;
        }
        // nearest known prefix in (j, lane) order
        int jstar = 4, lstar = 32;
#pragma unroll
        for (int j = 3; j >= 0; --j) {
            const uint32_t b = __ballot_sync(0xffffffffu, (w[j] >> 62) == 2);
            if (b) { jstar = j; lstar = __ffs(b) - 1; }
        }
        uint64_t v = 0;
#pragma unroll
        for (int j = 0; j < 4; ++j)
            if (j < jstar || (j == jstar && lane <= lstar)) v += w[j] & kValueMask;
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
        excl += v;
        if (jstar < 4) break;
        idx -= 128;
    }
    if (lane == 0) st_volatile_u64(status + tile, kFlagPrefix | (excl + count));
    return excl;
}

}  // namespace igl
