// device_utils.cuh — sm_100a device helpers: mbarrier, 1-D bulk async copies
// (TMA engine, SASS UBLKCP), named barriers, streaming loads/stores, hashing and
// the single-pass ordered scan ("decoupled look-back") used by every kernel that
// compacts rows.
#pragma once

#ifdef __CUDACC_RTC__
#include "stdint.h"   // supplied by fp_jit.cpp: NVRTC has no system headers
#else
#include <cuda_runtime.h>
#include <stdint.h>
#endif

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

// ---- ordered single-pass scan over tiles: block-prefix scheme ------------------
// Every tile publishes its own count early (a 32-bit word).  Tiles are grouped in
// blocks of kScanBlock; a tile's exclusive prefix = the prefix at its block's
// start + the counts of the tiles before it inside the block.  The LAST tile of a
// block publishes the next block's start prefix, so the only serial chain is one
// step per block, and a tile reads at most kScanBlock 4-byte counts (coalesced)
// plus one 8-byte block prefix.  Measured on B200: what bounds a decoupled
// look-back is the L2 request traffic of hundreds of warps polling the same few
// status lines, not latency — this layout issues ~6x fewer requests than a
// 64-bit-per-tile windowed walk and never re-polls a wide window.
//
// Scratch layout behind `status` (zeroed before the launch):
//   uint32 count[ntiles]                     bit 31 = published
//   uint64 block_prefix[ntiles / B + 2]      bit 63 = published   (8-byte aligned, after the counts)
// scan_scratch_words(ntiles) 64-bit words are needed in total.
constexpr int kScanBlock = 256;
constexpr uint32_t kCountFlag = 0x80000000u;
constexpr uint64_t kPrefixFlag = 1ull << 63;

__host__ __device__ inline long long scan_scratch_words(long long ntiles) { return (ntiles + 1) / 2 + ntiles / kScanBlock + 2; }

__device__ __forceinline__ uint32_t ld_volatile_u32(const uint32_t* p) {
    uint32_t v;
    asm volatile("ld.volatile.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
    return v;
}
__device__ __forceinline__ void st_volatile_u32(uint32_t* p, uint32_t v) {
    asm volatile("st.volatile.global.u32 [%0], %1;" ::"l"(p), "r"(v) : "memory");
}

// One lane: publish this tile's count (counts are < 2^31 per tile).
__device__ __forceinline__ void scan_publish(uint64_t* status, long long ntiles, long long tile, uint64_t count) {
    uint32_t* counts = reinterpret_cast<uint32_t*>(status);
    st_volatile_u32(counts + tile, kCountFlag | (uint32_t)count);
    if (tile == 0) st_volatile_u64(status + (ntiles + 1) / 2, kPrefixFlag);   // block 0 starts at 0
}

// One full warp, after scan_publish: returns the exclusive prefix of `tile` (same
// value in every lane) and, for the last tile of a block, publishes the next
// block's start prefix.
__device__ __forceinline__ uint64_t scan_exclusive(uint64_t* status, long long ntiles, long long tile, uint64_t count, int lane) {
    const uint32_t* counts = reinterpret_cast<const uint32_t*>(status);
    uint64_t* bprefix = status + (ntiles + 1) / 2;
    const long long blk = tile / kScanBlock;
    const int j = (int)(tile % kScanBlock);   // tiles before this one inside the block
    const uint32_t* cb = counts + blk * kScanBlock;
    uint64_t sum = 0;
#pragma unroll
    for (int w = 0; w < kScanBlock / 32; ++w) {
        const int i = w * 32 + lane;
        if (w * 32 < j) {   // warp-uniform
            uint32_t v = 0;
            unsigned ns = 32;
            while (true) {
                v = (i < j) ? ld_volatile_u32(cb + i) : kCountFlag;
                if (!__any_sync(0xffffffffu, (v & kCountFlag) == 0)) break;
                __nanosleep(ns);
                if (ns < 256) ns <<= 1;
            }
            sum += v & ~kCountFlag;
        }
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) sum += __shfl_xor_sync(0xffffffffu, sum, o);
    uint64_t bp = 0;
    if (lane == 0) {
        unsigned ns = 32;
        while (((bp = ld_volatile_u64(bprefix + blk)) & kPrefixFlag) == 0) {
            __nanosleep(ns);
            if (ns < 256) ns <<= 1;
        }
    }
    bp = __shfl_sync(0xffffffffu, bp, 0) & ~kPrefixFlag;
    const uint64_t excl = bp + sum;
    if (lane == 0 && j == kScanBlock - 1) st_volatile_u64(bprefix + blk + 1, kPrefixFlag | (excl + count));
    return excl;
}

}  // namespace igl
