// partition.cu — K4: hash repartition of a column set, the data-plane half of
// the multi-GPU join's "worker fragment exchange".
//
// The reference only declares this step (FragmentType::Shuffle is never built,
// crates/coordinator/src/fragment.rs:8-13; GetDataForTask returns empty bytes,
// crates/worker/src/service.rs:26-32; the join fragment is simply pinned to the
// coordinator, crates/coordinator/src/distributed_planner.rs:65-92).  Here both
// join sides are split by dest = hash(key) so that equal keys meet on one GPU.
//
// Two passes: a histogram (so every (source, destination) pair knows its exact
// slab offset after one tiny count exchange) and a scatter that writes rows
// straight to their final place.  Destination pointers may be peer device
// pointers mapped over NVLink (CUDA IPC): the scatter then IS the exchange — no
// send buffer, no second copy.
//
// dest uses the LOW 32 bits of mix64(key); the local hash table indexes with the
// TOP bits, so partitioning does not correlate with slot placement.  NULL keys
// all go to partition 0 (they must meet: NULL == NULL in the reference).
#include "device_utils.cuh"
#include "internal.h"
#include "kernels.h"

namespace igl {

constexpr int kPartMaxParts = 16;
constexpr int kPartMaxCols = 16;
constexpr int kPartThreads = 256;
constexpr int kPartRowsPerThread = 8;
constexpr int kPartTile = kPartThreads * kPartRowsPerThread;

__device__ __forceinline__ uint64_t part_canon(int type, uint64_t raw) {
    if (type == IGL_T_I32) return (uint64_t)(int64_t)(int32_t)raw;
    if (type == IGL_T_F64) return ((raw & 0x7FFFFFFFFFFFFFFFull) > 0x7FF0000000000000ull) ? 0x7FF8000000000000ull : raw;
    return raw;
}

__device__ __forceinline__ int part_dest(const void* keys, int type, const uint8_t* valid, long long i, int nparts) {
    if (valid && !((valid[i >> 3] >> (i & 7)) & 1)) return 0;
    const uint64_t raw = type == IGL_T_I32 ? (uint64_t)(uint32_t)((const int32_t*)keys)[i] : ((const uint64_t*)keys)[i];
    const uint64_t h = mix64(part_canon(type, raw)) & 0xFFFFFFFFull;
    return (int)((h * (uint64_t)nparts) >> 32);
}

__global__ void partition_hist_kernel(const void* __restrict__ keys, int type, const uint8_t* __restrict__ valid,
                                      long long n, int nparts, unsigned long long* hist) {
    __shared__ unsigned int s_hist[kPartMaxParts];
    if (threadIdx.x < kPartMaxParts) s_hist[threadIdx.x] = 0;
    __syncthreads();
    const long long stride = (long long)gridDim.x * blockDim.x;
    unsigned int local[kPartMaxParts];
#pragma unroll
    for (int p = 0; p < kPartMaxParts; ++p) local[p] = 0;
    for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += stride) {
        const int d = part_dest(keys, type, valid, i, nparts);
#pragma unroll
        for (int p = 0; p < kPartMaxParts; ++p) local[p] += (d == p);
    }
#pragma unroll
    for (int p = 0; p < kPartMaxParts; ++p) {
        unsigned int v = local[p];
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
        if ((threadIdx.x & 31) == 0 && v) atomicAdd(&s_hist[p], v);
    }
    __syncthreads();
    if (threadIdx.x < nparts && s_hist[threadIdx.x]) atomicAdd(&hist[threadIdx.x], (unsigned long long)s_hist[threadIdx.x]);
}

struct ScatterParams {
    const uint8_t* src[kPartMaxCols];
    uint8_t shift[kPartMaxCols];
    uint8_t* dst[kPartMaxParts * kPartMaxCols];   // [part][col], already offset to this source's slab
    int ncols, nparts, key_col, key_type;
    const uint8_t* key_valid;
    long long n;
    unsigned long long* cursor;   // [nparts] rows already written per destination
};

__global__ void __launch_bounds__(kPartThreads) partition_scatter_kernel(const __grid_constant__ ScatterParams P) {
    __shared__ unsigned int s_cnt[kPartMaxParts];
    __shared__ unsigned long long s_base[kPartMaxParts];
    const long long ntiles = (P.n + kPartTile - 1) / kPartTile;
    for (long long tile = blockIdx.x; tile < ntiles; tile += gridDim.x) {
        if (threadIdx.x < kPartMaxParts) s_cnt[threadIdx.x] = 0;
        __syncthreads();
        const long long row0 = tile * kPartTile;
        int dest[kPartRowsPerThread];
        unsigned int rank[kPartRowsPerThread];
#pragma unroll
        for (int r = 0; r < kPartRowsPerThread; ++r) {
            const long long i = row0 + r * kPartThreads + threadIdx.x;
            dest[r] = -1;
            if (i < P.n) {
                dest[r] = part_dest(P.src[P.key_col], P.key_type, P.key_valid, i, P.nparts);
                rank[r] = atomicAdd(&s_cnt[dest[r]], 1u);   // shared-memory atomic: position inside this tile's run
            }
        }
        __syncthreads();
        if (threadIdx.x < P.nparts)
            s_base[threadIdx.x] = s_cnt[threadIdx.x] ? atomicAdd(&P.cursor[threadIdx.x], (unsigned long long)s_cnt[threadIdx.x]) : 0ull;
        __syncthreads();
        for (int c = 0; c < P.ncols; ++c) {
            const int sh = P.shift[c];
#pragma unroll
            for (int r = 0; r < kPartRowsPerThread; ++r) {
                if (dest[r] < 0) continue;
                const long long i = row0 + r * kPartThreads + threadIdx.x;
                const unsigned long long o = s_base[dest[r]] + rank[r];
                uint8_t* d = P.dst[dest[r] * kPartMaxCols + c];
                if (sh == 2) reinterpret_cast<uint32_t*>(d)[o] = reinterpret_cast<const uint32_t*>(P.src[c])[i];
                else reinterpret_cast<uint64_t*>(d)[o] = reinterpret_cast<const uint64_t*>(P.src[c])[i];
            }
        }
        __syncthreads();
    }
}

// ---------------------------------------------------------------------------
// High-fanout radix partition (up to 1024 parts) by the TOP bits of mix64(key) —
// the same bits the join table is indexed with, so rows sorted by partition walk
// the table region by region and every region (16 MB) stays L2-resident while
// it is built and probed.  Each random 16-byte table access otherwise costs about
// a 128-byte line of HBM traffic (profiles/README.md).
//
// Pass 1: per-CTA shared-memory histogram -> global counts -> one-block scan.
// Pass 2: every tile is counting-sorted by partition in shared memory first, so
// the write-out stores runs of consecutive rows per partition (coalesced) instead
// of one scattered store per row.
// ---------------------------------------------------------------------------
constexpr int kRadixMaxParts = 1024;
#ifndef IGL_RADIX_THREADS
#define IGL_RADIX_THREADS 512
#endif
#ifndef IGL_RADIX_CTAS
#define IGL_RADIX_CTAS 2
#endif
constexpr int kRadixThreads = IGL_RADIX_THREADS;
constexpr int kRadixBinsPerThread = kRadixMaxParts / kRadixThreads;
constexpr unsigned kRadixMaxStage = (IGL_RADIX_CTAS == 1 ? 176u : 96u) << 10;   // dynamic shared memory for the staged tile
constexpr int kRadixRowsPerThread = 8;
constexpr int kRadixTile = kRadixThreads * kRadixRowsPerThread;   // 4096 rows per tile: long write runs per partition

// bin = destination * 2^log2_parts + partition: destination from the LOW 32 hash bits (as
// part_dest), partition from the TOP bits (as the join table).  nhash == 1: partition only.
//
// Hot keys (skewed probe sides): a few keys carry so many probe rows that the GPU
// their hash names would receive far more than its share.  Keys listed in `H` are routed to the
// fixed destination H.dest instead — the probe side keeps them where they are (H.dest = this
// rank), the build side sends them to one extra "every rank reads this" destination — so the
// hash decides only the rest.  H.table: open-addressing set of canonical keys, kHotEmpty = free.
constexpr unsigned long long kHotEmpty = 0xFFFFFFFFFFFFFFFFull;
template <bool HOT>
__device__ __forceinline__ int radix_bin_of(uint64_t canon, uint64_t h, int nhash, int log2_parts, const HotKeys& H) {
    const int part = log2_parts ? (int)(h >> (64 - log2_parts)) : 0;
    int dest = nhash > 1 ? (int)(((h & 0xFFFFFFFFull) * (uint64_t)nhash) >> 32) : 0;
    if (HOT) {
        uint32_t s = (uint32_t)(h >> 24) & (uint32_t)H.mask;
        while (true) {
            const unsigned long long k = __ldg(H.table + s);
            if (k == canon) { dest = H.dest; break; }
            if (k == kHotEmpty) break;
            s = (s + 1) & (uint32_t)H.mask;
        }
    }
    return (dest << log2_parts) | part;
}
template <bool HOT>
__device__ __forceinline__ int radix_part(const void* keys, int type, long long i, int nhash, int log2_parts, const HotKeys& H) {
    const uint64_t raw = type == IGL_T_I32 ? (uint64_t)(uint32_t)((const int32_t*)keys)[i] : ((const uint64_t*)keys)[i];
    const uint64_t c = part_canon(type, raw);
    return radix_bin_of<HOT>(c, mix64(c), nhash, log2_parts, H);
}

template <bool HOT>
__global__ void __launch_bounds__(512) radix_hist_kernel(const void* __restrict__ keys, int type, long long n, int ndest, int nhash,
                                                                   int log2_parts, unsigned long long* hist, HotKeys H) {
    __shared__ unsigned int s_hist[kRadixMaxParts];
    const int nparts = ndest << log2_parts;
    for (int p = threadIdx.x; p < nparts; p += blockDim.x) s_hist[p] = 0;
    __syncthreads();
    const long long stride = (long long)gridDim.x * blockDim.x;
    for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += stride)
        atomicAdd(&s_hist[radix_part<HOT>(keys, type, i, nhash, log2_parts, H)], 1u);
    __syncthreads();
    for (int p = threadIdx.x; p < nparts; p += blockDim.x)
        if (s_hist[p]) atomicAdd(&hist[p], (unsigned long long)s_hist[p]);
}

// Every (n / nsamples)-th key in canonical form: the host counts them to find hot keys.
__global__ void sample_keys_kernel(const void* __restrict__ keys, int type, long long n, int nsamples, unsigned long long* out) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= nsamples) return;
    const long long row = (long long)(((__int128)i * n) / nsamples);
    const uint64_t raw = type == IGL_T_I32 ? (uint64_t)(uint32_t)((const int32_t*)keys)[row] : ((const uint64_t*)keys)[row];
    out[i] = part_canon(type, raw);
}

// offsets[p] = sum of hist[0..p) ; offsets[nparts] = total ; cursor[] zeroed.
__global__ void radix_scan_kernel(const unsigned long long* hist, unsigned long long* offsets, unsigned long long* cursor, int nparts) {
    __shared__ unsigned long long s[kRadixMaxParts];
    const int t = threadIdx.x;
    unsigned long long v = t < nparts ? hist[t] : 0;
    s[t] = v;
    __syncthreads();
    for (int o = 1; o < kRadixMaxParts; o <<= 1) {
        unsigned long long add = t >= o ? s[t - o] : 0;
        __syncthreads();
        s[t] += add;
        __syncthreads();
    }
    if (t < nparts) { offsets[t] = s[t] - v; cursor[t] = 0; }
    if (t == nparts - 1) offsets[nparts] = s[t];
}

struct RadixParams {
    const uint8_t* src[kPartMaxCols];
    uint8_t* dst[kPartMaxParts][kPartMaxCols];   // per destination (one when partitioning locally)
    uint8_t shift[kPartMaxCols];
    uint32_t stage_off[kPartMaxCols];   // byte offset of each column's staging area in dynamic shared memory
    int ncols, key_col, key_type, log2_parts, ndest;
    int nhash;                          // range of the hashed destination (== ndest unless hot keys add a destination)
    HotKeys hot;
    long long n;
    const unsigned long long* offsets;  // [bins]: first row of the bin inside its destination's buffer
    unsigned long long* cursor;         // [bins]
    // second-level ("fine") pass over rows already ordered by the first `log2_parts` hash bits: the bin of a row is
    // mix64 >> fine_shift (fine_bins of them — far more than a tile can rank), a tile ranks the 1024 bins that
    // start at its first row's first-level partition and places the odd row beyond them one by one
    int fine_shift, fine_lp2;
    long long fine_bins;
};

// One tile.  FULL = every one of the 4096 rows exists (all tiles but the last):
// no per-row existence predicates, hence no divergence blocks in the hot loops.
template <bool FULL>
__device__ __forceinline__ void radix_tile(const RadixParams& P, uint8_t* stage, unsigned int* s_cnt, unsigned long long* s_gbase,
                                           unsigned short* s_part, unsigned int* s_warp_tot, long long row0, int rows) {
    const int nparts = P.ndest << P.log2_parts;   // bins
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const int tid = threadIdx.x;
    int part[kRadixRowsPerThread];
    unsigned int rank[kRadixRowsPerThread];
    uint64_t keyraw[kRadixRowsPerThread];
    {
        const uint8_t* kbase = P.src[P.key_col] + (P.key_type == IGL_T_I32 ? (row0 << 2) : (row0 << 3));   // uniform base, 32-bit offsets
#pragma unroll
        for (int r = 0; r < kRadixRowsPerThread; ++r) {   // all key loads first: 8 independent requests in flight per thread
            const int li = FULL ? r * kRadixThreads + tid : min(r * kRadixThreads + tid, rows - 1);
            keyraw[r] = P.key_type == IGL_T_I32 ? (uint64_t)((const uint32_t*)kbase)[li] : ((const uint64_t*)kbase)[li];
        }
    }
#pragma unroll
    for (int r = 0; r < kRadixRowsPerThread; ++r) {
        const bool ok = FULL || (r * kRadixThreads + tid) < rows;
        const uint64_t ck = part_canon(P.key_type, keyraw[r]);
        part[r] = radix_bin_of<false>(ck, mix64(ck), P.nhash, P.log2_parts, P.hot);
        rank[r] = 0;
        if (ok) rank[r] = atomicAdd(&s_cnt[part[r]], 1u);
        else part[r] = -1;
    }
    __syncthreads();
    // exclusive scan of s_cnt[0..nparts) by the CTA (nparts <= 1024 = 2 per thread) + global reservation
    {
        const int i0 = tid * kRadixBinsPerThread;
        unsigned int c[kRadixBinsPerThread];
        unsigned int sum = 0;
#pragma unroll
        for (int q = 0; q < kRadixBinsPerThread; ++q) { c[q] = (i0 + q < nparts) ? s_cnt[i0 + q] : 0; sum += c[q]; }
        unsigned int inc = sum;
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) {
            const unsigned int t = __shfl_up_sync(0xffffffffu, inc, o);
            if (lane >= o) inc += t;
        }
        if (lane == 31) s_warp_tot[warp] = inc;
        __syncthreads();
        unsigned int ex = inc - sum;
        for (int w = 0; w < warp; ++w) ex += s_warp_tot[w];
#pragma unroll
        for (int q = 0; q < kRadixBinsPerThread; ++q) {
            if (i0 + q < nparts) {
                s_cnt[i0 + q] = ex;
                s_gbase[i0 + q] = c[q] ? P.offsets[i0 + q] + atomicAdd(&P.cursor[i0 + q], (unsigned long long)c[q]) : 0ull;
            }
            ex += c[q];
        }
    }
    __syncthreads();
    // stage: rows land grouped by partition.  Column by column, so the width test is uniform.
    unsigned int slot[kRadixRowsPerThread];
#pragma unroll
    for (int r = 0; r < kRadixRowsPerThread; ++r) {
        slot[r] = 0;
        if (FULL || part[r] >= 0) { slot[r] = s_cnt[part[r]] + rank[r]; s_part[slot[r]] = (unsigned short)part[r]; }
    }
    for (int c = 0; c < P.ncols; ++c) {
        const bool is_key = c == P.key_col;
        if (P.shift[c] == 3) {
            uint64_t* sc = reinterpret_cast<uint64_t*>(stage + P.stage_off[c]);
            const uint64_t* src = reinterpret_cast<const uint64_t*>(P.src[c]) + row0;
            uint64_t v[kRadixRowsPerThread];
#pragma unroll
            for (int r = 0; r < kRadixRowsPerThread; ++r)
                v[r] = is_key ? keyraw[r] : src[FULL ? r * kRadixThreads + tid : min(r * kRadixThreads + tid, rows - 1)];
#pragma unroll
            for (int r = 0; r < kRadixRowsPerThread; ++r) if (FULL || part[r] >= 0) sc[slot[r]] = v[r];
        } else {
            uint32_t* sc = reinterpret_cast<uint32_t*>(stage + P.stage_off[c]);
            const uint32_t* src = reinterpret_cast<const uint32_t*>(P.src[c]) + row0;
            uint32_t v[kRadixRowsPerThread];
#pragma unroll
            for (int r = 0; r < kRadixRowsPerThread; ++r)
                v[r] = is_key ? (uint32_t)keyraw[r] : src[FULL ? r * kRadixThreads + tid : min(r * kRadixThreads + tid, rows - 1)];
#pragma unroll
            for (int r = 0; r < kRadixRowsPerThread; ++r) if (FULL || part[r] >= 0) sc[slot[r]] = v[r];
        }
    }
    __syncthreads();
    // write-out: consecutive threads -> consecutive slots -> runs of consecutive destination rows
    unsigned long long d[kRadixRowsPerThread];
    int dq[kRadixRowsPerThread];   // destination of each staged slot (0 when partitioning locally)
#pragma unroll
    for (int r = 0; r < kRadixRowsPerThread; ++r) {
        const int sl = r * kRadixThreads + tid;
        d[r] = 0; dq[r] = 0;
        if (FULL || sl < rows) { const int p = s_part[sl]; d[r] = s_gbase[p] + (unsigned)(sl - s_cnt[p]); dq[r] = p >> P.log2_parts; }
    }
    for (int c = 0; c < P.ncols; ++c) {
        if (P.shift[c] == 3) {
            const uint64_t* sc = reinterpret_cast<const uint64_t*>(stage + P.stage_off[c]) + tid;
#pragma unroll
            for (int r = 0; r < kRadixRowsPerThread; ++r)
                if (FULL || r * kRadixThreads + tid < rows) reinterpret_cast<uint64_t*>(P.dst[dq[r]][c])[d[r]] = sc[r * kRadixThreads];
        } else {
            const uint32_t* sc = reinterpret_cast<const uint32_t*>(stage + P.stage_off[c]) + tid;
#pragma unroll
            for (int r = 0; r < kRadixRowsPerThread; ++r)
                if (FULL || r * kRadixThreads + tid < rows) reinterpret_cast<uint32_t*>(P.dst[dq[r]][c])[d[r]] = sc[r * kRadixThreads];
        }
    }
    __syncthreads();
}

__global__ void __launch_bounds__(kRadixThreads, IGL_RADIX_CTAS) radix_scatter_kernel(const __grid_constant__ RadixParams P) {
    extern __shared__ __align__(16) uint8_t stage[];
    __shared__ unsigned int s_cnt[kRadixMaxParts];     // count, then exclusive start
    __shared__ unsigned long long s_gbase[kRadixMaxParts];
    __shared__ unsigned short s_part[kRadixTile];       // partition of each staged slot
    __shared__ unsigned int s_warp_tot[kRadixThreads / 32];
    const int nparts = P.ndest << P.log2_parts;
    const long long ntiles = (P.n + kRadixTile - 1) / kRadixTile;
    for (long long tile = blockIdx.x; tile < ntiles; tile += gridDim.x) {
        for (int p = threadIdx.x; p < nparts; p += kRadixThreads) s_cnt[p] = 0;
        __syncthreads();
        const long long row0 = tile * kRadixTile;
        const int rows = (int)min((long long)kRadixTile, P.n - row0);
        if (rows == kRadixTile) radix_tile<true>(P, stage, s_cnt, s_gbase, s_part, s_warp_tot, row0, rows);
        else radix_tile<false>(P, stage, s_cnt, s_gbase, s_part, s_warp_tot, row0, rows);
    }
}

// ---- scatter, bulk-copy form -----------------------------------------------------------
// One 1024-thread CTA per SM.  A tile's columns arrive in shared memory through bulk async
// copies (double-buffered: the next tile travels while this one is sorted), so no global
// load sits on the tile's critical path and no row value passes through registers twice:
// rows are ranked by bin with shared-memory atomics, a permutation (slot -> row) is built,
// and the write-out reads each column through the permutation and stores runs of
// consecutive destination rows.  Used when the tile fits (2 x 4096 rows x row bytes <= 190 KB,
// or one buffer for wider rows) and the columns are 16-byte aligned.
constexpr int kRtThreads = 1024;
constexpr int kRtRows = 4;
constexpr int kRtTile = kRtThreads * kRtRows;          // 4096 rows
constexpr unsigned kRtMaxDyn = 190u << 10;   // tile buffers; the two permutation arrays (16 KB) follow them

template <int ROWS, bool HOT, int THREADS, bool FINE = false>   // ROWS rows per thread; tile = THREADS * ROWS rows (4096, or 8192 for the long-run form)
__global__ void __launch_bounds__(THREADS, THREADS >= 1024 ? 1 : 2) radix_scatter_bulk_kernel(const __grid_constant__ RadixParams P, int n_buf, uint32_t tile_bytes) {
    constexpr int kTile = THREADS * ROWS;
    constexpr uint32_t kOffScale = kTile / kRtTile;   // P.stage_off is laid out for 4096-row tiles
    constexpr int BPT = kRadixMaxParts / THREADS;     // bins per thread in the scan
    static_assert(kTile % kRtTile == 0 && kRadixMaxParts % THREADS == 0, "tile and bin layout");
    extern __shared__ __align__(128) uint8_t in_tiles[];
    __shared__ unsigned int s_cnt[2][kRadixMaxParts];      // by tile parity: count, then exclusive start
    __shared__ unsigned long long s_gbase[kRadixMaxParts];
    unsigned short* s_perm = reinterpret_cast<unsigned short*>(in_tiles + (size_t)n_buf * tile_bytes);   // slot -> row of the tile
    unsigned short* s_part = s_perm + kTile;                                                               // slot -> bin
    __shared__ unsigned int s_warp_tot[THREADS / 32];
    __shared__ uint64_t s_full[2];
    __shared__ int s_total;   // FINE: rows of the tile that were ranked (the others were placed one by one)
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const int nbins = FINE ? kRadixMaxParts : P.ndest << P.log2_parts;
    const long long ntiles = (P.n + kTile - 1) / kTile;
    auto issue = [&](long long tile, int buf) {   // thread 0
        if (tile >= ntiles) return;
        const long long row0 = tile * kTile;
        const uint32_t rows = (uint32_t)min((long long)kTile, P.n - row0);
        uint32_t total = 0;
        for (int c = 0; c < P.ncols; ++c) total += ((rows << P.shift[c]) + 15u) & ~15u;
        mbar_arrive_expect_tx(&s_full[buf], total);
        uint8_t* dst = in_tiles + (size_t)buf * tile_bytes;
        for (int c = 0; c < P.ncols; ++c)
            bulk_g2s(dst + P.stage_off[c] * kOffScale, P.src[c] + ((size_t)row0 << P.shift[c]), ((rows << P.shift[c]) + 15u) & ~15u, &s_full[buf]);
    };
    if (tid == 0) {
        mbar_init(&s_full[0], 1); mbar_init(&s_full[1], 1);
        mbar_fence_init();
        for (int b = 0; b < n_buf; ++b) issue((long long)blockIdx.x + (long long)b * gridDim.x, b);
    }
#pragma unroll
    for (int q = 0; q < BPT; ++q) { s_cnt[0][tid * BPT + q] = 0; s_cnt[1][tid * BPT + q] = 0; }
    __syncthreads();
    int it = 0;
    for (long long tile = blockIdx.x; tile < ntiles; tile += gridDim.x, ++it) {
        const int buf = n_buf == 2 ? (it & 1) : 0;
        const uint32_t phase = n_buf == 2 ? ((uint32_t)(it >> 1) & 1u) : ((uint32_t)it & 1u);
        unsigned int* cnt = s_cnt[it & 1];
        const long long row0 = tile * kTile;
        const int rows = (int)min((long long)kTile, P.n - row0);
        const uint8_t* in = in_tiles + (size_t)buf * tile_bytes;
        mbar_wait(&s_full[buf], phase);
        int bin[ROWS];
        unsigned int rank[ROWS];
        long long fbase = 0;   // FINE: first bin of the window this tile ranks
        {
            const uint8_t* kc = in + P.stage_off[P.key_col] * kOffScale;
            if constexpr (FINE) {   // the tile's first row sits in its lowest first-level partition: the window starts there
                const uint64_t raw0 = P.key_type == IGL_T_I32 ? (uint64_t)((const uint32_t*)kc)[0] : ((const uint64_t*)kc)[0];
                fbase = (long long)((mix64(part_canon(P.key_type, raw0)) >> P.fine_shift) >> P.fine_lp2) << P.fine_lp2;
            }
#pragma unroll
            for (int r = 0; r < ROWS; ++r) {
                const int i = r * THREADS + tid;
                bin[r] = -1;
                if (i < rows) {
                    const uint64_t raw = P.key_type == IGL_T_I32 ? (uint64_t)((const uint32_t*)kc)[i] : ((const uint64_t*)kc)[i];
                    const uint64_t ck = part_canon(P.key_type, raw);
                    if constexpr (FINE) {
                        const long long f = (long long)(mix64(ck) >> P.fine_shift) - fbase;
                        if (f >= 0 && f < kRadixMaxParts) {
                            bin[r] = (int)f;
                            rank[r] = atomicAdd(&cnt[bin[r]], 1u);
                        } else {   // beyond the window (tiny first-level partitions): this row is placed on its own
                            const long long fb = f + fbase;
                            const unsigned long long pos = P.offsets[fb] + atomicAdd(&P.cursor[fb], 1ull);
                            for (int c = 0; c < P.ncols; ++c) {
                                const uint8_t* sc = in + P.stage_off[c] * kOffScale;
                                if (P.shift[c] == 3) reinterpret_cast<uint64_t*>(P.dst[0][c])[pos] = ((const uint64_t*)sc)[i];
                                else reinterpret_cast<uint32_t*>(P.dst[0][c])[pos] = ((const uint32_t*)sc)[i];
                            }
                        }
                    } else {
                        bin[r] = radix_bin_of<HOT>(ck, mix64(ck), P.nhash, P.log2_parts, P.hot);
                        rank[r] = atomicAdd(&cnt[bin[r]], 1u);
                    }
                }
            }
        }
#pragma unroll
        for (int q = 0; q < BPT; ++q) s_cnt[(it + 1) & 1][tid * BPT + q] = 0;   // the other parity's counters: last read before the previous tile's final barrier
        __syncthreads();
        {   // exclusive scan of the bin counts (BPT bins per thread) + global reservation
            // (keeping the reservation atomics in flight across the permutation phase was measured: no gain)
            unsigned int c[BPT], sum = 0;
#pragma unroll
            for (int q = 0; q < BPT; ++q) { c[q] = tid * BPT + q < nbins ? cnt[tid * BPT + q] : 0; sum += c[q]; }
            unsigned int inc = sum;
#pragma unroll
            for (int o = 1; o < 32; o <<= 1) {
                const unsigned int t = __shfl_up_sync(0xffffffffu, inc, o);
                if (lane >= o) inc += t;
            }
            if (lane == 31) s_warp_tot[warp] = inc;
            __syncthreads();
            unsigned int ex = inc - sum;
            for (int w = 0; w < warp; ++w) ex += s_warp_tot[w];
#pragma unroll
            for (int q = 0; q < BPT; ++q) {
                const int bq = tid * BPT + q;
                if (bq < nbins) {
                    cnt[bq] = ex;
                    s_gbase[bq] = c[q] ? P.offsets[fbase + bq] + atomicAdd(&P.cursor[fbase + bq], (unsigned long long)c[q]) : 0ull;
                }
                ex += c[q];
            }
            if (FINE && tid == THREADS - 1) s_total = (int)ex;
        }
        __syncthreads();
#pragma unroll
        for (int r = 0; r < ROWS; ++r) {
            if (bin[r] >= 0) {
                const unsigned int slot = cnt[bin[r]] + rank[r];
                s_perm[slot] = (unsigned short)(r * THREADS + tid);
                s_part[slot] = (unsigned short)bin[r];
            }
        }
        __syncthreads();
        // write-out: consecutive threads -> consecutive slots -> runs of consecutive destination rows
        unsigned long long d[ROWS];
        int src_row[ROWS], dq[ROWS];
        const int nslots = FINE ? s_total : rows;
#pragma unroll
        for (int r = 0; r < ROWS; ++r) {
            const int sl = r * THREADS + tid;
            d[r] = 0; src_row[r] = 0; dq[r] = 0;
            if (sl < nslots) {
                const int p = s_part[sl];
                d[r] = s_gbase[p] + (unsigned)(sl - cnt[p]);
                src_row[r] = s_perm[sl];
                dq[r] = FINE ? 0 : p >> P.log2_parts;
            }
        }
        for (int c = 0; c < P.ncols; ++c) {
            if (P.shift[c] == 3) {
                const uint64_t* sc = reinterpret_cast<const uint64_t*>(in + P.stage_off[c] * kOffScale);
#pragma unroll
                for (int r = 0; r < ROWS; ++r)
                    if (r * THREADS + tid < nslots) reinterpret_cast<uint64_t*>(P.dst[dq[r]][c])[d[r]] = sc[src_row[r]];
            } else {
                const uint32_t* sc = reinterpret_cast<const uint32_t*>(in + P.stage_off[c] * kOffScale);
#pragma unroll
                for (int r = 0; r < ROWS; ++r)
                    if (r * THREADS + tid < nslots) reinterpret_cast<uint32_t*>(P.dst[dq[r]][c])[d[r]] = sc[src_row[r]];
            }
        }
        __syncthreads();   // the tile buffer, s_perm, s_part and s_gbase are free again
        if (tid == 0) issue(tile + (long long)n_buf * gridDim.x, buf);
    }
}

// Whether a column set can go through radix_partition / igl_partition_scatter_radix: fixed-width
// non-null columns whose 4096-row tile fits one of the two scatter forms.
bool radix_partition_supported(const igl_dev_col* cols, int ncols) {
    if (ncols < 1 || ncols > kPartMaxCols) return false;
    uint32_t off = 0;
    bool aligned = true;
    for (int c = 0; c < ncols; ++c) {
        const int t = cols[c].type;
        if (t != IGL_T_I32 && t != IGL_T_I64 && t != IGL_T_F64) return false;
        if (cols[c].validity) return false;
        if ((uintptr_t)cols[c].data & 15) aligned = false;
        off += (uint32_t)kRtTile << (t == IGL_T_I32 ? 2 : 3);
    }
    return off <= kRadixMaxStage || (aligned && off <= kRtMaxDyn);
}

// Launches the scatter pass for `P` (offsets and cursor already set): bulk-copy form when it applies.
// `yielding`: one CTA per tile instead of a persistent grid, so that kernels of a higher-priority stream
// get onto the SMs between tiles (the two-CTA form only: it holds no state across tiles anyway).
static cudaError_t launch_radix_scatter(igl_ctx* ctx, const RadixParams& P, uint32_t tile_bytes_4096, cudaStream_t st, bool long_runs = false,
                                        bool yielding = false) {
    bool aligned = true;
    for (int c = 0; c < P.ncols; ++c) if ((uintptr_t)P.src[c] & 15) aligned = false;
    const char* env = getenv("IGL_RADIX_BULK");
    const bool hot = P.hot.table != nullptr;
    if (hot && !(aligned && tile_bytes_4096 <= kRtMaxDyn)) return cudaErrorNotSupported;   // only the bulk form routes hot keys
    if (aligned && tile_bytes_4096 <= kRtMaxDyn && (!(env && env[0] == '0') || tile_bytes_4096 > kRadixMaxStage || hot)) {
        // 8192-row tiles (one buffer) when the caller wants the longest runs and the tile fits beside 48 KB of static arrays
        const char* lenv = getenv("IGL_RADIX_LONG");
        if (long_runs && !hot && !(lenv && lenv[0] == '0') && 2u * tile_bytes_4096 <= (176u << 10)) {
            cudaError_t e = cudaFuncSetAttribute((const void*)radix_scatter_bulk_kernel<8, false, 1024>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)((176u << 10) + 8u * kRtTile));
            if (e != cudaSuccess) return e;
            radix_scatter_bulk_kernel<8, false, 1024><<<ctx->sm_count, kRtThreads, 2u * tile_bytes_4096 + 4u * 2u * kRtTile, st>>>(P, 1, 2u * tile_bytes_4096);
            return cudaGetLastError();
        }
        // Two 512-thread CTAs per SM, each sorting its own 4096-row tile in one buffer: while one waits at
        // a barrier or for its next tile, the other runs (IGL_RADIX_CTAS2=0: one 1024-thread CTA, two buffers).
        const char* cenv = getenv("IGL_RADIX_CTAS2");
        const bool two = !(cenv && cenv[0] == '0') && tile_bytes_4096 + 4u * kRtTile <= (92u << 10);
        if (two) {
            const size_t dyn = (size_t)tile_bytes_4096 + 4u * kRtTile;
            const void* k = hot ? (const void*)radix_scatter_bulk_kernel<8, true, 512> : (const void*)radix_scatter_bulk_kernel<8, false, 512>;
            cudaError_t e = cudaFuncSetAttribute(k, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)dyn);
            if (e != cudaSuccess) return e;
            const long long ntiles = (P.n + kRtTile - 1) / kRtTile;
            const int grid = yielding ? (int)std::min<long long>(ntiles, 1ll << 30) : ctx->sm_count * 2;
            if (hot) radix_scatter_bulk_kernel<8, true, 512><<<grid, 512, dyn, st>>>(P, 1, tile_bytes_4096);
            else radix_scatter_bulk_kernel<8, false, 512><<<grid, 512, dyn, st>>>(P, 1, tile_bytes_4096);
            return cudaGetLastError();
        }
        const int n_buf = 2u * tile_bytes_4096 <= kRtMaxDyn ? 2 : 1;
        const size_t dyn = (size_t)n_buf * tile_bytes_4096;
        cudaError_t e = cudaFuncSetAttribute(hot ? (const void*)radix_scatter_bulk_kernel<4, true, 1024> : (const void*)radix_scatter_bulk_kernel<4, false, 1024>,
                                             cudaFuncAttributeMaxDynamicSharedMemorySize, (int)(kRtMaxDyn + 4u * kRtTile));
        if (e != cudaSuccess) return e;
        if (hot) radix_scatter_bulk_kernel<4, true, 1024><<<ctx->sm_count, kRtThreads, dyn + 4u * kRtTile, st>>>(P, n_buf, tile_bytes_4096);
        else radix_scatter_bulk_kernel<4, false, 1024><<<ctx->sm_count, kRtThreads, dyn + 4u * kRtTile, st>>>(P, n_buf, tile_bytes_4096);
        return cudaGetLastError();
    }
    cudaError_t e = cudaFuncSetAttribute((const void*)radix_scatter_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)kRadixMaxStage);
    if (e != cudaSuccess) return e;
    static_assert(kRadixTile == kRtTile, "both scatter forms stage 4096-row tiles");
    radix_scatter_kernel<<<ctx->sm_count * IGL_RADIX_CTAS, kRadixThreads, tile_bytes_4096, st>>>(P);
    return cudaGetLastError();
}

// Partitions `cols` (no validity) into dst[c] (capacity nrows each) by the top
// log2_parts bits of the join hash.  `scratch` holds 3 * (nparts + 1) u64.
int32_t radix_partition(igl_ctx* ctx, const igl_dev_col* cols, int ncols, int key_col, long long nrows, int log2_parts,
                        void* const* dst, void* scratch, cudaStream_t st) {
    if (ncols < 1 || ncols > kPartMaxCols || log2_parts < 1 || (1 << log2_parts) > kRadixMaxParts)
        return fail(IGL_ERR_INVALID_ARG, "radix_partition: bad arguments");
    const int nparts = 1 << log2_parts;
    unsigned long long* hist = (unsigned long long*)scratch;
    unsigned long long* offsets = hist + (nparts + 1);
    unsigned long long* cursor = offsets + (nparts + 1);
    IGL_CUDA_OK(cudaMemsetAsync(hist, 0, sizeof(unsigned long long) * (nparts + 1), st));
    RadixParams P{};
    P.ncols = ncols; P.key_col = key_col; P.key_type = cols[key_col].type; P.log2_parts = log2_parts; P.ndest = 1; P.nhash = 1; P.n = nrows;
    P.offsets = offsets; P.cursor = cursor;
    uint32_t off = 0;
    for (int c = 0; c < ncols; ++c) {
        const int t = cols[c].type;
        if (t != IGL_T_I32 && t != IGL_T_I64 && t != IGL_T_F64) return fail(IGL_ERR_UNSUPPORTED, "radix_partition: fixed-width columns only");
        if (cols[c].validity) return fail(IGL_ERR_UNSUPPORTED, "radix_partition: nullable columns are not partitioned");
        P.src[c] = (const uint8_t*)cols[c].data; P.dst[0][c] = (uint8_t*)dst[c];
        P.shift[c] = t == IGL_T_I32 ? 2 : 3;
        P.stage_off[c] = off;
        off += (uint32_t)kRadixTile << P.shift[c];
    }
    if (!radix_partition_supported(cols, ncols)) return fail(IGL_ERR_UNSUPPORTED, "radix_partition: row too wide (%u bytes of staging)", off);
    if (nrows > 0) {
        HostTrace tr("partition.hist");
        radix_hist_kernel<false><<<ctx->sm_count * 4, 512, 0, st>>>(cols[key_col].data, P.key_type, nrows, 1, 1, log2_parts, hist, HotKeys{});
        IGL_CUDA_OK(cudaGetLastError());
    }
    {
        HostTrace tr("partition.scan");
        radix_scan_kernel<<<1, kRadixMaxParts, 0, st>>>(hist, offsets, cursor, nparts);
        IGL_CUDA_OK(cudaGetLastError());
    }
    ctx->launches += 2;
    if (nrows > 0) {
        HostTrace tr("partition.scatter");
        IGL_CUDA_OK(launch_radix_scatter(ctx, P, off, st));
        ctx->launches++;
    }
    return IGL_OK;
}

// ---- second-level ("fine") partition: rows already ordered by their first hash bits are ordered by more of
// them, down to the table sub-region a single CTA can build in shared memory (join.cu, join_build_fine_kernel).
__global__ void __launch_bounds__(512) radix_hist_fine_kernel(const void* __restrict__ keys, int type, long long n, int fine_shift, int fine_lp2,
                                                              unsigned long long* hist) {
    __shared__ unsigned int s_hist[kRadixMaxParts];
    const long long ntiles = (n + kRtTile - 1) / kRtTile;
    for (long long tile = blockIdx.x; tile < ntiles; tile += gridDim.x) {
        for (int p = threadIdx.x; p < kRadixMaxParts; p += 512) s_hist[p] = 0;
        __syncthreads();
        const long long row0 = tile * kRtTile;
        const uint64_t raw0 = type == IGL_T_I32 ? (uint64_t)(uint32_t)((const int32_t*)keys)[row0] : ((const uint64_t*)keys)[row0];
        const long long fbase = (long long)((mix64(part_canon(type, raw0)) >> fine_shift) >> fine_lp2) << fine_lp2;
#pragma unroll
        for (int r = 0; r < kRtTile / 512; ++r) {
            const long long i = row0 + r * 512 + threadIdx.x;
            if (i < n) {
                const uint64_t raw = type == IGL_T_I32 ? (uint64_t)(uint32_t)((const int32_t*)keys)[i] : ((const uint64_t*)keys)[i];
                const long long f = (long long)(mix64(part_canon(type, raw)) >> fine_shift) - fbase;
                if (f >= 0 && f < kRadixMaxParts) atomicAdd(&s_hist[f], 1u);
                else atomicAdd(&hist[f + fbase], 1ull);
            }
        }
        __syncthreads();
        for (int p = threadIdx.x; p < kRadixMaxParts; p += 512)
            if (s_hist[p]) atomicAdd(&hist[fbase + p], (unsigned long long)s_hist[p]);
        __syncthreads();
    }
}

// exclusive scan of up to 2^20 counters by one CTA: offsets[nbins] = total, cursor zeroed
__global__ void __launch_bounds__(1024) radix_scan_fine_kernel(const unsigned long long* __restrict__ hist, unsigned long long* __restrict__ offsets,
                                                               unsigned long long* __restrict__ cursor, long long nbins) {
    __shared__ unsigned long long s_warp[32];
    const int t = threadIdx.x, lane = t & 31, warp = t >> 5;
    const long long per = (nbins + 1023) / 1024, b0 = (long long)t * per, b1 = min(nbins, b0 + per);
    unsigned long long sum = 0;
    for (long long b = b0; b < b1; ++b) sum += hist[b];
    unsigned long long inc = sum;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
        const unsigned long long v = __shfl_up_sync(0xffffffffu, inc, o);
        if (lane >= o) inc += v;
    }
    if (lane == 31) s_warp[warp] = inc;
    __syncthreads();
    unsigned long long ex = inc - sum;
    for (int w = 0; w < warp; ++w) ex += s_warp[w];
    for (long long b = b0; b < b1; ++b) { offsets[b] = ex; cursor[b] = 0; ex += hist[b]; }
    if (t == 1023) offsets[nbins] = ex;
}

bool radix_partition_fine_supported(const igl_dev_col* cols, int ncols) {
    if (!radix_partition_supported(cols, ncols)) return false;
    uint32_t off = 0;
    for (int c = 0; c < ncols; ++c) {
        if ((uintptr_t)cols[c].data & 15) return false;
        off += (uint32_t)kRtTile << (cols[c].type == IGL_T_I32 ? 2 : 3);
    }
    return off + 4u * kRtTile <= (92u << 10);   // the two-CTA scatter form
}

// Orders `cols` (already ordered by the top `fine_bits - fine_lp2` bits of the join hash) by its top `fine_bits`
// bits into dst[c].  `scratch` holds 3 * (2^fine_bits + 1) u64: counts, offsets (kept: offsets[b] .. offsets[b + 1]
// are bin b's rows), cursors.
int32_t radix_partition_fine(igl_ctx* ctx, const igl_dev_col* cols, int ncols, int key_col, long long nrows, int fine_bits, int fine_lp2,
                             void* const* dst, void* scratch, cudaStream_t st) {
    if (fine_bits < 1 || fine_bits > 20 || fine_lp2 < 0 || fine_lp2 > fine_bits || !radix_partition_fine_supported(cols, ncols))
        return fail(IGL_ERR_INVALID_ARG, "radix_partition_fine: bad arguments");
    const long long nbins = 1ll << fine_bits;
    unsigned long long* hist = (unsigned long long*)scratch;
    unsigned long long* offsets = hist + (nbins + 1);
    unsigned long long* cursor = offsets + (nbins + 1);
    IGL_CUDA_OK(cudaMemsetAsync(hist, 0, sizeof(unsigned long long) * (nbins + 1), st));
    RadixParams P{};
    P.ncols = ncols; P.key_col = key_col; P.key_type = cols[key_col].type; P.log2_parts = 0; P.ndest = 1; P.nhash = 1; P.n = nrows;
    P.offsets = offsets; P.cursor = cursor;
    P.fine_shift = 64 - fine_bits; P.fine_lp2 = fine_lp2; P.fine_bins = nbins;
    uint32_t off = 0;
    for (int c = 0; c < ncols; ++c) {
        P.src[c] = (const uint8_t*)cols[c].data; P.dst[0][c] = (uint8_t*)dst[c];
        P.shift[c] = cols[c].type == IGL_T_I32 ? 2 : 3;
        P.stage_off[c] = off;
        off += (uint32_t)kRadixTile << P.shift[c];
    }
    if (nrows > 0) {
        radix_hist_fine_kernel<<<ctx->sm_count * 4, 512, 0, st>>>(cols[key_col].data, P.key_type, nrows, P.fine_shift, fine_lp2, hist);
        IGL_CUDA_OK(cudaGetLastError());
    }
    radix_scan_fine_kernel<<<1, 1024, 0, st>>>(hist, offsets, cursor, nbins);
    IGL_CUDA_OK(cudaGetLastError());
    ctx->launches += 2;
    if (nrows > 0) {
        const size_t dyn = (size_t)off + 4u * kRtTile;
        IGL_CUDA_OK(cudaFuncSetAttribute((const void*)radix_scatter_bulk_kernel<8, false, 512, true>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)dyn));
        radix_scatter_bulk_kernel<8, false, 512, true><<<ctx->sm_count * 2, 512, dyn, st>>>(P, 1, off);
        IGL_CUDA_OK(cudaGetLastError());
        ctx->launches++;
    }
    return IGL_OK;
}

// ---- the two halves of a (destination x partition) split, stream-ordered and without any
// host synchronisation (shuffle.cu drives them) --------------------------------------------
// dev_hist[ndest << log2_parts] is zeroed and filled.
int32_t radix_histogram_async(igl_ctx* ctx, const igl_dev_col* key, long long nrows, int ndest, int nhash, int log2_parts,
                              const HotKeys& hot, unsigned long long* dev_hist, cudaStream_t st) {
    if (ndest < 1 || ndest > kPartMaxParts || nhash < 1 || nhash > ndest || log2_parts < 0 || (ndest << log2_parts) > kRadixMaxParts)
        return fail(IGL_ERR_UNSUPPORTED, "ndest in [1, %d] and ndest << log2_parts <= %d", kPartMaxParts, kRadixMaxParts);
    if (key->type != IGL_T_I32 && key->type != IGL_T_I64 && key->type != IGL_T_F64)
        return fail(IGL_ERR_UNSUPPORTED, "partition key must be Int32/Int64/Float64");
    if (key->validity) return fail(IGL_ERR_UNSUPPORTED, "nullable join keys are not repartitioned across GPUs");
    IGL_CUDA_OK(cudaMemsetAsync(dev_hist, 0, sizeof(unsigned long long) * ((size_t)ndest << log2_parts), st));
    if (nrows == 0) return IGL_OK;
    if (hot.table) radix_hist_kernel<true><<<ctx->sm_count * 4, 512, 0, st>>>(key->data, key->type, nrows, ndest, nhash, log2_parts, dev_hist, hot);
    else radix_hist_kernel<false><<<ctx->sm_count * 4, 512, 0, st>>>(key->data, key->type, nrows, ndest, nhash, log2_parts, dev_hist, hot);
    IGL_CUDA_OK(cudaGetLastError());
    ctx->launches++;
    return IGL_OK;
}

int32_t sample_keys_async(igl_ctx* ctx, const igl_dev_col* key, long long nrows, int nsamples, unsigned long long* dev_out, cudaStream_t st) {
    if (nrows <= 0 || nsamples <= 0) return IGL_OK;
    sample_keys_kernel<<<(nsamples + 255) / 256, 256, 0, st>>>(key->data, key->type, nrows, nsamples, dev_out);
    IGL_CUDA_OK(cudaGetLastError());
    ctx->launches++;
    return IGL_OK;
}

// Rows of `cols` are written to dst[c] (one buffer per column, all bins inside it) at
// dev_offsets[bin] + arrival order; dev_cursor[bins] must be zero.  Bins are destination-major.
int32_t radix_scatter_bins_async(igl_ctx* ctx, const igl_dev_col* cols, int ncols, int key_col, long long nrows, int ndest, int nhash,
                                 int log2_parts, const HotKeys& hot, void* const* dst, const unsigned long long* dev_offsets,
                                 unsigned long long* dev_cursor, cudaStream_t st, bool yielding) {
    if (ndest < 1 || ndest > kPartMaxParts || nhash < 1 || nhash > ndest || log2_parts < 0 || (ndest << log2_parts) > kRadixMaxParts)
        return fail(IGL_ERR_UNSUPPORTED, "ndest in [1, %d] and ndest << log2_parts <= %d", kPartMaxParts, kRadixMaxParts);
    if (ncols < 1 || ncols > kPartMaxCols || key_col < 0 || key_col >= ncols) return fail(IGL_ERR_INVALID_ARG, "radix_scatter_bins: bad column set");
    if (!radix_partition_supported(cols, ncols)) return fail(IGL_ERR_UNSUPPORTED, "columns cannot be radix-partitioned (nullable, not fixed-width, or too wide)");
    if (nrows == 0) return IGL_OK;
    RadixParams P{};
    P.ncols = ncols; P.key_col = key_col; P.key_type = cols[key_col].type; P.log2_parts = log2_parts; P.ndest = ndest; P.nhash = nhash; P.hot = hot; P.n = nrows;
    P.offsets = dev_offsets; P.cursor = dev_cursor;
    uint32_t off = 0;
    for (int c = 0; c < ncols; ++c) {
        P.src[c] = (const uint8_t*)cols[c].data;
        P.shift[c] = cols[c].type == IGL_T_I32 ? 2 : 3;
        P.stage_off[c] = off;
        off += (uint32_t)kRadixTile << P.shift[c];
        for (int d = 0; d < ndest; ++d) P.dst[d][c] = (uint8_t*)dst[c];
    }
    IGL_CUDA_OK(launch_radix_scatter(ctx, P, off, st, false, yielding));
    ctx->launches++;
    return IGL_OK;
}

}  // namespace igl

using namespace igl;

extern "C" {

int32_t igl_partition_histogram(igl_ctx* ctx, const igl_dev_col* key, int64_t nrows, int32_t nparts, int64_t* dev_hist,
                                void* stream) {
    if (!ctx || !key || !dev_hist || nrows < 0) return fail(IGL_ERR_INVALID_ARG, "igl_partition_histogram: bad arguments");
    if (nparts < 1 || nparts > kPartMaxParts) return fail(IGL_ERR_UNSUPPORTED, "nparts must be in [1, %d]", kPartMaxParts);
    if (key->type != IGL_T_I32 && key->type != IGL_T_I64 && key->type != IGL_T_F64)
        return fail(IGL_ERR_UNSUPPORTED, "partition key must be Int32/Int64/Float64");
    DeviceGuard g(ctx->device);
    cudaStream_t st = pick_stream(ctx, stream);
    IGL_CUDA_OK(cudaMemsetAsync(dev_hist, 0, sizeof(int64_t) * nparts, st));
    if (nrows == 0) return IGL_OK;
    const int grid = (int)std::min<int64_t>((nrows + 255) / 256, (int64_t)ctx->sm_count * 8);
    partition_hist_kernel<<<grid, 256, 0, st>>>(key->data, key->type, key->validity, nrows, nparts, (unsigned long long*)dev_hist);
    IGL_CUDA_OK(cudaGetLastError());
    ctx->launches++;
    return IGL_OK;
}

int32_t igl_partition_scatter(igl_ctx* ctx, const igl_dev_col* cols, int32_t ncols, int32_t key_col, int64_t nrows,
                              int32_t nparts, void* const* dst, const int64_t* dst_row_offset, void* stream) {
    if (!ctx || !cols || !dst || !dst_row_offset || nrows < 0) return fail(IGL_ERR_INVALID_ARG, "igl_partition_scatter: bad arguments");
    if (nparts < 1 || nparts > kPartMaxParts) return fail(IGL_ERR_UNSUPPORTED, "nparts must be in [1, %d]", kPartMaxParts);
    if (ncols < 1 || ncols > kPartMaxCols || key_col < 0 || key_col >= ncols) return fail(IGL_ERR_INVALID_ARG, "igl_partition_scatter: bad column set");
    DeviceGuard g(ctx->device);
    cudaStream_t st = pick_stream(ctx, stream);
    ScatterParams P{};
    P.ncols = ncols; P.nparts = nparts; P.key_col = key_col; P.key_type = cols[key_col].type;
    P.key_valid = cols[key_col].validity;
    P.n = nrows;
    for (int c = 0; c < ncols; ++c) {
        const int t = cols[c].type;
        if (t != IGL_T_I32 && t != IGL_T_I64 && t != IGL_T_F64) return fail(IGL_ERR_UNSUPPORTED, "partitioned columns must be Int32/Int64/Float64");
        if (cols[c].validity && c != key_col) return fail(IGL_ERR_UNSUPPORTED, "nullable payload columns are not repartitioned yet");
        P.src[c] = (const uint8_t*)cols[c].data;
        P.shift[c] = t == IGL_T_I32 ? 2 : 3;
    }
    for (int p = 0; p < nparts; ++p)
        for (int c = 0; c < ncols; ++c) {
            if (!dst[p * ncols + c]) return fail(IGL_ERR_INVALID_ARG, "destination pointer [%d][%d] is NULL", p, c);
            P.dst[p * kPartMaxCols + c] = (uint8_t*)dst[p * ncols + c] + ((size_t)dst_row_offset[p] << P.shift[c]);
        }
    if (nrows == 0) return IGL_OK;
    void* cursor = nullptr;
    IGL_TRY(pool_alloc(ctx, sizeof(unsigned long long) * kPartMaxParts, &cursor));
    cudaError_t e = cudaMemsetAsync(cursor, 0, sizeof(unsigned long long) * kPartMaxParts, st);
    P.cursor = (unsigned long long*)cursor;
    if (e == cudaSuccess) {
        const int64_t ntiles = (nrows + kPartTile - 1) / kPartTile;
        const int grid = (int)std::min<int64_t>(ntiles, (int64_t)ctx->sm_count * 8);
        partition_scatter_kernel<<<grid, kPartThreads, 0, st>>>(P);
        e = cudaGetLastError();
        ctx->launches++;
    }
    // the cursor block returns to the pool only after the kernel has consumed it
    if (e == cudaSuccess) e = cudaStreamSynchronize(st);
    pool_free(ctx, cursor);
    if (e != cudaSuccess) return fail(IGL_ERR_CUDA, "igl_partition_scatter: %s", cudaGetErrorString(e));
    return IGL_OK;
}

// ---- exchange fused with the join's own partitioning ---------------------------------
// Bins = ndest * 2^log2_parts, destination-major.  A receiver lays its slab out partition by
// partition (inside a partition: source by source), so what arrives is already in the order
// igl_join_build / igl_join_probe want (igl_join_set_prepartitioned) and rows cross NVLink as
// runs staged in shared memory, not as single 8-byte stores.
int32_t igl_partition_histogram_radix(igl_ctx* ctx, const igl_dev_col* key, int64_t nrows, int32_t ndest, int32_t log2_parts,
                                      int64_t* dev_hist, void* stream) {
    if (!ctx || !key || !dev_hist || nrows < 0) return fail(IGL_ERR_INVALID_ARG, "igl_partition_histogram_radix: bad arguments");
    if (ndest < 1 || ndest > kPartMaxParts || log2_parts < 0 || (ndest << log2_parts) > kRadixMaxParts)
        return fail(IGL_ERR_UNSUPPORTED, "ndest in [1, %d] and ndest << log2_parts <= %d", kPartMaxParts, kRadixMaxParts);
    if (key->type != IGL_T_I32 && key->type != IGL_T_I64 && key->type != IGL_T_F64)
        return fail(IGL_ERR_UNSUPPORTED, "partition key must be Int32/Int64/Float64");
    if (key->validity) return fail(IGL_ERR_UNSUPPORTED, "nullable keys take igl_partition_histogram / igl_partition_scatter");
    DeviceGuard g(ctx->device);
    cudaStream_t st = pick_stream(ctx, stream);
    IGL_CUDA_OK(cudaMemsetAsync(dev_hist, 0, sizeof(int64_t) * ((size_t)ndest << log2_parts), st));
    if (nrows == 0) return IGL_OK;
    radix_hist_kernel<false><<<ctx->sm_count * 4, 512, 0, st>>>(key->data, key->type, nrows, ndest, ndest, log2_parts, (unsigned long long*)dev_hist, HotKeys{});
    IGL_CUDA_OK(cudaGetLastError());
    ctx->launches++;
    return IGL_OK;
}

int32_t igl_partition_radix_supported(const igl_dev_col* cols, int32_t ncols) {
    return cols && radix_partition_supported(cols, ncols) ? 1 : 0;
}

int32_t igl_partition_scatter_radix(igl_ctx* ctx, const igl_dev_col* cols, int32_t ncols, int32_t key_col, int64_t nrows,
                                    int32_t ndest, int32_t log2_parts, void* const* dst, const int64_t* bin_row_offset,
                                    void* stream) {
    if (!ctx || !cols || !dst || !bin_row_offset || nrows < 0) return fail(IGL_ERR_INVALID_ARG, "igl_partition_scatter_radix: bad arguments");
    if (ndest < 1 || ndest > kPartMaxParts || log2_parts < 0 || (ndest << log2_parts) > kRadixMaxParts)
        return fail(IGL_ERR_UNSUPPORTED, "ndest in [1, %d] and ndest << log2_parts <= %d", kPartMaxParts, kRadixMaxParts);
    if (ncols < 1 || ncols > kPartMaxCols || key_col < 0 || key_col >= ncols) return fail(IGL_ERR_INVALID_ARG, "igl_partition_scatter_radix: bad column set");
    DeviceGuard g(ctx->device);
    cudaStream_t st = pick_stream(ctx, stream);
    const int bins = ndest << log2_parts;
    RadixParams P{};
    P.ncols = ncols; P.key_col = key_col; P.key_type = cols[key_col].type; P.log2_parts = log2_parts; P.ndest = ndest; P.nhash = ndest; P.n = nrows;
    uint32_t off = 0;
    for (int c = 0; c < ncols; ++c) {
        const int t = cols[c].type;
        if (t != IGL_T_I32 && t != IGL_T_I64 && t != IGL_T_F64) return fail(IGL_ERR_UNSUPPORTED, "partitioned columns must be Int32/Int64/Float64");
        if (cols[c].validity) return fail(IGL_ERR_UNSUPPORTED, "nullable columns take igl_partition_scatter");
        P.src[c] = (const uint8_t*)cols[c].data;
        P.shift[c] = t == IGL_T_I32 ? 2 : 3;
        P.stage_off[c] = off;
        off += (uint32_t)kRadixTile << P.shift[c];
        for (int d = 0; d < ndest; ++d) {
            if (!dst[d * ncols + c]) return fail(IGL_ERR_INVALID_ARG, "destination pointer [%d][%d] is NULL", d, c);
            P.dst[d][c] = (uint8_t*)dst[d * ncols + c];
        }
    }
    if (!radix_partition_supported(cols, ncols)) return fail(IGL_ERR_UNSUPPORTED, "row too wide (%u bytes of staging)", off);
    if (nrows == 0) return IGL_OK;
    // device copy of the offsets + zeroed cursors; handed back to the pool once the kernel is done
    void* scratch = nullptr;
    IGL_TRY(pool_alloc(ctx, sizeof(unsigned long long) * 2 * kRadixMaxParts, &scratch));
    unsigned long long* offsets = (unsigned long long*)scratch;
    unsigned long long* cursor = offsets + kRadixMaxParts;
    cudaError_t e = cudaMemcpyAsync(offsets, bin_row_offset, sizeof(int64_t) * bins, cudaMemcpyHostToDevice, st);
    if (e == cudaSuccess) e = cudaMemsetAsync(cursor, 0, sizeof(unsigned long long) * bins, st);
    P.offsets = offsets; P.cursor = cursor;
    if (e == cudaSuccess) {
        e = launch_radix_scatter(ctx, P, off, st, /*long_runs=*/ndest > 1);   // run length is what NVLink efficiency hangs on
        ctx->launches++;
    }
    if (e == cudaSuccess) e = cudaStreamSynchronize(st);
    pool_free(ctx, scratch);
    if (e != cudaSuccess) return fail(IGL_ERR_CUDA, "igl_partition_scatter_radix: %s", cudaGetErrorString(e));
    return IGL_OK;
}

}  // extern "C"
