// join.cu — K5 build / K6 probe / K7 materialise for the hash join on
// device-resident columns.
//
// Reference semantics implemented (crates/engine/src/operators/hash_join.rs):
//   :35-76    output schema = left fields ++ right fields
//   :100-128  build: the right child is drained into the table, duplicates kept
//   :131-213  probe: per left row, one output row per stored right row with the
//             same key; Left/Full: an unmatched left row is emitted once with a
//             null per right field; Right == Inner, Full == Left (no unmatched
//             right rows are ever produced, :189, :263-276)
//   :117-122  key identity is the printed form: NULL == NULL, NaN == NaN,
//             -0.0 != 0.0, keys of different Arrow types never match.
// The reference keys a std HashMap by a heap string per row and emits 1-row
// batches; here the key is its canonical 64-bit identity and rows are emitted in
// bulk.
//
// Table: open addressing, linear probing, 16-byte slots {key, val}, claimed with
// atomicCAS on the key word.  Two layouts, chosen by what the build side holds:
//   unique keys (the optimistic first pass; any duplicate raises a flag and the
//   build is redone)   val = the 8-byte payload itself when the right side is
//                      (key, one fixed-width non-null column), else the row id
//   duplicate keys     val = {count, run end}: a CSR of build row ids grouped by
//                      key (count pass -> run allocation -> fill pass), so a
//                      heavy key's matches are one contiguous run that a whole
//                      warp walks together.
// NULL keys and the one key value that equals the empty-slot sentinel live in
// two side entries.
//
// Large build sides are ordered by the top bits of the join hash (partition.cu) so that
// the table is touched region by region while L2-resident; large UNIQUE ones are ordered
// once more and built in shared memory, one CTA per 64 KB of table (join_build_fine_kernel).
//
// Probe kernels.  General path (nulls, Left/Full, duplicate keys): the tile pipeline
// (tile_pipe.cuh) — left column tiles staged by bulk async copies, each lane looks up 4 keys
// with all first-slot loads in flight together, output offsets from the ordered look-back
// scan, output rows (left columns from shared memory ++ right payload) written once, to
// their final place.  Fast path (inner, unique keys, no nulls): join_probe_flat_kernel, one
// warp per CTA, matched rows written at their input position; its SEG form reads the probe
// rows from runs that may sit in other GPUs' memory (shuffle.cu).
#include "internal.h"
#include "kernels.h"
#include "tile_pipe.cuh"

#include <algorithm>
#include <cstdlib>
#include <cstring>
#include <memory>

namespace igl {

constexpr uint64_t kEmptyKey = 0x8000000000000000ull;
constexpr uint32_t kNoRow = 0xFFFFFFFFu;
constexpr int kJoinMaxCols = 16;

struct Slot {
    unsigned long long key;
    unsigned long long val;
};

struct JoinSide {          // side entries: [0] NULL key, [1] key == kEmptyKey
    unsigned int count[2];
    unsigned int run_end[2];
    unsigned long long val[2];
    unsigned int dup_flag;
    unsigned int pad;
    unsigned long long cursor;   // CSR run allocator
    unsigned int ticket;         // build kernels: chunk claims
    unsigned int pad2;
    unsigned long long ovf_count;   // shared-memory build: rows that ran past their sub-region's end
};

struct JoinTable {
    Slot* slots;
    uint64_t mask;
    int shift;               // 64 - log2(capacity)
    JoinSide* side;
    const uint32_t* rowids;  // CSR runs (duplicate layout)
};

__device__ __forceinline__ uint64_t canon_key(int type, uint64_t raw) {
    switch (type) {
        case IGL_T_I32: return (uint64_t)(int64_t)(int32_t)raw;
        case IGL_T_F64: {
            // every NaN prints as "NaN"; everything else round-trips its bits
            const bool nan = (raw & 0x7FFFFFFFFFFFFFFFull) > 0x7FF0000000000000ull;
            return nan ? 0x7FF8000000000000ull : raw;
        }
        case IGL_T_BOOL: return raw & 1;
        default: return raw;
    }
}

__device__ __forceinline__ uint64_t load_key_global(const void* keys, int type, long long i) {
    if (type == IGL_T_I32) return (uint32_t)((const int32_t*)keys)[i];
    if (type == IGL_T_BOOL) return (((const uint8_t*)keys)[i >> 3] >> (i & 7)) & 1;
    return ((const uint64_t*)keys)[i];
}

__device__ __forceinline__ bool bit_at(const uint8_t* bm, long long i) { return (bm[i >> 3] >> (i & 7)) & 1; }

// ---- table init ---------------------------------------------------------------------
__global__ void join_init_kernel(Slot* slots, uint64_t capacity, JoinSide* side) {
    const uint64_t stride = (uint64_t)gridDim.x * blockDim.x;
    ulonglong2 e; e.x = kEmptyKey; e.y = 0;
    for (uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x; i < capacity; i += stride)
        *reinterpret_cast<ulonglong2*>(slots + i) = e;
    if (blockIdx.x == 0 && threadIdx.x == 0) {
        side->count[0] = side->count[1] = 0; side->run_end[0] = side->run_end[1] = 0;
        side->val[0] = side->val[1] = 0; side->dup_flag = 0; side->cursor = 0; side->ticket = 0; side->ovf_count = 0;
    }
}

// ---- build in shared memory -------------------------------------------------------------------------
// Random 16-byte writes into an L2-resident table region — atomic or not — run at 28-37 G/s on B200
// (tools/micro/atomic_bench.cu: CAS 36.8, CAS + payload store 29.5, plain 16-byte stores 27.8 G/s), which
// is what join_build_unique_kernel gets (34 G inserts/s).  So for large unique build sides the rows are
// ordered once more (partition.cu, radix_partition_fine) down to sub-regions of kFineSlots slots, and one
// CTA builds each sub-region in shared memory and writes it out as 64 KB of consecutive stores — which
// also clears the empty slots, so no init pass.  A row whose walk runs past the sub-region's end (rare)
// goes to a short list and is inserted with global atomics afterwards.
constexpr int kFineSlotsLog2 = 12;
constexpr int kFineSlots = 1 << kFineSlotsLog2;   // 64 KB of table per CTA, three CTAs per SM
constexpr int kFineThreads = 1024;
constexpr int kFineU = 2;

__global__ void __launch_bounds__(kFineThreads, 2) join_build_fine_kernel(const void* __restrict__ keys, int key_type, const void* __restrict__ payload,
                                                                       int payload_shift, const unsigned long long* __restrict__ bin_offsets,
                                                                       JoinTable t, Slot* __restrict__ ovf, unsigned long long ovf_cap) {
    extern __shared__ __align__(16) uint8_t fine_smem[];
    Slot* tab = reinterpret_cast<Slot*>(fine_smem);
    const unsigned long long slot0 = (unsigned long long)blockIdx.x << kFineSlotsLog2;
    {
        ulonglong2 e; e.x = kEmptyKey; e.y = 0;
        for (int q = threadIdx.x; q < kFineSlots; q += kFineThreads) reinterpret_cast<ulonglong2*>(tab)[q] = e;
    }
    __syncthreads();
    const long long r0 = (long long)bin_offsets[blockIdx.x], r1 = (long long)bin_offsets[blockIdx.x + 1];
    for (long long base = r0; base < r1; base += kFineThreads * kFineU) {
        uint64_t k[kFineU];
        unsigned long long val[kFineU];
#pragma unroll
        for (int u = 0; u < kFineU; ++u) {   // all loads first
            const long long i = base + u * kFineThreads + threadIdx.x;
            k[u] = kEmptyKey; val[u] = 0;
            if (i < r1) {
                val[u] = payload_shift == 0xFF ? (unsigned long long)i
                         : payload_shift == 2 ? (unsigned long long)((const uint32_t*)payload)[i]
                                              : ((const unsigned long long*)payload)[i];
                k[u] = canon_key(key_type, load_key_global(keys, key_type, i));
                if (k[u] == kEmptyKey) {   // the one key equal to the empty-slot sentinel lives in a side entry
                    if (atomicAdd(&t.side->count[1], 1u) == 0) t.side->val[1] = val[u];
                    else t.side->dup_flag = 1;
                }
            }
        }
#pragma unroll
        for (int u = 0; u < kFineU; ++u) {
            if (k[u] == kEmptyKey) continue;
            unsigned long long h = ((mix64(k[u]) >> t.shift) & ~1ull) - slot0;   // < kFineSlots: that is what the fine order means
            bool placed = false;
            for (; h < (unsigned long long)kFineSlots; ++h) {
                const unsigned long long prev = atomicCAS(reinterpret_cast<unsigned long long*>(&tab[h].key), (unsigned long long)kEmptyKey, (unsigned long long)k[u]);
                if (prev == kEmptyKey) { tab[h].val = val[u]; placed = true; break; }
                if (prev == k[u]) { t.side->dup_flag = 1; placed = true; break; }
            }
            if (!placed) {
                const unsigned long long q = atomicAdd(&t.side->ovf_count, 1ull);
                if (q < ovf_cap) { ovf[q].key = k[u]; ovf[q].val = val[u]; }
            }
        }
    }
    __syncthreads();
    ulonglong2* g = reinterpret_cast<ulonglong2*>(t.slots + slot0);
    for (int q = threadIdx.x; q < kFineSlots; q += kFineThreads) g[q] = reinterpret_cast<const ulonglong2*>(tab)[q];
}

// the rows join_build_fine_kernel could not place: ordinary linear probing over the finished table
__global__ void join_build_overflow_kernel(const Slot* __restrict__ ovf, unsigned long long ovf_cap, JoinTable t) {
    const unsigned long long n = min(t.side->ovf_count, ovf_cap);
    const unsigned long long stride = (unsigned long long)gridDim.x * blockDim.x;
    for (unsigned long long i = (unsigned long long)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += stride) {
        const unsigned long long k = ovf[i].key;
        uint64_t h = (mix64(k) >> t.shift) & ~1ull;
        while (true) {
            const unsigned long long prev = atomicCAS(&t.slots[h].key, (unsigned long long)kEmptyKey, k);
            if (prev == kEmptyKey) { t.slots[h].val = ovf[i].val; break; }
            if (prev == k) { t.side->dup_flag = 1; break; }
            h = (h + 1) & t.mask;
        }
    }
}

// IGL_BUILD_CAS128=1: claim the slot and write the payload with ONE 128-bit compare-and-swap instead
// of a 64-bit CAS on the key followed by a store of the payload.  Measured at C3 (128 Mi inserts): 6.04 ms
// against 3.95 ms for CAS + store — the 128-bit atomic is the slower L2 operation; kept for the record.
#ifndef IGL_BUILD_CAS128
#define IGL_BUILD_CAS128 0
#endif
// 128-bit compare-and-swap on a 16-byte-aligned slot (SASS ATOMG.E.CAS.128); returns the old value.
__device__ __forceinline__ void cas_u128(void* addr, unsigned long long c0, unsigned long long c1, unsigned long long s0,
                                         unsigned long long s1, unsigned long long& o0, unsigned long long& o1) {
    asm volatile(
        "{\n\t.reg .b128 cmp, swp, old;\n\tmov.b128 cmp, {%2, %3};\n\tmov.b128 swp, {%4, %5};\n\t"
        "atom.global.cas.b128 old, [%6], cmp, swp;\n\tmov.b128 {%0, %1}, old;\n\t}"
        : "=l"(o0), "=l"(o1) : "l"(c0), "l"(c1), "l"(s0), "l"(s1), "l"(addr) : "memory");
}

// ---- K5 build, unique layout ---------------------------------------------------------
// payload_shift: 2 / 3 = width of the inlined payload column, 0xFF = store the row id.
constexpr int kBuildThreads = 256;
#ifndef IGL_BUILD_U
#define IGL_BUILD_U 4
#endif
constexpr int kBuildU = IGL_BUILD_U;                                   // rows per thread per chunk: 4 independent atomics in flight
constexpr int kBuildChunk = kBuildThreads * kBuildU;

// Chunks are claimed from an atomic ticket (not a static grid stride): all CTAs
// then work on one moving frontier of the input, so when the build side has been
// radix-partitioned the table region being filled stays L2-resident.
__global__ void __launch_bounds__(kBuildThreads) join_build_unique_kernel(const void* __restrict__ keys, int key_type,
                                                                           const uint8_t* __restrict__ key_valid,
                                                                           const void* __restrict__ payload, int payload_shift,
                                                                           long long n, JoinTable t, unsigned int* ticket) {
    constexpr int U = kBuildU;
    __shared__ long long s_chunk[2];
    const long long nchunks = (n + kBuildChunk - 1) / kBuildChunk;
    // the ticket for the NEXT chunk is drawn while this one is inserted: its L2 round trip is hidden
    if (threadIdx.x == 0) s_chunk[0] = (long long)atomicAdd(ticket, 1u);
    __syncthreads();
    for (int it = 0;; ++it) {
        const long long chunk = s_chunk[it & 1];
        if (chunk >= nchunks) break;
        if (threadIdx.x == 0) {
            long long nxt = (long long)atomicAdd(ticket, 1u);
            // a duplicate has been seen: the build will be redone as CSR runs, so stop claiming work —
            // every further insert of a heavy key would be one more atomic on the same slot (C5b: 29 M
            // of them).  Read beside the ticket, one chunk ahead: its latency hides behind the inserts.
            if (*reinterpret_cast<volatile unsigned int*>(&t.side->dup_flag)) nxt = nchunks;
            s_chunk[(it + 1) & 1] = nxt;
        }
        const long long i0 = chunk * kBuildChunk + threadIdx.x;
        uint64_t k[U], h[U];
        unsigned long long val[U];
        uint32_t pend = 0;
#pragma unroll
        for (int u = 0; u < U; ++u) {
            const long long i = i0 + u * kBuildThreads;
            if (i >= n) continue;
            val[u] = payload_shift == 0xFF ? (unsigned long long)i
                     : payload_shift == 2 ? (unsigned long long)((const uint32_t*)payload)[i]
                                          : ((const unsigned long long*)payload)[i];
            const bool valid = !key_valid || bit_at(key_valid, i);
            k[u] = valid ? canon_key(key_type, load_key_global(keys, key_type, i)) : 0;
            if (!valid || k[u] == kEmptyKey) {
                const int sidx = valid ? 1 : 0;
                if (atomicAdd(&t.side->count[sidx], 1u) == 0) t.side->val[sidx] = val[u];
                else t.side->dup_flag = 1;
                continue;
            }
            h[u] = (mix64(k[u]) >> t.shift) & ~1ull;   // homes are even: a bucket is one 32-byte sector
            pend |= 1u << u;
        }
        while (pend) {
            unsigned long long prev[U];
#if IGL_BUILD_CAS128
            // an empty slot is always {kEmptyKey, 0} (join_init_kernel), so the swap succeeds exactly on empty slots
#pragma unroll
            for (int u = 0; u < U; ++u) {
                unsigned long long pv;
                if ((pend >> u) & 1u) cas_u128(&t.slots[h[u]], (unsigned long long)kEmptyKey, 0ull, (unsigned long long)k[u], val[u], prev[u], pv);
            }
#pragma unroll
            for (int u = 0; u < U; ++u) {
                if (!((pend >> u) & 1u)) continue;
                if (prev[u] == kEmptyKey) pend &= ~(1u << u);
                else if (prev[u] == k[u]) { t.side->dup_flag = 1; pend &= ~(1u << u); }
                else h[u] = (h[u] + 1) & t.mask;
            }
#else
#pragma unroll
            for (int u = 0; u < U; ++u)
                if ((pend >> u) & 1u) prev[u] = atomicCAS(&t.slots[h[u]].key, (unsigned long long)kEmptyKey, (unsigned long long)k[u]);
#pragma unroll
            for (int u = 0; u < U; ++u) {
                if (!((pend >> u) & 1u)) continue;
                if (prev[u] == kEmptyKey) { t.slots[h[u]].val = val[u]; pend &= ~(1u << u); }
                else if (prev[u] == k[u]) { t.side->dup_flag = 1; pend &= ~(1u << u); }
                else h[u] = (h[u] + 1) & t.mask;
            }
#endif
        }
        __syncthreads();   // s_chunk[(it + 1) & 1] is written; s_chunk[it & 1] may be overwritten next round
    }
}

// ---- K5 build, duplicate layout: count, allocate runs, fill --------------------------------
// Lanes of a warp that hold the same key act as one: the first of them walks to the key's slot
// and adds the whole group's count.  With skewed build keys (C5b: one key owns a fifth of the
// rows) this cuts the atomics on the hot slot 32-fold.
__global__ void join_build_count_kernel(const void* __restrict__ keys, int key_type, const uint8_t* __restrict__ key_valid,
                                        long long n, JoinTable t) {
    const long long stride = (long long)gridDim.x * blockDim.x;
    const int lane = threadIdx.x & 31;
    for (long long i0 = (long long)blockIdx.x * blockDim.x + threadIdx.x - lane; i0 < n; i0 += stride) {   // warp-uniform trip count
        const long long i = i0 + lane;
        const bool in = i < n;
        const bool valid = in && (!key_valid || bit_at(key_valid, i));
        const uint64_t k = valid ? canon_key(key_type, load_key_global(keys, key_type, i)) : kEmptyKey;
        const bool normal = valid && k != kEmptyKey;
        if (in && !normal) atomicAdd(&t.side->count[valid ? 1 : 0], 1u);
        const unsigned peers = __match_any_sync(0xffffffffu, k);   // side-entry lanes all carry kEmptyKey and skip below
        if (normal && lane == __ffs(peers) - 1) {
            uint64_t h = (mix64(k) >> t.shift) & ~1ull;   // homes are even: a bucket is one 32-byte sector
            while (true) {
                const unsigned long long prev = atomicCAS(&t.slots[h].key, (unsigned long long)kEmptyKey, (unsigned long long)k);
                if (prev == kEmptyKey || prev == k) { atomicAdd(&t.slots[h].val, (unsigned long long)__popc(peers) << 32); break; }
                h = (h + 1) & t.mask;
            }
        }
    }
}

__global__ void join_build_runs_kernel(JoinTable t, uint64_t capacity) {
    const uint64_t stride = (uint64_t)gridDim.x * blockDim.x;
    const int lane = threadIdx.x & 31;
    for (uint64_t i0 = ((uint64_t)blockIdx.x * blockDim.x + threadIdx.x) - lane; i0 < capacity; i0 += stride) {
        const uint64_t i = i0 + lane;
        unsigned long long cnt = 0;
        if (i < capacity && t.slots[i].key != kEmptyKey) cnt = t.slots[i].val >> 32;
        // one atomic per warp: warp-scan the counts, lane 31 reserves the warp's block of runs
        unsigned long long inc = cnt;
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) {
            const unsigned long long nb = __shfl_up_sync(0xffffffffu, inc, o);
            if (lane >= o) inc += nb;
        }
        unsigned long long base = 0;
        if (lane == 31 && inc) base = atomicAdd(&t.side->cursor, inc);
        base = __shfl_sync(0xffffffffu, base, 31);
        if (cnt) t.slots[i].val = (cnt << 32) | (base + inc - cnt);   // low word: run start, advanced by the fill pass
    }
    if (blockIdx.x == 0 && threadIdx.x == 0) {
        for (int s = 0; s < 2; ++s)
            if (t.side->count[s]) t.side->run_end[s] = (unsigned int)atomicAdd(&t.side->cursor, (unsigned long long)t.side->count[s]);
    }
}

__global__ void join_build_fill_kernel(const void* __restrict__ keys, int key_type, const uint8_t* __restrict__ key_valid,
                                       long long n, JoinTable t, uint32_t* __restrict__ rowids) {
    const long long stride = (long long)gridDim.x * blockDim.x;
    const int lane = threadIdx.x & 31;
    for (long long i0 = (long long)blockIdx.x * blockDim.x + threadIdx.x - lane; i0 < n; i0 += stride) {   // warp-uniform trip count
        const long long i = i0 + lane;
        const bool in = i < n;
        const bool valid = in && (!key_valid || bit_at(key_valid, i));
        const uint64_t k = valid ? canon_key(key_type, load_key_global(keys, key_type, i)) : kEmptyKey;
        const bool normal = valid && k != kEmptyKey;
        const unsigned peers = __match_any_sync(0xffffffffu, k);
        const int leader = __ffs(peers) - 1;
        unsigned int base = 0;
        if (normal && lane == leader) {   // one reservation in the key's run for the whole group
            uint64_t h = (mix64(k) >> t.shift) & ~1ull;   // homes are even: a bucket is one 32-byte sector
            while (t.slots[h].key != k) h = (h + 1) & t.mask;
            base = atomicAdd(reinterpret_cast<unsigned int*>(&t.slots[h].val), (unsigned int)__popc(peers));   // little endian: the low word
        }
        base = __shfl_sync(0xffffffffu, base, leader);
        unsigned int pos = base + __popc(peers & ((1u << lane) - 1u));
        if (in && !normal) pos = atomicAdd(&t.side->run_end[valid ? 1 : 0], 1u);
        if (in) rowids[pos] = (uint32_t)i;   // after this pass the low word / run_end is the END of the run
    }
}

// The build columns in CSR order: dst[pos] = src[rowids[pos]].  Afterwards the row ids are the
// identity, so a probe row's matches (one run) are read from consecutive rows of the build columns
// instead of one random 8-byte gather per output row and column.
__global__ void gather_rows_kernel(const uint8_t* __restrict__ src, uint8_t* __restrict__ dst, const uint32_t* __restrict__ rowids,
                                   long long n, int shift) {
    const long long stride = (long long)gridDim.x * blockDim.x;
    for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += stride) {
        const uint32_t r = rowids[i];
        if (shift == 2) ((uint32_t*)dst)[i] = ((const uint32_t*)src)[r];
        else ((uint64_t*)dst)[i] = ((const uint64_t*)src)[r];
    }
}
__global__ void iota_u32_kernel(uint32_t* out, long long n) {
    const long long stride = (long long)gridDim.x * blockDim.x;
    for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += stride) out[i] = (uint32_t)i;
}

// Build rows as records: column c of row i at out + i * stride + off[c].
struct PackRowsParams {
    const uint8_t* src[kJoinMaxCols];
    uint8_t off[kJoinMaxCols], shift[kJoinMaxCols];
    int ncols, stride;
    long long n;
    uint8_t* out;
};
__global__ void pack_rows_kernel(const __grid_constant__ PackRowsParams P) {
    const long long strideg = (long long)gridDim.x * blockDim.x;
    for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < P.n; i += strideg) {
        // the record is assembled in registers and leaves as 16-byte stores (a store per column would
        // touch 32 sectors per warp instruction for 4 or 8 useful bytes each)
        unsigned long long w0 = 0, w1 = 0, w2 = 0, w3 = 0;
        for (int c = 0; c < P.ncols; ++c) {
            const unsigned long long v = P.shift[c] == 3 ? ((const unsigned long long*)P.src[c])[i]
                                                         : (unsigned long long)((const uint32_t*)P.src[c])[i] << ((P.off[c] & 4) * 8);
            const int wi = P.off[c] >> 3;
            if (wi == 0) w0 |= v; else if (wi == 1) w1 |= v; else if (wi == 2) w2 |= v; else w3 |= v;
        }
        ulonglong2* rec = reinterpret_cast<ulonglong2*>(P.out + (size_t)i * P.stride);
        rec[0] = make_ulonglong2(w0, w1);
        if (P.stride == 32) rec[1] = make_ulonglong2(w2, w3);
    }
}

// ---- probe ---------------------------------------------------------------------------------
struct JoinProbeParams {
    PipeCols T;                       // every left column, staged
    JoinTable table;
    int key_slot, key_type;           // left key column
    int never_match;                  // left / right key types differ: nothing can match
    int n_left, n_right;
    int right_key_col;                // index among the right columns
    int payload_col;                  // unique layout with inlined payload: that right column; else -1
    int left_join;
    void* out_data[2 * kJoinMaxCols];
    uint8_t* out_valid[2 * kJoinMaxCols];   // byte-per-row validity planes (packed afterwards) or NULL
    uint8_t out_shift[2 * kJoinMaxCols];
    const uint8_t* right_data[kJoinMaxCols];
    const uint8_t* right_valid[kJoinMaxCols];
    // duplicate layout: runs longer than kListRun are not walked by the probing warp (one warp
    // would emit millions of rows alone) but listed here as {left row, run end, count, output
    // position} and expanded afterwards by the whole grid (join_expand_giant_kernel)
    unsigned long long* giant_list;   // runs of kListRun+1 .. kHugeRun rows: expanded one CTA per run
    unsigned long long* giant_count;
    unsigned int giant_cap;
    unsigned long long* huge_list;    // runs longer than kHugeRun: expanded by the whole grid, one after another
    unsigned long long* huge_count;
    unsigned int huge_cap;
};

// A probing warp walks a run of up to kListRun rows itself (32 rows per step).  Longer runs would hold
// its tile — and, through the ordered output scan, every tile behind it — for as long as one warp
// needs to write them (measured at C5b/8: 120 ms per join with an 8192-row limit), so they are listed
// and expanded after the probe kernel: a CTA per listed run, the whole grid per run beyond kHugeRun.
constexpr uint32_t kListRun = 256;
constexpr uint32_t kHugeRun = 65536;

struct Match {
    unsigned long long val;   // unique: payload or row id;  duplicate: {count, run end}
    bool hit;
};

// Looks up this lane's R keys.  The table is read as 32-byte buckets of two slots
// (one sector, one request); homes are even slots, so a key sits in its home
// bucket or a following one.  All first-bucket loads are issued before any is
// used, and collisions are resolved in warp-wide rounds: a round reloads every
// still-pending row at once, so the cost is the longest probe chain in the warp,
// not the sum of the chains.
template <int R, bool HAS_NULLS>
__device__ __forceinline__ void lookup_rows(const JoinProbeParams& P, const uint8_t* stage, int row0, int rows_in_tile,
                                            Match (&m)[R]) {
    const uint8_t* kb = stage + P.T.smem_off[P.key_slot];
    const int ksh = P.T.shift[P.key_slot];
    const uint64_t bmask = P.table.mask >> 1;
    uint64_t k[R], b[R];
    ulonglong4 q[R];
    uint32_t pend = 0;
#pragma unroll
    for (int r = 0; r < R; ++r) {
        const int row = row0 + r * 32;
        const uint64_t raw = ksh == 2 ? (uint64_t) * reinterpret_cast<const uint32_t*>(kb + (row << 2))
                             : ksh == 3 ? *reinterpret_cast<const uint64_t*>(kb + (row << 3))
                                        : (uint64_t)((kb[row >> 3] >> (row & 7)) & 1);
        bool valid = true;
        if (HAS_NULLS && P.T.valid[P.key_slot]) valid = bit_at(stage + P.T.valid_off[P.key_slot], row);
        k[r] = canon_key(P.key_type, raw);
        const int kind = (row >= rows_in_tile || P.never_match) ? 2 : (!valid ? 1 : (k[r] == kEmptyKey ? 1 : 0));
        m[r].hit = false; m[r].val = 0;
        if (kind == 1) {   // NULL key / the key that equals the empty-slot sentinel: side entries
            const int sidx = valid ? 1 : 0;
            if (P.table.side->count[sidx]) {
                m[r].hit = true;
                m[r].val = P.table.rowids ? (((unsigned long long)P.table.side->count[sidx] << 32) | P.table.side->run_end[sidx])
                                          : P.table.side->val[sidx];
            }
        }
        b[r] = (mix64(k[r]) >> P.table.shift) >> 1;
        if (kind == 0) {
            q[r] = ld_nc_u64x4(P.table.slots + 2 * b[r]);
            pend |= 1u << r;
        }
    }
    while (__any_sync(0xffffffffu, pend != 0)) {
#pragma unroll
        for (int r = 0; r < R; ++r) {
            if (!((pend >> r) & 1u)) continue;
            if (q[r].x == k[r]) { m[r].hit = true; m[r].val = q[r].y; pend &= ~(1u << r); }
            else if (q[r].x == kEmptyKey) pend &= ~(1u << r);
            else if (q[r].z == k[r]) { m[r].hit = true; m[r].val = q[r].w; pend &= ~(1u << r); }
            else if (q[r].z == kEmptyKey) pend &= ~(1u << r);
            else {
                b[r] = (b[r] + 1) & bmask;
                q[r] = ld_nc_u64x4(P.table.slots + 2 * b[r]);
            }
        }
    }
}

// Writes one output row: left columns from the staged tile, right columns from
// the inlined payload (unique layout) or gathered by build row id.
template <bool HAS_NULLS>
__device__ __forceinline__ void emit_row(const JoinProbeParams& P, const uint8_t* stage, int row, long long pos,
                                         bool matched, unsigned long long val, uint64_t raw_key) {
    for (int c = 0; c < P.n_left; ++c) {
        if (!P.out_data[c]) continue;   // hidden column (a computed / combined key): joined on, not emitted
        const uint8_t* src = stage + P.T.smem_off[c];
        if (P.out_shift[c] == 2) st_cs_u32((uint8_t*)P.out_data[c] + (pos << 2), *reinterpret_cast<const uint32_t*>(src + (row << 2)));
        else st_cs_u64((uint8_t*)P.out_data[c] + (pos << 3), *reinterpret_cast<const uint64_t*>(src + (row << 3)));
        if (HAS_NULLS && P.out_valid[c]) P.out_valid[c][pos] = P.T.valid[c] ? (uint8_t)bit_at(stage + P.T.valid_off[c], row) : (uint8_t)1;
    }
    for (int c = 0; c < P.n_right; ++c) {
        const int oc = P.n_left + c;
        if (!P.out_data[oc]) continue;
        uint64_t v = 0;
        uint8_t ok = matched ? 1 : 0;
        if (matched) {
            if (P.payload_col >= 0) {
                v = c == P.payload_col ? val : raw_key;   // inlined payload; the key column is the probe key itself
            } else {
                const uint32_t rid = (uint32_t)val;
                v = P.out_shift[oc] == 2 ? (uint64_t)((const uint32_t*)P.right_data[c])[rid] : ((const uint64_t*)P.right_data[c])[rid];
                if (HAS_NULLS && P.right_valid[c]) ok = (uint8_t)bit_at(P.right_valid[c], rid);
            }
        }
        if (P.out_shift[oc] == 2) st_cs_u32((uint8_t*)P.out_data[oc] + (pos << 2), (uint32_t)v);
        else st_cs_u64((uint8_t*)P.out_data[oc] + (pos << 3), v);
        if (HAS_NULLS && P.out_valid[oc]) P.out_valid[oc][pos] = ok;
    }
}

// K6/K7, unique layout: at most one output row per probe row.  The lookups of
// tile t+1 are issued before the rows of tile t are written, so the scan of tile
// t and the table's DRAM latency overlap with useful work.
template <int R>
struct UniqueTile {
    const uint8_t* stage;
    long long row_begin;
    int rows;
    uint32_t keep[R], hitb[R];
    unsigned long long val[R];
    int count;
};

template <int R, bool HAS_NULLS>
__device__ __forceinline__ void unique_lookup(const JoinProbeParams& P, UniqueTile<R>& t, int warp, int lane) {
    const int row0 = warp * (32 * R) + lane;
    Match m[R];
    lookup_rows<R, HAS_NULLS>(P, t.stage, row0, warp < P.T.active_warps ? t.rows : 0, m);
    t.count = 0;
#pragma unroll
    for (int r = 0; r < R; ++r) {
        const bool inr = warp < P.T.active_warps && (row0 + r * 32) < t.rows;
        t.val[r] = m[r].val;
        t.hitb[r] = __ballot_sync(0xffffffffu, m[r].hit);
        t.keep[r] = __ballot_sync(0xffffffffu, inr && (m[r].hit || P.left_join));
        t.count += __popc(t.keep[r]);
    }
}

// Column by column: width / source tests are uniform per column, the row loop is
// straight predicated code.
template <int R, bool HAS_NULLS>
__device__ __forceinline__ void unique_emit(const JoinProbeParams& P, const UniqueTile<R>& t, long long wbase, int warp, int lane) {
    if (t.count == 0) return;
    const uint32_t lt_mask = (1u << lane) - 1u;
    const int row0 = warp * (32 * R) + lane;
    long long p[R];
    uint32_t mine = 0, hitm = 0;
    {
        long long pos = wbase;
#pragma unroll
        for (int r = 0; r < R; ++r) {
            p[r] = pos + __popc(t.keep[r] & lt_mask);
            pos += __popc(t.keep[r]);
            mine |= ((t.keep[r] >> lane) & 1u) << r;
            hitm |= ((t.hitb[r] >> lane) & 1u) << r;
        }
    }
    if (!mine) return;
    for (int c = 0; c < P.n_left; ++c) {
        const uint8_t* src = t.stage + P.T.smem_off[c];
        uint8_t* dst = (uint8_t*)P.out_data[c];
        if (!dst) continue;
        if (P.out_shift[c] == 3) {
#pragma unroll
            for (int r = 0; r < R; ++r)
                if ((mine >> r) & 1u) st_cs_u64(dst + (p[r] << 3), *reinterpret_cast<const uint64_t*>(src + ((row0 + r * 32) << 3)));
        } else {
#pragma unroll
            for (int r = 0; r < R; ++r)
                if ((mine >> r) & 1u) st_cs_u32(dst + (p[r] << 2), *reinterpret_cast<const uint32_t*>(src + ((row0 + r * 32) << 2)));
        }
        if (HAS_NULLS && P.out_valid[c]) {
#pragma unroll
            for (int r = 0; r < R; ++r)
                if ((mine >> r) & 1u)
                    P.out_valid[c][p[r]] = P.T.valid[c] ? (uint8_t)bit_at(t.stage + P.T.valid_off[c], row0 + r * 32) : (uint8_t)1;
        }
    }
    const uint8_t* kb = t.stage + P.T.smem_off[P.key_slot];
    const int ksh = P.T.shift[P.key_slot];
    for (int c = 0; c < P.n_right; ++c) {
        const int oc = P.n_left + c;
        uint8_t* dst = (uint8_t*)P.out_data[oc];
        if (!dst) continue;
        const int mode = P.payload_col < 0 ? 2 : (c == P.payload_col ? 0 : 1);   // 0 inlined payload, 1 the probe key, 2 gather by row id
#pragma unroll
        for (int r = 0; r < R; ++r) {
            if (!((mine >> r) & 1u)) continue;
            const bool matched = (hitm >> r) & 1u;
            uint64_t v = 0;
            uint8_t ok = matched ? 1 : 0;
            if (matched) {
                if (mode == 0) v = t.val[r];
                else if (mode == 1) {
                    const int row = row0 + r * 32;
                    v = ksh == 2 ? (uint64_t) * reinterpret_cast<const uint32_t*>(kb + (row << 2)) : *reinterpret_cast<const uint64_t*>(kb + (row << 3));
                } else {
                    const uint32_t rid = (uint32_t)t.val[r];
                    v = P.out_shift[oc] == 2 ? (uint64_t)((const uint32_t*)P.right_data[c])[rid] : ((const uint64_t*)P.right_data[c])[rid];
                    if (HAS_NULLS && P.right_valid[c]) ok = (uint8_t)bit_at(P.right_valid[c], rid);
                }
            }
            if (P.out_shift[oc] == 2) st_cs_u32(dst + (p[r] << 2), (uint32_t)v);
            else st_cs_u64(dst + (p[r] << 3), v);
            if (HAS_NULLS && P.out_valid[oc]) P.out_valid[oc][p[r]] = ok;
        }
    }
}

template <int R, bool HAS_NULLS>
__global__ void __launch_bounds__(kPipeThreads, kPipeMinCtas) join_probe_unique_kernel(const __grid_constant__ JoinProbeParams P) {
    extern __shared__ __align__(128) uint8_t smem[];
    __shared__ PipeShared sh;
    TilePipe<true, HAS_NULLS> pipe(P.T, sh, smem);
    pipe.init();
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    if (warp == kPipeComputeWarps) { pipe.control_loop(); return; }
    UniqueTile<R> cur;
    if (!pipe.acquire(0, cur.stage, cur.row_begin, cur.rows)) return;
    unique_lookup<R, HAS_NULLS>(P, cur, warp, lane);
    pipe.post_count(0, warp, cur.count);
    for (int it = 0;; ++it) {
        UniqueTile<R> nxt;
        const bool has_next = pipe.acquire(it + 1, nxt.stage, nxt.row_begin, nxt.rows);
        if (has_next) {
            unique_lookup<R, HAS_NULLS>(P, nxt, warp, lane);
            pipe.post_count(it + 1, warp, nxt.count);
        }
        const long long wbase = pipe.wait_base(it, warp);
        unique_emit<R, HAS_NULLS>(P, cur, wbase, warp, lane);
        pipe.release(it);
        if (!has_next) break;
        cur = nxt;
    }
}

// K6/K7, duplicate layout: a probe row emits `count` rows; runs longer than a
// warp are walked by the whole warp.
template <int R>
struct MultiTile {
    const uint8_t* stage;
    long long row_begin;
    int rows;
    uint32_t cnt[R], excl[R], rend[R], hitb[R];
    long long count;
};

template <int R, bool HAS_NULLS>
__device__ __forceinline__ void multi_lookup(const JoinProbeParams& P, MultiTile<R>& t, int warp, int lane) {
    const int row0 = warp * (32 * R) + lane;
    Match m[R];
    lookup_rows<R, HAS_NULLS>(P, t.stage, row0, warp < P.T.active_warps ? t.rows : 0, m);
    // sums in 64 bits: 128 probe rows that each hit a run of tens of millions of rows overflow 32;
    // a warp total of 2^31 or more is reported as such and take_counts raises the tile-overflow flag
    unsigned long long running = 0;
#pragma unroll
    for (int r = 0; r < R; ++r) {
        const bool inr = warp < P.T.active_warps && (row0 + r * 32) < t.rows;
        uint32_t c = m[r].hit ? (uint32_t)(m[r].val >> 32) : ((inr && P.left_join) ? 1u : 0u);
        if (!inr) c = 0;
        t.hitb[r] = __ballot_sync(0xffffffffu, m[r].hit);
        t.rend[r] = (uint32_t)m[r].val;
        unsigned long long inc = c;
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) {
            const unsigned long long nb = __shfl_up_sync(0xffffffffu, inc, o);
            if (lane >= o) inc += nb;
        }
        t.cnt[r] = c;
        t.excl[r] = (uint32_t)(running + inc - c);   // exact whenever the total below stays under 2^31 (else the call fails)
        running += __shfl_sync(0xffffffffu, inc, 31);
    }
    t.count = (long long)running;
}

// Per-warp scratch of the duplicate-layout probe: one entry per probe row of the warp's group.
struct MultiWarpScratch {
    uint32_t eexcl[32 * 4];   // exclusive prefix of the rows this warp writes itself (listed runs count 0)
    uint32_t excl[32 * 4];    // exclusive prefix of ALL output rows (where a probe row's matches start)
    uint32_t rend[32 * 4];    // end of the probe row's run in the CSR
    uint32_t cnt[32 * 4];     // bit 31: the row matched (else a null-extended Left-join row)
};

// Writes the warp's output rows.  Runs longer than kListRun are listed for the expansion kernel (a
// run of millions of rows must not hang on one warp).  Everything else is expanded load-balanced:
// the warp's output rows are numbered consecutively, lane l writes rows l, l + 32, ... and finds the
// probe row each belongs to by bisecting the prefix sums — whatever the mix of run lengths, all 32
// lanes write (a per-lane loop over its own row's matches leaves 31 lanes waiting for the longest).
template <int R, bool HAS_NULLS>
__device__ __forceinline__ void multi_emit(const JoinProbeParams& P, const MultiTile<R>& t, long long wbase, int warp, int lane,
                                           MultiWarpScratch& S) {
    if (t.count == 0) return;
    static_assert(R <= 4, "MultiWarpScratch holds 4 rows per lane");
    const int row0 = warp * (32 * R) + lane;
    uint32_t ec[R];
#pragma unroll
    for (int r = 0; r < R; ++r) {
        const int row = row0 + r * 32;
        const uint32_t c = t.cnt[r];
        ec[r] = c;
        const long long pos = wbase + t.excl[r];
        uint32_t hv = __ballot_sync(0xffffffffu, c > kListRun);
        while (hv) {   // list one long run (warp-uniform)
            const int src = __ffs(hv) - 1;
            hv &= hv - 1;
            const uint32_t hc = __shfl_sync(0xffffffffu, c, src);
            const uint32_t he = __shfl_sync(0xffffffffu, t.rend[r], src);
            const long long hp = __shfl_sync(0xffffffffu, pos, src);
            const int hrow = row - lane + src;
            bool listed = false;
            if (P.giant_list) {
                const bool huge = hc > kHugeRun;
                unsigned long long slot = 0;
                if (lane == 0) slot = atomicAdd(huge ? P.huge_count : P.giant_count, 1ull);
                slot = __shfl_sync(0xffffffffu, slot, 0);
                if (slot < (huge ? P.huge_cap : P.giant_cap)) {
                    if (lane == 0) {
                        unsigned long long* e = (huge ? P.huge_list : P.giant_list) + 4 * slot;
                        e[0] = (unsigned long long)(t.row_begin + hrow); e[1] = he; e[2] = hc; e[3] = (unsigned long long)hp;
                    }
                    listed = true;
                }
            }
            if (!listed)   // the list is full: the warp walks the run itself
                for (uint32_t j = lane; j < hc; j += 32)
                    emit_row<HAS_NULLS>(P, t.stage, hrow, hp + j, true, P.table.rowids[he - hc + j], 0);
            if (lane == src) ec[r] = 0;
        }
    }
    // prefix sums over the rows the warp writes itself, in the order of the output (r, then lane)
    uint32_t run = 0;
#pragma unroll
    for (int r = 0; r < R; ++r) {
        uint32_t inc = ec[r];
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) {
            const uint32_t nb = __shfl_up_sync(0xffffffffu, inc, o);
            if (lane >= o) inc += nb;
        }
        const int idx = r * 32 + lane;
        S.eexcl[idx] = run + inc - ec[r];
        S.excl[idx] = t.excl[r];
        S.rend[idx] = t.rend[r];
        S.cnt[idx] = t.cnt[r] | (((t.hitb[r] >> lane) & 1u) << 31);
        run += __shfl_sync(0xffffffffu, inc, 31);
    }
    __syncwarp();
    const uint32_t total = run;   // warp-uniform
    for (uint32_t o = lane; o < total; o += 32) {
        int lo = 0, hi = 32 * R - 1;   // last idx with eexcl[idx] <= o
        while (lo < hi) {
            const int mid = (lo + hi + 1) >> 1;
            if (S.eexcl[mid] <= o) lo = mid; else hi = mid - 1;
        }
        const uint32_t j = o - S.eexcl[lo];
        const uint32_t cw = S.cnt[lo], c = cw & 0x7FFFFFFFu;
        const bool hit = cw >> 31;
        const int row = warp * (32 * R) + (lo & 31) + (lo >> 5) * 32;
        emit_row<HAS_NULLS>(P, t.stage, row, wbase + S.excl[lo] + j, hit, hit ? P.table.rowids[S.rend[lo] - c + j] : 0u, 0);
    }
    __syncwarp();   // the scratch is rewritten by the next tile
}

template <int R, bool HAS_NULLS>
__global__ void __launch_bounds__(kPipeThreads, kPipeMinCtas) join_probe_multi_kernel(const __grid_constant__ JoinProbeParams P) {
    extern __shared__ __align__(128) uint8_t smem[];
    __shared__ PipeShared sh;
    __shared__ MultiWarpScratch wscratch[kPipeComputeWarps];
    TilePipe<true, HAS_NULLS> pipe(P.T, sh, smem);
    pipe.init();
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    if (warp == kPipeComputeWarps) { pipe.control_loop(); return; }
    MultiTile<R> cur;
    if (!pipe.acquire(0, cur.stage, cur.row_begin, cur.rows)) return;
    multi_lookup<R, HAS_NULLS>(P, cur, warp, lane);
    pipe.post_count(0, warp, cur.count);
    for (int it = 0;; ++it) {
        MultiTile<R> nxt;
        const bool has_next = pipe.acquire(it + 1, nxt.stage, nxt.row_begin, nxt.rows);
        if (has_next) {
            multi_lookup<R, HAS_NULLS>(P, nxt, warp, lane);
            pipe.post_count(it + 1, warp, nxt.count);
        }
        const long long wbase = pipe.wait_base(it, warp);
        multi_emit<R, HAS_NULLS>(P, cur, wbase, warp, lane, wscratch[warp]);
        pipe.release(it);
        if (!has_next) break;
        cur = nxt;
    }
}

// ---------------------------------------------------------------------------
// Fast path (inner join, unique layout, no validity anywhere): a flat kernel
// without staging, scan or control warp.  Each lane owns 4 rows, issues its 4
// bucket loads back to back, resolves them, and a CTA reserves its output range
// with one atomicAdd — output order is by CTA chunk, which a join does not
// define anyway.  ~10x fewer instructions per row than the generic tile
// pipeline, which matters once the table is L2-resident (radix-partitioned
// input) and HBM latency no longer hides them.
// ---------------------------------------------------------------------------
#ifndef IGL_FLAT_THREADS
#define IGL_FLAT_THREADS 32
#endif
constexpr int kFlatThreads = IGL_FLAT_THREADS;
#ifndef IGL_FLAT_ROWS
#define IGL_FLAT_ROWS 4
#endif
#ifndef IGL_FLAT_CTAS
#define IGL_FLAT_CTAS 24
#endif
constexpr int kFlatRows = IGL_FLAT_ROWS;
constexpr int kFlatChunk = kFlatThreads * kFlatRows;   // 1024 rows per CTA iteration
constexpr int kFlatClaim = 4;                          // consecutive chunks claimed per ticket
static_assert(kFlatChunk == kSegChunkRows, "ProbeSeg::chunk0 counts chunks of the flat probe kernel");

// Expands the listed runs.  The left row's values are read once per thread from the (global) left
// columns; right columns are gathered by row id (the identity when the build columns were put into
// CSR order).  `first`/`step` select the rows of the run this thread writes.
template <bool HAS_NULLS>
__device__ __forceinline__ void expand_run(const JoinProbeParams& P, const unsigned long long* ent, long long first, long long step) {
    const long long lrow = (long long)ent[0];
    const uint32_t rend = (uint32_t)ent[1], cnt = (uint32_t)ent[2];
    const long long pos0 = (long long)ent[3];
    uint64_t lv[kJoinMaxCols];
    uint8_t lok[kJoinMaxCols];
    for (int c = 0; c < P.n_left; ++c) {
        lv[c] = P.out_shift[c] == 2 ? (uint64_t)((const uint32_t*)P.T.data[c])[lrow] : ((const uint64_t*)P.T.data[c])[lrow];
        lok[c] = (HAS_NULLS && P.T.valid[c]) ? (uint8_t)bit_at(P.T.valid[c], lrow) : (uint8_t)1;
    }
    for (long long j = first; j < (long long)cnt; j += step) {
        const long long pos = pos0 + j;
        const uint32_t rid = P.table.rowids[rend - cnt + (uint32_t)j];
        for (int c = 0; c < P.n_left; ++c) {
            if (!P.out_data[c]) continue;
            if (P.out_shift[c] == 2) st_cs_u32((uint8_t*)P.out_data[c] + (pos << 2), (uint32_t)lv[c]);
            else st_cs_u64((uint8_t*)P.out_data[c] + (pos << 3), lv[c]);
            if (HAS_NULLS && P.out_valid[c]) P.out_valid[c][pos] = lok[c];
        }
        for (int c = 0; c < P.n_right; ++c) {
            const int oc = P.n_left + c;
            if (!P.out_data[oc]) continue;
            const uint64_t v = P.out_shift[oc] == 2 ? (uint64_t)((const uint32_t*)P.right_data[c])[rid] : ((const uint64_t*)P.right_data[c])[rid];
            if (P.out_shift[oc] == 2) st_cs_u32((uint8_t*)P.out_data[oc] + (pos << 2), (uint32_t)v);
            else st_cs_u64((uint8_t*)P.out_data[oc] + (pos << 3), v);
            if (HAS_NULLS && P.out_valid[oc]) P.out_valid[oc][pos] = P.right_valid[c] ? (uint8_t)bit_at(P.right_valid[c], rid) : (uint8_t)1;
        }
    }
}

template <bool HAS_NULLS>
__global__ void __launch_bounds__(256) join_expand_giant_kernel(const __grid_constant__ JoinProbeParams P) {
    // listed runs: one CTA per run
    const unsigned long long nm = min(*P.giant_count, (unsigned long long)P.giant_cap);
    for (unsigned long long e = blockIdx.x; e < nm; e += gridDim.x) expand_run<HAS_NULLS>(P, P.giant_list + 4 * e, threadIdx.x, blockDim.x);
    // huge runs: every CTA takes a slice of every run
    const unsigned long long nh = min(*P.huge_count, (unsigned long long)P.huge_cap);
    for (unsigned long long e = 0; e < nh; ++e)
        expand_run<HAS_NULLS>(P, P.huge_list + 4 * e, (long long)blockIdx.x * blockDim.x + threadIdx.x, (long long)gridDim.x * blockDim.x);
}

struct FlatProbeParams {
    const uint8_t* left[kJoinMaxCols];
    const uint8_t* right_data[kJoinMaxCols];
    void* out[2 * kJoinMaxCols];
    uint8_t lshift[kJoinMaxCols], rshift[kJoinMaxCols];
    int n_left, n_right, key_col, key_type, right_key_col, payload_col;
    JoinTable table;
    long long nrows;
    unsigned long long* out_counter;
    unsigned int* ticket;   // chunks are claimed in order, so all CTAs work on one moving frontier of the
                            // (partitioned) input and the table region they touch stays L2-resident
    uint32_t stage_off[kJoinMaxCols];   // byte offset of each left column's chunk slice inside a stage
    uint32_t stage_bytes;
    // DENSE form: matched rows keep their input position, the warps leave a hit bitmap
    uint32_t* hit_words;                // bit i = probe row i matched; padded to whole 1024-row tiles, zero-filled
    unsigned long long* miss_counter;
    // Build rows as packed records (unique layout with row-id payload, rows of at most 32 bytes): one
    // 16- or 32-byte load per match fetches every right column, instead of one random sector per column.
    const uint8_t* packed;
    int packed_stride;                  // 0 = none, else 16 or 32
    uint8_t pk_off[kJoinMaxCols];       // byte offset of each right column inside a record
    // SEG form (multi-GPU join): the probe rows are a list of runs spread over several source
    // buffers — this GPU's staging buffer and the peers', read over NVLink by the same bulk
    // copies.  `nrows` is the total over all runs.
    const ProbeSeg* segs;
    int nsegs;
    long long seg_nchunks;
    const uint8_t* seg_left[kSegMaxSrc][kJoinMaxCols];
};

#ifndef IGL_FLAT_COMPACT
#define IGL_FLAT_COMPACT 1
#endif
constexpr int kFlatStages = 3;   // chunks in flight per CTA: the one being probed + two on their way from HBM

// The left columns of every chunk reach shared memory through a ring of bulk async
// copies issued by thread 0 two chunks ahead (the threads themselves only have four
// rows each in flight, far too few bytes to cover HBM latency); table buckets are
// read directly (L2-resident region) and the output is stored straight from registers.
//
// One warp per CTA (24 CTAs per SM): measured, the probe's speed follows the number of
// independent chunk chains per SM, not the rows in flight (256-thread CTAs 19.5 ms, 128: 17.6,
// 64: 16.0, 32: 14.9 for C3).  With one output reservation per 128-row chunk the shared
// cursor then becomes the limit (same-address atomics top out near 0.55 G/s), so the
// DENSE form drops the cursor: an inner join against a unique table yields one row or
// none per probe row — nearly always one in the foreign-key joins this path exists for —
// and matched rows are written at their INPUT position.  If any row missed, dense_fix
// afterwards moves matched rows from the tail into the holes (output order carries no
// meaning here: the partitioned probe is already unordered).
//
// SEG form: the chunk a CTA claims is looked up in a list of runs (ProbeSeg) that may sit in
// another GPU's memory; the bulk copy then crosses NVLink and lands in shared memory directly,
// so the exchange of the probe side and the probe are one kernel: no receive buffer, and the
// transfer of a chunk overlaps the lookups of the chunks before it.  (Plain bulk copies, never
// multicast: TMA multicast loads here read only this GPU's memory, since on B200 a TMA
// multicast load from another GPU's memory hung the GPU.)
template <int KEY_TYPE, bool DENSE, bool SEG>
__global__ void __launch_bounds__(kFlatThreads, IGL_FLAT_CTAS) join_probe_flat_kernel(const __grid_constant__ FlatProbeParams P) {
    static_assert(!SEG || DENSE, "segmented input is only written densely");
    extern __shared__ __align__(128) uint8_t fsmem[];
    __shared__ uint64_t s_full[kFlatStages];
    __shared__ long long s_chunk_of[kFlatStages];
    __shared__ int s_rows_of[kFlatStages];   // SEG: rows of the chunk held by each stage
    __shared__ unsigned long long s_obase[kFlatStages];   // SEG: output row of the chunk's first row
    __shared__ unsigned int s_wcnt[kFlatThreads / 32];
    __shared__ unsigned long long s_base;
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const uint32_t lt_mask = (1u << lane) - 1u;
    const uint64_t bmask = P.table.mask >> 1;
    const long long nchunks = SEG ? P.seg_nchunks : (P.nrows + kFlatChunk - 1) / kFlatChunk;
    const Slot* slots = P.table.slots;
    const int tshift = P.table.shift + 1;   // slot index -> bucket index
    const int lrow = warp * (32 * kFlatRows) + lane;   // this lane's first row inside a chunk
    unsigned int misses = 0;
    // ---- producer state (thread 0 only) ----
    long long grp_base = 0;
    int grp_used = kFlatClaim;
    unsigned int pending = 0;
    int cur_seg = 0;   // SEG: claims only grow, so the run holding a chunk is found by walking forward
    auto refill = [&](int s) {   // claim the next chunk and start its copies into stage s
        if (grp_used == kFlatClaim) {
            grp_base = (long long)pending * kFlatClaim;
            pending = atomicAdd(P.ticket, 1u);   // for the group after this one: its round trip is hidden
            grp_used = 0;
        }
        const long long chunk = grp_base + grp_used++;
        if (chunk >= nchunks) { s_chunk_of[s] = -1; mbar_arrive(&s_full[s]); return; }
        s_chunk_of[s] = chunk;
        uint8_t* st = fsmem + (size_t)s * P.stage_bytes;
        if (SEG) {
            while (cur_seg + 1 < P.nsegs && (long long)__ldg(&P.segs[cur_seg + 1].chunk0) <= chunk) ++cur_seg;
            const uint32_t c0 = __ldg(&P.segs[cur_seg].chunk0), srows = __ldg(&P.segs[cur_seg].rows);
            const unsigned long long rs = __ldg(&P.segs[cur_seg].row0_src);
            const uint32_t done = (uint32_t)(chunk - c0) * kFlatChunk;
            const uint32_t rows = min((uint32_t)kFlatChunk, srows - done);
            const int src = (int)(rs >> 56);
            const size_t srow = (size_t)(rs & 0x00FFFFFFFFFFFFFFull) + done;
            s_rows_of[s] = (int)rows;
            s_obase[s] = __ldg(&P.segs[cur_seg].out0) + done;
            uint32_t total = 0;
            for (int c = 0; c < P.n_left; ++c) total += ((rows << P.lshift[c]) + 15u) & ~15u;
            mbar_arrive_expect_tx(&s_full[s], total);
            for (int c = 0; c < P.n_left; ++c)
                bulk_g2s(st + P.stage_off[c], P.seg_left[src][c] + (srow << P.lshift[c]), ((rows << P.lshift[c]) + 15u) & ~15u, &s_full[s]);
            return;
        }
        const long long cb = chunk * kFlatChunk;
        const uint32_t rows = (uint32_t)min((long long)kFlatChunk, P.nrows - cb);
        uint32_t total = 0;
        for (int c = 0; c < P.n_left; ++c) total += ((rows << P.lshift[c]) + 15u) & ~15u;
        mbar_arrive_expect_tx(&s_full[s], total);
        for (int c = 0; c < P.n_left; ++c)
            bulk_g2s(st + P.stage_off[c], P.left[c] + ((size_t)cb << P.lshift[c]), ((rows << P.lshift[c]) + 15u) & ~15u, &s_full[s]);
    };
    if (threadIdx.x == 0) {
        for (int s = 0; s < kFlatStages; ++s) mbar_init(&s_full[s], 1);
        mbar_fence_init();
        pending = atomicAdd(P.ticket, 1u);
        for (int s = 0; s < kFlatStages; ++s) refill(s);
    }
    __syncthreads();
    for (int it = 0;; ++it) {
        const int sidx = it % kFlatStages;
        mbar_wait(&s_full[sidx], (uint32_t)(it / kFlatStages) & 1u);
        const long long chunk = s_chunk_of[sidx];
        if (chunk < 0) break;   // claims only grow: nothing is left for this CTA
        const uint8_t* stage = fsmem + (size_t)sidx * P.stage_bytes;
        const long long cbase = chunk * kFlatChunk;                       // uniform
        const int nvalid = SEG ? s_rows_of[sidx] : (int)min((long long)kFlatChunk, P.nrows - cbase);   // rows of this chunk that exist
        uint64_t raw[kFlatRows], k[kFlatRows];
        unsigned long long val[kFlatRows];
        ulonglong4 q[kFlatRows];
        uint32_t live = 0;
        {
            const uint8_t* kc = stage + P.stage_off[P.key_col];
#pragma unroll
            for (int r = 0; r < kFlatRows; ++r) {
                const int i = lrow + r * 32;   // rows past nvalid read stale shared memory and do not count
                raw[r] = KEY_TYPE == IGL_T_I32 ? (uint64_t)((const uint32_t*)kc)[i] : ((const uint64_t*)kc)[i];
                live |= ((lrow + r * 32) < nvalid ? 1u : 0u) << r;
            }
        }
#pragma unroll
        for (int r = 0; r < kFlatRows; ++r) {
            k[r] = canon_key(KEY_TYPE, raw[r]);
            const uint64_t b = mix64(k[r]) >> tshift;
            q[r] = ld_nc_u64x4(slots + 2 * b);
        }
        // first bucket, branch-free: most rows are decided here
        uint32_t hit = 0, pend = 0;
#pragma unroll
        for (int r = 0; r < kFlatRows; ++r) {
            const bool h0 = q[r].x == k[r], h1 = q[r].z == k[r];
            const bool e0 = q[r].x == kEmptyKey, e1 = q[r].z == kEmptyKey;
            const bool found = h0 || (h1 && !e0);
            val[r] = h0 ? q[r].y : q[r].w;
            hit |= (found ? 1u : 0u) << r;
            pend |= ((found || e0 || e1) ? 0u : 1u) << r;
        }
        uint32_t special = 0;
#pragma unroll
        for (int r = 0; r < kFlatRows; ++r) special |= (k[r] == kEmptyKey ? 1u : 0u) << r;
        hit &= live & ~special;
        pend &= live & ~special;
        if (special & live) {   // the key equal to the empty-slot sentinel lives in a side entry (rare)
            if (P.table.side->count[1]) {
#pragma unroll
                for (int r = 0; r < kFlatRows; ++r)
                    if (((special & live) >> r) & 1u) { hit |= 1u << r; val[r] = P.table.side->val[1]; }
            }
        }
#if IGL_FLAT_COMPACT
        // About one row in ten is not decided by its home bucket (load factor 0.5, two slots per
        // bucket), so nearly every 128-row chunk has some — on average a dozen, spread over the
        // lanes.  Walking them where they sit costs every lane the predicated code of all four of
        // its rows per round (a third of this kernel's instructions, measured); instead the
        // undecided keys of the whole warp are gathered into a dense list in shared memory, each
        // lane walks ONE of them to its end, and the owners read the answers back.
        if (__any_sync(0xffffffffu, pend != 0)) {
            static_assert(kFlatThreads == 32, "the walk list is per warp");
            __shared__ unsigned long long s_walk[kFlatChunk];   // key in, value out
            __shared__ unsigned char s_walk_hit[kFlatChunk];
            int pos[kFlatRows], total = 0;
#pragma unroll
            for (int r = 0; r < kFlatRows; ++r) {
                const uint32_t pb = __ballot_sync(0xffffffffu, (pend >> r) & 1u);
                pos[r] = total + __popc(pb & lt_mask);
                total += __popc(pb);
                if ((pend >> r) & 1u) s_walk[pos[r]] = k[r];
            }
            __syncwarp();
            for (int i = lane; i < total; i += 32) {
                const uint64_t key = s_walk[i];
                uint64_t b = ((mix64(key) >> tshift) + 1) & bmask;
                unsigned long long v = 0;
                bool found = false;
                while (true) {
                    const ulonglong4 qq = ld_nc_u64x4(slots + 2 * b);
                    if (qq.x == key) { v = qq.y; found = true; break; }
                    if (qq.x == kEmptyKey) break;
                    if (qq.z == key) { v = qq.w; found = true; break; }
                    if (qq.z == kEmptyKey) break;
                    b = (b + 1) & bmask;
                }
                s_walk[i] = v;
                s_walk_hit[i] = found ? 1 : 0;
            }
            __syncwarp();
#pragma unroll
            for (int r = 0; r < kFlatRows; ++r)
                if (((pend >> r) & 1u) && s_walk_hit[pos[r]]) { hit |= 1u << r; val[r] = s_walk[pos[r]]; }
            pend = 0;
        }
#else
        if (pend) {
            // About one row in ten is not decided by its home bucket (load factor 0.5, two
            // slots per bucket), so nearly every warp comes here: fetch the next bucket of
            // ALL this lane's undecided rows at once — one L2 round trip, not one per row.
            uint64_t b[kFlatRows];
#pragma unroll
            for (int r = 0; r < kFlatRows; ++r) {
                b[r] = ((mix64(k[r]) >> tshift) + 1) & bmask;
                if ((pend >> r) & 1u) q[r] = ld_nc_u64x4(slots + 2 * b[r]);
            }
#pragma unroll
            for (int r = 0; r < kFlatRows; ++r) {
                if ((pend >> r) & 1u) {
                    if (q[r].x == k[r]) { hit |= 1u << r; val[r] = q[r].y; pend &= ~(1u << r); }
                    else if (q[r].x == kEmptyKey) pend &= ~(1u << r);
                    else if (q[r].z == k[r]) { hit |= 1u << r; val[r] = q[r].w; pend &= ~(1u << r); }
                    else if (q[r].z == kEmptyKey) pend &= ~(1u << r);
                }
            }
            while (pend) {   // longer walks (about one row in a hundred): again all undecided rows per round trip
#pragma unroll
                for (int r = 0; r < kFlatRows; ++r) {
                    b[r] = (b[r] + 1) & bmask;
                    if ((pend >> r) & 1u) q[r] = ld_nc_u64x4(slots + 2 * b[r]);
                }
#pragma unroll
                for (int r = 0; r < kFlatRows; ++r) {
                    if ((pend >> r) & 1u) {
                        if (q[r].x == k[r]) { hit |= 1u << r; val[r] = q[r].y; pend &= ~(1u << r); }
                        else if (q[r].x == kEmptyKey) pend &= ~(1u << r);
                        else if (q[r].z == k[r]) { hit |= 1u << r; val[r] = q[r].w; pend &= ~(1u << r); }
                        else if (q[r].z == kEmptyKey) pend &= ~(1u << r);
                    }
                }
            }
        }
#endif
        uint32_t hb[kFlatRows];
        unsigned int wcnt = 0;
#pragma unroll
        for (int r = 0; r < kFlatRows; ++r) { hb[r] = __ballot_sync(0xffffffffu, (hit >> r) & 1u); wcnt += __popc(hb[r]); }
        unsigned long long obase;
        unsigned int p[kFlatRows];   // 32-bit output offsets from obase
        if (DENSE) {
            uint32_t mine = 0;   // the ballots are the bitmap words of this warp's rows
#pragma unroll
            for (int r = 0; r < kFlatRows; ++r) {
                misses += __popc(__ballot_sync(0xffffffffu, (live >> r) & 1u)) - __popc(hb[r]);
                mine = lane == r ? hb[r] : mine;
                p[r] = (unsigned int)(lrow + r * 32);
            }
            if (SEG) {
                // Runs are packed in the output, so a chunk starts at any row: its four ballot words
                // land on five bitmap words.  Words wholly inside a full chunk are stored; the two
                // edge words (and every word of a run's short last chunk) are shared with the
                // neighbouring chunks and OR-ed in (the bitmap starts zeroed).
                static_assert(!SEG || kFlatThreads == 32, "one warp per CTA");
                obase = s_obase[sidx];
                const uint32_t sh = (uint32_t)obase & 31u;
                uint32_t prev = 0;   // hb[lane - 1]
#pragma unroll
                for (int r = 0; r < kFlatRows; ++r) prev = lane == r + 1 ? hb[r] : prev;
                const uint32_t word = (lane < kFlatRows ? mine << sh : 0u) | ((sh && lane >= 1 && lane <= kFlatRows) ? prev >> (32u - sh) : 0u);
                if (lane <= kFlatRows && word) {
                    uint32_t* wp = P.hit_words + (obase >> 5) + lane;
                    const bool own = nvalid == kFlatChunk && (sh == 0 ? lane < kFlatRows : (lane >= 1 && lane < kFlatRows));
                    if (own) *wp = word; else atomicOr(wp, word);
                }
            } else {
                if (lane < kFlatRows) P.hit_words[(cbase >> 5) + warp * kFlatRows + lane] = mine;
                obase = (unsigned long long)cbase;
            }
            __syncthreads();   // every thread is done with the previous chunk's stage
            if (threadIdx.x == 0 && it > 0) refill((it - 1) % kFlatStages);
        } else {
            if (lane == 0) s_wcnt[warp] = wcnt;
            __syncthreads();   // every thread is also done with the previous chunk's stage
            if (threadIdx.x == 0) {
                if (it > 0) refill((it - 1) % kFlatStages);
                unsigned int tot = 0;
#pragma unroll
                for (int w = 0; w < kFlatThreads / 32; ++w) tot += s_wcnt[w];
                s_base = tot ? atomicAdd(P.out_counter, (unsigned long long)tot) : 0ull;
            }
            __syncthreads();
            obase = s_base;   // uniform
            unsigned int lpos = 0;
            for (int w = 0; w < warp; ++w) lpos += s_wcnt[w];
            __syncthreads();   // s_wcnt / s_base are rewritten by the next chunk
#pragma unroll
            for (int r = 0; r < kFlatRows; ++r) { p[r] = lpos + __popc(hb[r] & lt_mask); lpos += __popc(hb[r]); }
        }
        // column by column: uniform width test, unconditional (clamped) loads, predicated stores
        for (int c = 0; c < P.n_left; ++c) {
            const bool is_key = c == P.key_col;
            if (!P.out[c]) continue;   // hidden column (a computed / combined key): joined on, not emitted
            if (P.lshift[c] == 3) {
                const uint64_t* src = (const uint64_t*)(stage + P.stage_off[c]);
                uint64_t* dst = (uint64_t*)P.out[c] + obase;
                uint64_t v[kFlatRows];
#pragma unroll
                for (int r = 0; r < kFlatRows; ++r) v[r] = is_key ? raw[r] : src[lrow + r * 32];
#pragma unroll
                for (int r = 0; r < kFlatRows; ++r) if ((hit >> r) & 1u) st_cs_u64(dst + p[r], v[r]);
            } else {
                const uint32_t* src = (const uint32_t*)(stage + P.stage_off[c]);
                uint32_t* dst = (uint32_t*)P.out[c] + obase;
                uint32_t v[kFlatRows];
#pragma unroll
                for (int r = 0; r < kFlatRows; ++r) v[r] = is_key ? (uint32_t)raw[r] : src[lrow + r * 32];
#pragma unroll
                for (int r = 0; r < kFlatRows; ++r) if ((hit >> r) & 1u) st_cs_u32(dst + p[r], v[r]);
            }
        }
        if (P.packed_stride) {
            // row-id payload, packed records: all four record loads first, then every column from registers
            ulonglong4 rec[kFlatRows];
#pragma unroll
            for (int r = 0; r < kFlatRows; ++r) {
                rec[r] = ulonglong4{0, 0, 0, 0};
                if ((hit >> r) & 1u) {
                    const uint8_t* a = P.packed + (size_t)(uint32_t)val[r] * (size_t)P.packed_stride;
                    if (P.packed_stride == 32) rec[r] = ld_nc_u64x4(a);
                    else { const ulonglong2 h = ld_nc_u64x2(a); rec[r].x = h.x; rec[r].y = h.y; }
                }
            }
            for (int c = 0; c < P.n_right; ++c) {
                if (!P.out[P.n_left + c]) continue;
                const int wi = P.pk_off[c] >> 3, sh = (P.pk_off[c] & 4) * 8;
#pragma unroll
                for (int r = 0; r < kFlatRows; ++r) {
                    if (!((hit >> r) & 1u)) continue;
                    const unsigned long long w = wi == 0 ? rec[r].x : wi == 1 ? rec[r].y : wi == 2 ? rec[r].z : rec[r].w;
                    if (P.rshift[c] == 3) st_cs_u64((uint64_t*)P.out[P.n_left + c] + obase + p[r], w);
                    else st_cs_u32((uint32_t*)P.out[P.n_left + c] + obase + p[r], (uint32_t)(w >> sh));
                }
            }
        } else
        for (int c = 0; c < P.n_right; ++c) {
            const int mode = P.payload_col < 0 ? 2 : (c == P.payload_col ? 0 : 1);   // 0 inlined payload, 1 the key itself, 2 gather by row id
            if (!P.out[P.n_left + c]) continue;
            if (P.rshift[c] == 3) {
                const uint64_t* src = (const uint64_t*)P.right_data[c];
                uint64_t* dst = (uint64_t*)P.out[P.n_left + c] + obase;
#pragma unroll
                for (int r = 0; r < kFlatRows; ++r)
                    if ((hit >> r) & 1u) st_cs_u64(dst + p[r], mode == 0 ? val[r] : mode == 1 ? raw[r] : src[(uint32_t)val[r]]);
            } else {
                const uint32_t* src = (const uint32_t*)P.right_data[c];
                uint32_t* dst = (uint32_t*)P.out[P.n_left + c] + obase;
#pragma unroll
                for (int r = 0; r < kFlatRows; ++r)
                    if ((hit >> r) & 1u) st_cs_u32(dst + p[r], mode == 0 ? (uint32_t)val[r] : mode == 1 ? (uint32_t)raw[r] : src[(uint32_t)val[r]]);
            }
        }
    }
    if (DENSE && lane == 0 && misses) atomicAdd(P.miss_counter, (unsigned long long)misses);
}

// Hole filling after a dense probe with misses.  H = matched rows in total; the output is
// dense once every unmatched position below H holds a matched row taken from at or above H.
// One warp per 32-row bitmap word; tile_off[] (per `tile_rows` rows) gives the matched rows
// before each tile.  The j-th hole (from the front) receives the matched tail row that has
// j matched rows after it.
struct DenseFixParams {
    const uint32_t* hit_words;
    const unsigned long long* tile_off;   // [ntiles + 1]
    int tile_rows;
    long long nrows;
    uint32_t* holes;                      // [misses]
    void* out[2 * kJoinMaxCols];
    uint8_t shift[2 * kJoinMaxCols];
    int ncols;
};

template <bool FILL>
__global__ void __launch_bounds__(256) dense_fix_kernel(const __grid_constant__ DenseFixParams P) {
    const int lane = threadIdx.x & 31;
    const long long nwords = (P.nrows + 31) >> 5;
    const long long ntiles = (P.nrows + P.tile_rows - 1) / P.tile_rows;
    const unsigned long long H = P.tile_off[ntiles];
    const int wpt = P.tile_rows >> 5;
    const long long warps = ((long long)gridDim.x * blockDim.x) >> 5;
    // the fill pass walks the words at or above H's word, the hole pass the words up to it
    const long long w_lo = FILL ? (long long)(H >> 5) : 0, w_hi = FILL ? nwords : (long long)((H + 31) >> 5);
    for (long long w = w_lo + (((long long)blockIdx.x * blockDim.x + threadIdx.x) >> 5); w < w_hi; w += warps) {
        const long long tile = w / wpt;
        const int j = (int)(w - tile * wpt);
        int c = 0;
        for (int i = lane; i < j; i += 32) c += __popc(P.hit_words[tile * wpt + i]);
        c = __reduce_add_sync(0xffffffffu, c);
        const uint32_t word = P.hit_words[w];
        const long long i = (w << 5) + lane;
        const bool bit = (word >> lane) & 1u;
        const unsigned long long hb = P.tile_off[tile] + (unsigned)c + __popc(word & ((1u << lane) - 1u));   // matched rows before row i
        if (!FILL) {
            if (i < (long long)H && !bit) P.holes[(unsigned long long)i - hb] = (uint32_t)i;
        } else if (i >= (long long)H && i < P.nrows && bit) {
            const uint32_t dst = P.holes[H - hb - 1];
            for (int col = 0; col < P.ncols; ++col) {
                if (!P.out[col]) continue;
                if (P.shift[col] == 3) ((uint64_t*)P.out[col])[dst] = ((const uint64_t*)P.out[col])[i];
                else ((uint32_t*)P.out[col])[dst] = ((const uint32_t*)P.out[col])[i];
            }
        }
    }
}

// Exact output row count of a probe call (needed to size the output of the
// duplicate layout; also serves igl_join_probe_count for the unique layout).
__global__ void join_probe_count_kernel(const void* __restrict__ keys, int key_type, const uint8_t* __restrict__ key_valid,
                                        long long n, JoinTable t, int csr, int left_join, int never_match,
                                        unsigned long long* total) {
    unsigned long long local = 0;
    const long long stride = (long long)gridDim.x * blockDim.x;
    for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += stride) {
        unsigned long long c = 0;
        if (!never_match) {
            const bool valid = !key_valid || bit_at(key_valid, i);
            const uint64_t k = valid ? canon_key(key_type, load_key_global(keys, key_type, i)) : 0;
            if (!valid || k == kEmptyKey) c = csr ? t.side->count[valid ? 1 : 0] : (t.side->count[valid ? 1 : 0] ? 1 : 0);
            else {
                uint64_t h = (mix64(k) >> t.shift) & ~1ull;   // homes are even: a bucket is one 32-byte sector
                while (true) {
                    const ulonglong2 v = ld_nc_u64x2(t.slots + h);
                    if (v.x == k) { c = csr ? (v.y >> 32) : 1; break; }
                    if (v.x == kEmptyKey) break;
                    h = (h + 1) & t.mask;
                }
            }
        }
        if (c == 0 && left_join) c = 1;
        local += c;
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) local += __shfl_xor_sync(0xffffffffu, local, o);
    if ((threadIdx.x & 31) == 0 && local) atomicAdd(total, local);
}

}  // namespace igl

using namespace igl;

struct igl_join {
    igl_ctx* ctx = nullptr;
    int join_type = 0;
    bool left_join = false;
    std::vector<int> lt, rt;
    std::vector<uint8_t> ln, rn;
    int lkey = 0, rkey = 0;
    bool never_match = false;
    std::vector<uint8_t> lhide, rhide;       // hidden columns: joined on (computed or combined keys) but not emitted
    void* ex = nullptr;                      // join_ex.cu: this handle fronts a wider join (several / computed keys, Boolean columns)
    // build state
    bool built = false, csr = false;
    int payload_col = -1;
    JoinTable table{};
    uint64_t capacity = 0;
    void* d_slots = nullptr;
    void* d_side = nullptr;
    void* d_rowids = nullptr;
    std::vector<igl_dev_col> right;          // what the table's row ids index: the caller's columns, or their partitioned copy
    int64_t n_right_rows = 0;
    // radix partitioning (large build sides): both sides are reordered by the top
    // bits of the hash so each 16 MB table region is built and probed while L2-resident
    int log2_parts = 0;
    bool prepartitioned = false;             // both sides arrive ordered by the top hash bits: skip the radix passes
    int prepart_bits = 0;                    // then: by how many bits (0 / 1 = not told)
    void* fine_scratch = nullptr;            // shared-memory build: counts / offsets / cursors of the fine bins
    size_t fine_scratch_bytes = 0;
    void* fine_ovf = nullptr;                // rows that ran past their sub-region
    size_t fine_ovf_cap = 0;
    bool fine_built = false;                 // the last build went through join_build_fine_kernel
    std::vector<void*> part_right;           // partitioned build columns (owned)
    std::vector<void*> csr_right;            // duplicate layout: the build columns in CSR order (owned)
    void* packed = nullptr;                  // unique layout with row-id payload: the build rows as 16/32-byte records (owned)
    int packed_stride = 0;
    uint8_t pk_off[kJoinMaxCols] = {};
    std::vector<void*> part_left;            // partitioned probe columns (owned, grow-only)
    int64_t part_left_rows = 0;
    void* part_scratch = nullptr;
    // probe scratch
    void* scratch = nullptr;
    size_t scratch_bytes = 0;
    std::vector<void*> valid_planes;
    int64_t plane_rows = 0;
    PipeHeader* host_hdr = nullptr;
    // device-timeline marks of the last build / probe (igl_join_phase_ms): [0] build start, [1] after its
    // partitioning, [2] build end, [3] probe start, [4] after its partitioning, [5] probe kernel end
    cudaEvent_t ev[6] = {};
    bool ev_build = false, ev_probe = false;
    // a probe that has been queued but not finished (join_probe_begin / join_probe_end)
    // the last igl_join_probe_count (a caller sizes its output with it; igl_join_probe then needs the same number)
    const void* counted_keys = nullptr;
    int64_t counted_nrows = -1, counted_rows = 0;
    int pend_kind = 0;                       // 0 none, 1 flat kernel queued, 2 already complete (pend_rows)
    int64_t pend_rows = 0, pend_nrows = 0;
    bool pend_dense = false;
    std::unique_ptr<FlatProbeParams> pend_F;
    JoinSide* host_side = nullptr;
    PipeCols T{};
    PipeGeometry geo;
    bool has_nulls = false;
    size_t smem_bytes = 0;
    int grid_cap_u = 0, grid_cap_m = 0;
};

namespace {

typedef void (*probe_kernel_t)(const JoinProbeParams);
probe_kernel_t pick_probe(bool csr, int rows_per_lane, bool nulls) {
    if (rows_per_lane == 4) {
        if (csr) return nulls ? join_probe_multi_kernel<4, true> : join_probe_multi_kernel<4, false>;
        return nulls ? join_probe_unique_kernel<4, true> : join_probe_unique_kernel<4, false>;
    }
    if (csr) return nulls ? join_probe_multi_kernel<1, true> : join_probe_multi_kernel<1, false>;
    return nulls ? join_probe_unique_kernel<1, true> : join_probe_unique_kernel<1, false>;
}

bool fixed_width(int t) { return t == IGL_T_I32 || t == IGL_T_I64 || t == IGL_T_F64; }

const void* flat_kernel_for(int kt, bool dense, bool seg) {
    if (seg)
        return kt == IGL_T_I32 ? (const void*)join_probe_flat_kernel<IGL_T_I32, true, true>
               : kt == IGL_T_F64 ? (const void*)join_probe_flat_kernel<IGL_T_F64, true, true>
                                 : (const void*)join_probe_flat_kernel<IGL_T_I64, true, true>;
    if (dense)
        return kt == IGL_T_I32 ? (const void*)join_probe_flat_kernel<IGL_T_I32, true, false>
               : kt == IGL_T_F64 ? (const void*)join_probe_flat_kernel<IGL_T_F64, true, false>
                                 : (const void*)join_probe_flat_kernel<IGL_T_I64, true, false>;
    return kt == IGL_T_I32 ? (const void*)join_probe_flat_kernel<IGL_T_I32, false, false>
           : kt == IGL_T_F64 ? (const void*)join_probe_flat_kernel<IGL_T_F64, false, false>
                             : (const void*)join_probe_flat_kernel<IGL_T_I64, false, false>;
}

uint32_t flat_stage_bytes(const igl_join* j) {
    uint32_t b = 0;
    for (int t : j->lt) b += (uint32_t)kFlatChunk << (t == IGL_T_I32 ? 2 : 3);
    return b;
}

// The flat kernel's preconditions that do not depend on the probe columns themselves.
bool flat_eligible(const igl_join* j) {
    if (!j->built || j->csr || j->left_join || j->never_match) return false;
    for (const auto& rc : j->right) if (rc.validity) return false;
    return (size_t)flat_stage_bytes(j) * kFlatStages <= (200u << 10);   // very wide probe rows take the tiled kernels
}

// Everything of FlatProbeParams but the probe-side source pointers.
int32_t flat_fill_params(igl_join* j, FlatProbeParams& F, igl_dev_col* out, int64_t nrows) {
    const int nl = (int)j->lt.size(), nr = (int)j->rt.size();
    F.stage_bytes = flat_stage_bytes(j);
    for (uint32_t c = 0, off = 0; c < (uint32_t)nl; ++c) { F.stage_off[c] = off; off += (uint32_t)kFlatChunk << (j->lt[c] == IGL_T_I32 ? 2 : 3); }
    F.n_left = nl; F.n_right = nr; F.key_col = j->lkey; F.key_type = j->lt[j->lkey];
    F.right_key_col = j->rkey; F.payload_col = j->payload_col; F.table = j->table; F.nrows = nrows;
    F.packed = (const uint8_t*)j->packed; F.packed_stride = j->packed_stride;
    memcpy(F.pk_off, j->pk_off, sizeof F.pk_off);
    for (int c = 0; c < nl; ++c) {
        F.lshift[c] = j->lt[c] == IGL_T_I32 ? 2 : 3;
        if (j->lhide[c]) { F.out[c] = nullptr; continue; }
        if (!out[c].data) return fail(IGL_ERR_INVALID_ARG, "output column %d has no data buffer", c);
        F.out[c] = out[c].data;
    }
    for (int c = 0; c < nr; ++c) {
        F.right_data[c] = (const uint8_t*)j->right[c].data; F.rshift[c] = j->rt[c] == IGL_T_I32 ? 2 : 3;
        if (j->rhide[c]) { F.out[nl + c] = nullptr; continue; }
        if (!out[nl + c].data) return fail(IGL_ERR_INVALID_ARG, "output column %d has no data buffer", nl + c);
        F.out[nl + c] = out[nl + c].data;
    }
    return IGL_OK;
}

// First half: scratch, launch, header copy — everything is queued on `st`, nothing waits.
int32_t flat_probe_launch(igl_join* j, FlatProbeParams& F, bool seg, int64_t nrows, long long nchunks, int64_t out_capacity,
                          cudaStream_t st) {
    HostTrace tr_all("probe.flat");
    igl_ctx* ctx = j->ctx;
    const int nl = F.n_left, nr = F.n_right;
    // DENSE form (see the kernel): needs room for every probe row and 32-bit row numbers
    const char* denv = getenv("IGL_JOIN_DENSE");
    const bool dense = seg || (!(denv && denv[0] == '0') && out_capacity >= nrows && nrows < (1ll << 32));
    if (seg && !(out_capacity >= nrows && nrows < (1ll << 32)))
        return fail(IGL_ERR_INVALID_ARG, "segmented probe needs an output capacity of %lld rows (< 2^32)", (long long)nrows);
    constexpr int kFixTile = 1024;   // rows per tile of the hole-filling pass
    const long long ntiles = (nrows + kFixTile - 1) / kFixTile;
    const size_t scan_bytes = (((size_t)scan_scratch_words(tile_offsets_scan_tiles(ntiles)) * 8) + 255) & ~(size_t)255;
    const size_t off_bytes = (((size_t)(ntiles + 1) * 8) + 255) & ~(size_t)255;
    const size_t bm_bytes = (size_t)ntiles * (kFixTile / 8) + 256;
    // scratch: [header][offsets-kernel header][scan words][tile offsets][hit bitmap]
    const size_t need = dense ? 512 + scan_bytes + off_bytes + bm_bytes : 256;
    if (need > j->scratch_bytes) {
        if (j->scratch) pool_free(ctx, j->scratch);
        j->scratch = nullptr; j->scratch_bytes = 0;
        IGL_TRY(pool_alloc(ctx, need, &j->scratch));
        j->scratch_bytes = need;
    }
    uint8_t* sb = (uint8_t*)j->scratch;
    IGL_CUDA_OK(cudaMemsetAsync(sb, 0, dense ? 512 + scan_bytes : 256, st));
    if (dense) IGL_CUDA_OK(cudaMemsetAsync(sb + 512 + scan_bytes + off_bytes, 0, bm_bytes, st));
    F.out_counter = &((PipeHeader*)j->scratch)->aux[1];
    F.ticket = &((PipeHeader*)j->scratch)->ticket;
    F.hit_words = dense ? (uint32_t*)(sb + 512 + scan_bytes + off_bytes) : nullptr;
    F.miss_counter = &((PipeHeader*)j->scratch)->aux[2];
    const void* kfn = flat_kernel_for(F.key_type, dense, seg);
    const size_t fsmem = (size_t)F.stage_bytes * kFlatStages;
    int per_sm = 0;
    {
        HostTrace tr("probe.flat.attr");
        IGL_CUDA_OK(cudaFuncSetAttribute(kfn, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)fsmem));
        IGL_CUDA_OK(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, kfn, kFlatThreads, fsmem));
    }
    HostTrace tr_launch("probe.flat.launch");
    if (const char* e = getenv("IGL_FLAT_CTAS_PER_SM")) per_sm = std::min(per_sm, std::max(1, atoi(e)));   // experiments
    const int grid = (int)std::min<long long>(nchunks, (long long)ctx->sm_count * std::max(per_sm, 1));
    void* kargs[] = {&F};
    cudaEventRecord(j->ev[4], st);
    IGL_CUDA_OK(cudaLaunchKernel(kfn, dim3(grid), dim3(kFlatThreads), kargs, fsmem, st));
    cudaEventRecord(j->ev[5], st);
    j->ev_probe = true;
    ctx->launches++;
    IGL_CUDA_OK(cudaMemcpyAsync(j->host_hdr, j->scratch, sizeof(PipeHeader), cudaMemcpyDeviceToHost, st));
    if (!j->pend_F) j->pend_F = std::make_unique<FlatProbeParams>();
    *j->pend_F = F;
    j->pend_kind = 1; j->pend_nrows = nrows; j->pend_dense = dense;
    return IGL_OK;
}

// Second half: waits for the queued flat probe, fills the holes of a dense output, reports the rows.
int32_t flat_probe_finish(igl_join* j, int64_t* out_rows, cudaStream_t st) {
    igl_ctx* ctx = j->ctx;
    const FlatProbeParams& F = *j->pend_F;
    const int64_t nrows = j->pend_nrows;
    const bool dense = j->pend_dense;
    const int nl = F.n_left, nr = F.n_right;
    j->pend_kind = 0;
    constexpr int kFixTile = 1024;
    const long long ntiles = (nrows + kFixTile - 1) / kFixTile;
    const size_t scan_bytes = (((size_t)scan_scratch_words(tile_offsets_scan_tiles(ntiles)) * 8) + 255) & ~(size_t)255;
    uint8_t* sb = (uint8_t*)j->scratch;
    IGL_CUDA_OK(cudaStreamSynchronize(st));
    if (dense) {
        const int64_t misses = (int64_t)j->host_hdr->aux[2];
        if (misses > 0 && misses < nrows) {   // fill the holes below the matched-row count from the tail
            void* holes = nullptr;
            IGL_TRY(pool_alloc(ctx, (size_t)misses * 4 + 256, &holes));
            unsigned long long* tile_off = (unsigned long long*)(sb + 512 + scan_bytes);
            cudaError_t e = launch_tile_offsets(F.hit_words, kFixTile, ntiles, tile_off, (PipeHeader*)sb,
                                                &((PipeHeader*)(sb + 256))->ticket, (uint64_t*)(sb + 512), ctx->sm_count, st);
            DenseFixParams X{};
            X.hit_words = F.hit_words; X.tile_off = tile_off; X.tile_rows = kFixTile; X.nrows = nrows;
            X.holes = (uint32_t*)holes; X.ncols = nl + nr;
            for (int c = 0; c < nl + nr; ++c) { X.out[c] = F.out[c]; X.shift[c] = c < nl ? F.lshift[c] : F.rshift[c - nl]; }
            const int fgrid = (int)std::min<long long>(((nrows + 31) / 32 + 7) / 8, (long long)ctx->sm_count * 8);
            if (e == cudaSuccess) { dense_fix_kernel<false><<<fgrid, 256, 0, st>>>(X); e = cudaGetLastError(); }
            if (e == cudaSuccess) { dense_fix_kernel<true><<<fgrid, 256, 0, st>>>(X); e = cudaGetLastError(); }
            if (e == cudaSuccess) e = cudaStreamSynchronize(st);
            pool_free(ctx, holes);
            if (e != cudaSuccess) return fail(IGL_ERR_CUDA, "probe hole filling failed: %s", cudaGetErrorString(e));
            ctx->launches += 3;
        }
        if (out_rows) *out_rows = nrows - misses;
        return IGL_OK;
    }
    if (out_rows) *out_rows = (int64_t)j->host_hdr->aux[1];
    return IGL_OK;
}

}  // namespace

namespace igl {

bool join_segments_eligible(const igl_join* j) { return j && flat_eligible(j); }

igl_ctx* join_ctx(const igl_join* j) { return j->ctx; }
void** join_ex_slot(igl_join* j) { return &j->ex; }
igl_join* join_new_shell(igl_ctx* ctx) { igl_join* j = new igl_join(); j->ctx = ctx; return j; }
void join_set_hidden(igl_join* j, const uint8_t* lhide, const uint8_t* rhide) {
    if (lhide) j->lhide.assign(lhide, lhide + j->lt.size());
    if (rhide) j->rhide.assign(rhide, rhide + j->rt.size());
}

// The probe of a segmented probe side (kernels.h): this GPU's and its peers' staging buffers,
// walked run by run.  The output is dense, `out_rows` rows; it needs S.total_rows rows of room.
int32_t join_probe_segments(igl_join* j, const SegSource& S, igl_dev_col* out, int64_t out_capacity, int64_t* out_rows,
                            cudaStream_t st) {
    if (!j || !out || S.nsrc < 1 || S.nsrc > kSegMaxSrc) return fail(IGL_ERR_INVALID_ARG, "join_probe_segments: bad arguments");
    if (!flat_eligible(j)) return fail(IGL_ERR_UNSUPPORTED, "join_probe_segments: this join takes the general probe");
    DeviceGuard g(j->ctx->device);
    cudaEventRecord(j->ev[3], st);
    j->ev_probe = false;
    if (S.nchunks == 0 || S.nsegs == 0) { if (out_rows) *out_rows = 0; return IGL_OK; }
    FlatProbeParams F{};
    const int nl = (int)j->lt.size();
    for (int s = 0; s < S.nsrc; ++s)
        for (int c = 0; c < nl; ++c) {
            if ((uintptr_t)S.cols[s][c] & 15) return fail(IGL_ERR_INVALID_ARG, "segment source %d column %d is not 16-byte aligned", s, c);
            F.seg_left[s][c] = S.cols[s][c];
        }
    F.segs = S.dev_segs; F.nsegs = S.nsegs; F.seg_nchunks = S.nchunks;
    IGL_TRY(flat_fill_params(j, F, out, S.total_rows));
    IGL_TRY(flat_probe_launch(j, F, /*seg=*/true, S.total_rows, S.nchunks, out_capacity, st));
    return flat_probe_finish(j, out_rows, st);
}

}  // namespace igl

extern "C" {

int32_t igl_join_create(igl_ctx* ctx, int32_t join_type, const int32_t* left_types, const uint8_t* left_nullable,
                        int32_t n_left, const int32_t* right_types, const uint8_t* right_nullable, int32_t n_right,
                        const igl_expr* left_key, const igl_expr* right_key, igl_join** out) {
    if (!ctx || !out || !left_types || !right_types || !left_key || !right_key || n_left <= 0 || n_right <= 0)
        return fail(IGL_ERR_INVALID_ARG, "igl_join_create: bad arguments");
    *out = nullptr;
    if (join_type < IGL_JOIN_INNER || join_type > IGL_JOIN_FULL)
        return fail(IGL_ERR_UNSUPPORTED, "Join type %d not implemented", join_type);   // physical_planner.rs:127-129
    if (n_left > kJoinMaxCols || n_right > kJoinMaxCols) return fail(IGL_ERR_UNSUPPORTED, "more than %d columns on one join side", kJoinMaxCols);
    if (!left_key->is_plain_col() || !right_key->is_plain_col())
        return fail(IGL_ERR_UNSUPPORTED, "device join keys must be column references; project computed keys first");
    auto j = std::make_unique<igl_join>();
    j->ctx = ctx;
    j->join_type = join_type;
    // Right == Inner and Full == Left: the reference never emits unmatched right rows.
    j->left_join = join_type == IGL_JOIN_LEFT || join_type == IGL_JOIN_FULL;
    j->lt.assign(left_types, left_types + n_left);
    j->rt.assign(right_types, right_types + n_right);
    j->ln.assign(n_left, 0); j->rn.assign(n_right, 0);
    j->lhide.assign(n_left, 0); j->rhide.assign(n_right, 0);
    if (left_nullable) j->ln.assign(left_nullable, left_nullable + n_left);
    if (right_nullable) j->rn.assign(right_nullable, right_nullable + n_right);
    j->lkey = left_key->nodes[0].col;
    j->rkey = right_key->nodes[0].col;
    if (j->lkey >= n_left || j->rkey >= n_right) return fail(IGL_ERR_INVALID_ARG, "join key column out of range");
    for (int t : j->lt) if (!fixed_width(t)) return fail(IGL_ERR_UNSUPPORTED, "device join columns must be Int32/Int64/Float64 (got type %d)", t);
    for (int t : j->rt) if (!fixed_width(t)) return fail(IGL_ERR_UNSUPPORTED, "device join columns must be Int32/Int64/Float64 (got type %d)", t);
    j->never_match = j->lt[j->lkey] != j->rt[j->rkey];   // the Arrow type name is part of the printed key
    j->has_nulls = j->left_join;
    for (auto v : j->ln) if (v) j->has_nulls = true;
    for (auto v : j->rn) if (v) j->has_nulls = true;
    uint8_t shifts[kJoinMaxCols];
    for (int c = 0; c < n_left; ++c) shifts[c] = j->lt[c] == IGL_T_I32 ? 2 : 3;
    j->geo = pipe_geometry(shifts, n_left, j->T);
    if (!j->geo.ok) return fail(IGL_ERR_UNSUPPORTED, "left side too wide for one tile (%u bytes)", j->geo.stage_bytes);
    j->smem_bytes = (size_t)j->geo.stage_bytes * j->geo.stages;
    DeviceGuard g(ctx->device);
    for (int csr = 0; csr < 2; ++csr) {
        probe_kernel_t k = pick_probe(csr, j->geo.rows_per_lane, j->has_nulls);
        IGL_CUDA_OK(cudaFuncSetAttribute((const void*)k, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)j->smem_bytes));
        int per_sm = 0;
        IGL_CUDA_OK(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, (const void*)k, kPipeThreads, j->smem_bytes));
        if (per_sm < 1) return fail(IGL_ERR_INTERNAL, "join probe kernel does not fit on an SM");
        (csr ? j->grid_cap_m : j->grid_cap_u) = per_sm * ctx->sm_count;
    }
    IGL_CUDA_OK(cudaHostAlloc((void**)&j->host_hdr, sizeof(PipeHeader), cudaHostAllocDefault));
    for (auto& e : j->ev) IGL_CUDA_OK(cudaEventCreate(&e));
    IGL_CUDA_OK(cudaHostAlloc((void**)&j->host_side, sizeof(JoinSide), cudaHostAllocDefault));
    j->valid_planes.assign(n_left + n_right, nullptr);
    *out = j.release();
    return IGL_OK;
}

// The radix fan-out igl_join_build would pick for a build side of `build_rows` rows (0 = it
// would not partition): callers that partition during the exchange use the same value.
int32_t igl_join_partition_bits(igl_ctx* ctx, int64_t build_rows) {
    if (!ctx || build_rows <= 0) return 0;
    const char* env = getenv("IGL_JOIN_PARTITION");
    if (env && env[0] == '0') return 0;
    uint64_t cap = 1024;
    while ((double)cap * ctx->cfg.hash_load_factor < (double)build_rows) cap <<= 1;
    if (cap * sizeof(Slot) < (64ull << 20)) return 0;
    const char* pmb = getenv("IGL_PART_MB");
    const uint64_t target = (uint64_t)(pmb ? atoi(pmb) : 16) << 20;
    int lp = 0;
    while (((cap * sizeof(Slot)) >> lp) > target && lp < 10) ++lp;
    return lp < 4 ? 4 : lp;
}

int32_t igl_join_set_prepartitioned(igl_join* j, int32_t on) {
    if (!j) return fail(IGL_ERR_INVALID_ARG, "igl_join_set_prepartitioned: NULL join");
    if (j->ex) return igl_join_set_prepartitioned(join_ex_inner(j), on);
    j->prepartitioned = on != 0;
    j->prepart_bits = on;
    return IGL_OK;
}

int32_t igl_join_build(igl_join* j, const igl_dev_col* right, int64_t nrows, void* stream) {
    if (!j || !right || nrows < 0) return fail(IGL_ERR_INVALID_ARG, "igl_join_build: bad arguments");
    if (j->ex) return join_ex_build(j, right, nrows, stream);
    if (nrows >= (int64_t)kNoRow) return fail(IGL_ERR_UNSUPPORTED, "build side limited to 2^32-2 rows per table");
    igl_ctx* ctx = j->ctx;
    DeviceGuard g(ctx->device);
    cudaStream_t st = pick_stream(ctx, stream);
    const int nr = (int)j->rt.size();
    for (int c = 0; c < nr; ++c)
        if (right[c].type != j->rt[c]) return fail(IGL_ERR_INVALID_ARG, "right column %d has type %d, join expects %d", c, right[c].type, j->rt[c]);
    j->right.assign(right, right + nr);
    j->n_right_rows = nrows;
    j->counted_keys = nullptr;
    // capacity: power of two >= rows / load factor
    uint64_t cap = 1024;
    while ((double)cap * ctx->cfg.hash_load_factor < (double)nrows) cap <<= 1;
    if (j->d_slots && cap != j->capacity) { pool_free(ctx, j->d_slots); j->d_slots = nullptr; }
    if (!j->d_slots) IGL_TRY(pool_alloc(ctx, cap * sizeof(Slot), &j->d_slots));
    if (!j->d_side) IGL_TRY(pool_alloc(ctx, sizeof(JoinSide), &j->d_side));
    if (j->d_rowids) { pool_free(ctx, j->d_rowids); j->d_rowids = nullptr; }
    j->capacity = cap;
    int lg = 0; while ((1ull << lg) < cap) ++lg;
    j->table.slots = (Slot*)j->d_slots;
    j->table.mask = cap - 1;
    j->table.shift = 64 - lg;
    j->table.side = (JoinSide*)j->d_side;
    j->table.rowids = nullptr;
    const int grid = ctx->sm_count * 8;
    cudaEventRecord(j->ev[0], st);
    // Large tables: reorder the build side by the top hash bits (partition.cu).
    for (void* p : j->part_right) pool_free(ctx, p);
    j->part_right.clear();
    for (void* p : j->csr_right) pool_free(ctx, p);
    j->csr_right.clear();
    j->log2_parts = 0;
    {
        // IGL_JOIN_PARTITION=0 switches it off (A/B measurements in profiles/README.md).
        const char* env = getenv("IGL_JOIN_PARTITION");
        bool want = !(env && env[0] == '0') && cap * sizeof(Slot) >= (64ull << 20) && nrows > 0 && !j->prepartitioned;
        if (!radix_partition_supported(right, nr)) want = false;   // nullable or very wide rows: one big table, unordered probes
        if (want) {
            int lp = 0;
            const char* pmb = getenv("IGL_PART_MB");
            const uint64_t target = (uint64_t)(pmb ? atoi(pmb) : 16) << 20;          // ~16 MB of table per partition
            while (((cap * sizeof(Slot)) >> lp) > target && lp < 10) ++lp;
            if (lp < 4) lp = 4;
            if (!j->part_scratch) IGL_TRY(pool_alloc(ctx, 3 * 1025 * 8, &j->part_scratch));
            std::vector<void*> dst(nr, nullptr);
            for (int c = 0; c < nr; ++c) {
                IGL_TRY(pool_alloc(ctx, (size_t)nrows * (j->rt[c] == IGL_T_I32 ? 4 : 8) + 256, &dst[c]));
                j->part_right.push_back(dst[c]);
            }
            IGL_TRY(radix_partition(ctx, right, nr, j->rkey, nrows, lp, dst.data(), j->part_scratch, st));
            for (int c = 0; c < nr; ++c) j->right[c].data = dst[c];
            j->log2_parts = lp;
        }
    }
    // Large unique build sides are built in shared memory, sub-region by sub-region (join_build_fine_kernel):
    // the rows, ordered by their first hash bits above (or by the exchange), are first ordered down to sub-regions.
    bool fine = false;
    std::vector<void*> level1;   // the first-level copies: not needed once the fine order exists (freed after the stream has drained)
    {
        const int lp1 = j->log2_parts > 0 ? j->log2_parts : (j->prepartitioned && j->prepart_bits >= 2 ? j->prepart_bits : 0);
        const int fine_bits = lg - kFineSlotsLog2;
        const char* fenv = getenv("IGL_JOIN_FINE");
        if (!(fenv && fenv[0] == '0') && lp1 > 0 && nrows > 0 && fine_bits > lp1 && fine_bits - lp1 <= 10 && fine_bits <= 20 &&
            radix_partition_fine_supported(j->right.data(), nr)) {
            const size_t need = 3 * ((size_t(1) << fine_bits) + 1) * 8;
            if (need > j->fine_scratch_bytes) {
                if (j->fine_scratch) pool_free(ctx, j->fine_scratch);
                j->fine_scratch = nullptr; j->fine_scratch_bytes = 0;
                IGL_TRY(pool_alloc(ctx, need, &j->fine_scratch));
                j->fine_scratch_bytes = need;
            }
            std::vector<void*> dst(nr, nullptr);
            level1 = j->part_right;
            j->part_right.clear();
            for (int c = 0; c < nr; ++c) {
                IGL_TRY(pool_alloc(ctx, (size_t)nrows * (j->rt[c] == IGL_T_I32 ? 4 : 8) + 256, &dst[c]));
                j->part_right.push_back(dst[c]);
            }
            IGL_TRY(radix_partition_fine(ctx, j->right.data(), nr, j->rkey, nrows, fine_bits, fine_bits - lp1, dst.data(), j->fine_scratch, st));
            for (int c = 0; c < nr; ++c) j->right[c].data = dst[c];
            fine = true;
        }
    }
    j->fine_built = fine;
    right = j->right.data();
    cudaEventRecord(j->ev[1], st);
    const igl_dev_col& kc = right[j->rkey];
    // payload inlining: (key, one non-null fixed-width column), integer key (exact round trip)
    j->payload_col = -1;
    if (nr == 2 && !right[0].validity && !right[1].validity &&   // what is there decides, not the schema's nullable flags
        (j->rt[j->rkey] == IGL_T_I32 || j->rt[j->rkey] == IGL_T_I64))
        j->payload_col = 1 - j->rkey;
    // (clearing the table on a side stream under the build side's partitioning was measured: both are
    // streaming passes over HBM, the sum of the two phases did not change — 5.87 vs 5.83 ms at C3)
    const void* payload = j->payload_col >= 0 ? right[j->payload_col].data : nullptr;
    const int pshift = j->payload_col >= 0 ? (j->rt[j->payload_col] == IGL_T_I32 ? 2 : 3) : 0xFF;
    auto plain_build = [&]() -> int32_t {
        join_init_kernel<<<grid, 256, 0, st>>>((Slot*)j->d_slots, cap, (JoinSide*)j->d_side);
        IGL_CUDA_OK(cudaGetLastError());
        ctx->launches++;
        if (nrows > 0) {
            join_build_unique_kernel<<<grid, kBuildThreads, 0, st>>>(kc.data, j->rt[j->rkey], kc.validity, payload, pshift, nrows, j->table,
                                                                      &((JoinSide*)j->d_side)->ticket);
            IGL_CUDA_OK(cudaGetLastError());
            ctx->launches++;
        }
        return IGL_OK;
    };
    if (fine) {
        const size_t ovf_cap = (size_t)std::max<int64_t>(4096, nrows / 32);
        if (ovf_cap > j->fine_ovf_cap) {
            if (j->fine_ovf) pool_free(ctx, j->fine_ovf);
            j->fine_ovf = nullptr; j->fine_ovf_cap = 0;
            IGL_TRY(pool_alloc(ctx, ovf_cap * sizeof(Slot), &j->fine_ovf));
            j->fine_ovf_cap = ovf_cap;
        }
        static bool attr_set = false;
        if (!attr_set) {
            IGL_CUDA_OK(cudaFuncSetAttribute((const void*)join_build_fine_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)(kFineSlots * sizeof(Slot))));
            attr_set = true;
        }
        IGL_CUDA_OK(cudaMemsetAsync(j->d_side, 0, sizeof(JoinSide), st));
        const unsigned long long* bin_offsets = (const unsigned long long*)j->fine_scratch + ((size_t(1) << (lg - kFineSlotsLog2)) + 1);
        join_build_fine_kernel<<<(unsigned)(cap >> kFineSlotsLog2), kFineThreads, kFineSlots * sizeof(Slot), st>>>(
            kc.data, j->rt[j->rkey], payload, pshift, bin_offsets, j->table, (Slot*)j->fine_ovf, (unsigned long long)j->fine_ovf_cap);
        join_build_overflow_kernel<<<ctx->sm_count, 256, 0, st>>>((const Slot*)j->fine_ovf, (unsigned long long)j->fine_ovf_cap, j->table);
        IGL_CUDA_OK(cudaGetLastError());
        ctx->launches += 2;
    } else {
        IGL_TRY(plain_build());
    }
    cudaEventRecord(j->ev[2], st);
    j->ev_build = true;
    IGL_CUDA_OK(cudaMemcpyAsync(j->host_side, j->d_side, sizeof(JoinSide), cudaMemcpyDeviceToHost, st));
    IGL_CUDA_OK(cudaStreamSynchronize(st));
    if (fine && j->host_side->ovf_count > (unsigned long long)j->fine_ovf_cap) {
        // more rows ran past their sub-regions than the list holds (a pathological key set): the plain build, over the same rows
        IGL_TRY(plain_build());
        cudaEventRecord(j->ev[2], st);
        IGL_CUDA_OK(cudaMemcpyAsync(j->host_side, j->d_side, sizeof(JoinSide), cudaMemcpyDeviceToHost, st));
        IGL_CUDA_OK(cudaStreamSynchronize(st));
        j->fine_built = false;
    }
    for (void* p : level1) pool_free(ctx, p);
    level1.clear();
    j->csr = j->host_side->dup_flag != 0;
    if (j->packed) { pool_free(ctx, j->packed); j->packed = nullptr; }
    j->packed_stride = 0;
    if (!j->csr && j->payload_col < 0 && nrows > 0 && !(getenv("IGL_JOIN_PACKED") && getenv("IGL_JOIN_PACKED")[0] == '0')) {
        // Row-id payload: the probe would gather every right column of a match from its own array — one
        // random 32-byte sector per column (C4: four per output row).  Rows of up to 32 bytes are packed
        // into records, 8-byte columns first, so one aligned load brings the whole row.
        bool ok = true;
        int bytes = 0;
        PackRowsParams K{};
        for (int c = 0; c < nr; ++c) if (right[c].validity) ok = false;
        for (int pass = 0; pass < 2 && ok; ++pass)
            for (int c = 0; c < nr; ++c) {
                const int sh = j->rt[c] == IGL_T_I32 ? 2 : 3;
                if ((pass == 0) != (sh == 3)) continue;
                K.off[c] = (uint8_t)bytes; K.shift[c] = (uint8_t)sh; K.src[c] = (const uint8_t*)right[c].data;
                bytes += 1 << sh;
            }
        if (ok && bytes <= 32) {
            const int stride = bytes <= 16 ? 16 : 32;
            IGL_TRY(pool_alloc(ctx, (size_t)nrows * stride + 256, &j->packed));
            K.ncols = nr; K.stride = stride; K.n = nrows; K.out = (uint8_t*)j->packed;
            pack_rows_kernel<<<grid, 256, 0, st>>>(K);
            IGL_CUDA_OK(cudaGetLastError());
            ctx->launches++;
            j->packed_stride = stride;
            memcpy(j->pk_off, K.off, sizeof j->pk_off);
        }
    }
    if (j->csr) {   // duplicates: rebuild as CSR runs
        j->payload_col = -1;
        IGL_TRY(pool_alloc(ctx, (size_t)nrows * 4, &j->d_rowids));
        join_init_kernel<<<grid, 256, 0, st>>>((Slot*)j->d_slots, cap, (JoinSide*)j->d_side);
        join_build_count_kernel<<<grid, 256, 0, st>>>(kc.data, j->rt[j->rkey], kc.validity, nrows, j->table);
        join_build_runs_kernel<<<grid, 256, 0, st>>>(j->table, cap);
        join_build_fill_kernel<<<grid, 256, 0, st>>>(kc.data, j->rt[j->rkey], kc.validity, nrows, j->table, (uint32_t*)j->d_rowids);
        IGL_CUDA_OK(cudaGetLastError());
        ctx->launches += 4;
        j->table.rowids = (const uint32_t*)j->d_rowids;
        // Heavy buckets: put the build columns themselves into CSR order (one gather pass over the build
        // side), so that expanding a run reads consecutive rows.  Columns with a validity bitmap keep
        // the row-id indirection.
        bool reorder = nrows > 0 && !(getenv("IGL_JOIN_CSR_REORDER") && getenv("IGL_JOIN_CSR_REORDER")[0] == '0');
        for (int c = 0; c < nr && reorder; ++c) if (right[c].validity) reorder = false;
        for (void* p : j->csr_right) pool_free(ctx, p);
        j->csr_right.clear();
        if (reorder) {
            for (int c = 0; c < nr; ++c) {
                void* p = nullptr;
                IGL_TRY(pool_alloc(ctx, (size_t)nrows * (j->rt[c] == IGL_T_I32 ? 4 : 8) + 256, &p));
                j->csr_right.push_back(p);
                gather_rows_kernel<<<grid, 256, 0, st>>>((const uint8_t*)right[c].data, (uint8_t*)p, (const uint32_t*)j->d_rowids, nrows, j->rt[c] == IGL_T_I32 ? 2 : 3);
            }
            iota_u32_kernel<<<grid, 256, 0, st>>>((uint32_t*)j->d_rowids, nrows);
            IGL_CUDA_OK(cudaGetLastError());
            ctx->launches += nr + 1;
            for (int c = 0; c < nr; ++c) j->right[c].data = j->csr_right[c];
        }
        IGL_CUDA_OK(cudaStreamSynchronize(st));
    }
    j->built = true;
    return IGL_OK;
}

int32_t igl_join_has_duplicates(const igl_join* j) {
    if (j && j->ex) return igl_join_has_duplicates(join_ex_inner(const_cast<igl_join*>(j)));
    return j && j->csr ? 1 : 0;
}

int32_t igl_join_probe_count(igl_join* j, const igl_dev_col* left, int64_t nrows, int64_t* out_rows, void* stream) {
    if (!j || !left || !out_rows || nrows < 0) return fail(IGL_ERR_INVALID_ARG, "igl_join_probe_count: bad arguments");
    if (j->ex) return join_ex_probe_count(j, left, nrows, out_rows, stream);
    if (!j->built) return fail(IGL_ERR_INVALID_ARG, "igl_join_probe_count: build side not finished");
    igl_ctx* ctx = j->ctx;
    DeviceGuard g(ctx->device);
    cudaStream_t st = pick_stream(ctx, stream);
    if (nrows == 0) { *out_rows = 0; return IGL_OK; }
    if (!j->scratch) { IGL_TRY(pool_alloc(ctx, 256, &j->scratch)); j->scratch_bytes = 256; }
    IGL_CUDA_OK(cudaMemsetAsync(j->scratch, 0, 256, st));
    PipeHeader* hdr = (PipeHeader*)j->scratch;
    const igl_dev_col& kc = left[j->lkey];
    join_probe_count_kernel<<<ctx->sm_count * 8, 256, 0, st>>>(kc.data, j->lt[j->lkey], kc.validity, nrows, j->table, j->csr ? 1 : 0,
                                                               j->left_join ? 1 : 0, j->never_match ? 1 : 0, &hdr->aux[0]);
    IGL_CUDA_OK(cudaGetLastError());
    ctx->launches++;
    IGL_CUDA_OK(cudaMemcpyAsync(j->host_hdr, j->scratch, sizeof(PipeHeader), cudaMemcpyDeviceToHost, st));
    IGL_CUDA_OK(cudaStreamSynchronize(st));
    *out_rows = (int64_t)j->host_hdr->aux[0];
    j->counted_keys = kc.data; j->counted_nrows = nrows; j->counted_rows = *out_rows;
    return IGL_OK;
}

}  // extern "C"

namespace igl {

// igl_join_probe in two halves (kernels.h): `begin` queues everything it can on `st` — for the flat
// fast path that is all of the work, nothing waits — and `end` waits, finishes (hole filling) and
// reports the row count.  One probe per join may be pending.  The streaming operator puts the
// staging of its next chunk between the two.
int32_t join_probe_begin(igl_join* j, const igl_dev_col* left, int64_t nrows, igl_dev_col* out, int64_t out_capacity,
                         cudaStream_t st) {
    void* stream = (void*)st;
    if (!j || !left || !out || nrows < 0) return fail(IGL_ERR_INVALID_ARG, "igl_join_probe: bad arguments");
    if (j->ex) {   // wider joins are not split: the whole probe runs here, `end` only reports
        int64_t n = 0;
        IGL_TRY(join_ex_probe(j, left, nrows, out, out_capacity, &n, stream));
        j->pend_kind = 2; j->pend_rows = n;
        return IGL_OK;
    }
    if (!j->built) return fail(IGL_ERR_INVALID_ARG, "igl_join_probe: build side not finished");
    igl_ctx* ctx = j->ctx;
    DeviceGuard g(ctx->device);
    const int nl = (int)j->lt.size(), nr = (int)j->rt.size();
    if (j->pend_kind != 0) return fail(IGL_ERR_INVALID_ARG, "igl_join_probe: the previous probe of this join has not been finished");
    if (nrows == 0) { j->pend_kind = 2; j->pend_rows = 0; return IGL_OK; }
    // capacity check: unique layout emits at most one row per probe row
    if (!j->csr && out_capacity < nrows) {
        int64_t need = 0;
        IGL_TRY(igl_join_probe_count(j, left, nrows, &need, stream));
        if (out_capacity < need) return fail(IGL_ERR_INVALID_ARG, "output capacity %lld < %lld rows", (long long)out_capacity, (long long)need);
    } else if (j->csr) {
        int64_t need = 0;
        // the caller has usually just asked (to size `out`): same key column, same rows, same table
        if (j->counted_keys == left[j->lkey].data && j->counted_nrows == nrows) need = j->counted_rows;
        else IGL_TRY(igl_join_probe_count(j, left, nrows, &need, stream));
        j->counted_keys = nullptr;
        if (out_capacity < need) return fail(IGL_ERR_INVALID_ARG, "output capacity %lld < %lld rows", (long long)out_capacity, (long long)need);
    }
    std::vector<igl_dev_col> pleft;
    cudaEventRecord(j->ev[3], st);
    j->ev_probe = false;
    if (j->log2_parts > 0) {   // reorder the probe chunk the same way, so it walks the table region by region
        HostTrace tr("probe.partition");
        if (radix_partition_supported(left, nl)) {
            if (nrows > j->part_left_rows) {
                for (void* p : j->part_left) pool_free(ctx, p);
                j->part_left.assign(nl, nullptr);
                for (int c = 0; c < nl; ++c) IGL_TRY(pool_alloc(ctx, (size_t)nrows * (j->lt[c] == IGL_T_I32 ? 4 : 8) + 256, &j->part_left[c]));
                j->part_left_rows = nrows;
            }
            IGL_TRY(radix_partition(ctx, left, nl, j->lkey, nrows, j->log2_parts, j->part_left.data(), j->part_scratch, st));
            pleft.assign(left, left + nl);
            for (int c = 0; c < nl; ++c) pleft[c].data = j->part_left[c];
            left = pleft.data();
        }
    }
    // ---- fast path: inner join, unique layout, no validity anywhere ----
    {
        const char* env = getenv("IGL_JOIN_FLAT");
        bool flat = !(env && env[0] == '0') && flat_eligible(j);
        for (int c = 0; c < nl && flat; ++c) if (left[c].validity || ((uintptr_t)left[c].data & 15)) flat = false;
        if (flat) {
            FlatProbeParams F{};
            for (int c = 0; c < nl; ++c) {
                if (left[c].type != j->lt[c]) return fail(IGL_ERR_INVALID_ARG, "left column %d has type %d, join expects %d", c, left[c].type, j->lt[c]);
                F.left[c] = (const uint8_t*)left[c].data;
            }
            IGL_TRY(flat_fill_params(j, F, out, nrows));
            return flat_probe_launch(j, F, /*seg=*/false, nrows, (nrows + kFlatChunk - 1) / kFlatChunk, out_capacity, st);
        }
    }
    JoinProbeParams P{};
    P.T = j->T;
    P.T.nrows = nrows;
    P.T.ntiles = (nrows + P.T.tile_rows - 1) / P.T.tile_rows;
    for (int c = 0; c < nl; ++c) {
        if (left[c].type != j->lt[c]) return fail(IGL_ERR_INVALID_ARG, "left column %d has type %d, join expects %d", c, left[c].type, j->lt[c]);
        if (((uintptr_t)left[c].data & 15) || ((uintptr_t)left[c].validity & 15)) return fail(IGL_ERR_INVALID_ARG, "left column %d is not 16-byte aligned", c);
        P.T.data[c] = (const uint8_t*)left[c].data;
        P.T.valid[c] = j->has_nulls ? left[c].validity : nullptr;
    }
    // listed runs per probe call (more are walked in place): sized with the call — a skewed build side lists
    // a run for every probe row that hits one of its heavy keys
    const unsigned int kGiantCap = (unsigned int)std::min<int64_t>(4 << 20, std::max<int64_t>(64 << 10, nrows / 256)), kHugeCap = 4096;
    const size_t scan_part = (256 + (size_t)scan_scratch_words(P.T.ntiles) * 8 + 255) & ~(size_t)255;
    size_t need = scan_part + (j->csr ? ((size_t)kGiantCap + kHugeCap) * 32 : 0);
    if (need > j->scratch_bytes) {
        if (j->scratch) pool_free(ctx, j->scratch);
        j->scratch = nullptr; j->scratch_bytes = 0;
        IGL_TRY(pool_alloc(ctx, need, &j->scratch));
        j->scratch_bytes = need;
    }
    P.T.hdr = (PipeHeader*)j->scratch;
    P.T.status = (uint64_t*)((uint8_t*)j->scratch + 256);
    IGL_CUDA_OK(cudaMemsetAsync(j->scratch, 0, scan_part, st));
    if (j->csr) {
        P.giant_list = (unsigned long long*)((uint8_t*)j->scratch + scan_part);
        P.giant_count = &P.T.hdr->aux[3];
        P.giant_cap = kGiantCap;
        P.huge_list = P.giant_list + 4ull * kGiantCap;
        P.huge_count = &P.T.hdr->aux[2];
        P.huge_cap = kHugeCap;
    }
    P.table = j->table;
    P.key_slot = j->lkey;
    P.key_type = j->lt[j->lkey];
    P.never_match = j->never_match ? 1 : 0;
    P.n_left = nl; P.n_right = nr;
    P.right_key_col = j->rkey;
    P.payload_col = j->payload_col;
    P.left_join = j->left_join ? 1 : 0;
    // validity planes: a left column that has a bitmap; every right column when it
    // can be null (source nullable, or null-extended by a Left/Full join)
    const int64_t cap_rows = std::max<int64_t>(out_capacity, 1);
    if (cap_rows > j->plane_rows) {
        for (auto& p : j->valid_planes) if (p) { pool_free(ctx, p); p = nullptr; }
        j->plane_rows = cap_rows;
    }
    for (int c = 0; c < nl + nr; ++c) {
        const bool is_left = c < nl;
        const int t = is_left ? j->lt[c] : j->rt[c - nl];
        P.out_shift[c] = t == IGL_T_I32 ? 2 : 3;
        const bool hidden = is_left ? j->lhide[c] != 0 : j->rhide[c - nl] != 0;
        if (!hidden && !out[c].data) return fail(IGL_ERR_INVALID_ARG, "output column %d has no data buffer", c);
        P.out_data[c] = hidden ? nullptr : out[c].data;
        bool want = is_left ? (left[c].validity != nullptr) : (j->left_join || j->right[c - nl].validity != nullptr);
        if (!j->has_nulls || hidden) want = false;
        if (want && !out[c].validity) return fail(IGL_ERR_INVALID_ARG, "output column %d can hold nulls but has no validity buffer", c);
        if (want) {
            if (!j->valid_planes[c]) IGL_TRY(pool_alloc(ctx, (size_t)j->plane_rows, &j->valid_planes[c]));
            P.out_valid[c] = (uint8_t*)j->valid_planes[c];
        } else P.out_valid[c] = nullptr;
        if (!is_left) {
            P.right_data[c - nl] = (const uint8_t*)j->right[c - nl].data;
            P.right_valid[c - nl] = j->right[c - nl].validity;
        }
    }
    probe_kernel_t k = pick_probe(j->csr, j->geo.rows_per_lane, j->has_nulls);
    const int grid = (int)std::min<int64_t>(P.T.ntiles, j->csr ? j->grid_cap_m : j->grid_cap_u);
    k<<<grid, kPipeThreads, j->smem_bytes, st>>>(P);
    IGL_CUDA_OK(cudaGetLastError());
    ctx->launches++;
    if (j->csr) {
        if (j->has_nulls) join_expand_giant_kernel<true><<<ctx->sm_count * 8, 256, 0, st>>>(P);
        else join_expand_giant_kernel<false><<<ctx->sm_count * 8, 256, 0, st>>>(P);
        IGL_CUDA_OK(cudaGetLastError());
        ctx->launches++;
    }
    for (int c = 0; c < nl + nr; ++c) {
        if (!P.out_valid[c]) continue;
        IGL_CUDA_OK(launch_pack_bits(P.out_valid[c], (uint32_t*)out[c].validity, &P.T.hdr->out_count, 0, out_capacity, ctx->sm_count, st));
        ctx->launches++;
    }
    IGL_CUDA_OK(cudaMemcpyAsync(j->host_hdr, j->scratch, sizeof(PipeHeader), cudaMemcpyDeviceToHost, st));
    IGL_CUDA_OK(cudaStreamSynchronize(st));
    if (j->host_hdr->err_flags & 0x80000000u)
        return fail(IGL_ERR_UNSUPPORTED, "one probe tile expands to more than 2^31 output rows (extreme key skew on both sides); probe in smaller chunks");
    j->pend_kind = 2; j->pend_rows = (int64_t)j->host_hdr->out_count;
    return IGL_OK;
}

int32_t join_probe_end(igl_join* j, int64_t* out_rows, cudaStream_t st) {
    if (!j) return fail(IGL_ERR_INVALID_ARG, "join_probe_end: NULL join");
    DeviceGuard g(j->ctx->device);
    if (j->pend_kind == 1) return flat_probe_finish(j, out_rows, st);
    if (j->pend_kind == 2) { j->pend_kind = 0; if (out_rows) *out_rows = j->pend_rows; return IGL_OK; }
    return fail(IGL_ERR_INVALID_ARG, "join_probe_end: no probe is pending");
}

}  // namespace igl

extern "C" {

int32_t igl_join_probe(igl_join* j, const igl_dev_col* left, int64_t nrows, igl_dev_col* out, int64_t out_capacity,
                       int64_t* out_rows, void* stream) {
    if (!j || !left || !out || nrows < 0) return fail(IGL_ERR_INVALID_ARG, "igl_join_probe: bad arguments");
    if (j->ex) return join_ex_probe(j, left, nrows, out, out_capacity, out_rows, stream);
    cudaStream_t st = pick_stream(j->ctx, stream);
    IGL_TRY(join_probe_begin(j, left, nrows, out, out_capacity, st));
    return join_probe_end(j, out_rows, st);
}


// Device time of the last igl_join_build / igl_join_probe on this join, in milliseconds:
// out[0] build-side partitioning, out[1] table init + build kernels, out[2] probe-side
// partitioning, out[3] the probe kernel (fast path only; -1 where nothing was recorded).
int32_t igl_join_phase_ms(igl_join* j, double out[4]) {
    if (!j || !out) return fail(IGL_ERR_INVALID_ARG, "igl_join_phase_ms: NULL argument");
    if (j->ex) return igl_join_phase_ms(join_ex_inner(j), out);
    DeviceGuard g(j->ctx->device);
    for (int i = 0; i < 4; ++i) out[i] = -1.0;
    float ms = 0.f;
    if (j->ev_build && cudaEventSynchronize(j->ev[2]) == cudaSuccess) {
        if (cudaEventElapsedTime(&ms, j->ev[0], j->ev[1]) == cudaSuccess) out[0] = ms;
        if (cudaEventElapsedTime(&ms, j->ev[1], j->ev[2]) == cudaSuccess) out[1] = ms;
    }
    if (j->ev_probe && cudaEventSynchronize(j->ev[5]) == cudaSuccess) {
        if (cudaEventElapsedTime(&ms, j->ev[3], j->ev[4]) == cudaSuccess) out[2] = ms;
        if (cudaEventElapsedTime(&ms, j->ev[4], j->ev[5]) == cudaSuccess) out[3] = ms;
    }
    return IGL_OK;
}

void igl_join_destroy(igl_join* j) {
    if (!j) return;
    if (j->ex) { join_ex_destroy(j); delete j; return; }
    DeviceGuard g(j->ctx->device);
    pool_free(j->ctx, j->d_slots);
    pool_free(j->ctx, j->d_side);
    pool_free(j->ctx, j->d_rowids);
    pool_free(j->ctx, j->scratch);
    pool_free(j->ctx, j->part_scratch);
    pool_free(j->ctx, j->fine_scratch);
    pool_free(j->ctx, j->fine_ovf);
    for (void* p : j->part_right) pool_free(j->ctx, p);
    for (void* p : j->csr_right) pool_free(j->ctx, p);
    pool_free(j->ctx, j->packed);
    for (void* p : j->part_left) pool_free(j->ctx, p);
    for (void* p : j->valid_planes) pool_free(j->ctx, p);
    if (j->host_hdr) cudaFreeHost(j->host_hdr);
    if (j->host_side) cudaFreeHost(j->host_side);
    for (auto& e : j->ev) if (e) cudaEventDestroy(e);
    delete j;
}

}  // extern "C"
