// tile_pipe.cuh — the CTA skeleton shared by every row-streaming kernel
// (filter/project, join probe):
//
//   control warp  : takes tile tickets in order, streams each tile's column
//                   slices into a shared-memory ring with 1-D bulk async copies
//                   (cp.async.bulk -> mbarrier complete_tx, the TMA engine), and
//                   — when the kernel compacts or expands rows — turns per-warp
//                   output counts into global output offsets with an ordered
//                   single-pass scan (decoupled look-back over tile status words)
//   compute warps : wait for a stage, do the kernel's work on 4 rows per lane
//                   (lane i owns rows i, i+32, i+64, i+96 of its warp's 128-row
//                   group), hand their counts to the control warp, receive their
//                   output base, write, release the stage.
//
// Scanning kernels are software-pipelined: a compute warp runs the counting
// phase of tile t+1 BEFORE the writing phase of tile t, so the look-back of
// tile t (a chain of L2 round trips behind every tile in flight on the GPU)
// overlaps useful work instead of stalling the CTA at a barrier.  The named
// barriers alternate by tile parity so a warp that is one tile ahead cannot
// alias the previous tile's handshake.
//
// Tiles are claimed with an atomic ticket, so a tile's predecessors are always
// owned by CTAs that are already running: the look-back cannot deadlock however
// many CTAs are resident.
#pragma once

#include "device_utils.cuh"

namespace igl {

#ifndef IGL_PIPE_COMPUTE_WARPS
#define IGL_PIPE_COMPUTE_WARPS 15
#endif
#ifndef IGL_PIPE_MIN_CTAS
#define IGL_PIPE_MIN_CTAS 2
#endif
constexpr int kPipeComputeWarps = IGL_PIPE_COMPUTE_WARPS;   // + 1 control warp
constexpr int kPipeThreads = (kPipeComputeWarps + 1) * 32;
// rows per tile = compute warps * 32 lanes * R rows per lane (R = 4, or 1 for very wide
// tables); always a multiple of 128 rows so bitmap slices stay 16-byte aligned
constexpr int kPipeMinCtas = IGL_PIPE_MIN_CTAS;
constexpr int kPipeMaxStages = 8;
constexpr int kPipeMaxCols = 16;
static_assert(kPipeComputeWarps <= 32, "per-tile warp counts are scanned by one warp");

struct PipeHeader {   // first 256 bytes of a kernel's scratch block; zeroed before every launch
    unsigned int ticket;
    unsigned int err_flags;
    long long out_count;
    unsigned long long aux[4];
};

struct PipeCols {
    const uint8_t* data[kPipeMaxCols];
    const uint8_t* valid[kPipeMaxCols];
    uint32_t smem_off[kPipeMaxCols];
    uint32_t valid_off[kPipeMaxCols];
    uint8_t shift[kPipeMaxCols];   // log2(bytes per row); 0xFF = bit-packed boolean
    uint8_t n_cols, n_stages, sub_tiles, pad;
    uint32_t stage_bytes;
    int sub_rows;                  // active warps * 32 * rows-per-lane, a multiple of 128: what the CTA works on at a time
    int tile_rows;                 // sub_tiles * sub_rows: what one stage (one round of bulk copies) holds
    int active_warps;              // compute warps that own rows (all of them at 4 rows per lane)
    long long nrows, ntiles;
    // optional per-sub-tile table (8 bytes per sub-tile, e.g. output offsets) whose entries for a
    // tile are copied into the stage with the columns: aux_smem = stage offset of the copy, which
    // starts at the even entry at or below the tile's first one
    const unsigned long long* aux_table;
    uint32_t aux_smem;
    PipeHeader* hdr;
    uint64_t* status;              // one word per tile (only when SCAN)
};

struct PipeShared {
    uint64_t full_bar[kPipeMaxStages], empty_bar[kPipeMaxStages];
    long long tile_of_stage[kPipeMaxStages];
    long long counts[2][kPipeComputeWarps];
    long long local[2][kPipeComputeWarps];
    unsigned long long base[2];
};

enum { kBarCounts = 1, kBarBase = 3 };   // named barrier ids (+ tile parity)

template <bool SCAN, bool HAS_NULLS>
struct TilePipe {
    const PipeCols& T;
    PipeShared& sh;
    uint8_t* smem;
    int lane;

    __device__ TilePipe(const PipeCols& t, PipeShared& s, uint8_t* dyn_smem)
        : T(t), sh(s), smem(dyn_smem), lane(threadIdx.x & 31) {}

    __device__ void init() {   // whole CTA
        if (threadIdx.x == 0) {
            for (int s = 0; s < T.n_stages; ++s) { mbar_init(&sh.full_bar[s], 1); mbar_init(&sh.empty_bar[s], kPipeComputeWarps); }
            mbar_fence_init();
        }
        __syncthreads();
    }

    // ---- control warp ---------------------------------------------------------
    // Tile claims.  The first n_stages tiles of a CTA are taken round-robin
    // (stage s of CTA c = tile s * gridDim + c) and later ones from an atomic
    // ticket as stages free up: consecutive tiles then always sit on different
    // CTAs that reach them at about the same time.  (Claiming a CTA's first
    // stages as consecutive tickets would make its neighbour's first tile wait
    // for this CTA's LAST prefetched tile — a scan chain as long as the grid.)
    // The ticket for the NEXT refill is drawn as soon as the current one is used, so the
    // atomic's round trip to L2 overlaps the wait for a free stage.
    unsigned int next_ticket = 0;

    __device__ void issue(int s, bool prologue) {   // one lane
        const long long t = prologue ? (long long)s * gridDim.x + blockIdx.x
                                     : (long long)T.n_stages * gridDim.x + (long long)next_ticket;
        if (!prologue || s == T.n_stages - 1) next_ticket = atomicAdd(&T.hdr->ticket, 1u);
        if (t >= T.ntiles) {
            sh.tile_of_stage[s] = -1;
            mbar_arrive(&sh.full_bar[s]);
            return;
        }
        sh.tile_of_stage[s] = t;
        const long long row_begin = t * T.tile_rows;
        const long long rows = min((long long)T.tile_rows, T.nrows - row_begin);
        uint8_t* stage = smem + (size_t)s * T.stage_bytes;
        const uint32_t bit_bytes = ((uint32_t)((rows + 7) >> 3) + 15u) & ~15u;
        auto col_bytes = [&](int c) {
            const int shf = T.shift[c];
            return shf == 0xFF ? bit_bytes : (((uint32_t)(rows << shf) + 15u) & ~15u);
        };
        uint32_t total = 0;
        for (int c = 0; c < T.n_cols; ++c) total += col_bytes(c) + ((HAS_NULLS && T.valid[c]) ? bit_bytes : 0u);
        const long long aux_first = (t * T.sub_tiles) & ~1ll;
        const uint32_t aux_bytes = T.aux_table ? ((uint32_t)(T.sub_tiles + 1) * 8u + 15u) & ~15u : 0u;
        total += aux_bytes;
        mbar_arrive_expect_tx(&sh.full_bar[s], total);
        if (T.aux_table) bulk_g2s(stage + T.aux_smem, T.aux_table + aux_first, aux_bytes, &sh.full_bar[s]);
        for (int c = 0; c < T.n_cols; ++c) {
            const int shf = T.shift[c];
            const size_t goff = shf == 0xFF ? (size_t)(row_begin >> 3) : ((size_t)row_begin << shf);
            bulk_g2s(stage + T.smem_off[c], T.data[c] + goff, col_bytes(c), &sh.full_bar[s]);
            if (HAS_NULLS && T.valid[c])
                bulk_g2s(stage + T.valid_off[c], T.valid[c] + (row_begin >> 3), bit_bytes, &sh.full_bar[s]);
        }
    }

    // Scans this tile's per-warp counts (in sh.counts[it & 1]) and publishes the
    // tile's own count; returns {inclusive per-warp scan, tile total}.
    __device__ void take_counts(int it, long long tile, long long& c, long long& inc, long long& total) {
        bar_sync(kBarCounts + (it & 1), kPipeThreads);
        c = lane < kPipeComputeWarps ? sh.counts[it & 1][lane] : 0;
        inc = c;
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) {
            const long long n = __shfl_up_sync(0xffffffffu, inc, o);
            if (lane >= o) inc += n;
        }
        total = __shfl_sync(0xffffffffu, inc, 31);
        if (lane == 0) {
            if (total >> 31) atomicOr(&T.hdr->err_flags, 0x80000000u);   // tile counts are published in 31 bits
            scan_publish(T.status, T.ntiles, tile, (uint32_t)total);
        }
    }

    __device__ void control_loop() {   // the whole control warp
        const int S = T.n_stages;
        if (lane == 0)
            for (int s = 0; s < S; ++s) issue(s, true);
        __syncwarp();
        long long c = 0, inc = 0, total = 0;
        long long tile = *reinterpret_cast<volatile long long*>(&sh.tile_of_stage[0]);
        if (SCAN && tile >= 0) take_counts(0, tile, c, inc, total);
        for (int it = 0; tile >= 0; ++it) {
            const int s = it % S;
            const uint32_t ph = (uint32_t)(it / S) & 1u;
            if (SCAN) {
                const uint64_t base = scan_exclusive(T.status, T.ntiles, tile, (uint32_t)total, lane);
                if (lane < kPipeComputeWarps) sh.local[it & 1][lane] = inc - c;
                if (lane == 0) {
                    sh.base[it & 1] = base;
                    if (tile == T.ntiles - 1) T.hdr->out_count = (long long)(base + (uint64_t)total);
                }
                __syncwarp();
                bar_arrive(kBarBase + (it & 1), kPipeThreads);
            }
            // The compute warps post the counts of tile it+1 before they write tile
            // it: publish that count now, a whole write phase before its turn, so
            // the neighbours' look-backs rarely find it missing.
            const long long next_tile = *reinterpret_cast<volatile long long*>(&sh.tile_of_stage[(it + 1) % S]);
            if (SCAN && next_tile >= 0) take_counts(it + 1, next_tile, c, inc, total);
            if (lane == 0) {
                mbar_wait(&sh.empty_bar[s], ph);   // every compute warp has finished reading stage s
                issue(s, false);
            }
            __syncwarp();
            tile = next_tile;
        }
    }

    // ---- compute warps ----------------------------------------------------------
    // Waits for iteration `it`'s stage.  Returns false at end of input.
    __device__ bool acquire(int it, const uint8_t*& stage, long long& row_begin, int& rows_in_tile) {
        const int s = it % T.n_stages;
        mbar_wait(&sh.full_bar[s], (uint32_t)(it / T.n_stages) & 1u);
        const long long tile = sh.tile_of_stage[s];
        if (tile < 0) return false;
        stage = smem + (size_t)s * T.stage_bytes;
        row_begin = tile * T.tile_rows;
        rows_in_tile = (int)min((long long)T.tile_rows, T.nrows - row_begin);
        return true;
    }

    // This warp's output count for tile `it` (does not block).
    __device__ void post_count(int it, int warp, long long wcount) {
        if (lane == 0) sh.counts[it & 1][warp] = wcount;
        __syncwarp();
        bar_arrive(kBarCounts + (it & 1), kPipeThreads);
    }

    // This warp's global output base for tile `it` (ordered).
    __device__ long long wait_base(int it, int warp) {
        bar_sync(kBarBase + (it & 1), kPipeThreads);
        return (long long)sh.base[it & 1] + sh.local[it & 1][warp];
    }

    __device__ void release(int it) {
        __syncwarp();
        if (lane == 0) mbar_arrive(&sh.empty_bar[it % T.n_stages]);
    }
};

#ifndef __CUDACC_RTC__
// ---- host-side ring geometry ------------------------------------------------------
struct PipeGeometry {
    int stages = 2;
    int rows_per_lane = 4;
    uint32_t stage_bytes = 128;
    bool ok = true;
};

// shift[c] per staged column; every column reserves a validity slice too.
// `extra` = other dynamic shared memory the kernel needs per CTA.  Picks 4 rows
// per lane when two stages fit (two CTAs per SM if possible), else 1 row per lane.
// `max_sub` > 1 lets a narrow stage hold several sub-tiles, so the fixed cost of a stage
// hand-over (ticket, copies, barriers) is paid once per sub_tiles * sub_rows rows.
inline PipeGeometry pipe_geometry(const uint8_t* shift, int ncols, PipeCols& T, uint32_t extra = 0, int force_r = 0,
                                  int max_sub = 1) {
    PipeGeometry g;
    const uint32_t budget2 = (226u << 10) / kPipeMinCtas - 2048u;   // per CTA when kPipeMinCtas share an SM
    const uint32_t budget1 = 220u << 10;                             // one CTA per SM
    for (int R : {4, 1}) {
        if (force_r && R != force_r) continue;
        // 1 row per lane: only a multiple of 4 warps own rows, so the tile stays a multiple of 128 rows
        const int active = R == 4 ? kPipeComputeWarps : (kPipeComputeWarps / 4) * 4;
        const uint32_t sub_rows = (uint32_t)active * 32u * R;
        T.active_warps = active;
        T.sub_rows = (int)sub_rows;
        auto layout = [&](uint32_t rows) {
            uint32_t off = 0;
            for (int c = 0; c < ncols; ++c) {
                const uint32_t b = shift[c] == 0xFF ? rows / 8 : rows << shift[c];
                T.shift[c] = shift[c];
                T.smem_off[c] = off; off += (b + 127u) & ~127u;
                T.valid_off[c] = off; off += ((rows / 8) + 127u) & ~127u;
            }
            return off ? off : 128u;
        };
        int sub = 1;
        for (int k = max_sub; k > 1; k >>= 1)   // the most sub-tiles that still leave three stages for two CTAs per SM
            if ((layout(sub_rows * k) + 128u) * 3 + extra <= budget2) { sub = k; break; }
        g.stage_bytes = layout(sub_rows * sub);
        T.aux_smem = g.stage_bytes;      // 128 bytes behind the columns, used when aux_table is set
        g.stage_bytes += 128u;
        g.rows_per_lane = R;
        T.sub_tiles = (uint8_t)sub;
        T.tile_rows = (int)(sub_rows * sub);
        if (g.stage_bytes * 2 + extra <= budget2) {
            const uint32_t st = (budget2 - extra) / g.stage_bytes;
            g.stages = (int)(st < (uint32_t)kPipeMaxStages ? st : kPipeMaxStages);
            g.ok = true;
            break;
        }
        if (g.stage_bytes * 2 + extra <= budget1) { g.stages = 2; g.ok = true; break; }
        g.ok = false;
    }
    T.n_cols = (uint8_t)ncols;
    T.n_stages = (uint8_t)g.stages;
    T.stage_bytes = g.stage_bytes;
    return g;
}

#endif  // !__CUDACC_RTC__

}  // namespace igl
