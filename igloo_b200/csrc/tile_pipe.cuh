// tile_pipe.cuh — the CTA skeleton shared by every row-streaming kernel
// (filter/project, join probe, join probe-count):
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
// Tiles are claimed with an atomic ticket, so a tile's predecessors are always
// owned by CTAs that are already running: the look-back cannot deadlock however
// many CTAs are resident.
#pragma once

#include "device_utils.cuh"

namespace igl {

constexpr int kPipeComputeWarps = 7;   // + 1 control warp = 256 threads: 2 CTAs/SM at <= 128 registers
constexpr int kPipeThreads = (kPipeComputeWarps + 1) * 32;
constexpr int kPipeRowsPerSub = kPipeComputeWarps * 128;   // 896: a multiple of 128 keeps bitmap slices 16-B aligned
constexpr int kPipeMaxStages = 4;
constexpr int kPipeMaxCols = 16;

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
    uint8_t n_cols, n_stages, pad[2];
    uint32_t stage_bytes;
    long long nrows, ntiles;
    PipeHeader* hdr;
    uint64_t* status;              // one word per tile (only when SCAN)
};

template <int SUBS>
struct PipeShared {
    uint64_t full_bar[kPipeMaxStages], empty_bar[kPipeMaxStages];
    long long tile_of_stage[kPipeMaxStages];
    long long counts[2][SUBS * kPipeComputeWarps];
    long long local[2][SUBS * kPipeComputeWarps];
    unsigned long long base[2];
};

template <int SUBS, bool SCAN, bool HAS_NULLS>
struct TilePipe {
    static constexpr int TILE = SUBS * kPipeRowsPerSub;
    static constexpr int NC = SUBS * kPipeComputeWarps;
    static_assert(NC <= 32, "per-tile warp counts are scanned by one warp");

    const PipeCols& T;
    PipeShared<SUBS>& sh;
    uint8_t* smem;
    int lane;

    __device__ TilePipe(const PipeCols& t, PipeShared<SUBS>& s, uint8_t* dyn_smem)
        : T(t), sh(s), smem(dyn_smem), lane(threadIdx.x & 31) {}

    __device__ void init() {   // whole CTA
        if (threadIdx.x == 0) {
            for (int s = 0; s < T.n_stages; ++s) { mbar_init(&sh.full_bar[s], 1); mbar_init(&sh.empty_bar[s], kPipeComputeWarps); }
            mbar_fence_init();
        }
        __syncthreads();
    }

    // ---- control warp ---------------------------------------------------------
    __device__ void issue(int s) {   // one lane
        const long long t = (long long)atomicAdd(&T.hdr->ticket, 1u);
        if (t >= T.ntiles) {
            sh.tile_of_stage[s] = -1;
            mbar_arrive(&sh.full_bar[s]);
            return;
        }
        sh.tile_of_stage[s] = t;
        const long long row_begin = t * TILE;
        const long long rows = min((long long)TILE, T.nrows - row_begin);
        uint8_t* stage = smem + (size_t)s * T.stage_bytes;
        const uint32_t bit_bytes = ((uint32_t)((rows + 7) >> 3) + 15u) & ~15u;
        auto col_bytes = [&](int c) {
            const int shf = T.shift[c];
            return shf == 0xFF ? bit_bytes : (((uint32_t)(rows << shf) + 15u) & ~15u);
        };
        uint32_t total = 0;
        for (int c = 0; c < T.n_cols; ++c) total += col_bytes(c) + ((HAS_NULLS && T.valid[c]) ? bit_bytes : 0u);
        mbar_arrive_expect_tx(&sh.full_bar[s], total);
        for (int c = 0; c < T.n_cols; ++c) {
            const int shf = T.shift[c];
            const size_t goff = shf == 0xFF ? (size_t)(row_begin >> 3) : ((size_t)row_begin << shf);
            bulk_g2s(stage + T.smem_off[c], T.data[c] + goff, col_bytes(c), &sh.full_bar[s]);
            if (HAS_NULLS && T.valid[c])
                bulk_g2s(stage + T.valid_off[c], T.valid[c] + (row_begin >> 3), bit_bytes, &sh.full_bar[s]);
        }
    }

    __device__ void control_loop() {   // the whole control warp
        const int S = T.n_stages;
        if (lane == 0)
            for (int s = 0; s < S; ++s) issue(s);
        __syncwarp();
        for (int it = 0;; ++it) {
            const int s = it % S;
            const uint32_t ph = (uint32_t)(it / S) & 1u;
            const long long tile = *reinterpret_cast<volatile long long*>(&sh.tile_of_stage[s]);
            if (tile < 0) break;
            if (SCAN) {
                bar_sync(1, kPipeThreads);   // this tile's per-warp counts are in sh.counts[it & 1]
                const long long c = lane < NC ? sh.counts[it & 1][lane] : 0;
                long long inc = c;
#pragma unroll
                for (int o = 1; o < 32; o <<= 1) {
                    const long long n = __shfl_up_sync(0xffffffffu, inc, o);
                    if (lane >= o) inc += n;
                }
                const long long total = __shfl_sync(0xffffffffu, inc, 31);
                const uint64_t base = lookback_exclusive(T.status, tile, (uint64_t)total, lane);
                if (lane < NC) sh.local[it & 1][lane] = inc - c;
                if (lane == 0) {
                    sh.base[it & 1] = base;
                    if (tile == T.ntiles - 1) T.hdr->out_count = (long long)(base + (uint64_t)total);
                }
                __syncwarp();
                bar_arrive(2, kPipeThreads);
            }
            if (lane == 0) {
                mbar_wait(&sh.empty_bar[s], ph);   // every compute warp has finished reading stage s
                issue(s);
            }
            __syncwarp();
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
        row_begin = tile * TILE;
        rows_in_tile = (int)min((long long)TILE, T.nrows - row_begin);
        return true;
    }

    // Per-warp output counts in, per-warp global output base out (ordered).
    __device__ void offsets(int it, int warp, const long long (&wcount)[SUBS], long long (&wbase)[SUBS]) {
        if (lane == 0) {
#pragma unroll
            for (int sub = 0; sub < SUBS; ++sub) sh.counts[it & 1][sub * kPipeComputeWarps + warp] = wcount[sub];
        }
        __syncwarp();
        bar_arrive(1, kPipeThreads);
        bar_sync(2, kPipeThreads);
        const unsigned long long base = sh.base[it & 1];
#pragma unroll
        for (int sub = 0; sub < SUBS; ++sub) wbase[sub] = (long long)base + sh.local[it & 1][sub * kPipeComputeWarps + warp];
    }

    __device__ void release(int it) {
        __syncwarp();
        if (lane == 0) mbar_arrive(&sh.empty_bar[it % T.n_stages]);
    }
};

// ---- host-side ring geometry ------------------------------------------------------
struct PipeGeometry {
    int subs = 2, stages = 2;
    uint32_t stage_bytes = 128;
    bool ok = true;
};

// shift[c] per staged column; every column reserves a validity slice too.
inline PipeGeometry pipe_geometry(const uint8_t* shift, int ncols, PipeCols& T, uint32_t budget = 100u << 10) {
    auto stage_bytes_for = [&](int subs) {
        uint32_t off = 0;
        for (int c = 0; c < ncols; ++c) {
            const uint32_t rows = subs * kPipeRowsPerSub;
            const uint32_t b = shift[c] == 0xFF ? rows / 8 : rows << shift[c];
            off += (b + 127u) & ~127u;
            off += ((rows / 8) + 127u) & ~127u;
        }
        return off ? off : 128u;
    };
    PipeGeometry g;
    g.stage_bytes = stage_bytes_for(2);
    if (g.stage_bytes * 3 <= budget) g.stages = (int)(budget / g.stage_bytes < (uint32_t)kPipeMaxStages ? budget / g.stage_bytes : kPipeMaxStages);
    else {
        g.subs = 1;
        g.stage_bytes = stage_bytes_for(1);
        if (g.stage_bytes * 2 <= budget) g.stages = (int)(budget / g.stage_bytes < (uint32_t)kPipeMaxStages ? budget / g.stage_bytes : kPipeMaxStages);
        else if (g.stage_bytes * 2 <= (200u << 10)) g.stages = 2;
        else g.ok = false;
    }
    uint32_t off = 0;
    for (int c = 0; c < ncols; ++c) {
        const uint32_t rows = g.subs * kPipeRowsPerSub;
        const uint32_t b = shift[c] == 0xFF ? rows / 8 : rows << shift[c];
        T.shift[c] = shift[c];
        T.smem_off[c] = off; off += (b + 127u) & ~127u;
        T.valid_off[c] = off; off += ((rows / 8) + 127u) & ~127u;
    }
    T.n_cols = (uint8_t)ncols;
    T.n_stages = (uint8_t)g.stages;
    T.stage_bytes = g.stage_bytes;
    return g;
}

}  // namespace igl
