// kernels.h — launchers shared between translation units.
#pragma once

#include <cuda_runtime.h>
#include <stdint.h>

#include "../../include/igloo_cuda.h"

namespace igl {

// byte-per-row {0,1} -> Arrow bitmap (LSB first).  The row count is read from
// *n_ptr on the device when n_ptr != NULL (a count produced by the same stream),
// else n_fixed.  `cap_rows` bounds the grid.
cudaError_t launch_pack_bits(const uint8_t* bytes, uint32_t* bits, const long long* n_ptr, long long n_fixed,
                             long long cap_rows, int sm_count, cudaStream_t st);

// Popcount of every tile's slice of a row bitmap -> exclusive offsets (filter_project.cu):
// off[t] = set bits before tile t, off[ntiles] = hdr->out_count = all set bits; resets
// hdr->ticket.  `status` = scan_scratch_words(tile_offsets_scan_tiles(ntiles)) zeroed words,
// *ticket zero.  rows_per_tile is a multiple of 128; the bitmap is padded to whole tiles.
struct PipeHeader;
cudaError_t launch_tile_offsets(const void* bitmap, int rows_per_tile, long long ntiles, unsigned long long* off,
                                PipeHeader* hdr, unsigned int* ticket, uint64_t* status, int sm_count, cudaStream_t st);
long long tile_offsets_scan_tiles(long long ntiles);

// igl_fp_create with two internal extras used by the streaming operators: `skip`
// hides input columns from a pass-through filter, `with_rowid` appends an Int32
// output holding the input row index of each surviving row.
int32_t fp_create_ex(igl_ctx* ctx, const int32_t* in_types, const uint8_t* in_nullable, int32_t n_in,
                     const igl_expr* pred, const igl_expr* const* proj, int32_t nproj, const uint8_t* skip,
                     bool with_rowid, igl_fp** out);

// Utf8 gather by row ids (utf8.cu): offsets (lengths + ordered scan), then bytes.
int32_t utf8_gather_offsets(igl_ctx* ctx, const int32_t* src_off, const uint8_t* src_valid, const int32_t* rowid,
                            long long n, int32_t* out_off, uint8_t* out_valid_bytes, void* scratch, size_t scratch_bytes,
                            cudaStream_t st);
int32_t utf8_gather_bytes(igl_ctx* ctx, const int32_t* src_off, const uint8_t* src_bytes, const int32_t* rowid,
                          long long n, const int32_t* out_off, uint8_t* out_bytes, cudaStream_t st);

int32_t launch_iota_i32(igl_ctx* ctx, int32_t* out, long long n, cudaStream_t st);
int32_t launch_mask_rowids(igl_ctx* ctx, int32_t* rowid, const uint8_t* valid_bits, long long n, cudaStream_t st);
// Utf8 join keys: hash of each row's bytes as an Int64 key column; byte-for-byte check of the joined pairs
int32_t utf8_hash_keys(igl_ctx* ctx, const int32_t* off, const uint8_t* bytes, long long n, int bits, int64_t* out, cudaStream_t st);
int32_t utf8_verify_keys(igl_ctx* ctx, const int32_t* lrow, const int32_t* rrow, const uint8_t* r_out_valid, long long n,
                         const int32_t* loff, const uint8_t* lbytes, const uint8_t* lvalid,
                         const int32_t* roff, const uint8_t* rbytes, const uint8_t* rvalid,
                         int32_t* keep, unsigned long long* mismatches, cudaStream_t st);
int32_t launch_fill_i32(igl_ctx* ctx, int32_t* out, long long n, int32_t v, cudaStream_t st);

// Radix partition of a column set by the top bits of the join hash (partition.cu).
// dst[c]: device buffers of nrows rows; scratch: 3 * ((1 << log2_parts) + 1) * 8 bytes.
bool radix_partition_supported(const igl_dev_col* cols, int ncols);
int32_t radix_partition(igl_ctx* ctx, const igl_dev_col* cols, int ncols, int key_col, long long nrows, int log2_parts,
                        void* const* dst, void* scratch, cudaStream_t st);

// (destination x partition) split without host synchronisation (partition.cu; driven by shuffle.cu).
// HotKeys: a small device-resident set of canonical keys that go to the fixed destination `dest`
// instead of the one their hash names (table == NULL: none); `ndest` destinations exist, the hash
// picks among the first `nhash`.
struct HotKeys { const unsigned long long* table = nullptr; int mask = 0; int dest = 0; };
int32_t radix_histogram_async(igl_ctx* ctx, const igl_dev_col* key, long long nrows, int ndest, int nhash, int log2_parts,
                              const HotKeys& hot, unsigned long long* dev_hist, cudaStream_t st);
bool radix_partition_fine_supported(const igl_dev_col* cols, int ncols);
int32_t radix_partition_fine(igl_ctx* ctx, const igl_dev_col* cols, int ncols, int key_col, long long nrows, int fine_bits, int fine_lp2,
                             void* const* dst, void* scratch, cudaStream_t st);
int32_t radix_scatter_bins_async(igl_ctx* ctx, const igl_dev_col* cols, int ncols, int key_col, long long nrows, int ndest, int nhash,
                                 int log2_parts, const HotKeys& hot, void* const* dst, const unsigned long long* dev_offsets,
                                 unsigned long long* dev_cursor, cudaStream_t st, bool yielding = false);
// canonical form of every (nrows / nsamples)-th key
int32_t sample_keys_async(igl_ctx* ctx, const igl_dev_col* key, long long nrows, int nsamples, unsigned long long* dev_out, cudaStream_t st);

// ---- segmented probe input (join.cu; used by the multi-GPU join, shuffle.cu) ---------------------
// The probe side as a list of contiguous runs that may live in different buffers — the staging
// buffers of this and the peer GPUs.  Runs are walked in list order; each starts on a 128-row
// boundary of the (virtual) output, so `chunk0` is its first 128-row chunk.
constexpr int kSegMaxSrc = 16;       // up to 8 ranks' staging buffers + a local prefetched copy of each
constexpr int kShufMaxRanks = 8;
constexpr int kSegChunkRows = 128;   // rows per chunk of the flat probe kernel (join.cu asserts it)
constexpr int kSegAlignRows = 4;     // a run starts on a multiple of this many rows of its source: 16 bytes for 4-byte columns
struct ProbeSeg {
    uint32_t chunk0;    // first chunk of this run in the chunk space (strictly increasing, run 0 starts at 0)
    uint32_t rows;      // > 0
    uint64_t row0_src;  // low 56 bits: first row inside the source's columns (a multiple of 4); top 8 bits: source index
    uint64_t out0;      // rows of all earlier runs: where this run's rows go in the (dense) output
    uint64_t pad;
};
struct SegSource {
    const uint8_t* cols[kSegMaxSrc][16];   // [source][left column]
    int nsrc;
    const ProbeSeg* dev_segs;
    int nsegs;
    long long nchunks;                     // chunks in total (every run rounded up to whole chunks)
    long long total_rows;                  // rows in total: the output capacity needed
};
// igl_join_probe in two halves: `begin` queues the probe on `st` (the flat fast path does not
// wait for anything), `end` waits, finishes and reports the rows.  One pending probe per join.
int32_t join_probe_begin(igl_join* j, const igl_dev_col* left, int64_t nrows, igl_dev_col* out, int64_t out_capacity,
                         cudaStream_t st);
int32_t join_probe_end(igl_join* j, int64_t* out_rows, cudaStream_t st);
// True when `j` (built) can take segmented input: inner join, unique build keys, no validity.
bool join_segments_eligible(const igl_join* j);
int32_t join_probe_segments(igl_join* j, const SegSource& S, igl_dev_col* out, int64_t out_capacity, int64_t* out_rows,
                            cudaStream_t st);

// ---- wider joins (join_ex.cu): several key expressions, computed keys, Boolean columns — all
// reduced to the core join over an augmented column set with a hidden key column.
igl_ctx* join_ctx(const igl_join* j);
void** join_ex_slot(igl_join* j);
igl_join* join_new_shell(igl_ctx* ctx);
void join_set_hidden(igl_join* j, const uint8_t* lhide, const uint8_t* rhide);
igl_join* join_ex_inner(igl_join* j);
int32_t join_ex_build(igl_join* j, const igl_dev_col* right, int64_t nrows, void* stream);
int32_t join_ex_probe_count(igl_join* j, const igl_dev_col* left, int64_t nrows, int64_t* out_rows, void* stream);
int32_t join_ex_probe(igl_join* j, const igl_dev_col* left, int64_t nrows, igl_dev_col* out, int64_t out_capacity,
                      int64_t* out_rows, void* stream);
void join_ex_destroy(igl_join* j);

}  // namespace igl
