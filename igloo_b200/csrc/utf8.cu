// utf8.cu — K2s: Utf8 pass-through for Filter / Projection.
//
// filter_record_batch (filter.rs:57) compacts EVERY column, including strings
// (C1: `SELECT name, age FROM users WHERE age > 25`, README.md:27).  The fused
// kernel emits the surviving rows' input indices; a string column is then
// gathered in three steps: lengths -> exclusive scan (new offsets) -> bytes.
#include "device_utils.cuh"
#include "internal.h"
#include "kernels.h"

namespace igl {

// out_off[i + 1] = length of source row rowid[i]  (out_off[0] = 0); validity gathered to bytes.
__global__ void utf8_lengths_kernel(const int32_t* __restrict__ src_off, const uint8_t* __restrict__ src_valid,
                                    const int32_t* __restrict__ rowid, long long n, int32_t* __restrict__ out_off,
                                    uint8_t* __restrict__ out_valid_bytes) {
    const long long stride = (long long)gridDim.x * blockDim.x;
    for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += stride) {
        const int32_t r = rowid ? rowid[i] : (int32_t)i;
        if (r < 0) {   // no source row (the null-extended side of a Left join): an empty, NULL string
            out_off[i + 1] = 0;
            if (out_valid_bytes) out_valid_bytes[i] = 0;
            continue;
        }
        out_off[i + 1] = src_off[r + 1] - src_off[r];
        if (out_valid_bytes) out_valid_bytes[i] = src_valid ? ((src_valid[r >> 3] >> (r & 7)) & 1) : 1;
    }
    if (blockIdx.x == 0 && threadIdx.x == 0) out_off[0] = 0;
}

// In-place inclusive scan of v[1..n] (so v becomes the offsets array), one tile
// of 2048 elements per CTA iteration, ordered by the same look-back as the
// row-compaction kernels.
constexpr int kScanThreads = 256;
constexpr int kScanPerThread = 8;
constexpr int kScanTile = kScanThreads * kScanPerThread;

__global__ void __launch_bounds__(kScanThreads) scan_offsets_kernel(int32_t* v, long long n, uint64_t* status,
                                                                    unsigned int* ticket) {
    __shared__ long long warp_sums[kScanThreads / 32];
    __shared__ long long s_base;
    __shared__ long long s_tile;
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const long long ntiles = (n + kScanTile - 1) / kScanTile;
    while (true) {
        if (threadIdx.x == 0) s_tile = (long long)atomicAdd(ticket, 1u);
        __syncthreads();
        const long long tile = s_tile;
        if (tile >= ntiles) break;
        const long long base_i = tile * kScanTile + (long long)threadIdx.x * kScanPerThread;
        long long x[kScanPerThread];
        long long run = 0;
#pragma unroll
        for (int k = 0; k < kScanPerThread; ++k) {
            const long long i = base_i + k;
            run += (i < n) ? (long long)v[1 + i] : 0;
            x[k] = run;
        }
        long long inc = run;
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) {
            const long long t = __shfl_up_sync(0xffffffffu, inc, o);
            if (lane >= o) inc += t;
        }
        if (lane == 31) warp_sums[warp] = inc;
        __syncthreads();
        long long wprefix = 0, total = 0;
#pragma unroll
        for (int w = 0; w < kScanThreads / 32; ++w) {
            if (w < warp) wprefix += warp_sums[w];
            total += warp_sums[w];
        }
        if (warp == 0) {
            if (lane == 0) scan_publish(status, ntiles, tile, (uint64_t)total);
            __syncwarp();
            const uint64_t b = scan_exclusive(status, ntiles, tile, (uint64_t)total, lane);
            if (lane == 0) s_base = (long long)b;
        }
        __syncthreads();
        const long long off = s_base + wprefix + inc - run;
#pragma unroll
        for (int k = 0; k < kScanPerThread; ++k) {
            const long long i = base_i + k;
            if (i < n) v[1 + i] = (int32_t)(off + x[k]);
        }
        __syncthreads();
    }
}

__global__ void utf8_gather_kernel(const int32_t* __restrict__ src_off, const uint8_t* __restrict__ src_bytes,
                                   const int32_t* __restrict__ rowid, long long n, const int32_t* __restrict__ out_off,
                                   uint8_t* __restrict__ out_bytes) {
    // one warp per output string: short strings dominate, a warp copies 32 bytes per step
    const int lane = threadIdx.x & 31;
    const long long warps = ((long long)gridDim.x * blockDim.x) >> 5;
    for (long long i = ((long long)blockIdx.x * blockDim.x + threadIdx.x) >> 5; i < n; i += warps) {
        const int32_t r = rowid ? rowid[i] : (int32_t)i;
        if (r < 0) continue;
        const int32_t s = src_off[r], len = src_off[r + 1] - s, d = out_off[i];
        for (int32_t b = lane; b < len; b += 32) out_bytes[d + b] = src_bytes[s + b];
    }
}

__global__ void iota_i32_kernel(int32_t* out, long long n) {
    const long long stride = (long long)gridDim.x * blockDim.x;
    for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += stride) out[i] = (int32_t)i;
}
__global__ void mask_rowids_kernel(int32_t* rowid, const uint8_t* __restrict__ valid_bits, long long n) {
    const long long stride = (long long)gridDim.x * blockDim.x;
    for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += stride)
        if (!((valid_bits[i >> 3] >> (i & 7)) & 1)) rowid[i] = -1;
}
int32_t launch_iota_i32(igl_ctx* ctx, int32_t* out, long long n, cudaStream_t st) {
    if (n <= 0) return IGL_OK;
    iota_i32_kernel<<<(int)std::min<long long>((n + 255) / 256, (long long)ctx->sm_count * 8), 256, 0, st>>>(out, n);
    IGL_CUDA_OK(cudaGetLastError());
    ctx->launches++;
    return IGL_OK;
}
// rowid[i] = -1 wherever bit i of `valid_bits` is clear (row ids of null-extended join rows)
int32_t launch_mask_rowids(igl_ctx* ctx, int32_t* rowid, const uint8_t* valid_bits, long long n, cudaStream_t st) {
    if (n <= 0) return IGL_OK;
    mask_rowids_kernel<<<(int)std::min<long long>((n + 255) / 256, (long long)ctx->sm_count * 8), 256, 0, st>>>(rowid, valid_bits, n);
    IGL_CUDA_OK(cudaGetLastError());
    ctx->launches++;
    return IGL_OK;
}

// ---- Utf8 join keys (hash_join.rs:117-122: the key is the printed string; here: a 64-bit hash of
// the bytes joined on as an Int64 key, every joined pair then checked byte for byte) ----

// out[i] = hash of row i's bytes (FNV-1a over the bytes, length folded in, finished with splitmix64).
// `bits` < 64 keeps only that many low bits: a test switch that makes different strings collide.
__global__ void utf8_hash_kernel(const int32_t* __restrict__ off, const uint8_t* __restrict__ bytes, long long n, int bits,
                                 int64_t* __restrict__ out) {
    const long long stride = (long long)gridDim.x * blockDim.x;
    const uint64_t mask = bits >= 64 ? ~0ull : ((1ull << bits) - 1);
    for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += stride) {
        const int32_t s = off[i], e = off[i + 1];
        uint64_t h = 0xCBF29CE484222325ull ^ ((uint64_t)(uint32_t)(e - s) * 0x9E3779B97F4A7C15ull);
        for (int32_t b = s; b < e; ++b) h = (h ^ bytes[b]) * 0x100000001B3ull;
        out[i] = (int64_t)(splitmix64(h) & mask);
    }
}

// keep[i] &= (left string of joined row i == right string of joined row i), NULL == NULL
// (hash_join.rs:120 prints NULL as "null" on both sides); null-extended rows of a Left join
// (bit i of r_out_valid clear) are kept.  *mismatches counts the rows cleared.
__global__ void utf8_verify_kernel(const int32_t* __restrict__ lrow, const int32_t* __restrict__ rrow,
                                   const uint8_t* __restrict__ r_out_valid, long long n,
                                   const int32_t* __restrict__ loff, const uint8_t* __restrict__ lbytes, const uint8_t* __restrict__ lvalid,
                                   const int32_t* __restrict__ roff, const uint8_t* __restrict__ rbytes, const uint8_t* __restrict__ rvalid,
                                   int32_t* __restrict__ keep, unsigned long long* __restrict__ mismatches) {
    const long long stride = (long long)gridDim.x * blockDim.x;
    unsigned long long bad = 0;
    for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += stride) {
        if (r_out_valid && !((r_out_valid[i >> 3] >> (i & 7)) & 1)) continue;
        const int32_t l = lrow[i], r = rrow[i];
        const bool lnull = lvalid && !((lvalid[l >> 3] >> (l & 7)) & 1);
        const bool rnull = rvalid && !((rvalid[r >> 3] >> (r & 7)) & 1);
        bool same;
        if (lnull || rnull) same = lnull && rnull;
        else {
            const int32_t ls = loff[l], len = loff[l + 1] - ls, rs = roff[r];
            same = len == roff[r + 1] - rs;
            for (int32_t b = 0; same && b < len; ++b) same = lbytes[ls + b] == rbytes[rs + b];
        }
        if (!same && keep[i]) { keep[i] = 0; ++bad; }
    }
    if (bad) atomicAdd(mismatches, bad);
}

__global__ void fill_i32_kernel(int32_t* out, long long n, int32_t v) {
    const long long stride = (long long)gridDim.x * blockDim.x;
    for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += stride) out[i] = v;
}

int32_t utf8_hash_keys(igl_ctx* ctx, const int32_t* off, const uint8_t* bytes, long long n, int bits, int64_t* out, cudaStream_t st) {
    if (n <= 0) return IGL_OK;
    utf8_hash_kernel<<<(int)std::min<long long>((n + 255) / 256, (long long)ctx->sm_count * 8), 256, 0, st>>>(off, bytes, n, bits, out);
    IGL_CUDA_OK(cudaGetLastError());
    ctx->launches++;
    return IGL_OK;
}

int32_t launch_fill_i32(igl_ctx* ctx, int32_t* out, long long n, int32_t v, cudaStream_t st) {
    if (n <= 0) return IGL_OK;
    fill_i32_kernel<<<(int)std::min<long long>((n + 255) / 256, (long long)ctx->sm_count * 8), 256, 0, st>>>(out, n, v);
    IGL_CUDA_OK(cudaGetLastError());
    ctx->launches++;
    return IGL_OK;
}

int32_t utf8_verify_keys(igl_ctx* ctx, const int32_t* lrow, const int32_t* rrow, const uint8_t* r_out_valid, long long n,
                         const int32_t* loff, const uint8_t* lbytes, const uint8_t* lvalid,
                         const int32_t* roff, const uint8_t* rbytes, const uint8_t* rvalid,
                         int32_t* keep, unsigned long long* mismatches, cudaStream_t st) {
    if (n <= 0) return IGL_OK;
    utf8_verify_kernel<<<(int)std::min<long long>((n + 255) / 256, (long long)ctx->sm_count * 8), 256, 0, st>>>(
        lrow, rrow, r_out_valid, n, loff, lbytes, lvalid, roff, rbytes, rvalid, keep, mismatches);
    IGL_CUDA_OK(cudaGetLastError());
    ctx->launches++;
    return IGL_OK;
}

// Gathers a Utf8 column by row ids.  out_off must hold n+1 entries, scratch
// 8 * ceil(n / 2048) + 256 bytes (zeroed here).  Returns after enqueueing; the
// total byte count is out_off[n] on the device.
int32_t utf8_gather_offsets(igl_ctx* ctx, const int32_t* src_off, const uint8_t* src_valid, const int32_t* rowid,
                            long long n, int32_t* out_off, uint8_t* out_valid_bytes, void* scratch, size_t scratch_bytes,
                            cudaStream_t st) {
    const long long ntiles = (n + kScanTile - 1) / kScanTile;
    const size_t need = 256 + (size_t)scan_scratch_words(ntiles) * 8;
    if (scratch_bytes < need) return fail(IGL_ERR_INTERNAL, "utf8 gather scratch too small");
    IGL_CUDA_OK(cudaMemsetAsync(scratch, 0, need, st));
    const int grid = (int)std::min<long long>((n + 255) / 256 + 1, (long long)ctx->sm_count * 8);
    utf8_lengths_kernel<<<grid, 256, 0, st>>>(src_off, src_valid, rowid, n, out_off, out_valid_bytes);
    IGL_CUDA_OK(cudaGetLastError());
    ctx->launches++;
    if (n > 0) {
        const int sgrid = (int)std::min<long long>(ntiles, (long long)ctx->sm_count * 4);
        scan_offsets_kernel<<<sgrid, kScanThreads, 0, st>>>(out_off, n, (uint64_t*)((uint8_t*)scratch + 256), (unsigned int*)scratch);
        IGL_CUDA_OK(cudaGetLastError());
        ctx->launches++;
    }
    return IGL_OK;
}

int32_t utf8_gather_bytes(igl_ctx* ctx, const int32_t* src_off, const uint8_t* src_bytes, const int32_t* rowid,
                          long long n, const int32_t* out_off, uint8_t* out_bytes, cudaStream_t st) {
    if (n == 0) return IGL_OK;
    const int grid = (int)std::min<long long>((n * 32 + 255) / 256, (long long)ctx->sm_count * 16);
    utf8_gather_kernel<<<grid, 256, 0, st>>>(src_off, src_bytes, rowid, n, out_off, out_bytes);
    IGL_CUDA_OK(cudaGetLastError());
    ctx->launches++;
    return IGL_OK;
}

}  // namespace igl
