// gen.cu — device-side synthetic data generator and order-independent row
// digest.  Counter based, so any row range of any column can be
// produced on any rank:  u = splitmix64((seed ^ col_id * GOLDEN) + row).
// The digest folds every row to one 64-bit hash and accumulates (sum, xor,
// sum of a second mix); the CPU oracle implements the same arithmetic
// independently (oracle/igloo_oracle.cpp `digest`).
#include "device_utils.cuh"
#include "internal.h"

#include <cmath>

namespace igl {

__device__ __forceinline__ uint64_t gen_u64(uint64_t seed, uint64_t col, uint64_t row) {
    return splitmix64((seed ^ (col * 0x9E3779B97F4A7C15ull)) + row);
}

__host__ __device__ __forceinline__ uint64_t inv_odd(uint64_t a) {
    uint64_t x = a;
    for (int i = 0; i < 6; ++i) x *= 2 - a * x;
    return x;
}
__device__ __forceinline__ uint64_t unxorshift(uint64_t y, int s) {
    uint64_t x = y;
    for (int i = s; i < 64; i += s) x = y ^ (x >> s);
    return x;
}
__device__ __forceinline__ uint64_t perm_pow2(uint64_t x, int bits) {
    const uint64_t M = bits >= 64 ? ~0ull : ((1ull << bits) - 1);
    const int s1 = bits / 2 + 1, s2 = bits / 2, s3 = bits / 2 + 2;
    x = (x * 0x9E3779B97F4A7C15ull) & M; x ^= x >> s1;
    x = (x * 0xBF58476D1CE4E5B9ull) & M; x ^= x >> s2;
    x = (x * 0x94D049BB133111EBull) & M; x ^= x >> s3;
    return x;
}
__device__ __forceinline__ uint64_t perm_pow2_inv(uint64_t x, int bits, uint64_t i1, uint64_t i2, uint64_t i3) {
    const uint64_t M = bits >= 64 ? ~0ull : ((1ull << bits) - 1);
    const int s1 = bits / 2 + 1, s2 = bits / 2, s3 = bits / 2 + 2;
    x = unxorshift(x, s3); x = (x * i3) & M;
    x = unxorshift(x, s2); x = (x * i2) & M;
    x = unxorshift(x, s1); x = (x * i1) & M;
    return x;
}

// C4: lineitem row -> index of its order, 1-7 lines per order, addressable by
// row.  Orders come in blocks of 7 whose line counts are a permutation of 1..7 picked by the
// block's hash (28 lines per block, 4 per order on average); rows past the last whole block go
// round-robin to the remaining orders and then wrap from order 0.
__device__ __forceinline__ uint64_t line_order_index(uint64_t seed, uint64_t col, uint64_t row, uint64_t n_orders) {
    const uint64_t nblocks = n_orders / 7;
    const uint64_t blk = row / 28;
    if (blk >= nblocks) return (7 * nblocks + (row - 28 * nblocks)) % n_orders;
    const uint64_t u = gen_u64(seed, col, blk);
    const uint32_t r = (uint32_t)(u % 7), mul = 1 + (uint32_t)((u >> 8) % 6);   // t -> (t * mul + r) mod 7 permutes 0..6
    uint32_t q = (uint32_t)(row - blk * 28), t = 0;
    for (; t < 6; ++t) {
        const uint32_t c = 1 + (t * mul + r) % 7;
        if (q < c) break;
        q -= c;
    }
    return blk * 7 + t;
}

// C5: Zipf(1.1) rank in [1, m] by inverse CDF.  H(k) = sum_{i<=k} i^-s is
// tabulated exactly for k <= kZipfHead and continued by the midpoint rule
// H(k) = H(K) + ((K + .5)^(1-s) - (k + .5)^(1-s)) / (s - 1) (relative error < 1e-8 here).
constexpr int kZipfHead = 1024;
constexpr double kZipfS = 1.1;
struct ZipfParams { const double* head; double h_total, tail_base; };   // head[k-1] = H(k); tail_base = (K + .5)^(1-s)
__device__ __forceinline__ uint64_t zipf_rank(double u01, uint64_t m, const ZipfParams& Z) {
    const double t = u01 * Z.h_total;
    const int K = (int)min((uint64_t)kZipfHead, m);
    if (t < Z.head[K - 1]) {   // first k with H(k) > t
        int lo = 0, hi = K - 1;
        while (lo < hi) { const int mid = (lo + hi) >> 1; if (Z.head[mid] > t) hi = mid; else lo = mid + 1; }
        return (uint64_t)lo + 1;
    }
    const double y = Z.tail_base - (t - Z.head[K - 1]) * (kZipfS - 1.0);
    double k = floor(pow(y, 1.0 / (1.0 - kZipfS)) + 0.5);
    if (!(k >= (double)(K + 1))) k = (double)(K + 1);
    if (k > (double)m) k = (double)m;
    return (uint64_t)k;
}

__global__ void gen_kernel(int kind, uint64_t seed, uint64_t col, long long row0, long long n, uint64_t m, int bits,
                           const long long* __restrict__ keys, void* __restrict__ out, uint64_t i1, uint64_t i2,
                           uint64_t i3, ZipfParams Z) {
    const long long stride = (long long)gridDim.x * blockDim.x;
    for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += stride) {
        // kinds 1-3 and 6 take their row number from keys[i] when keys are given ("the value the
        // generator gives row keys[i]": expected join payloads without a hash table)
        const uint64_t row = (keys && kind != IGL_GEN_I64_LOOKUP) ? (uint64_t)keys[i] : (uint64_t)(row0 + i);
        switch (kind) {
            case IGL_GEN_I64_TPCH_ORDERKEY:
                ((int64_t*)out)[i] = (int64_t)((((row >> 3) << 5) | (row & 7)) + 1);
                break;
            case IGL_GEN_I64_LINE_ORDER:
                ((int64_t*)out)[i] = (int64_t)line_order_index(seed, col, row, m);
                break;
            case IGL_GEN_I64_ZIPF: {
                const double u01 = (double)(gen_u64(seed, col, row) >> 11) * 0x1.0p-53;
                const uint64_t r = zipf_rank(u01, m, Z) - 1;
                ((int64_t*)out)[i] = (int64_t)(bits > 0 ? perm_pow2(r, bits) : r);
                break;
            }
            case IGL_GEN_I32_UNIFORM:
                ((int32_t*)out)[i] = (int32_t)__umul64hi(gen_u64(seed, col, row), m);
                break;
            case IGL_GEN_I64_UNIFORM: {
                uint64_t u = gen_u64(seed, col, row);
                ((int64_t*)out)[i] = (int64_t)(m ? __umul64hi(u, m) : u);
                break;
            }
            case IGL_GEN_F64_UNIT:
                ((double*)out)[i] = (double)(gen_u64(seed, col, row) >> 11) * 0x1.0p-53;
                break;
            case IGL_GEN_I64_PERM:
                ((int64_t*)out)[i] = (int64_t)perm_pow2(row, bits);
                break;
            case IGL_GEN_I64_LOOKUP: {
                uint64_t src = perm_pow2_inv((uint64_t)keys[i], bits, i1, i2, i3);
                uint64_t u = gen_u64(seed, col, src);
                ((int64_t*)out)[i] = (int64_t)(m ? __umul64hi(u, m) : u);
                break;
            }
        }
    }
}

struct DigestCols {
    const uint8_t* data[32];
    const uint8_t* valid[32];
    uint8_t type[32];
    int n;
};

__global__ void digest_kernel(const __grid_constant__ DigestCols C, long long nrows, unsigned long long* acc) {
    uint64_t s = 0, x = 0, s2 = 0;
    const long long stride = (long long)gridDim.x * blockDim.x;
    for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < nrows; i += stride) {
        uint64_t h = 0x243F6A8885A308D3ull;
        for (int j = 0; j < C.n; ++j) {
            uint64_t bits;
            const bool ok = !C.valid[j] || ((C.valid[j][i >> 3] >> (i & 7)) & 1);
            if (!ok) bits = 0x9E3779B97F4A7C15ull ^ (uint64_t)j;
            else switch (C.type[j]) {
                case IGL_T_BOOL: bits = (C.data[j][i >> 3] >> (i & 7)) & 1; break;
                case IGL_T_I32: bits = (uint64_t)(int64_t)((const int32_t*)C.data[j])[i]; break;
                default: bits = ((const uint64_t*)C.data[j])[i]; break;
            }
            h = splitmix64(h ^ (bits + (uint64_t)j * 0xD6E8FEB86659FD93ull));
        }
        s += h; x ^= h; s2 += splitmix64(h ^ 0xA5A5A5A5A5A5A5A5ull);
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
        s += __shfl_xor_sync(0xffffffffu, s, o);
        x ^= __shfl_xor_sync(0xffffffffu, x, o);
        s2 += __shfl_xor_sync(0xffffffffu, s2, o);
    }
    if ((threadIdx.x & 31) == 0) {
        atomicAdd(acc + 0, (unsigned long long)s);
        atomicXor(acc + 1, (unsigned long long)x);
        atomicAdd(acc + 2, (unsigned long long)s2);
    }
}

}  // namespace igl

using namespace igl;

extern "C" {

int32_t igl_gen_column(igl_ctx* ctx, int32_t kind, uint64_t seed, uint64_t col_id, int64_t row0, int64_t n,
                       uint64_t m, int32_t bits, const int64_t* dev_keys, void* dev_out, void* stream) {
    if (!ctx || !dev_out || n < 0) return fail(IGL_ERR_INVALID_ARG, "igl_gen_column: bad arguments");
    if (kind < IGL_GEN_I32_UNIFORM || kind > IGL_GEN_I64_ZIPF) return fail(IGL_ERR_INVALID_ARG, "igl_gen_column: unknown kind %d", kind);
    if ((kind == IGL_GEN_I64_PERM || kind == IGL_GEN_I64_LOOKUP) && (bits < 2 || bits > 62))
        return fail(IGL_ERR_INVALID_ARG, "igl_gen_column: bits must be in [2, 62]");
    if (kind == IGL_GEN_I64_LOOKUP && !dev_keys) return fail(IGL_ERR_INVALID_ARG, "igl_gen_column: LOOKUP needs keys");
    if (kind == IGL_GEN_I64_LINE_ORDER && m < 1) return fail(IGL_ERR_INVALID_ARG, "igl_gen_column: LINE_ORDER needs the order count in m");
    if (kind == IGL_GEN_I64_ZIPF && (m < 1 || bits < 0 || bits > 62 || (bits > 0 && m > (1ull << bits))))
        return fail(IGL_ERR_INVALID_ARG, "igl_gen_column: ZIPF needs a domain size m >= 1 (and m <= 2^bits when permuted)");
    if (n == 0) return IGL_OK;
    DeviceGuard g(ctx->device);
    cudaStream_t st = pick_stream(ctx, stream);
    ZipfParams Z{};
    void* head = nullptr;
    if (kind == IGL_GEN_I64_ZIPF) {
        std::vector<double> h(kZipfHead);
        double acc = 0;
        for (int k = 1; k <= kZipfHead; ++k) { acc += pow((double)k, -kZipfS); h[k - 1] = acc; }
        const int K = (int)std::min<uint64_t>(kZipfHead, m);
        Z.tail_base = pow((double)K + 0.5, 1.0 - kZipfS);
        Z.h_total = m <= (uint64_t)kZipfHead ? h[m - 1] : h[K - 1] + (Z.tail_base - pow((double)m + 0.5, 1.0 - kZipfS)) / (kZipfS - 1.0);
        IGL_TRY(pool_alloc(ctx, sizeof(double) * kZipfHead, &head));
        cudaError_t e = cudaMemcpyAsync(head, h.data(), sizeof(double) * kZipfHead, cudaMemcpyHostToDevice, st);
        if (e == cudaSuccess) e = cudaStreamSynchronize(st);   // `h` is pageable and goes out of scope
        if (e != cudaSuccess) { pool_free(ctx, head); return fail(IGL_ERR_CUDA, "igl_gen_column: %s", cudaGetErrorString(e)); }
        Z.head = (const double*)head;
    }
    int grid = (int)std::min<int64_t>((n + 255) / 256, (int64_t)ctx->sm_count * 16);
    gen_kernel<<<grid, 256, 0, st>>>(kind, seed, col_id, row0, n, m, bits, (const long long*)dev_keys,
                                     dev_out, inv_odd(0x9E3779B97F4A7C15ull),
                                     inv_odd(0xBF58476D1CE4E5B9ull), inv_odd(0x94D049BB133111EBull), Z);
    cudaError_t le = cudaGetLastError();
    if (head) { if (le == cudaSuccess) le = cudaStreamSynchronize(st); pool_free(ctx, head); }
    if (le != cudaSuccess) return fail(IGL_ERR_CUDA, "igl_gen_column: %s", cudaGetErrorString(le));
    ctx->launches++;
    return IGL_OK;
}

int32_t igl_digest_rows(igl_ctx* ctx, const igl_dev_col* cols, int32_t ncols, int64_t nrows, uint64_t out[3],
                        void* stream) {
    if (!ctx || !cols || !out || ncols < 1 || ncols > 32 || nrows < 0) return fail(IGL_ERR_INVALID_ARG, "igl_digest_rows: bad arguments");
    DeviceGuard g(ctx->device);
    cudaStream_t st = pick_stream(ctx, stream);
    DigestCols C{};
    C.n = ncols;
    for (int j = 0; j < ncols; ++j) {
        if (cols[j].type < IGL_T_BOOL || cols[j].type > IGL_T_F64) return fail(IGL_ERR_UNSUPPORTED, "igl_digest_rows: fixed-width columns only");
        C.data[j] = (const uint8_t*)cols[j].data;
        C.valid[j] = cols[j].validity;
        C.type[j] = (uint8_t)cols[j].type;
    }
    void* acc = nullptr;
    IGL_TRY(pool_alloc(ctx, 32, &acc));
    cudaError_t e = cudaMemsetAsync(acc, 0, 32, st);
    if (e == cudaSuccess && nrows > 0) {
        int grid = (int)std::min<int64_t>((nrows + 255) / 256, (int64_t)ctx->sm_count * 16);
        digest_kernel<<<grid, 256, 0, st>>>(C, nrows, (unsigned long long*)acc);
        e = cudaGetLastError();
        ctx->launches++;
    }
    if (e == cudaSuccess) e = cudaMemcpyAsync(out, acc, 24, cudaMemcpyDeviceToHost, st);
    if (e == cudaSuccess) e = cudaStreamSynchronize(st);
    pool_free(ctx, acc);
    if (e != cudaSuccess) return fail(IGL_ERR_CUDA, "igl_digest_rows: %s", cudaGetErrorString(e));
    return IGL_OK;
}

}  // extern "C"
