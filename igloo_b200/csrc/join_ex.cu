// join_ex.cu — igl_join_create_ex: everything HashJoinExec::new accepts beyond one key column
// over fixed-width columns, reduced to the core join (join.cu).
//
// Reference semantics (crates/engine/src/operators/hash_join.rs):
//   :35-41    new(left, right, left_keys: Vec<PhysicalExpr>, right_keys: Vec<..>, join_type)
//   :109-112  build: every right key expression is evaluated per batch        (computed keys)
//   :117-122  the row's key = the printed forms of ALL key values, concatenated (several keys)
//   :125,229  whole batch rows are stored and emitted: any column type rides along
//   physical_planner.rs:106-120 pushes one (left, right) pair per `join.on` entry.
//
// How each case reaches the core join, which wants ONE key column and 4/8-byte columns:
//   computed key        the key expressions are evaluated by the fused projection kernel
//                       (filter_project.cu) into temporary columns
//   one key             the (possibly computed) key column is appended as a HIDDEN column: joined on,
//                       never emitted (join.cu skips output columns without a buffer)
//   two Int32 keys,     packed exactly into one hidden Int64: (a << 32) | b — no collisions, every
//   never null          join type
//   other key lists     a 64-bit hash of the canonical key values (type tags and NULL markers
//                       included) is the hidden key; since a hash can collide, the joined rows are
//                       then filtered by the real condition, key by key and null-safe
//                       (a = b OR (a IS NULL AND b IS NULL)) with the fused filter kernel.  A Left /
//                       Full join must also know when ALL matches of a probe row were false ones:
//                       both sides then carry a row-number column through the join, true matches are
//                       counted per probe row, and the first false pair of a row without any true
//                       match is turned into its null-extended row (the others are dropped).
//   Boolean columns     widened to Int32 on the way in, packed back to bits on the way out
//   key types differ    the printed key carries the Arrow type: such a pair never matches (the core
//                       join's own rule, so one mismatching pair is handed to it as the key)
// Utf8 columns stay with the streaming operator (igl_op_hash_join_create), which owns its
// output buffers and gathers strings by row id.
#include "internal.h"
#include "kernels.h"
#include "device_utils.cuh"

#include <algorithm>
#include <memory>

namespace igl {

constexpr int kExMaxKeys = 4;

__global__ void bool_to_i32_kernel(const uint8_t* __restrict__ bits, long long n, int32_t* __restrict__ out) {
    const long long stride = (long long)gridDim.x * blockDim.x;
    for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += stride) out[i] = (bits[i >> 3] >> (i & 7)) & 1;
}

// one thread per 32-row word
__global__ void i32_to_bool_kernel(const int32_t* __restrict__ v, long long n, uint32_t* __restrict__ bits) {
    const long long stride = (long long)gridDim.x * blockDim.x;
    const long long nwords = (n + 31) >> 5;
    for (long long w = (long long)blockIdx.x * blockDim.x + threadIdx.x; w < nwords; w += stride) {
        uint32_t x = 0;
        const long long base = w << 5;
        for (int b = 0; b < 32 && base + b < n; ++b) x |= (uint32_t)(v[base + b] & 1) << b;
        bits[w] = x;
    }
}

__global__ void pack2_i32_kernel(const int32_t* __restrict__ a, const int32_t* __restrict__ b, long long n, int64_t* __restrict__ out) {
    const long long stride = (long long)gridDim.x * blockDim.x;
    for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += stride)
        out[i] = (int64_t)(((uint64_t)(uint32_t)a[i] << 32) | (uint32_t)b[i]);
}

struct HashKeyCols {
    const uint8_t* data[kExMaxKeys];
    const uint8_t* valid[kExMaxKeys];
    int type[kExMaxKeys];
    int n;
};

// Canonical value per key (join.cu canon_key: Int32 sign-extended, every NaN one NaN, Boolean 0/1),
// a marker for NULL, the type and the position folded in; equal key lists give equal hashes.
__global__ void hash_keys_kernel(const __grid_constant__ HashKeyCols K, long long n, int bits, int64_t* __restrict__ out) {
    const long long stride = (long long)gridDim.x * blockDim.x;
    const uint64_t mask = bits >= 64 ? ~0ull : ((1ull << bits) - 1);   // IGL_JOIN_KEY_HASH_BITS: a test switch that forces collisions
    for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += stride) {
        uint64_t h = 0x9E3779B97F4A7C15ull;
        for (int k = 0; k < K.n; ++k) {
            const bool ok = !K.valid[k] || ((K.valid[k][i >> 3] >> (i & 7)) & 1);
            uint64_t v;
            if (!ok) v = 0xA24BAED4963EE407ull;
            else switch (K.type[k]) {
                case IGL_T_BOOL: v = (K.data[k][i >> 3] >> (i & 7)) & 1; break;
                case IGL_T_I32: v = (uint64_t)(int64_t)((const int32_t*)K.data[k])[i]; break;
                case IGL_T_F64: {
                    v = ((const uint64_t*)K.data[k])[i];
                    if ((v & 0x7FFFFFFFFFFFFFFFull) > 0x7FF0000000000000ull) v = 0x7FF8000000000000ull;
                    break;
                }
                default: v = ((const uint64_t*)K.data[k])[i]; break;
            }
            h = splitmix64(h ^ (v + (uint64_t)(K.type[k] * 16 + k + (ok ? 0 : 8)) * 0xD6E8FEB86659FD93ull));
        }
        out[i] = (int64_t)(h & mask);
    }
}

// ---- Left / Full join on a hashed key list: which joined pairs survive the real condition ----
__device__ __forceinline__ bool bit_at(const uint8_t* b, long long i) { return (b[i >> 3] >> (i & 7)) & 1; }

__global__ void lj_count_true_kernel(const int32_t* __restrict__ lrow, const uint8_t* __restrict__ matched, const uint8_t* __restrict__ ok_bits,
                                     const uint8_t* __restrict__ ok_valid, long long n, int32_t* __restrict__ true_cnt) {
    const long long stride = (long long)gridDim.x * blockDim.x;
    for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += stride)
        if (bit_at(matched, i) && bit_at(ok_bits, i) && (!ok_valid || bit_at(ok_valid, i))) atomicAdd(&true_cnt[lrow[i]], 1);
}

// keep[i]: the row stays.  nullify[i]: it stays as the null-extended row of its probe row (the first false pair of a
// probe row none of whose matches was a true one).
__global__ void lj_decide_kernel(const int32_t* __restrict__ lrow, const uint8_t* __restrict__ matched, const uint8_t* __restrict__ ok_bits,
                                 const uint8_t* __restrict__ ok_valid, long long n, const int32_t* __restrict__ true_cnt,
                                 int32_t* __restrict__ claimed, int32_t* __restrict__ keep, uint8_t* __restrict__ nullify) {
    const long long stride = (long long)gridDim.x * blockDim.x;
    for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += stride) {
        int k = 1, z = 0;
        if (bit_at(matched, i) && !(bit_at(ok_bits, i) && (!ok_valid || bit_at(ok_valid, i)))) {
            const int32_t l = lrow[i];
            if (true_cnt[l] == 0 && atomicExch(&claimed[l], 1) == 0) z = 1; else k = 0;
        }
        keep[i] = k; nullify[i] = (uint8_t)z;
    }
}

// one thread per 32-row word of a validity bitmap: clears the bits of the rows to null-extend
__global__ void lj_nullify_kernel(const uint8_t* __restrict__ nullify, long long n, uint32_t* __restrict__ valid_words) {
    const long long stride = (long long)gridDim.x * blockDim.x;
    const long long nwords = (n + 31) >> 5;
    for (long long w = (long long)blockIdx.x * blockDim.x + threadIdx.x; w < nwords; w += stride) {
        uint32_t m = 0;
        const long long base = w << 5;
        for (int b = 0; b < 32 && base + b < n; ++b) m |= (uint32_t)(nullify[base + b] & 1) << b;
        if (m) valid_words[w] &= ~m;
    }
}

enum { EX_SINGLE = 1, EX_PACKED = 2, EX_HASHED = 3 };

struct JoinEx {
    igl_ctx* ctx = nullptr;
    int join_type = 0;
    bool left_join = false;
    int nl = 0, nr = 0, nkeys = 0;
    std::vector<int> lt, rt;
    std::vector<uint8_t> ln, rn;
    std::vector<igl_expr> lkeys, rkeys;        // copies of the caller's expressions
    int mode = EX_SINGLE;
    bool hidden_key = false;                   // false: the single key is a fixed-width user column already
    int plain_l = -1, plain_r = -1;            // then: its index on each side
    igl_fp* lfp = nullptr;                     // projections of the key expressions that are not plain columns
    igl_fp* rfp = nullptr;
    std::vector<int> lfp_of_key, rfp_of_key;   // key i -> output # of the projection (-1: plain column)
    int key_type = IGL_T_I64;                  // type of the hidden key column
    igl_join* inner = nullptr;
    igl_fp* verify = nullptr;                  // EX_HASHED: the real join condition over the output schema
    bool track = false;                        // EX_HASHED Left / Full: both sides carry an emitted row-number column
    igl_fp* verify_ok = nullptr;               // then: the condition as a Boolean output column,
    igl_fp* compact = nullptr;                 //       and (output columns, keep flag) -> output columns where keep != 0
    int hash_bits = 64;
    std::vector<void*> rtemp;                  // build-side temporaries (widened booleans, key column): live as long as the table
    std::vector<uint8_t> r_has_valid;          // which build columns came with a validity bitmap
    int64_t build_rows = 0;
};

static int grid_for(igl_ctx* ctx, long long n) { return (int)std::min<long long>((n + 255) / 256 + 1, (long long)ctx->sm_count * 16); }
static size_t pad_bytes(size_t b) { return (b + 63) / 64 * 64 + 256; }

// Key value columns of one side: plain columns as they are, the others evaluated by the projection.
static int32_t key_columns(JoinEx* E, bool left, const igl_dev_col* cols, int64_t n, cudaStream_t st, std::vector<igl_dev_col>& kv,
                           std::vector<void*>& temps) {
    const auto& keys = left ? E->lkeys : E->rkeys;
    igl_fp* fp = left ? E->lfp : E->rfp;
    const auto& of = left ? E->lfp_of_key : E->rfp_of_key;
    std::vector<igl_dev_col> fout;
    if (fp) {
        const int nf = igl_fp_num_outputs(fp);
        fout.assign(nf, igl_dev_col{});
        for (int q = 0; q < nf; ++q) {
            int32_t t = 0, nl = 0;
            igl_fp_output_type(fp, q, &t, &nl);
            fout[q].type = t;
            void* p = nullptr;
            IGL_TRY(pool_alloc(E->ctx, pad_bytes(t == IGL_T_BOOL ? (size_t)(n + 7) / 8 : (size_t)n * type_width(t)), &p));
            temps.push_back(p); fout[q].data = p;
            if (nl) { IGL_TRY(pool_alloc(E->ctx, pad_bytes((size_t)(n + 7) / 8), &p)); temps.push_back(p); fout[q].validity = (uint8_t*)p; }
        }
        int64_t rows = 0;
        if (n > 0) IGL_TRY(igl_fp_run(fp, cols, n, fout.data(), &rows, (void*)st));   // key expression errors (division by zero, casts) surface here
    }
    kv.assign(E->nkeys, igl_dev_col{});
    for (int k = 0; k < E->nkeys; ++k) kv[k] = of[k] >= 0 ? fout[of[k]] : cols[keys[k].nodes[0].col];
    return IGL_OK;
}

// The hidden key column of one side (EX_SINGLE: the key value column itself).
static int32_t hidden_key_column(JoinEx* E, const std::vector<igl_dev_col>& kv, int64_t n, cudaStream_t st, igl_dev_col& out,
                                 std::vector<void*>& temps) {
    igl_ctx* ctx = E->ctx;
    if (E->mode == EX_SINGLE) {
        out = kv[0];
        if (out.type == IGL_T_BOOL) {   // a Boolean key: widened like any Boolean column
            void* p = nullptr;
            IGL_TRY(pool_alloc(ctx, pad_bytes((size_t)n * 4), &p));
            temps.push_back(p);
            if (n > 0) bool_to_i32_kernel<<<grid_for(ctx, n), 256, 0, st>>>((const uint8_t*)kv[0].data, n, (int32_t*)p);
            ctx->launches++;
            out.type = IGL_T_I32; out.data = p;
        }
        return IGL_OK;
    }
    void* p = nullptr;
    IGL_TRY(pool_alloc(ctx, pad_bytes((size_t)n * 8), &p));
    temps.push_back(p);
    out = igl_dev_col{IGL_T_I64, 0, p, nullptr, nullptr};
    if (n == 0) return IGL_OK;
    if (E->mode == EX_PACKED) {
        if (kv[0].validity || kv[1].validity) return fail(IGL_ERR_INVALID_ARG, "join key declared non-nullable carries a validity bitmap");
        pack2_i32_kernel<<<grid_for(ctx, n), 256, 0, st>>>((const int32_t*)kv[0].data, (const int32_t*)kv[1].data, n, (int64_t*)p);
    } else {
        HashKeyCols K{};
        K.n = E->nkeys;
        for (int k = 0; k < E->nkeys; ++k) { K.data[k] = (const uint8_t*)kv[k].data; K.valid[k] = kv[k].validity; K.type[k] = kv[k].type; }
        hash_keys_kernel<<<grid_for(ctx, n), 256, 0, st>>>(K, n, E->hash_bits, (int64_t*)p);
    }
    IGL_CUDA_OK(cudaGetLastError());
    ctx->launches++;
    return IGL_OK;
}

// User columns -> the core join's columns: Booleans widened, the hidden key appended.
static int32_t augment(JoinEx* E, bool left, const igl_dev_col* cols, int64_t n, cudaStream_t st, std::vector<igl_dev_col>& aug,
                       std::vector<void*>& temps) {
    igl_ctx* ctx = E->ctx;
    const auto& types = left ? E->lt : E->rt;
    const int nc = left ? E->nl : E->nr;
    aug.clear();
    for (int c = 0; c < nc; ++c) {
        if (cols[c].type != types[c]) return fail(IGL_ERR_INVALID_ARG, "%s column %d has type %d, join expects %d", left ? "left" : "right", c, cols[c].type, types[c]);
        igl_dev_col a = cols[c];
        if (types[c] == IGL_T_BOOL) {
            void* p = nullptr;
            IGL_TRY(pool_alloc(ctx, pad_bytes((size_t)n * 4), &p));
            temps.push_back(p);
            if (n > 0) bool_to_i32_kernel<<<grid_for(ctx, n), 256, 0, st>>>((const uint8_t*)cols[c].data, n, (int32_t*)p);
            ctx->launches++;
            a.type = IGL_T_I32; a.data = p;
        }
        aug.push_back(a);
    }
    if (E->hidden_key) {
        std::vector<igl_dev_col> kv;
        IGL_TRY(key_columns(E, left, cols, n, st, kv, temps));
        igl_dev_col hk{};
        IGL_TRY(hidden_key_column(E, kv, n, st, hk, temps));
        aug.push_back(hk);
    }
    if (E->track) {   // this side's row numbers, carried through the join
        if (n >= ((int64_t)1 << 31)) return fail(IGL_ERR_UNSUPPORTED, "a Left/Full join on a hashed key list is limited to 2^31 rows per side and call");
        void* p = nullptr;
        IGL_TRY(pool_alloc(ctx, pad_bytes((size_t)n * 4), &p));
        temps.push_back(p);
        IGL_TRY(launch_iota_i32(ctx, (int32_t*)p, n, st));
        aug.push_back(igl_dev_col{IGL_T_I32, 0, p, nullptr, nullptr});
    }
    IGL_CUDA_OK(cudaGetLastError());
    return IGL_OK;
}

igl_join* join_ex_inner(igl_join* j) { return ((JoinEx*)*join_ex_slot(j))->inner; }

int32_t join_ex_build(igl_join* j, const igl_dev_col* right, int64_t nrows, void* stream) {
    JoinEx* E = (JoinEx*)*join_ex_slot(j);
    DeviceGuard g(E->ctx->device);
    cudaStream_t st = pick_stream(E->ctx, stream);
    IGL_CUDA_OK(cudaStreamSynchronize(st));   // nothing may still read the previous build's temporaries
    for (void* p : E->rtemp) pool_free(E->ctx, p);
    E->rtemp.clear();
    std::vector<igl_dev_col> aug;
    IGL_TRY(augment(E, false, right, nrows, st, aug, E->rtemp));
    E->build_rows = nrows;
    E->r_has_valid.assign(E->nr, 0);
    for (int c = 0; c < E->nr; ++c) E->r_has_valid[c] = right[c].validity != nullptr;
    return igl_join_build(E->inner, aug.data(), nrows, (void*)st);
}

int32_t join_ex_probe_count(igl_join* j, const igl_dev_col* left, int64_t nrows, int64_t* out_rows, void* stream) {
    JoinEx* E = (JoinEx*)*join_ex_slot(j);
    DeviceGuard g(E->ctx->device);
    cudaStream_t st = pick_stream(E->ctx, stream);
    std::vector<igl_dev_col> aug;
    std::vector<void*> temps;
    int32_t rc = augment(E, true, left, nrows, st, aug, temps);
    if (rc >= 0) rc = igl_join_probe_count(E->inner, aug.data(), nrows, out_rows, (void*)st);   // hashed keys: an upper bound (collisions are filtered afterwards)
    cudaStreamSynchronize(st);
    for (void* p : temps) pool_free(E->ctx, p);
    return rc;
}

int32_t join_ex_probe(igl_join* j, const igl_dev_col* left, int64_t nrows, igl_dev_col* out, int64_t out_capacity,
                      int64_t* out_rows, void* stream) {
    JoinEx* E = (JoinEx*)*join_ex_slot(j);
    igl_ctx* ctx = E->ctx;
    DeviceGuard g(ctx->device);
    cudaStream_t st = pick_stream(ctx, stream);
    std::vector<void*> temps;
    auto done = [&](int32_t rc) { cudaStreamSynchronize(st); for (void* p : temps) pool_free(ctx, p); return rc; };
    std::vector<igl_dev_col> aug;
    int32_t rc = augment(E, true, left, nrows, st, aug, temps);
    if (rc < 0) return done(rc);
    const int nu = E->nl + E->nr;                       // user-visible output columns
    const int64_t cap = std::max<int64_t>(out_capacity, 1);
    const bool staged = E->mode == EX_HASHED;           // the core writes to temporaries, the verify filter to `out`
    // the core join's output columns: user columns (Booleans as Int32 temporaries), hidden keys without a buffer
    std::vector<igl_dev_col> iout, mid(nu);
    auto out_type = [&](int c) { return c < E->nl ? E->lt[c] : E->rt[c - E->nl]; };
    for (int c = 0; c < nu; ++c) {
        const int t = out_type(c);
        igl_dev_col m = out[c];
        m.type = t;
        if (staged) {   // same shape as the caller's column, in pool memory
            void* p = nullptr;
            rc = pool_alloc(ctx, pad_bytes(t == IGL_T_BOOL ? (size_t)(cap + 7) / 8 : (size_t)cap * type_width(t)), &p);
            if (rc < 0) return done(rc);
            temps.push_back(p); m.data = p;
            // a bitmap only where the core join writes one (the verify filter reads what it is given)
            const bool has_valid = c < E->nl ? left[c].validity != nullptr : (E->left_join || (!E->r_has_valid.empty() && E->r_has_valid[c - E->nl]));
            m.validity = nullptr;
            if (has_valid) {
                rc = pool_alloc(ctx, pad_bytes((size_t)(cap + 7) / 8), &p);
                if (rc < 0) return done(rc);
                temps.push_back(p); m.validity = (uint8_t*)p;
            }
        }
        mid[c] = m;
    }
    auto push_side = [&](int c0, int c1) -> int32_t {
        for (int c = c0; c < c1; ++c) {
            igl_dev_col o = mid[c];
            if (out_type(c) == IGL_T_BOOL) {
                void* p = nullptr;
                IGL_TRY(pool_alloc(ctx, pad_bytes((size_t)cap * 4), &p));
                temps.push_back(p);
                o.type = IGL_T_I32; o.data = p;
            }
            iout.push_back(o);
        }
        if (E->hidden_key) iout.push_back(igl_dev_col{E->key_type, 0, nullptr, nullptr, nullptr});
        if (E->track) {
            void *p = nullptr, *v = nullptr;
            IGL_TRY(pool_alloc(ctx, pad_bytes((size_t)cap * 4), &p));
            temps.push_back(p);
            IGL_TRY(pool_alloc(ctx, pad_bytes((size_t)(cap + 7) / 8), &v));
            temps.push_back(v);
            iout.push_back(igl_dev_col{IGL_T_I32, 0, p, (uint8_t*)v, nullptr});
        }
        return IGL_OK;
    };
    rc = push_side(0, E->nl);
    if (rc >= 0) rc = push_side(E->nl, nu);
    if (rc < 0) return done(rc);
    int64_t n = 0;
    rc = igl_join_probe(E->inner, aug.data(), nrows, iout.data(), out_capacity, &n, (void*)st);
    if (rc < 0) return done(rc);
    // Booleans back to bits
    for (int c = 0, ic = 0; c < nu; ++c, ++ic) {
        if (c == E->nl) ic += (E->hidden_key ? 1 : 0) + (E->track ? 1 : 0);
        if (out_type(c) != IGL_T_BOOL || n == 0) continue;
        i32_to_bool_kernel<<<grid_for(ctx, (n + 31) / 32), 256, 0, st>>>((const int32_t*)iout[ic].data, n, (uint32_t*)mid[c].data);
        ctx->launches++;
    }
    if (cudaGetLastError() != cudaSuccess) return done(fail(IGL_ERR_CUDA, "boolean repack failed"));
    if (staged && n > 0 && E->track) {
        // Left / Full: the condition as a column, true matches counted per probe row, false pairs dropped —
        // except the first one of a probe row without any true match, which becomes its null-extended row
        const int extra = (E->hidden_key ? 1 : 0) + 1;
        const int32_t* lrow = (const int32_t*)iout[E->nl + extra - 1].data;
        const uint8_t* matched = iout[E->nl + extra + E->nr + extra - 1].validity;   // the right row number is valid = the pair is a match
        int32_t okt = 0, oknl = 0;
        igl_fp_output_type(E->verify_ok, 0, &okt, &oknl);
        void *ok_bits = nullptr, *ok_valid = nullptr, *true_cnt = nullptr, *claimed = nullptr, *keep = nullptr, *nullify = nullptr;
        auto get = [&](size_t bytes, void** p) -> int32_t { int32_t r = pool_alloc(ctx, pad_bytes(bytes), p); if (r >= 0) temps.push_back(*p); return r; };
        if ((rc = get((size_t)(n + 7) / 8, &ok_bits)) < 0 || (oknl && (rc = get((size_t)(n + 7) / 8, &ok_valid)) < 0) ||
            (rc = get((size_t)std::max<int64_t>(nrows, 1) * 4, &true_cnt)) < 0 || (rc = get((size_t)std::max<int64_t>(nrows, 1) * 4, &claimed)) < 0 ||
            (rc = get((size_t)n * 4, &keep)) < 0 || (rc = get((size_t)n, &nullify)) < 0) return done(rc);
        igl_dev_col okcol{IGL_T_BOOL, 0, ok_bits, (uint8_t*)ok_valid, nullptr};
        int64_t rows = 0;
        rc = igl_fp_run(E->verify_ok, mid.data(), n, &okcol, &rows, (void*)st);
        if (rc < 0) return done(rc);
        if (cudaMemsetAsync(true_cnt, 0, (size_t)std::max<int64_t>(nrows, 1) * 4, st) != cudaSuccess ||
            cudaMemsetAsync(claimed, 0, (size_t)std::max<int64_t>(nrows, 1) * 4, st) != cudaSuccess) return done(fail(IGL_ERR_CUDA, "memset failed"));
        lj_count_true_kernel<<<grid_for(ctx, n), 256, 0, st>>>(lrow, matched, (const uint8_t*)ok_bits, (const uint8_t*)ok_valid, n, (int32_t*)true_cnt);
        lj_decide_kernel<<<grid_for(ctx, n), 256, 0, st>>>(lrow, matched, (const uint8_t*)ok_bits, (const uint8_t*)ok_valid, n, (const int32_t*)true_cnt,
                                                          (int32_t*)claimed, (int32_t*)keep, (uint8_t*)nullify);
        for (int c = E->nl; c < nu; ++c)
            lj_nullify_kernel<<<grid_for(ctx, (n + 31) / 32), 256, 0, st>>>((const uint8_t*)nullify, n, (uint32_t*)mid[c].validity);
        ctx->launches += 2 + E->nr;
        if (cudaGetLastError() != cudaSuccess) return done(fail(IGL_ERR_CUDA, "left-join verification failed to launch"));
        std::vector<igl_dev_col> in2(mid), fin(nu);
        in2.push_back(igl_dev_col{IGL_T_I32, 0, keep, nullptr, nullptr});
        for (int c = 0; c < nu; ++c) { fin[c] = out[c]; fin[c].type = out_type(c); }
        int64_t kept = 0;
        rc = igl_fp_run(E->compact, in2.data(), n, fin.data(), &kept, (void*)st);
        if (rc < 0) return done(rc);
        n = kept;
    } else if (staged && n > 0) {
        // the real join condition, key by key: drops rows whose 64-bit key hashes collided
        std::vector<igl_dev_col> fin(nu);
        for (int c = 0; c < nu; ++c) { fin[c] = out[c]; fin[c].type = out_type(c); }
        int64_t kept = 0;
        rc = igl_fp_run(E->verify, mid.data(), n, fin.data(), &kept, (void*)st);
        if (rc < 0) return done(rc);
        n = kept;
    }
    if (out_rows) *out_rows = n;
    return done(IGL_OK);
}

void join_ex_destroy(igl_join* j) {
    JoinEx* E = (JoinEx*)*join_ex_slot(j);
    if (!E) return;
    DeviceGuard g(E->ctx->device);
    cudaStreamSynchronize(E->ctx->stream);
    igl_join_destroy(E->inner);
    if (E->lfp) igl_fp_destroy(E->lfp);
    if (E->rfp) igl_fp_destroy(E->rfp);
    if (E->verify) igl_fp_destroy(E->verify);
    if (E->verify_ok) igl_fp_destroy(E->verify_ok);
    if (E->compact) igl_fp_destroy(E->compact);
    for (void* p : E->rtemp) pool_free(E->ctx, p);
    delete E;
    *join_ex_slot(j) = nullptr;
}

}  // namespace igl

using namespace igl;

namespace {

igl_expr plain_col_expr(int col, int type, bool nullable, const std::vector<int>& types, const std::vector<uint8_t>& nulls) {
    igl_expr e;
    ExprNode n;
    n.op = IGL_OP_COL; n.col = col; n.type = type; n.nullable = nullable;
    e.nodes.push_back(n);
    e.type = type; e.nullable = nullable; e.ncols = (int)types.size();
    e.col_types.assign(types.begin(), types.end());
    e.col_nullable = nulls;
    return e;
}

void push_col(std::vector<uint8_t>& bc, int c) { bc.push_back(IGL_OP_COL); bc.push_back((uint8_t)(c & 0xFF)); bc.push_back((uint8_t)(c >> 8)); }

}  // namespace

extern "C" int32_t igl_join_create_ex(igl_ctx* ctx, int32_t join_type, const int32_t* left_types, const uint8_t* left_nullable,
                                      int32_t n_left, const int32_t* right_types, const uint8_t* right_nullable, int32_t n_right,
                                      const igl_expr* const* left_keys, const igl_expr* const* right_keys, int32_t nkeys,
                                      igl_join** out) {
    if (!ctx || !out || !left_types || !right_types || !left_keys || !right_keys || n_left <= 0 || n_right <= 0)
        return fail(IGL_ERR_INVALID_ARG, "igl_join_create_ex: bad arguments");
    *out = nullptr;
    if (nkeys < 1 || nkeys > kExMaxKeys) return fail(IGL_ERR_UNSUPPORTED, "1 to %d join key pairs", kExMaxKeys);
    if (join_type < IGL_JOIN_INNER || join_type > IGL_JOIN_FULL)
        return fail(IGL_ERR_UNSUPPORTED, "Join type %d not implemented", join_type);   // physical_planner.rs:127-129
    for (int k = 0; k < nkeys; ++k) {
        if (!left_keys[k] || !right_keys[k]) return fail(IGL_ERR_INVALID_ARG, "join key %d is NULL", k);
        if (left_keys[k]->ncols != n_left || right_keys[k]->ncols != n_right) return fail(IGL_ERR_INVALID_ARG, "join key %d was compiled against a different schema", k);
        if (left_keys[k]->type == IGL_T_UTF8 || right_keys[k]->type == IGL_T_UTF8)
            return fail(IGL_ERR_UNSUPPORTED, "Utf8 join keys are not supported on device-resident columns");
    }
    bool any_wide = false;
    for (int c = 0; c < n_left; ++c) {
        if (left_types[c] == IGL_T_UTF8) return fail(IGL_ERR_UNSUPPORTED, "Utf8 columns ride through the streaming operator (igl_op_hash_join_create), not the device-level join");
        if (left_types[c] == IGL_T_BOOL) any_wide = true;
    }
    for (int c = 0; c < n_right; ++c) {
        if (right_types[c] == IGL_T_UTF8) return fail(IGL_ERR_UNSUPPORTED, "Utf8 columns ride through the streaming operator (igl_op_hash_join_create), not the device-level join");
        if (right_types[c] == IGL_T_BOOL) any_wide = true;
    }
    const bool simple_key = nkeys == 1 && left_keys[0]->is_plain_col() && right_keys[0]->is_plain_col();
    if (simple_key && !any_wide)   // what the core join takes as it is
        return igl_join_create(ctx, join_type, left_types, left_nullable, n_left, right_types, right_nullable, n_right, left_keys[0], right_keys[0], out);

    auto E = std::make_unique<JoinEx>();
    E->ctx = ctx; E->join_type = join_type;
    E->left_join = join_type == IGL_JOIN_LEFT || join_type == IGL_JOIN_FULL;
    E->nl = n_left; E->nr = n_right; E->nkeys = nkeys;
    E->lt.assign(left_types, left_types + n_left); E->rt.assign(right_types, right_types + n_right);
    E->ln.assign(n_left, 0); E->rn.assign(n_right, 0);
    if (left_nullable) E->ln.assign(left_nullable, left_nullable + n_left);
    if (right_nullable) E->rn.assign(right_nullable, right_nullable + n_right);
    for (int k = 0; k < nkeys; ++k) { E->lkeys.push_back(*left_keys[k]); E->rkeys.push_back(*right_keys[k]); }
    DeviceGuard g(ctx->device);
    struct Cleanup { JoinEx* e; ~Cleanup() { if (e) { if (e->inner) igl_join_destroy(e->inner); if (e->lfp) igl_fp_destroy(e->lfp); if (e->rfp) igl_fp_destroy(e->rfp); if (e->verify) igl_fp_destroy(e->verify); if (e->verify_ok) igl_fp_destroy(e->verify_ok); if (e->compact) igl_fp_destroy(e->compact); } } } guard{E.get()};

    // a pair whose types differ can never match (the printed key names its Arrow type): hand that pair
    // to the core join as the one key — its own rule then yields no match / all null-extended rows
    int use_only = -1;
    for (int k = 0; k < nkeys; ++k) if (left_keys[k]->type != right_keys[k]->type) { use_only = k; break; }
    if (use_only >= 0) {
        E->lkeys = {*left_keys[use_only]}; E->rkeys = {*right_keys[use_only]};
        E->nkeys = 1;
    }
    // key mode
    bool all_plain = true, any_nullable = false;
    for (int k = 0; k < E->nkeys; ++k) {
        all_plain = all_plain && E->lkeys[k].is_plain_col() && E->rkeys[k].is_plain_col();
        any_nullable = any_nullable || E->lkeys[k].nullable || E->rkeys[k].nullable;
    }
    if (E->nkeys == 1) {
        E->mode = EX_SINGLE;
        const bool plain_fixed = all_plain && E->lkeys[0].type != IGL_T_BOOL && E->rkeys[0].type != IGL_T_BOOL;
        E->hidden_key = !plain_fixed;
        if (plain_fixed) { E->plain_l = E->lkeys[0].nodes[0].col; E->plain_r = E->rkeys[0].nodes[0].col; }
        E->key_type = E->lkeys[0].type == IGL_T_BOOL ? IGL_T_I32 : E->lkeys[0].type;
    } else if (E->nkeys == 2 && !any_nullable && E->lkeys[0].type == IGL_T_I32 && E->lkeys[1].type == IGL_T_I32) {
        E->mode = EX_PACKED; E->hidden_key = true; E->key_type = IGL_T_I64;
    } else {
        E->mode = EX_HASHED; E->hidden_key = true; E->key_type = IGL_T_I64;
        E->track = E->left_join;
        if (const char* hb = getenv("IGL_JOIN_KEY_HASH_BITS")) E->hash_bits = std::min(64, std::max(1, atoi(hb)));
        if (!all_plain) return fail(IGL_ERR_UNSUPPORTED, "computed keys in a join on a hashed key list are not supported; project them first");
    }
    // projections for the computed keys
    auto make_fp = [&](bool left, igl_fp** fp, std::vector<int>& of) -> int32_t {
        const auto& keys = left ? E->lkeys : E->rkeys;
        std::vector<const igl_expr*> pe;
        of.assign(E->nkeys, -1);
        for (int k = 0; k < E->nkeys; ++k) if (!keys[k].is_plain_col()) { of[k] = (int)pe.size(); pe.push_back(&keys[k]); }
        if (pe.empty()) return IGL_OK;
        return fp_create_ex(ctx, left ? E->lt.data() : E->rt.data(), left ? E->ln.data() : E->rn.data(), left ? n_left : n_right,
                            nullptr, pe.data(), (int32_t)pe.size(), nullptr, false, fp);
    };
    IGL_TRY(make_fp(true, &E->lfp, E->lfp_of_key));
    IGL_TRY(make_fp(false, &E->rfp, E->rfp_of_key));
    // the core join over the augmented columns
    std::vector<int32_t> ilt, irt;
    std::vector<uint8_t> iln = E->ln, irn = E->rn, lhide(n_left, 0), rhide(n_right, 0);
    for (int t : E->lt) ilt.push_back(t == IGL_T_BOOL ? IGL_T_I32 : t);
    for (int t : E->rt) irt.push_back(t == IGL_T_BOOL ? IGL_T_I32 : t);
    int ilkey = E->plain_l, irkey = E->plain_r;
    if (E->hidden_key) {
        const int lkt = E->mode == EX_SINGLE ? (E->lkeys[0].type == IGL_T_BOOL ? IGL_T_I32 : E->lkeys[0].type) : IGL_T_I64;
        const int rkt = E->mode == EX_SINGLE ? (E->rkeys[0].type == IGL_T_BOOL ? IGL_T_I32 : E->rkeys[0].type) : IGL_T_I64;
        ilkey = n_left; irkey = n_right;
        ilt.push_back(lkt); irt.push_back(rkt);
        iln.push_back(E->mode == EX_SINGLE && E->lkeys[0].nullable ? 1 : 0);
        irn.push_back(E->mode == EX_SINGLE && E->rkeys[0].nullable ? 1 : 0);
        lhide.push_back(1); rhide.push_back(1);
    }
    if (E->track) {   // row numbers of both sides: emitted, never NULL on their own side
        ilt.push_back(IGL_T_I32); irt.push_back(IGL_T_I32);
        iln.push_back(0); irn.push_back(0);
        lhide.push_back(0); rhide.push_back(0);
    }
    std::vector<int> ilt_i(ilt.begin(), ilt.end()), irt_i(irt.begin(), irt.end());
    igl_expr lk = plain_col_expr(ilkey, ilt[ilkey], iln[ilkey] != 0, ilt_i, iln);
    igl_expr rk = plain_col_expr(irkey, irt[irkey], irn[irkey] != 0, irt_i, irn);
    IGL_TRY(igl_join_create(ctx, join_type, ilt.data(), iln.data(), (int32_t)ilt.size(), irt.data(), irn.data(), (int32_t)irt.size(), &lk, &rk, &E->inner));
    join_set_hidden(E->inner, lhide.data(), rhide.data());
    if (E->mode == EX_HASHED) {
        // AND over the key pairs of (l = r OR (l IS NULL AND r IS NULL)), on the output schema (right columns follow the left ones)
        std::vector<int32_t> ot(E->lt.begin(), E->lt.end());
        ot.insert(ot.end(), E->rt.begin(), E->rt.end());
        std::vector<uint8_t> on(E->ln);
        on.insert(on.end(), E->rn.begin(), E->rn.end());
        if (E->left_join) for (int c = n_left; c < n_left + n_right; ++c) on[c] = 1;   // null-extended rows
        std::vector<uint8_t> bc;
        for (int k = 0; k < E->nkeys; ++k) {
            const int a = E->lkeys[k].nodes[0].col, b = n_left + E->rkeys[k].nodes[0].col;
            push_col(bc, a); push_col(bc, b); bc.push_back(IGL_OP_EQ);
            push_col(bc, a); bc.push_back(IGL_OP_IS_NULL); push_col(bc, b); bc.push_back(IGL_OP_IS_NULL); bc.push_back(IGL_OP_AND);
            bc.push_back(IGL_OP_OR);
            if (k > 0) bc.push_back(IGL_OP_AND);
        }
        igl_expr* pred = nullptr;
        IGL_TRY(igl_expr_compile(ctx, bc.data(), bc.size(), ot.data(), on.data(), (int32_t)ot.size(), &pred));
        int32_t rc;
        if (!E->track) rc = fp_create_ex(ctx, ot.data(), on.data(), (int32_t)ot.size(), pred, nullptr, 0, nullptr, false, &E->verify);
        else {
            const igl_expr* pe[1] = {pred};
            rc = fp_create_ex(ctx, ot.data(), on.data(), (int32_t)ot.size(), nullptr, pe, 1, nullptr, false, &E->verify_ok);
        }
        igl_expr_destroy(pred);
        IGL_TRY(rc);
        if (E->track) {
            const int kc = (int)ot.size();
            ot.push_back(IGL_T_I32); on.push_back(0);
            std::vector<uint8_t> skip(ot.size(), 0);
            skip[kc] = 1;
            const uint8_t kb[9] = {IGL_OP_COL, (uint8_t)(kc & 0xFF), (uint8_t)(kc >> 8), IGL_OP_LIT_I32, 0, 0, 0, 0, IGL_OP_NE};
            igl_expr* kp = nullptr;
            IGL_TRY(igl_expr_compile(ctx, kb, sizeof kb, ot.data(), on.data(), (int32_t)ot.size(), &kp));
            rc = fp_create_ex(ctx, ot.data(), on.data(), (int32_t)ot.size(), kp, nullptr, 0, skip.data(), false, &E->compact);
            igl_expr_destroy(kp);
            IGL_TRY(rc);
        }
    }
    igl_join* shell = join_new_shell(ctx);
    guard.e = nullptr;
    *join_ex_slot(shell) = E.release();
    *out = shell;
    return IGL_OK;
}
