// fp_device.cuh — device code of the fused Filter / Projection operator (see
// filter_project.cu for the reference semantics and the kernel shape).  Compiled twice:
//   * by nvcc into the library: the program is read from the kernel parameter
//     (an interpreter with specialised handlers);
//   * by NVRTC at run time (fp_jit.cpp) with IGL_JIT_PROGRAM defined: the jit_*()
//     accessors generated for ONE operator return constants, the loops over instructions
//     and outputs are fully unrolled, and the interpreter folds away into straight-line
//     code for that operator.
// Must compile without host-only headers.
#pragma once

#include "expr_isa.h"
#include "tile_pipe.cuh"

#ifdef IGL_JIT_PROGRAM
#define FP_INSTR_DECL(in, P, pc) const Instr in = jit_instr(pc)
#define FP_EXPR_FIRST(P, e) jit_expr_first(e)
#define FP_EXPR_COUNT(P, e) jit_expr_count(e)
#define FP_N_OUT(P) jit_n_out()
#define FP_OUT_SHIFT(P, j) jit_out_shift(j)
#define FP_SMEM_OFF(P, slot) jit_smem_off(slot)
#define FP_VALID_OFF(P, slot) jit_valid_off(slot)
#define FP_SHIFT(P, slot) jit_shift(slot)
#define FP_UNROLL _Pragma("unroll")
#define FP_SLOW_INLINE __forceinline__
#else
#define FP_INSTR_DECL(in, P, pc) const Instr& in = (P).code[pc]
#define FP_EXPR_FIRST(P, e) (P).expr_first[e]
#define FP_EXPR_COUNT(P, e) (P).expr_count[e]
#define FP_N_OUT(P) (P).n_out
#define FP_OUT_SHIFT(P, j) (P).out_shift[j]
#define FP_SMEM_OFF(P, slot) (P).T.smem_off[slot]
#define FP_VALID_OFF(P, slot) (P).T.valid_off[slot]
#define FP_SHIFT(P, slot) (P).T.shift[slot]
#define FP_UNROLL _Pragma("unroll 1")
#define FP_SLOW_INLINE __noinline__
#endif

namespace igl {

constexpr int kMaxOut = 17;   // 16 expressions + the optional row-id column

enum : uint32_t { ERRF_DIV_ZERO = 1, ERRF_OVERFLOW = 2, ERRF_CAST = 4 };

struct FpParams {
    Instr code[kMaxInstr];
    uint8_t expr_first[kMaxExprs], expr_count[kMaxExprs];
    uint8_t n_out, n_tmp, pad0[2];
    uint32_t tmp_off;              // byte offset of the per-warp temporaries in dynamic shared memory
    PipeCols T;                    // staged input columns + scan state
    void* out_data[kMaxOut];
    uint8_t* out_valid[kMaxOut];   // byte-per-row validity (packed afterwards) or NULL
    uint8_t out_shift[kMaxOut];    // 2 / 3, or 0 = byte per row (boolean, packed afterwards)
    // two-pass filter (FP_MASK writes, FP_APPLY reads): survivor bitmap in input row order, one
    // bit per row, padded to whole tiles; exclusive output offset of every tile
    uint8_t mask_slot;             // FP_APPLY: staged slot holding the tile's slice of the bitmap
    uint32_t* mask_words;
};

// What one launch of fp_kernel does.
//   FP_PROJECT  no predicate: output position == input position
//   FP_SCAN     predicate + outputs in one pass, ordered by the in-kernel tile scan
//   FP_MASK     predicate only: writes the survivor bitmap          (pass 1 of 2)
//   FP_APPLY    outputs only: reads the bitmap and the tile offsets  (pass 2 of 2)
enum FpMode : int { FP_PROJECT = 0, FP_SCAN = 1, FP_MASK = 2, FP_APPLY = 3 };

template <int R>
struct V {
    uint64_t v[R];
    uint32_t ok;  // bit r = row r is valid (non-null)
};
template <int R> struct FullMask { static constexpr uint32_t value = (1u << R) - 1u; };

__device__ __forceinline__ int64_t total_order_key(uint64_t bits) {
    int64_t b = (int64_t)bits;
    return b ^ (int64_t)(((uint64_t)(b >> 63)) >> 1);
}

template <int R, bool HAS_NULLS>
__device__ __forceinline__ V<R> load_col(const FpParams& P, int slot, const uint8_t* stage, int row0) {
    V<R> x;
    x.ok = FullMask<R>::value;
    const uint8_t* base = stage + FP_SMEM_OFF(P, slot);
    const int sh = FP_SHIFT(P, slot);
    if (sh == 2) {
#pragma unroll
        for (int r = 0; r < R; ++r) x.v[r] = *reinterpret_cast<const uint32_t*>(base + ((row0 + r * 32) << 2));
    } else if (sh == 3) {
#pragma unroll
        for (int r = 0; r < R; ++r) x.v[r] = *reinterpret_cast<const uint64_t*>(base + ((row0 + r * 32) << 3));
    } else {
#pragma unroll
        for (int r = 0; r < R; ++r) {
            int row = row0 + r * 32;
            x.v[r] = (base[row >> 3] >> (row & 7)) & 1;
        }
    }
    if (HAS_NULLS && P.T.valid[slot] != nullptr) {
        const uint8_t* vb = stage + FP_VALID_OFF(P, slot);
        uint32_t ok = 0;
#pragma unroll
        for (int r = 0; r < R; ++r) {
            int row = row0 + r * 32;
            ok |= ((vb[row >> 3] >> (row & 7)) & 1u) << r;
        }
        x.ok = ok;
    }
    return x;
}

// ---- one IR instruction on R rows ---------------------------------------------
// Everything that is uniform across rows (operator, type, operand kinds) is
// resolved ONCE per instruction, outside the per-row loops: the type selects a
// template instantiation, comparison operators become three uniform masks, and
// the operand kinds select a `step` instantiation so operands are read straight
// into the registers the operation consumes (no struct copies, no per-row
// switches).  Per-row code is then the arithmetic itself.
template <int TY> struct TypeTag {};

template <int R, int TY>
__device__ __forceinline__ void cmp_rows(int op, const uint64_t (&a)[R], const uint64_t (&b)[R], uint64_t (&d)[R]) {
    const bool m_lt = op == IGL_OP_LT || op == IGL_OP_LE || op == IGL_OP_NE;
    const bool m_eq = op == IGL_OP_EQ || op == IGL_OP_LE || op == IGL_OP_GE;
    const bool m_gt = op == IGL_OP_GT || op == IGL_OP_GE || op == IGL_OP_NE;
#pragma unroll
    for (int r = 0; r < R; ++r) {
        bool lt, eq;
        if (TY == IGL_T_I32) { const int32_t x = (int32_t)a[r], y = (int32_t)b[r]; lt = x < y; eq = x == y; }
        else if (TY == IGL_T_F64) { const int64_t x = total_order_key(a[r]), y = total_order_key(b[r]); lt = x < y; eq = x == y; }
        else if (TY == IGL_T_BOOL) { const uint32_t x = (uint32_t)a[r] & 1u, y = (uint32_t)b[r] & 1u; lt = x < y; eq = x == y; }
        else { const int64_t x = (int64_t)a[r], y = (int64_t)b[r]; lt = x < y; eq = x == y; }
        d[r] = ((lt && m_lt) || (eq && m_eq) || (!lt && !eq && m_gt)) ? 1 : 0;
    }
}

template <int R, int TY>
__device__ __forceinline__ void arith_rows(int op, const uint64_t (&a)[R], const uint64_t (&b)[R], uint64_t (&d)[R]) {
    if (TY == IGL_T_F64) {
        // explicit round-to-nearest intrinsics: never contracted into an FMA
        if (op == IGL_OP_ADD) {
#pragma unroll
            for (int r = 0; r < R; ++r) d[r] = (uint64_t)__double_as_longlong(__dadd_rn(__longlong_as_double((long long)a[r]), __longlong_as_double((long long)b[r])));
        } else if (op == IGL_OP_SUB) {
#pragma unroll
            for (int r = 0; r < R; ++r) d[r] = (uint64_t)__double_as_longlong(__dsub_rn(__longlong_as_double((long long)a[r]), __longlong_as_double((long long)b[r])));
        } else {
#pragma unroll
            for (int r = 0; r < R; ++r) d[r] = (uint64_t)__double_as_longlong(__dmul_rn(__longlong_as_double((long long)a[r]), __longlong_as_double((long long)b[r])));
        }
    } else if (TY == IGL_T_I32) {   // two's-complement wrap; I32 lives in the low 32 bits
        if (op == IGL_OP_ADD) {
#pragma unroll
            for (int r = 0; r < R; ++r) d[r] = (uint32_t)a[r] + (uint32_t)b[r];
        } else if (op == IGL_OP_SUB) {
#pragma unroll
            for (int r = 0; r < R; ++r) d[r] = (uint32_t)a[r] - (uint32_t)b[r];
        } else {
#pragma unroll
            for (int r = 0; r < R; ++r) d[r] = (uint32_t)a[r] * (uint32_t)b[r];
        }
    } else {
        if (op == IGL_OP_ADD) {
#pragma unroll
            for (int r = 0; r < R; ++r) d[r] = a[r] + b[r];
        } else if (op == IGL_OP_SUB) {
#pragma unroll
            for (int r = 0; r < R; ++r) d[r] = a[r] - b[r];
        } else {
#pragma unroll
            for (int r = 0; r < R; ++r) d[r] = a[r] * b[r];
        }
    }
}

// The rare operators (division, casts, boolean logic, null tests, negation).
template <int R, bool HAS_NULLS>
__device__ FP_SLOW_INLINE void slow_rows(const Instr& in, const uint64_t (&a)[R], uint32_t aok, const uint64_t (&b)[R], uint32_t bok,
                                       uint32_t inrange, unsigned int* err_flags, uint64_t (&d)[R], uint32_t& dok) {
    const int op = in.op, ty = in.type;
    constexpr uint32_t FULL = FullMask<R>::value;
    dok = HAS_NULLS ? (aok & bok) : FULL;
    switch (op) {
        case IGL_OP_DIV: case IGL_OP_MOD: {
            if (ty == IGL_T_F64) {
#pragma unroll
                for (int r = 0; r < R; ++r) {
                    const double x = __longlong_as_double((long long)a[r]), y = __longlong_as_double((long long)b[r]);
                    d[r] = (uint64_t)__double_as_longlong(op == IGL_OP_DIV ? __ddiv_rn(x, y) : fmod(x, y));
                }
            } else {
                uint32_t ef = 0;
#pragma unroll
                for (int r = 0; r < R; ++r) {
                    const int64_t x = ty == IGL_T_I32 ? (int64_t)(int32_t)a[r] : (int64_t)a[r];
                    const int64_t y = ty == IGL_T_I32 ? (int64_t)(int32_t)b[r] : (int64_t)b[r];
                    const bool live = ((dok & inrange) >> r) & 1;  // arrow evaluates valid slots only; rows past the end do not exist
                    const int64_t mn = ty == IGL_T_I32 ? (int64_t)INT32_MIN : INT64_MIN;
                    int64_t z = 0;
                    if (y == 0) { if (live) ef |= ERRF_DIV_ZERO; }
                    else if (y == -1) {
                        if (op == IGL_OP_DIV) { if (x == mn) { if (live) ef |= ERRF_OVERFLOW; } else z = -x; }
                    } else z = op == IGL_OP_DIV ? x / y : x % y;
                    d[r] = (uint64_t)z;
                }
                if (ef) atomicOr(err_flags, ef);
            }
            break;
        }
        case IGL_OP_NEG:
            dok = aok;
#pragma unroll
            for (int r = 0; r < R; ++r) d[r] = ty == IGL_T_F64 ? (a[r] ^ 0x8000000000000000ull) : (0ull - a[r]);
            break;
        case IGL_OP_AND: case IGL_OP_OR: {  // Kleene
            uint32_t ok = 0;
#pragma unroll
            for (int r = 0; r < R; ++r) {
                const bool av = (aok >> r) & 1, bv = (bok >> r) & 1;
                const bool x = av && (a[r] & 1), y = bv && (b[r] & 1);
                bool val, known;
                if (op == IGL_OP_AND) { const bool kf = (av && !x) || (bv && !y); known = kf || (av && bv); val = !kf && av && bv; }
                else { const bool kt = x || y; known = kt || (av && bv); val = kt; }
                d[r] = val ? 1 : 0;
                ok |= (known ? 1u : 0u) << r;
            }
            dok = HAS_NULLS ? ok : FULL;
            break;
        }
        case IGL_OP_NOT:
            dok = aok;
#pragma unroll
            for (int r = 0; r < R; ++r) d[r] = (a[r] & 1) ^ 1;
            break;
        case IGL_OP_IS_NULL: case IGL_OP_IS_NOT_NULL:
            dok = FULL;
#pragma unroll
            for (int r = 0; r < R; ++r) d[r] = (((aok >> r) & 1) != 0) == (op == IGL_OP_IS_NOT_NULL) ? 1 : 0;
            break;
        case IGL_OP_CAST: {
            dok = aok;
            const int to = in.aux;
            uint32_t ef = 0;
#pragma unroll
            for (int r = 0; r < R; ++r) {
                const bool live = ((aok & inrange) >> r) & 1;
                uint64_t z = 0;
                if (ty == IGL_T_I32 && to == IGL_T_I64) z = (uint64_t)(int64_t)(int32_t)a[r];
                else if (ty == IGL_T_I32 && to == IGL_T_F64) z = (uint64_t)__double_as_longlong((double)(int32_t)a[r]);
                else if (ty == IGL_T_I64 && to == IGL_T_F64) z = (uint64_t)__double_as_longlong(__ll2double_rn((long long)a[r]));
                else if (ty == IGL_T_I64 && to == IGL_T_I32) {
                    const int64_t x = (int64_t)a[r];
                    if (x < INT32_MIN || x > INT32_MAX) { if (live) ef |= ERRF_CAST; } else z = (uint64_t)(uint32_t)(int32_t)x;
                } else if (ty == IGL_T_F64 && to == IGL_T_I32) {
                    const double t = trunc(__longlong_as_double((long long)a[r]));
                    if (!(t >= -2147483648.0 && t <= 2147483647.0)) { if (live) ef |= ERRF_CAST; } else z = (uint64_t)(uint32_t)(int32_t)t;
                } else if (ty == IGL_T_F64 && to == IGL_T_I64) {
                    const double t = trunc(__longlong_as_double((long long)a[r]));
                    if (!(t >= -9223372036854775808.0 && t < 9223372036854775808.0)) { if (live) ef |= ERRF_CAST; } else z = (uint64_t)(int64_t)t;
                } else z = a[r];
                d[r] = z;
            }
            if (ef) atomicOr(err_flags, ef);
            break;
        }
        default:   // IOP_MOV
            dok = aok;
#pragma unroll
            for (int r = 0; r < R; ++r) d[r] = a[r];
            break;
    }
}

// Per-warp temporaries: kMaxTmp x 32R rows x 8 bytes, then one validity nibble per lane.
struct WarpTmp {
    uint64_t* vals;   // [n_tmp][32 * R]
    uint8_t* oks;     // [n_tmp][32]
};

// Reads one operand of statically known kind into (v, ok).
template <int R, bool HAS_NULLS, int KIND>
__device__ __forceinline__ void read_operand(const FpParams& P, int idx, uint64_t lit, bool lit_null, const uint8_t* stage, int row0,
                                             const uint64_t (&acc)[R], uint32_t acc_ok, const WarpTmp& tmp, int lane,
                                             uint64_t (&v)[R], uint32_t& ok) {
    constexpr uint32_t FULL = FullMask<R>::value;
    if (KIND == OPK_ACC) {
#pragma unroll
        for (int r = 0; r < R; ++r) v[r] = acc[r];
        ok = acc_ok;
    } else if (KIND == OPK_COL) {
        const V<R> x = load_col<R, HAS_NULLS>(P, idx, stage, row0);
#pragma unroll
        for (int r = 0; r < R; ++r) v[r] = x.v[r];
        ok = x.ok;
    } else if (KIND == OPK_TMP) {
#pragma unroll
        for (int r = 0; r < R; ++r) v[r] = tmp.vals[idx * (32 * R) + r * 32 + lane];
        ok = HAS_NULLS ? tmp.oks[idx * 32 + lane] : FULL;
    } else {
#pragma unroll
        for (int r = 0; r < R; ++r) v[r] = lit;
        ok = (HAS_NULLS && lit_null) ? 0 : FULL;
    }
}

template <int R, bool HAS_NULLS, int AK, int BK>
__device__ __forceinline__ void step(const FpParams& P, const Instr& in, const uint8_t* stage, int row0, uint32_t inrange,
                                     const WarpTmp& tmp, int lane, uint64_t (&acc)[R], uint32_t& acc_ok) {
    constexpr uint32_t FULL = FullMask<R>::value;
    const bool is_cast = in.op == IGL_OP_CAST;
    uint64_t a[R], b[R], d[R];
    uint32_t aok, bok, dok;
    read_operand<R, HAS_NULLS, AK>(P, in.a_idx, in.a_lit, !is_cast && (in.aux & AUX_A_NULL), stage, row0, acc, acc_ok, tmp, lane, a, aok);
    if (BK == OPK_NONE) {
#pragma unroll
        for (int r = 0; r < R; ++r) b[r] = a[r];
        bok = aok;
    } else {
        read_operand<R, HAS_NULLS, BK>(P, in.b_idx, in.b_lit, !is_cast && (in.aux & AUX_B_NULL), stage, row0, acc, acc_ok, tmp, lane, b, bok);
    }
    const int op = in.op, ty = in.type;
    if (op >= IGL_OP_EQ && op <= IGL_OP_GE) {
        dok = HAS_NULLS ? (aok & bok) : FULL;
        if (ty == IGL_T_I32) cmp_rows<R, IGL_T_I32>(op, a, b, d);
        else if (ty == IGL_T_F64) cmp_rows<R, IGL_T_F64>(op, a, b, d);
        else if (ty == IGL_T_I64) cmp_rows<R, IGL_T_I64>(op, a, b, d);
        else cmp_rows<R, IGL_T_BOOL>(op, a, b, d);
    } else if (op >= IGL_OP_ADD && op <= IGL_OP_MUL) {
        dok = HAS_NULLS ? (aok & bok) : FULL;
        if (ty == IGL_T_I32) arith_rows<R, IGL_T_I32>(op, a, b, d);
        else if (ty == IGL_T_F64) arith_rows<R, IGL_T_F64>(op, a, b, d);
        else arith_rows<R, IGL_T_I64>(op, a, b, d);
    } else if (op == IOP_MOV) {
        dok = aok;
#pragma unroll
        for (int r = 0; r < R; ++r) d[r] = a[r];
    } else {
        // out-of-line: hand it copies, so a / b / d themselves never need an address
        uint64_t sa[R], sb[R], sd[R];
#pragma unroll
        for (int r = 0; r < R; ++r) { sa[r] = a[r]; sb[r] = b[r]; }
        slow_rows<R, HAS_NULLS>(in, sa, aok, sb, bok, inrange, &P.T.hdr->err_flags, sd, dok);
#pragma unroll
        for (int r = 0; r < R; ++r) d[r] = sd[r];
    }
    if (in.dst_kind == OPK_TMP) {
#pragma unroll
        for (int r = 0; r < R; ++r) tmp.vals[in.dst_idx * (32 * R) + r * 32 + lane] = d[r];
        if (HAS_NULLS) tmp.oks[in.dst_idx * 32 + lane] = (uint8_t)dok;
    } else {
#pragma unroll
        for (int r = 0; r < R; ++r) acc[r] = d[r];
        acc_ok = dok;
    }
}

// ---- fast handlers ----------------------------------------------------------------
// The common instructions (column/accumulator against column/literal; compare, add,
// subtract, multiply; Int32 / Int64 / Float64; no nulls anywhere) get one fully
// specialised handler each, selected by a dense id computed on the host: one
// jump-table branch per instruction, native-width operands, no per-row tests.
template <int TY> struct NativeOf { using type = uint64_t; };
template <> struct NativeOf<FP_TY_I32> { using type = uint32_t; };

template <int R, int PAIR, int CLS, int TY>
__device__ __forceinline__ void fast_step(const FpParams& P, const Instr& in, const uint8_t* stage, int row0, uint64_t (&acc)[R]) {
    using T = typename NativeOf<TY>::type;
    T a[R], b[R];
    if (PAIR == FP_PAIR_COL_LIT || PAIR == FP_PAIR_COL_COL) {
        const T* ca = reinterpret_cast<const T*>(stage + FP_SMEM_OFF(P, in.a_idx)) + row0;
#pragma unroll
        for (int r = 0; r < R; ++r) a[r] = ca[r * 32];
    } else {
#pragma unroll
        for (int r = 0; r < R; ++r) a[r] = (T)acc[r];
    }
    if (PAIR == FP_PAIR_COL_COL || PAIR == FP_PAIR_ACC_COL) {
        const T* cb = reinterpret_cast<const T*>(stage + FP_SMEM_OFF(P, in.b_idx)) + row0;
#pragma unroll
        for (int r = 0; r < R; ++r) b[r] = cb[r * 32];
    } else {
        const T lit = (T)in.b_lit;
#pragma unroll
        for (int r = 0; r < R; ++r) b[r] = lit;
    }
    if (CLS == FP_CLS_CMP) {
        const int op = in.op;
        const bool m_lt = op == IGL_OP_LT || op == IGL_OP_LE || op == IGL_OP_NE;
        const bool m_eq = op == IGL_OP_EQ || op == IGL_OP_LE || op == IGL_OP_GE;
        const bool m_gt = op == IGL_OP_GT || op == IGL_OP_GE || op == IGL_OP_NE;
#pragma unroll
        for (int r = 0; r < R; ++r) {
            bool lt, eq;
            if (TY == FP_TY_I32) { lt = (int32_t)a[r] < (int32_t)b[r]; eq = a[r] == b[r]; }
            else if (TY == FP_TY_I64) { lt = (int64_t)a[r] < (int64_t)b[r]; eq = a[r] == b[r]; }
            else { const int64_t x = total_order_key((uint64_t)a[r]), y = total_order_key((uint64_t)b[r]); lt = x < y; eq = x == y; }
            acc[r] = ((lt && m_lt) || (eq && m_eq) || (!lt && !eq && m_gt)) ? 1 : 0;
        }
    } else if (TY == FP_TY_F64) {
#pragma unroll
        for (int r = 0; r < R; ++r) {
            const double x = __longlong_as_double((long long)a[r]), y = __longlong_as_double((long long)b[r]);
            const double z = CLS == FP_CLS_ADD ? __dadd_rn(x, y) : CLS == FP_CLS_SUB ? __dsub_rn(x, y) : __dmul_rn(x, y);   // never an FMA
            acc[r] = (uint64_t)__double_as_longlong(z);
        }
    } else {
#pragma unroll
        for (int r = 0; r < R; ++r) acc[r] = (uint64_t)(T)(CLS == FP_CLS_ADD ? a[r] + b[r] : CLS == FP_CLS_SUB ? a[r] - b[r] : a[r] * b[r]);
    }
}

template <int R, class T>
__device__ __forceinline__ void fast_mov(const FpParams& P, const Instr& in, const uint8_t* stage, int row0, uint64_t (&acc)[R]) {
    const T* ca = reinterpret_cast<const T*>(stage + FP_SMEM_OFF(P, in.a_idx)) + row0;
#pragma unroll
    for (int r = 0; r < R; ++r) acc[r] = (uint64_t)ca[r * 32];
}

// Runs expression `e` of the fused program on this lane's R rows.
template <int R, bool HAS_NULLS>
__device__ __forceinline__ V<R> run_expr(const FpParams& P, int e, const uint8_t* stage, int row0, uint32_t inrange,
                                         long long grow0, const WarpTmp& tmp, int lane) {
    uint64_t acc[R];
    uint32_t acc_ok = FullMask<R>::value;
#pragma unroll
    for (int r = 0; r < R; ++r) acc[r] = 0;
    const int first = FP_EXPR_FIRST(P, e), last = first + FP_EXPR_COUNT(P, e);
    FP_UNROLL
    for (int pc = first; pc < last; ++pc) {
        FP_INSTR_DECL(in, P, pc);
        if (in.op == IOP_ROWID) {
#pragma unroll
            for (int r = 0; r < R; ++r) acc[r] = (uint64_t)(grow0 + r * 32);
            acc_ok = FullMask<R>::value;
            continue;
        }
        if (!HAS_NULLS && in.fast) {   // dense id -> jump table -> specialised straight-line handler
#define IGL_FAST3(PAIR, CLS, BASE)                                                              \
            case BASE + 0: fast_step<R, PAIR, CLS, FP_TY_I32>(P, in, stage, row0, acc); break;   \
            case BASE + 1: fast_step<R, PAIR, CLS, FP_TY_I64>(P, in, stage, row0, acc); break;   \
            case BASE + 2: fast_step<R, PAIR, CLS, FP_TY_F64>(P, in, stage, row0, acc); break;
#define IGL_FAST_PAIR(PAIR)                                  \
            IGL_FAST3(PAIR, FP_CLS_CMP, 1 + (PAIR * 4 + 0) * 3) \
            IGL_FAST3(PAIR, FP_CLS_ADD, 1 + (PAIR * 4 + 1) * 3) \
            IGL_FAST3(PAIR, FP_CLS_SUB, 1 + (PAIR * 4 + 2) * 3) \
            IGL_FAST3(PAIR, FP_CLS_MUL, 1 + (PAIR * 4 + 3) * 3)
            switch (in.fast) {
                IGL_FAST_PAIR(FP_PAIR_COL_LIT)
                IGL_FAST_PAIR(FP_PAIR_COL_COL)
                IGL_FAST_PAIR(FP_PAIR_ACC_LIT)
                IGL_FAST_PAIR(FP_PAIR_ACC_COL)
                case kFastMov32: fast_mov<R, uint32_t>(P, in, stage, row0, acc); break;
                default: fast_mov<R, uint64_t>(P, in, stage, row0, acc); break;
            }
#undef IGL_FAST_PAIR
#undef IGL_FAST3
            continue;
        }
#define IGL_STEP(AK, BK) step<R, HAS_NULLS, AK, BK>(P, in, stage, row0, inrange, tmp, lane, acc, acc_ok)
        switch (in.a_kind * 8 + in.b_kind) {
            case OPK_COL * 8 + OPK_LIT: IGL_STEP(OPK_COL, OPK_LIT); break;
            case OPK_COL * 8 + OPK_COL: IGL_STEP(OPK_COL, OPK_COL); break;
            case OPK_ACC * 8 + OPK_LIT: IGL_STEP(OPK_ACC, OPK_LIT); break;
            case OPK_ACC * 8 + OPK_COL: IGL_STEP(OPK_ACC, OPK_COL); break;
            case OPK_COL * 8 + OPK_NONE: IGL_STEP(OPK_COL, OPK_NONE); break;
            case OPK_ACC * 8 + OPK_NONE: IGL_STEP(OPK_ACC, OPK_NONE); break;
            case OPK_LIT * 8 + OPK_NONE: IGL_STEP(OPK_LIT, OPK_NONE); break;
            case OPK_TMP * 8 + OPK_ACC: IGL_STEP(OPK_TMP, OPK_ACC); break;
            case OPK_TMP * 8 + OPK_COL: IGL_STEP(OPK_TMP, OPK_COL); break;
            case OPK_TMP * 8 + OPK_LIT: IGL_STEP(OPK_TMP, OPK_LIT); break;
            case OPK_TMP * 8 + OPK_TMP: IGL_STEP(OPK_TMP, OPK_TMP); break;
            case OPK_TMP * 8 + OPK_NONE: IGL_STEP(OPK_TMP, OPK_NONE); break;
            case OPK_COL * 8 + OPK_ACC: IGL_STEP(OPK_COL, OPK_ACC); break;
            case OPK_COL * 8 + OPK_TMP: IGL_STEP(OPK_COL, OPK_TMP); break;
            case OPK_LIT * 8 + OPK_COL: IGL_STEP(OPK_LIT, OPK_COL); break;
            case OPK_LIT * 8 + OPK_ACC: IGL_STEP(OPK_LIT, OPK_ACC); break;
            case OPK_LIT * 8 + OPK_LIT: IGL_STEP(OPK_LIT, OPK_LIT); break;
            case OPK_LIT * 8 + OPK_TMP: IGL_STEP(OPK_LIT, OPK_TMP); break;
            case OPK_ACC * 8 + OPK_TMP: IGL_STEP(OPK_ACC, OPK_TMP); break;
            default: IGL_STEP(OPK_ACC, OPK_ACC); break;
        }
#undef IGL_STEP
    }
    V<R> out;
#pragma unroll
    for (int r = 0; r < R; ++r) out.v[r] = acc[r];
    out.ok = acc_ok;
    return out;
}

template <int R>
struct FpTile {
    const uint8_t* stage;
    long long row_begin;   // first row of the staged tile
    int rows;              // rows in the staged tile
    int row_off;           // first row of the sub-tile being worked on, relative to row_begin
    uint32_t keep[R];   // ballots: bit = lane whose row (r*32 + lane) survives
    int count;
};

// Bit r of the result = row (row0 + 32 r) exists in this tile.
template <int R>
__device__ __forceinline__ uint32_t rows_in_range(int row0, int rows_in_tile) {
    uint32_t m = 0;
#pragma unroll
    for (int r = 0; r < R; ++r) m |= ((row0 + r * 32) < rows_in_tile ? 1u : 0u) << r;
    return m;
}

template <int R, bool HAS_PRED, bool HAS_NULLS>
__device__ __forceinline__ void fp_count(const FpParams& P, FpTile<R>& t, int warp, int lane, const WarpTmp& tmp) {
    const int row0 = t.row_off + warp * (32 * R) + lane;
    V<R> m;
    const bool owner = warp < P.T.active_warps;   // warp-uniform
    const uint32_t inrange = rows_in_range<R>(row0, t.rows);
    if (HAS_PRED && owner) m = run_expr<R, HAS_NULLS>(P, 0, t.stage, row0, inrange, t.row_begin + row0, tmp, lane);
    t.count = 0;
#pragma unroll
    for (int r = 0; r < R; ++r) {
        bool k = owner && ((inrange >> r) & 1u);
        if (HAS_PRED) k = k && ((m.ok >> r) & 1) && (m.v[r] & 1);   // null mask => row dropped
        t.keep[r] = __ballot_sync(0xffffffffu, k);
        t.count += __popc(t.keep[r]);
    }
}

// FP_APPLY: the ballots come from the staged slice of the survivor bitmap (word w of the
// slice = rows 32w .. 32w+31 of the tile = row group w % R of warp w / R).  Returns this
// warp's offset inside the tile's output run.
template <int R>
__device__ __forceinline__ int fp_take_mask(const FpParams& P, FpTile<R>& t, int warp, int lane) {
    const uint32_t* mw = reinterpret_cast<const uint32_t*>(t.stage + P.T.smem_off[P.mask_slot]) + (t.row_off >> 5);
    // a short last tile copies only the bitmap bytes that exist: clip every word to the rows present
    auto word = [&](int w) -> uint32_t {
        const int left = t.rows - t.row_off - w * 32;
        if (left <= 0) return 0u;
        const uint32_t x = mw[w];
        return left >= 32 ? x : (x & ((1u << left) - 1u));
    };
    const int first = warp * R;       // this warp's first word
    int before = 0;
    for (int w = lane; w < first; w += 32) before += __popc(word(w));
    before = __reduce_add_sync(0xffffffffu, before);
    t.count = 0;
    const bool owner = warp < P.T.active_warps;
#pragma unroll
    for (int r = 0; r < R; ++r) {
        t.keep[r] = owner ? word(first + r) : 0u;
        t.count += __popc(t.keep[r]);
    }
    return before;
}

// E0 = index of the first output expression in the program.
template <int R, int E0, bool HAS_NULLS>
__device__ __forceinline__ void fp_write(const FpParams& P, const FpTile<R>& t, long long wbase, int warp, int lane,
                                         const WarpTmp& tmp) {
    if (t.count == 0) return;   // warp-uniform
    const uint32_t lt_mask = (1u << lane) - 1u;
    const int row0 = t.row_off + warp * (32 * R) + lane;
    // 32-bit offsets inside this warp's output run, computed once for all output columns
    uint32_t off[R];
    uint32_t mine = 0;
    {
        uint32_t run = 0;
#pragma unroll
        for (int r = 0; r < R; ++r) {
            off[r] = run + __popc(t.keep[r] & lt_mask);
            run += __popc(t.keep[r]);
            mine |= ((t.keep[r] >> lane) & 1u) << r;
        }
    }
    FP_UNROLL
    for (int j = 0; j < FP_N_OUT(P); ++j) {
        const int osh = FP_OUT_SHIFT(P, j);
        // ProjectionExec only ever sees the rows FilterExec kept (physical_planner.rs:52-94):
        // an output expression cannot fail on a row the predicate dropped
        const V<R> val = run_expr<R, HAS_NULLS>(P, E0 + j, t.stage, row0, mine, t.row_begin + row0, tmp, lane);
        if (osh == 3) {
            uint64_t* od = reinterpret_cast<uint64_t*>(P.out_data[j]) + wbase;
#pragma unroll
            for (int r = 0; r < R; ++r) if ((mine >> r) & 1u) st_cs_u64(od + off[r], val.v[r]);
        } else if (osh == 2) {
            uint32_t* od = reinterpret_cast<uint32_t*>(P.out_data[j]) + wbase;
#pragma unroll
            for (int r = 0; r < R; ++r) if ((mine >> r) & 1u) st_cs_u32(od + off[r], (uint32_t)val.v[r]);
        } else {
            uint8_t* od = reinterpret_cast<uint8_t*>(P.out_data[j]) + wbase;
#pragma unroll
            for (int r = 0; r < R; ++r) if ((mine >> r) & 1u) od[off[r]] = (uint8_t)(val.v[r] & 1);
        }
        if (HAS_NULLS && P.out_valid[j]) {
            uint8_t* ov = P.out_valid[j] + wbase;
#pragma unroll
            for (int r = 0; r < R; ++r) if ((mine >> r) & 1u) ov[off[r]] = (uint8_t)((val.ok >> r) & 1u);
        }
    }
}

template <int R, int MODE, bool HAS_NULLS>
__device__ __forceinline__ void fp_body(const FpParams& P) {
    extern __shared__ __align__(128) uint8_t smem[];
    __shared__ PipeShared sh;
    TilePipe<MODE == FP_SCAN, HAS_NULLS> pipe(P.T, sh, smem);
    pipe.init();
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    if (warp == kPipeComputeWarps) { pipe.control_loop(); return; }
    WarpTmp tmp;
    tmp.vals = reinterpret_cast<uint64_t*>(smem + P.tmp_off) + (size_t)warp * P.n_tmp * (32 * R);
    tmp.oks = smem + P.tmp_off + (size_t)kPipeComputeWarps * P.n_tmp * 1024 + (size_t)warp * P.n_tmp * 32;

    if (MODE == FP_PROJECT) {   // nothing to scan: output position == input position
        for (int it = 0;; ++it) {
            FpTile<R> t;
            if (!pipe.acquire(it, t.stage, t.row_begin, t.rows)) break;
            for (t.row_off = 0; t.row_off < t.rows; t.row_off += P.T.sub_rows) {
                fp_count<R, false, HAS_NULLS>(P, t, warp, lane, tmp);
                fp_write<R, 0, HAS_NULLS>(P, t, t.row_begin + t.row_off + warp * (32 * R), warp, lane, tmp);
            }
            pipe.release(it);
        }
        return;
    }
    if (MODE == FP_MASK) {      // pass 1: the predicate's ballots ARE the bitmap words
        for (int it = 0;; ++it) {
            FpTile<R> t;
            if (!pipe.acquire(it, t.stage, t.row_begin, t.rows)) break;
            for (t.row_off = 0; t.row_off < t.rows; t.row_off += P.T.sub_rows) {
                fp_count<R, true, HAS_NULLS>(P, t, warp, lane, tmp);
                if (warp < P.T.active_warps && lane < R) {
                    uint32_t w = t.keep[0];
#pragma unroll
                    for (int r = 1; r < R; ++r) w = lane == r ? t.keep[r] : w;
                    P.mask_words[((t.row_begin + t.row_off) >> 5) + warp * R + lane] = w;
                }
            }
            pipe.release(it);
        }
        return;
    }
    if (MODE == FP_APPLY) {     // pass 2: every tile knows where its output starts
        for (int it = 0;; ++it) {
            FpTile<R> t;
            if (!pipe.acquire(it, t.stage, t.row_begin, t.rows)) break;
            // the tile's output offsets arrived with its columns (PipeCols::aux_table)
            const unsigned long long* offs = reinterpret_cast<const unsigned long long*>(t.stage + P.T.aux_smem) +
                                             ((t.row_begin / P.T.sub_rows) & 1);
            int k = 0;
            for (t.row_off = 0; t.row_off < t.rows; t.row_off += P.T.sub_rows, ++k) {
                const unsigned long long tbase = offs[k];
                const int before = fp_take_mask<R>(P, t, warp, lane);
                fp_write<R, 0, HAS_NULLS>(P, t, (long long)tbase + before, warp, lane, tmp);
            }
            pipe.release(it);
        }
        return;
    }
    // FP_SCAN, software pipeline: count(t+1) is issued before write(t)
    FpTile<R> cur;
    cur.row_off = 0;
    if (!pipe.acquire(0, cur.stage, cur.row_begin, cur.rows)) return;
    fp_count<R, true, HAS_NULLS>(P, cur, warp, lane, tmp);
    pipe.post_count(0, warp, cur.count);
    for (int it = 0;; ++it) {
        FpTile<R> nxt;
        nxt.row_off = 0;
        const bool has_next = pipe.acquire(it + 1, nxt.stage, nxt.row_begin, nxt.rows);
        if (has_next) {
            fp_count<R, true, HAS_NULLS>(P, nxt, warp, lane, tmp);
            pipe.post_count(it + 1, warp, nxt.count);
        }
        const long long wbase = pipe.wait_base(it, warp);
        fp_write<R, 1, HAS_NULLS>(P, cur, wbase, warp, lane, tmp);
        pipe.release(it);
        if (!has_next) break;
        cur = nxt;
    }
}

}  // namespace igl
