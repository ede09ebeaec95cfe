// igloo_oracle.cpp — CPU ORACLE.  TEST INFRASTRUCTURE ONLY.
//
// A plain C++17 restatement of the reference's Filter / Projection / HashJoin
// operator semantics, used as the checker by tests/, __graft_entry__.smoke()
// and bench.py's cpu_baseline / --impl reference legs.  Nothing in the product
// path (igloo_b200/, libigloo_cuda.so) may include, link or call this file.
//
// Pinning: Filter / Projection are checked against the reference's own known answers
// (integration_test.rs:21-29,59,72-75 and `SELECT 42`; tests/golden/users.json, and the CSV
// scan fixture tests/golden/scan_fixture.json).  HashJoin: PARITY UNPINNED — the reference
// holds no test, fixture or golden vector for HashJoinExec, and it cannot be built here
// (Rust, no toolchain); the join below follows hash_join.rs line by line, with Arrow C++
// (Acero) as a second opinion in tests/test_oracle.py.
//
// Reference followed (all under /root/reference/crates/engine/src/):
//   operators/filter.rs:39-68        FilterExec::execute
//   operators/projection.rs:48-72    ProjectionExec::execute
//   operators/hash_join.rs:35-76     HashJoinExec::new (output schema)
//   operators/hash_join.rs:100-128   build phase
//   operators/hash_join.rs:131-213   probe phase
//   operators/hash_join.rs:221-280   combine_batches / combine_with_nulls
//   operators/parquet_scan.rs:54     1024-row upstream batches
//
// Arithmetic the operators reach lives in crates that are NOT vendored under
// /root/reference (Cargo.lock pins arrow 55.1.0 and datafusion 48.0.0:
// arrow-select filter_record_batch, arrow-ord cmp, arrow-arith numeric,
// arrow-cast, arrow-array Debug, datafusion-physical-expr BinaryExpr/CastExpr).
// Their published behaviour is restated here:
//   * filter keeps rows whose mask is valid AND true; null mask => dropped
//   * int + - * and unary minus wrap; int / and % by zero => error
//     ("Divide by zero error"); MIN / -1 => overflow error; MIN % -1 => 0
//   * f64 arithmetic is IEEE-754, never contracted (this file is built with
//     -ffp-contract=off); f64 % is fmod; f64 / 0 gives inf/NaN, no error
//   * comparisons on f64 use IEEE-754 totalOrder (NaN == NaN, -0.0 < +0.0)
//   * any null operand => null result, except Kleene AND / OR
//   * CAST (DataFusion CastExpr, safe=false): out-of-range => error
//   * literals broadcast to the batch length (ColumnarValue::into_array)
//   * join key identity = the Debug text of a 1-row slice of the key array
//     (type name + printed value): NULL == NULL, NaN == NaN, -0.0 != 0.0,
//     Int32(1) != Int64(1).
//
// PARITY PINNING: Filter/Projection are pinned by the reference's one
// known-answer fixture (crates/engine/tests/integration_test.rs:21-29,59,72-75,
// transcribed into tests/golden/users.json).  HashJoin is PARITY UNPINNED: the
// reference holds no test, fixture or golden vector for HashJoinExec, and its
// Rust sources cannot be built here (no cargo/rustc, and they do not compile
// against their pinned dependencies — SURVEY.md §0.3).
//
// Two join modes: FAITHFUL (printed-string keys, std::unordered_map, one output
// row per emit — the reference's algorithm, used for small inputs and as the
// timed CPU baseline) and FAST (typed open addressing, multi-threaded probe —
// same row multiset and order, for multi-million-row checks).

#include <algorithm>
#include <cinttypes>
#include <cmath>
#include <cstdint>
#include <cstdio>
#include <cstring>
#include <memory>
#include <stdexcept>
#include <string>
#include <thread>
#include <unordered_map>
#include <vector>

#include "../include/igloo_expr.h"

namespace {

// ----------------------------------------------------------------------------
// Columns
// ----------------------------------------------------------------------------
int width_of(int type) {
    switch (type) {
        case IGL_T_BOOL: return 1;  // one byte per row inside the oracle
        case IGL_T_I32: return 4;
        case IGL_T_I64: return 8;
        case IGL_T_F64: return 8;
        default: return 0;
    }
}

struct Arr {
    int type = 0;
    int64_t n = 0;
    std::vector<uint8_t> data;   // fixed width values (BOOL: 1 byte/row)
    std::vector<uint8_t> valid;  // 1 byte/row; empty => all valid
    std::vector<int32_t> off;    // UTF8 only: n+1 offsets
    std::string name;
    bool nullable = false;

    bool is_valid(int64_t i) const { return valid.empty() || valid[i] != 0; }
    template <class T> const T* as() const { return reinterpret_cast<const T*>(data.data()); }
    template <class T> T* as() { return reinterpret_cast<T*>(data.data()); }
    void alloc(int t, int64_t rows) {
        type = t; n = rows;
        data.assign(static_cast<size_t>(rows) * width_of(t), 0);
    }
    void ensure_valid() { if (valid.empty()) valid.assign(n, 1); }
};

struct Table {
    std::vector<Arr> cols;
    int64_t nrows = 0;
    std::vector<int64_t> batch_sizes;  // stream granularity
    bool unit_batches = false;         // join output: one 1-row batch per row
    int64_t n_batches() const { return unit_batches ? nrows : (int64_t)batch_sizes.size(); }
};

// Append rows [s, s+len) of src to dst (same type).
void append_rows(Arr& dst, const Arr& src, int64_t s, int64_t len) {
    if (len == 0) return;
    if (src.type == IGL_T_UTF8) {
        if (dst.off.empty()) dst.off.push_back(0);
        int32_t b0 = src.off[s], b1 = src.off[s + len];
        int32_t base = dst.off.back();
        dst.data.insert(dst.data.end(), src.data.begin() + b0, src.data.begin() + b1);
        for (int64_t i = 1; i <= len; ++i) dst.off.push_back(base + (src.off[s + i] - b0));
    } else {
        int w = width_of(src.type);
        dst.data.insert(dst.data.end(), src.data.begin() + s * w, src.data.begin() + (s + len) * w);
    }
    if (!src.valid.empty()) {
        if (dst.valid.empty()) dst.valid.assign(dst.n, 1);
        dst.valid.insert(dst.valid.end(), src.valid.begin() + s, src.valid.begin() + s + len);
    } else if (!dst.valid.empty()) {
        dst.valid.insert(dst.valid.end(), len, 1);
    }
    dst.n += len;
}

void append_null(Arr& dst) {  // arrow::array::new_null_array(type, 1)  (hash_join.rs:260)
    if (dst.type == IGL_T_UTF8) {
        if (dst.off.empty()) dst.off.push_back(0);
        dst.off.push_back(dst.off.back());
    } else {
        dst.data.insert(dst.data.end(), width_of(dst.type), 0);
    }
    if (dst.valid.empty()) dst.valid.assign(dst.n, 1);
    dst.valid.push_back(0);
    dst.n += 1;
}

// ----------------------------------------------------------------------------
// Expressions: postfix bytecode -> tree; each node materialises an array per
// batch, the way DataFusion's PhysicalExpr::evaluate does.
// ----------------------------------------------------------------------------
struct Node {
    int op = 0;
    int col = 0;
    int lit_type = 0;
    uint64_t lit_bits = 0;
    bool lit_null = false;
    int cast_to = 0;
    std::unique_ptr<Node> a, b;
};

[[noreturn]] void fail(const std::string& m) { throw std::runtime_error(m); }

std::unique_ptr<Node> parse(const uint8_t* bc, size_t len) {
    std::vector<std::unique_ptr<Node>> st;
    size_t p = 0;
    auto need = [&](size_t k) { if (p + k > len) fail("truncated expression bytecode"); };
    auto pop = [&]() {
        if (st.empty()) fail("expression stack underflow");
        auto n = std::move(st.back()); st.pop_back(); return n;
    };
    while (p < len) {
        int op = bc[p++];
        auto n = std::make_unique<Node>();
        n->op = op;
        switch (op) {
            case IGL_OP_COL: { need(2); uint16_t c; memcpy(&c, bc + p, 2); p += 2; n->col = c; break; }
            case IGL_OP_LIT_BOOL: { need(1); n->lit_type = IGL_T_BOOL; n->lit_bits = bc[p++] ? 1 : 0; break; }
            case IGL_OP_LIT_I32: { need(4); int32_t v; memcpy(&v, bc + p, 4); p += 4; n->lit_type = IGL_T_I32; n->lit_bits = (uint64_t)(int64_t)v; break; }
            case IGL_OP_LIT_I64: { need(8); memcpy(&n->lit_bits, bc + p, 8); p += 8; n->lit_type = IGL_T_I64; break; }
            case IGL_OP_LIT_F64: { need(8); memcpy(&n->lit_bits, bc + p, 8); p += 8; n->lit_type = IGL_T_F64; break; }
            case IGL_OP_LIT_NULL: { need(1); n->lit_type = bc[p++]; n->lit_null = true; break; }
            case IGL_OP_EQ: case IGL_OP_NE: case IGL_OP_LT: case IGL_OP_LE: case IGL_OP_GT: case IGL_OP_GE:
            case IGL_OP_ADD: case IGL_OP_SUB: case IGL_OP_MUL: case IGL_OP_DIV: case IGL_OP_MOD:
            case IGL_OP_AND: case IGL_OP_OR:
                n->b = pop(); n->a = pop(); break;
            case IGL_OP_NOT: case IGL_OP_NEG: case IGL_OP_IS_NULL: case IGL_OP_IS_NOT_NULL:
                n->a = pop(); break;
            case IGL_OP_CAST: { need(1); n->cast_to = bc[p++]; n->a = pop(); break; }
            default: fail("unknown opcode in expression bytecode");
        }
        st.push_back(std::move(n));
    }
    if (st.size() != 1) fail("expression must leave exactly one value");
    return std::move(st.back());
}

// Lit -> array of n (ColumnarValue::into_array, filter.rs:48 / projection.rs:62)
Arr broadcast(const Node& nd, int64_t n) {
    Arr r; r.alloc(nd.lit_type, n);
    if (nd.lit_null) { r.valid.assign(n, 0); return r; }
    switch (nd.lit_type) {
        case IGL_T_BOOL: std::fill(r.data.begin(), r.data.end(), (uint8_t)nd.lit_bits); break;
        case IGL_T_I32: std::fill_n(r.as<int32_t>(), n, (int32_t)nd.lit_bits); break;
        case IGL_T_I64: std::fill_n(r.as<int64_t>(), n, (int64_t)nd.lit_bits); break;
        case IGL_T_F64: { double d; memcpy(&d, &nd.lit_bits, 8); std::fill_n(r.as<double>(), n, d); break; }
        default: fail("unsupported literal type");
    }
    return r;
}

// IEEE-754 totalOrder key (as Rust f64::total_cmp, used by arrow-ord cmp).
inline int64_t total_order_key(double d) {
    int64_t b; memcpy(&b, &d, 8);
    b ^= (int64_t)(((uint64_t)(b >> 63)) >> 1);
    return b;
}

template <class T> inline bool cmp_op(int op, T a, T b) {
    switch (op) {
        case IGL_OP_EQ: return a == b;
        case IGL_OP_NE: return a != b;
        case IGL_OP_LT: return a < b;
        case IGL_OP_LE: return a <= b;
        case IGL_OP_GT: return a > b;
        default: return a >= b;
    }
}

void null_union(Arr& r, const Arr& a, const Arr& b) {
    if (a.valid.empty() && b.valid.empty()) return;
    r.valid.assign(r.n, 1);
    for (int64_t i = 0; i < r.n; ++i) r.valid[i] = a.is_valid(i) && b.is_valid(i);
}

template <class T, class U> Arr arith_int(int op, const Arr& a, const Arr& b) {
    Arr r; r.alloc(a.type, a.n);
    null_union(r, a, b);
    const T* x = a.as<T>(); const T* y = b.as<T>(); T* o = r.as<T>();
    for (int64_t i = 0; i < a.n; ++i) {
        if (!r.is_valid(i)) { o[i] = 0; continue; }  // arrow evaluates valid slots only
        U ux = (U)x[i], uy = (U)y[i];
        switch (op) {
            case IGL_OP_ADD: o[i] = (T)(ux + uy); break;
            case IGL_OP_SUB: o[i] = (T)(ux - uy); break;
            case IGL_OP_MUL: o[i] = (T)(ux * uy); break;
            case IGL_OP_DIV:
                if (y[i] == 0) fail("Divide by zero error");
                if (x[i] == std::numeric_limits<T>::min() && y[i] == (T)-1) fail("Arithmetic overflow: Overflow happened on division");
                o[i] = x[i] / y[i]; break;
            default:  // MOD
                if (y[i] == 0) fail("Divide by zero error");
                o[i] = (y[i] == (T)-1) ? 0 : x[i] % y[i]; break;
        }
    }
    return r;
}

Arr arith_f64(int op, const Arr& a, const Arr& b) {
    Arr r; r.alloc(IGL_T_F64, a.n);
    null_union(r, a, b);
    const double* x = a.as<double>(); const double* y = b.as<double>(); double* o = r.as<double>();
    for (int64_t i = 0; i < a.n; ++i) {
        if (!r.is_valid(i)) { o[i] = 0; continue; }
        switch (op) {
            case IGL_OP_ADD: o[i] = x[i] + y[i]; break;
            case IGL_OP_SUB: o[i] = x[i] - y[i]; break;
            case IGL_OP_MUL: o[i] = x[i] * y[i]; break;
            case IGL_OP_DIV: o[i] = x[i] / y[i]; break;
            default: o[i] = std::fmod(x[i], y[i]); break;
        }
    }
    return r;
}

struct BatchView { const std::vector<Arr>* cols; int64_t start; int64_t len; };

Arr slice_col(const Arr& c, int64_t s, int64_t len) {
    Arr r; r.type = c.type; r.name = c.name; r.nullable = c.nullable;
    append_rows(r, c, s, len);
    if (c.type == IGL_T_UTF8 && r.off.empty()) r.off.push_back(0);
    return r;
}

Arr eval(const Node& nd, const BatchView& bv) {
    const int64_t n = bv.len;
    switch (nd.op) {
        case IGL_OP_COL: {
            if (nd.col < 0 || nd.col >= (int)bv.cols->size()) fail("column index out of range");
            return slice_col((*bv.cols)[nd.col], bv.start, n);
        }
        case IGL_OP_LIT_BOOL: case IGL_OP_LIT_I32: case IGL_OP_LIT_I64: case IGL_OP_LIT_F64: case IGL_OP_LIT_NULL:
            return broadcast(nd, n);
        case IGL_OP_EQ: case IGL_OP_NE: case IGL_OP_LT: case IGL_OP_LE: case IGL_OP_GT: case IGL_OP_GE: {
            Arr a = eval(*nd.a, bv), b = eval(*nd.b, bv);
            if (a.type != b.type) fail("comparison operands must have the same type");
            Arr r; r.alloc(IGL_T_BOOL, n); null_union(r, a, b);
            uint8_t* o = r.data.data();
            for (int64_t i = 0; i < n; ++i) {
                if (!r.is_valid(i)) { o[i] = 0; continue; }
                switch (a.type) {
                    case IGL_T_BOOL: o[i] = cmp_op<int>(nd.op, a.data[i], b.data[i]); break;
                    case IGL_T_I32: o[i] = cmp_op<int32_t>(nd.op, a.as<int32_t>()[i], b.as<int32_t>()[i]); break;
                    case IGL_T_I64: o[i] = cmp_op<int64_t>(nd.op, a.as<int64_t>()[i], b.as<int64_t>()[i]); break;
                    case IGL_T_F64: o[i] = cmp_op<int64_t>(nd.op, total_order_key(a.as<double>()[i]), total_order_key(b.as<double>()[i])); break;
                    default: fail("unsupported comparison type");
                }
            }
            return r;
        }
        case IGL_OP_ADD: case IGL_OP_SUB: case IGL_OP_MUL: case IGL_OP_DIV: case IGL_OP_MOD: {
            Arr a = eval(*nd.a, bv), b = eval(*nd.b, bv);
            if (a.type != b.type) fail("arithmetic operands must have the same type");
            switch (a.type) {
                case IGL_T_I32: return arith_int<int32_t, uint32_t>(nd.op, a, b);
                case IGL_T_I64: return arith_int<int64_t, uint64_t>(nd.op, a, b);
                case IGL_T_F64: return arith_f64(nd.op, a, b);
                default: fail("unsupported arithmetic type");
            }
        }
        case IGL_OP_NEG: {
            Arr a = eval(*nd.a, bv);
            Arr r; r.alloc(a.type, n); r.valid = a.valid;
            for (int64_t i = 0; i < n; ++i) {
                if (!r.is_valid(i)) continue;
                switch (a.type) {
                    case IGL_T_I32: r.as<int32_t>()[i] = (int32_t)(0u - (uint32_t)a.as<int32_t>()[i]); break;
                    case IGL_T_I64: r.as<int64_t>()[i] = (int64_t)(0ull - (uint64_t)a.as<int64_t>()[i]); break;
                    case IGL_T_F64: r.as<double>()[i] = -a.as<double>()[i]; break;
                    default: fail("unsupported type for unary minus");
                }
            }
            return r;
        }
        case IGL_OP_AND: case IGL_OP_OR: {  // Kleene (arrow and_kleene / or_kleene)
            Arr a = eval(*nd.a, bv), b = eval(*nd.b, bv);
            if (a.type != IGL_T_BOOL || b.type != IGL_T_BOOL) fail("AND/OR operands must be boolean");
            Arr r; r.alloc(IGL_T_BOOL, n);
            bool anynull = !a.valid.empty() || !b.valid.empty();
            if (anynull) r.valid.assign(n, 1);
            for (int64_t i = 0; i < n; ++i) {
                bool av = a.is_valid(i), bvv = b.is_valid(i);
                bool x = av && a.data[i], y = bvv && b.data[i];
                if (nd.op == IGL_OP_AND) {
                    bool known_false = (av && !x) || (bvv && !y);
                    if (known_false) { r.data[i] = 0; }
                    else if (av && bvv) { r.data[i] = 1; }
                    else { r.data[i] = 0; r.valid[i] = 0; }
                } else {
                    bool known_true = x || y;
                    if (known_true) { r.data[i] = 1; }
                    else if (av && bvv) { r.data[i] = 0; }
                    else { r.data[i] = 0; r.valid[i] = 0; }
                }
            }
            return r;
        }
        case IGL_OP_NOT: {
            Arr a = eval(*nd.a, bv);
            if (a.type != IGL_T_BOOL) fail("NOT operand must be boolean");
            Arr r; r.alloc(IGL_T_BOOL, n); r.valid = a.valid;
            for (int64_t i = 0; i < n; ++i) r.data[i] = r.is_valid(i) ? !a.data[i] : 0;
            return r;
        }
        case IGL_OP_IS_NULL: case IGL_OP_IS_NOT_NULL: {
            Arr a = eval(*nd.a, bv);
            Arr r; r.alloc(IGL_T_BOOL, n);
            for (int64_t i = 0; i < n; ++i) r.data[i] = (nd.op == IGL_OP_IS_NULL) ? !a.is_valid(i) : a.is_valid(i);
            return r;
        }
        case IGL_OP_CAST: {
            Arr a = eval(*nd.a, bv);
            if (a.type == nd.cast_to) return a;
            Arr r; r.alloc(nd.cast_to, n); r.valid = a.valid;
            for (int64_t i = 0; i < n; ++i) {
                if (!r.is_valid(i)) continue;
                if (a.type == IGL_T_I32 && nd.cast_to == IGL_T_I64) r.as<int64_t>()[i] = a.as<int32_t>()[i];
                else if (a.type == IGL_T_I32 && nd.cast_to == IGL_T_F64) r.as<double>()[i] = (double)a.as<int32_t>()[i];
                else if (a.type == IGL_T_I64 && nd.cast_to == IGL_T_F64) r.as<double>()[i] = (double)a.as<int64_t>()[i];
                else if (a.type == IGL_T_I64 && nd.cast_to == IGL_T_I32) {
                    int64_t v = a.as<int64_t>()[i];
                    if (v < INT32_MIN || v > INT32_MAX) fail("Cast error: Can't cast value to type Int32");
                    r.as<int32_t>()[i] = (int32_t)v;
                } else if (a.type == IGL_T_F64 && nd.cast_to == IGL_T_I32) {
                    double d = std::trunc(a.as<double>()[i]);  // num::cast: truncate, error if unrepresentable
                    if (!(d >= -2147483648.0 && d <= 2147483647.0)) fail("Cast error: Can't cast value to type Int32");
                    r.as<int32_t>()[i] = (int32_t)d;
                } else if (a.type == IGL_T_F64 && nd.cast_to == IGL_T_I64) {
                    double d = std::trunc(a.as<double>()[i]);
                    if (!(d >= -9223372036854775808.0 && d < 9223372036854775808.0)) fail("Cast error: Can't cast value to type Int64");
                    r.as<int64_t>()[i] = (int64_t)d;
                } else fail("unsupported cast");
            }
            return r;
        }
    }
    fail("bad expression node");
}

// Static result type (PhysicalExpr::data_type): the schema of a stream that
// carries no batch at all still comes from the planner (projection.rs:19-29).
int infer_type(const Node& nd, const std::vector<Arr>& cols) {
    switch (nd.op) {
        case IGL_OP_COL: return (nd.col >= 0 && nd.col < (int)cols.size()) ? cols[nd.col].type : 0;
        case IGL_OP_LIT_BOOL: case IGL_OP_LIT_I32: case IGL_OP_LIT_I64: case IGL_OP_LIT_F64: case IGL_OP_LIT_NULL: return nd.lit_type;
        case IGL_OP_EQ: case IGL_OP_NE: case IGL_OP_LT: case IGL_OP_LE: case IGL_OP_GT: case IGL_OP_GE:
        case IGL_OP_AND: case IGL_OP_OR: case IGL_OP_NOT: case IGL_OP_IS_NULL: case IGL_OP_IS_NOT_NULL: return IGL_T_BOOL;
        case IGL_OP_CAST: return nd.cast_to;
        default: return infer_type(*nd.a, cols);
    }
}

// ----------------------------------------------------------------------------
// Operators
// ----------------------------------------------------------------------------
Table make_like(const Table& in) {
    Table t; t.cols.resize(in.cols.size());
    for (size_t j = 0; j < in.cols.size(); ++j) {
        t.cols[j].type = in.cols[j].type; t.cols[j].name = in.cols[j].name; t.cols[j].nullable = in.cols[j].nullable;
    }
    return t;
}

// filter.rs:39-68
Table* op_filter(const Table& in, const Node& pred) {
    auto out = std::make_unique<Table>(make_like(in));
    int64_t start = 0;
    for (int64_t bs : in.batch_sizes) {
        BatchView bv{&in.cols, start, bs};
        Arr mask = eval(pred, bv);                                  // filter.rs:47-48
        if (mask.type != IGL_T_BOOL) fail("Filter predicate must return boolean array");  // :50-55
        int64_t kept = 0;
        for (int64_t i = 0; i < bs;) {                              // filter_record_batch, :57
            if (!(mask.is_valid(i) && mask.data[i])) { ++i; continue; }
            int64_t e = i + 1;                                       // copy maximal runs, column by column
            while (e < bs && mask.is_valid(e) && mask.data[e]) ++e;
            for (size_t j = 0; j < in.cols.size(); ++j) append_rows(out->cols[j], in.cols[j], start + i, e - i);
            kept += e - i;
            i = e;
        }
        if (kept > 0) { out->batch_sizes.push_back(kept); out->nrows += kept; }  // :59-63 drop empty
        start += bs;
    }
    for (auto& c : out->cols) if (c.type == IGL_T_UTF8 && c.off.empty()) c.off.push_back(0);
    return out.release();
}

// projection.rs:48-72
Table* op_project(const Table& in, const std::vector<std::unique_ptr<Node>>& exprs, const std::vector<std::string>& names) {
    auto out = std::make_unique<Table>();
    out->cols.resize(exprs.size());
    int64_t start = 0;
    bool first = true;
    for (int64_t bs : in.batch_sizes) {
        BatchView bv{&in.cols, start, bs};
        for (size_t e = 0; e < exprs.size(); ++e) {
            Arr a = eval(*exprs[e], bv);                            // projection.rs:60-64
            if (first) { out->cols[e].type = a.type; }
            if (out->cols[e].type != a.type) fail("projection type changed between batches");
            append_rows(out->cols[e], a, 0, a.n);
        }
        first = false;
        out->batch_sizes.push_back(bs);                             // never drops a batch (:67)
        out->nrows += bs;
        start += bs;
    }
    for (size_t e = 0; e < exprs.size(); ++e) {
        out->cols[e].name = names[e];
        if (first) out->cols[e].type = infer_type(*exprs[e], in.cols);
        if (out->cols[e].type == IGL_T_UTF8 && out->cols[e].off.empty()) out->cols[e].off.push_back(0);
    }
    return out.release();
}

// Debug text of a one-row slice: what `format!("{:?}", key_array.slice(i, 1))`
// renders (hash_join.rs:120,168; arrow-array 55.1.0 `impl Debug`).
void debug_slice(std::string& s, const Arr& a, int64_t i) {
    char buf[64];
    switch (a.type) {
        case IGL_T_BOOL: s += "BooleanArray\n[\n  "; break;
        case IGL_T_I32: s += "PrimitiveArray<Int32>\n[\n  "; break;
        case IGL_T_I64: s += "PrimitiveArray<Int64>\n[\n  "; break;
        case IGL_T_F64: s += "PrimitiveArray<Float64>\n[\n  "; break;
        case IGL_T_UTF8: s += "StringArray\n[\n  "; break;
        default: fail("unsupported join key type");
    }
    if (!a.is_valid(i)) s += "null";
    else switch (a.type) {
        case IGL_T_BOOL: s += a.data[i] ? "true" : "false"; break;
        case IGL_T_I32: snprintf(buf, sizeof buf, "%d", a.as<int32_t>()[i]); s += buf; break;
        case IGL_T_I64: snprintf(buf, sizeof buf, "%" PRId64, a.as<int64_t>()[i]); s += buf; break;
        case IGL_T_F64: {
            double d = a.as<double>()[i];
            // Rust prints every NaN as "NaN" and round-trips everything else, so
            // identity of the text == identity of the bits, NaNs collapsed.
            if (std::isnan(d)) s += "NaN";
            else { snprintf(buf, sizeof buf, "%.17g", d); s += buf; }
            break;
        }
        case IGL_T_UTF8: s += '"'; s.append((const char*)a.data.data() + a.off[i], a.off[i + 1] - a.off[i]); s += '"'; break;
    }
    s += ",\n]";
}

enum { JOIN_INNER = 0, JOIN_LEFT = 1, JOIN_RIGHT = 2, JOIN_FULL = 3 };

// hash_join.rs:35-76
Table join_schema(const Table& left, const Table& right) {
    Table t;
    for (auto& c : left.cols) { Arr a; a.type = c.type; a.name = c.name; a.nullable = c.nullable; t.cols.push_back(std::move(a)); }
    for (auto& c : right.cols) {
        Arr a; a.type = c.type; a.nullable = c.nullable; a.name = c.name;
        for (auto& l : left.cols) if (l.name == c.name) { a.name = "right_" + c.name; break; }   // :56-61
        t.cols.push_back(std::move(a));
    }
    return t;
}

void emit_pairs(Table& out, const Table& left, const Table& right,
                const std::vector<int64_t>& li, const std::vector<int64_t>& ri) {
    const size_t nl = left.cols.size();
    for (size_t p = 0; p < li.size(); ++p) {
        for (size_t j = 0; j < nl; ++j) append_rows(out.cols[j], left.cols[j], li[p], 1);
        for (size_t j = 0; j < right.cols.size(); ++j) {
            if (ri[p] >= 0) append_rows(out.cols[nl + j], right.cols[j], ri[p], 1);      // combine_batches :221-240
            else append_null(out.cols[nl + j]);                                           // combine_with_nulls :242-262
        }
    }
    out.nrows += (int64_t)li.size();
}

std::vector<Arr> eval_keys(const Table& t, const std::vector<std::unique_ptr<Node>>& keys) {
    // Evaluated batch by batch as the reference does (hash_join.rs:106-113,
    // :145-162) and concatenated; a per-batch error aborts the stream.
    std::vector<Arr> out(keys.size());
    int64_t start = 0;
    bool first = true;
    for (int64_t bs : t.batch_sizes) {
        BatchView bv{&t.cols, start, bs};
        for (size_t k = 0; k < keys.size(); ++k) {
            Arr a = eval(*keys[k], bv);
            if (first) out[k].type = a.type;
            append_rows(out[k], a, 0, a.n);
        }
        first = false;
        start += bs;
    }
    for (auto& a : out) if (a.type == IGL_T_UTF8 && a.off.empty()) a.off.push_back(0);
    return out;
}

// FAITHFUL: hash_join.rs:100-213 as written.
Table* op_join_faithful(const Table& left, const Table& right,
                        const std::vector<std::unique_ptr<Node>>& lkeys,
                        const std::vector<std::unique_ptr<Node>>& rkeys, int join_type) {
    auto out = std::make_unique<Table>(join_schema(left, right));
    out->unit_batches = true;
    // build: drain the right child (:100-128)
    std::unordered_map<std::string, std::vector<int64_t>> table;
    {
        std::vector<Arr> ka = eval_keys(right, rkeys);
        std::string key;
        for (int64_t i = 0; i < right.nrows; ++i) {
            key.clear();
            for (auto& a : ka) debug_slice(key, a, i);               // :117-122
            table[key].push_back(i);                                 // :125-126 arrival order
        }
    }
    // probe (:131-213)
    std::vector<Arr> ka = eval_keys(left, lkeys);
    std::vector<int64_t> li, ri;
    std::string key;
    for (int64_t i = 0; i < left.nrows; ++i) {
        key.clear();
        for (auto& a : ka) debug_slice(key, a, i);                   // :165-170
        auto it = table.find(key);                                   // :174
        if (it != table.end()) {
            for (int64_t r : it->second) { li.push_back(i); ri.push_back(r); }   // :176-181
        } else if (join_type == JOIN_LEFT || join_type == JOIN_FULL) {           // :189
            li.push_back(i); ri.push_back(-1);
        }
        // Right / Full never emit unmatched right rows: no visited tracking exists
        // and combine_with_nulls(left_side=false) is unreachable (:263-276).
    }
    emit_pairs(*out, left, right, li, ri);
    for (auto& c : out->cols) if (c.type == IGL_T_UTF8 && c.off.empty()) c.off.push_back(0);
    return out.release();
}

// ---- FAST mode -------------------------------------------------------------
inline uint64_t splitmix64(uint64_t x) {
    x += 0x9E3779B97F4A7C15ull;
    x = (x ^ (x >> 30)) * 0xBF58476D1CE4E5B9ull;
    x = (x ^ (x >> 27)) * 0x94D049BB133111EBull;
    return x ^ (x >> 31);
}

// Canonical 64-bit identity of one key value == identity of its Debug text.
inline bool canon_key(const Arr& a, int64_t i, uint64_t& k) {
    if (!a.is_valid(i)) return false;
    switch (a.type) {
        case IGL_T_BOOL: k = a.data[i] ? 1 : 0; break;
        case IGL_T_I32: k = (uint64_t)(int64_t)a.as<int32_t>()[i]; break;
        case IGL_T_I64: k = (uint64_t)a.as<int64_t>()[i]; break;
        case IGL_T_F64: { double d = a.as<double>()[i]; if (std::isnan(d)) k = 0x7FF8000000000000ull; else memcpy(&k, &d, 8); break; }
        default: fail("fast join supports BOOL/I32/I64/F64 keys");
    }
    return true;
}

Table* op_join_fast(const Table& left, const Table& right,
                    const std::vector<std::unique_ptr<Node>>& lkeys,
                    const std::vector<std::unique_ptr<Node>>& rkeys, int join_type, int threads) {
    if (lkeys.size() != 1 || rkeys.size() != 1) fail("fast join supports one key column");
    auto out = std::make_unique<Table>(join_schema(left, right));
    out->unit_batches = true;
    std::vector<Arr> rk = eval_keys(right, rkeys), lk = eval_keys(left, lkeys);
    const bool same_type = rk[0].type == lk[0].type;   // the type name is part of the printed key
    const int64_t nr = right.nrows, nl = left.nrows;
    uint64_t cap = 16; while (cap < (uint64_t)nr * 2) cap <<= 1;
    const uint64_t mask = cap - 1;
    std::vector<uint64_t> skey(cap);
    std::vector<int64_t> head(cap, -1), tail(cap, -1), next((size_t)nr, -1);
    int64_t null_head = -1, null_tail = -1;
    for (int64_t i = 0; i < nr; ++i) {
        uint64_t k;
        if (!canon_key(rk[0], i, k)) {                               // NULL == NULL (printed "null")
            if (null_head < 0) null_head = i; else next[null_tail] = i;
            null_tail = i; continue;
        }
        uint64_t h = splitmix64(k) & mask;
        while (head[h] >= 0 && skey[h] != k) h = (h + 1) & mask;
        if (head[h] < 0) { skey[h] = k; head[h] = i; } else next[tail[h]] = i;
        tail[h] = i;
    }
    if (threads < 1) threads = 1;
    std::vector<std::vector<int64_t>> tli(threads), tri(threads);
    auto work = [&](int t) {
        int64_t lo = nl * t / threads, hi = nl * (t + 1) / threads;
        auto& li = tli[t]; auto& ri = tri[t];
        for (int64_t i = lo; i < hi; ++i) {
            int64_t r = -1;
            uint64_t k;
            if (same_type) {
                if (!canon_key(lk[0], i, k)) r = null_head;
                else {
                    uint64_t h = splitmix64(k) & mask;
                    while (head[h] >= 0 && skey[h] != k) h = (h + 1) & mask;
                    r = head[h];
                }
            }
            if (r >= 0) { for (; r >= 0; r = next[r]) { li.push_back(i); ri.push_back(r); } }
            else if (join_type == JOIN_LEFT || join_type == JOIN_FULL) { li.push_back(i); ri.push_back(-1); }
        }
    };
    std::vector<std::thread> th;
    for (int t = 1; t < threads; ++t) th.emplace_back(work, t);
    work(0);
    for (auto& x : th) x.join();
    for (int t = 0; t < threads; ++t) emit_pairs(*out, left, right, tli[t], tri[t]);
    for (auto& c : out->cols) if (c.type == IGL_T_UTF8 && c.off.empty()) c.off.push_back(0);
    return out.release();
}

// ----------------------------------------------------------------------------
// Order-independent multiset digest of a table's rows (SURVEY.md §8d "parity
// check at scale").  The same arithmetic is implemented separately on the
// device (igloo_b200/csrc/gen.cu); UTF8 columns are folded by FNV-1a of bytes.
// ----------------------------------------------------------------------------
inline uint64_t col_bits(const Arr& a, int64_t i) {
    switch (a.type) {
        case IGL_T_BOOL: return a.data[i] ? 1 : 0;
        case IGL_T_I32: return (uint64_t)(int64_t)a.as<int32_t>()[i];
        case IGL_T_I64: return (uint64_t)a.as<int64_t>()[i];
        case IGL_T_F64: { uint64_t b; memcpy(&b, a.as<double>() + i, 8); return b; }
        case IGL_T_UTF8: {
            uint64_t h = 0xcbf29ce484222325ull;
            for (int32_t p = a.off[i]; p < a.off[i + 1]; ++p) { h ^= a.data[p]; h *= 0x100000001b3ull; }
            return h;
        }
    }
    return 0;
}

void digest(const Table& t, uint64_t out[3]) {
    uint64_t s = 0, x = 0, s2 = 0;
    for (int64_t i = 0; i < t.nrows; ++i) {
        uint64_t h = 0x243F6A8885A308D3ull;
        for (size_t j = 0; j < t.cols.size(); ++j) {
            const Arr& a = t.cols[j];
            uint64_t bits = a.is_valid(i) ? col_bits(a, i) : (0x9E3779B97F4A7C15ull ^ (uint64_t)j);
            h = splitmix64(h ^ (bits + (uint64_t)j * 0xD6E8FEB86659FD93ull));
        }
        s += h; x ^= h; s2 += splitmix64(h ^ 0xA5A5A5A5A5A5A5A5ull);
    }
    out[0] = s; out[1] = x; out[2] = s2;
}

// ----------------------------------------------------------------------------
// Deterministic generator
// ----------------------------------------------------------------------------
inline uint64_t gen_u64(uint64_t seed, uint64_t col, uint64_t row) {
    return splitmix64((seed ^ (col * 0x9E3779B97F4A7C15ull)) + row);
}
inline uint64_t mulhi64(uint64_t a, uint64_t b) { return (uint64_t)(((unsigned __int128)a * b) >> 64); }

// Bijection on [0, 2^bits): odd multiplies and xor-shifts, each invertible.
inline uint64_t perm_pow2(uint64_t x, int bits) {
    const uint64_t M = (bits >= 64) ? ~0ull : ((1ull << bits) - 1);
    const int s1 = bits / 2 + 1, s2 = bits / 2, s3 = bits / 2 + 2;
    x = (x * 0x9E3779B97F4A7C15ull) & M; x ^= x >> s1;
    x = (x * 0xBF58476D1CE4E5B9ull) & M; x ^= x >> s2;
    x = (x * 0x94D049BB133111EBull) & M; x ^= x >> s3;
    return x;
}
inline uint64_t inv_odd(uint64_t a) { uint64_t x = a; for (int i = 0; i < 6; ++i) x *= 2 - a * x; return x; }
inline uint64_t unxorshift(uint64_t y, int s) { uint64_t x = y; for (int i = s; i < 64; i += s) x = y ^ (x >> s); return x; }
inline uint64_t perm_pow2_inv(uint64_t x, int bits) {
    const uint64_t M = (bits >= 64) ? ~0ull : ((1ull << bits) - 1);
    const int s1 = bits / 2 + 1, s2 = bits / 2, s3 = bits / 2 + 2;
    x = unxorshift(x, s3); x = (x * inv_odd(0x94D049BB133111EBull)) & M;
    x = unxorshift(x, s2); x = (x * inv_odd(0xBF58476D1CE4E5B9ull)) & M;
    x = unxorshift(x, s1); x = (x * inv_odd(0x9E3779B97F4A7C15ull)) & M;
    return x;
}

}  // namespace

// ----------------------------------------------------------------------------
// C interface (ctypes from tests/ and bench.py only)
// ----------------------------------------------------------------------------
extern "C" {

typedef struct {
    int32_t type;
    int32_t nullable;
    int64_t n;
    const void* data;        // values (BOOL: one byte per row); UTF8: bytes
    const uint8_t* valid;    // one byte per row, or NULL
    const int32_t* offsets;  // UTF8 only
    const char* name;
} orc_col;

typedef struct Table orc_table;

static thread_local std::string g_err;
const char* orc_last_error(void) { return g_err.c_str(); }

#define ORC_TRY(...) try { __VA_ARGS__ } catch (const std::exception& e) { g_err = e.what(); return nullptr; }

// "Scan": wrap host columns as a stream of batch_rows-row batches
// (parquet_scan.rs:54 uses 1024).
orc_table* orc_scan(const orc_col* cols, int ncols, int64_t nrows, int64_t batch_rows) {
    ORC_TRY(
        auto t = std::make_unique<Table>();
        t->nrows = nrows;
        for (int j = 0; j < ncols; ++j) {
            Arr a; a.type = cols[j].type; a.n = nrows; a.nullable = cols[j].nullable != 0;
            a.name = cols[j].name ? cols[j].name : "";
            if (a.type == IGL_T_UTF8) {
                a.off.assign(cols[j].offsets, cols[j].offsets + nrows + 1);
                a.data.assign((const uint8_t*)cols[j].data, (const uint8_t*)cols[j].data + a.off[nrows]);
            } else {
                int w = width_of(a.type);
                if (!w) fail("unsupported column type");
                a.data.assign((const uint8_t*)cols[j].data, (const uint8_t*)cols[j].data + nrows * w);
            }
            if (cols[j].valid) a.valid.assign(cols[j].valid, cols[j].valid + nrows);
            t->cols.push_back(std::move(a));
        }
        if (batch_rows <= 0) batch_rows = 1024;
        for (int64_t s = 0; s < nrows; s += batch_rows) t->batch_sizes.push_back(std::min(batch_rows, nrows - s));
        return t.release();
    )
}

orc_table* orc_filter(const orc_table* in, const uint8_t* pred, size_t len) {
    ORC_TRY( auto p = parse(pred, len); return op_filter(*in, *p); )
}

orc_table* orc_project(const orc_table* in, const uint8_t* const* exprs, const size_t* lens,
                       const char* const* names, int nexpr) {
    ORC_TRY(
        std::vector<std::unique_ptr<Node>> e; std::vector<std::string> nm;
        for (int i = 0; i < nexpr; ++i) { e.push_back(parse(exprs[i], lens[i])); nm.push_back(names && names[i] ? names[i] : ""); }
        return op_project(*in, e, nm);
    )
}

// mode: 0 = faithful (printed-string keys), 1 = fast (typed, `threads` threads)
orc_table* orc_hash_join(const orc_table* left, const orc_table* right,
                         const uint8_t* const* lkeys, const size_t* llens,
                         const uint8_t* const* rkeys, const size_t* rlens, int nkeys,
                         int join_type, int mode, int threads) {
    ORC_TRY(
        std::vector<std::unique_ptr<Node>> lk, rk;
        for (int i = 0; i < nkeys; ++i) { lk.push_back(parse(lkeys[i], llens[i])); rk.push_back(parse(rkeys[i], rlens[i])); }
        if (mode == 0) return op_join_faithful(*left, *right, lk, rk, join_type);
        return op_join_fast(*left, *right, lk, rk, join_type, threads);
    )
}

int32_t orc_table_ncols(const orc_table* t) { return (int32_t)t->cols.size(); }
int64_t orc_table_nrows(const orc_table* t) { return t->nrows; }
int64_t orc_table_nbatches(const orc_table* t) { return t->n_batches(); }
int64_t orc_table_batch_size(const orc_table* t, int64_t i) { return t->unit_batches ? 1 : t->batch_sizes[i]; }
void orc_table_col(const orc_table* t, int j, orc_col* out) {
    const Arr& a = t->cols[j];
    out->type = a.type; out->nullable = a.nullable; out->n = a.n;
    out->data = a.data.data(); out->valid = a.valid.empty() ? nullptr : a.valid.data();
    out->offsets = a.off.empty() ? nullptr : a.off.data(); out->name = a.name.c_str();
}
void orc_table_digest(const orc_table* t, uint64_t out[3]) { digest(*t, out); }
void orc_table_free(orc_table* t) { delete t; }

// generator
uint64_t orc_splitmix64(uint64_t x) { return splitmix64(x); }
void orc_gen_u64(uint64_t seed, uint64_t col, int64_t row0, int64_t n, uint64_t* out) {
    for (int64_t i = 0; i < n; ++i) out[i] = gen_u64(seed, col, (uint64_t)(row0 + i));
}
void orc_gen_i32_uniform(uint64_t seed, uint64_t col, int64_t row0, int64_t n, uint64_t m, int32_t* out) {
    for (int64_t i = 0; i < n; ++i) out[i] = (int32_t)mulhi64(gen_u64(seed, col, (uint64_t)(row0 + i)), m);
}
void orc_gen_i64_uniform(uint64_t seed, uint64_t col, int64_t row0, int64_t n, uint64_t m, int64_t* out) {
    for (int64_t i = 0; i < n; ++i) {
        uint64_t u = gen_u64(seed, col, (uint64_t)(row0 + i));
        out[i] = (int64_t)(m ? mulhi64(u, m) : u);
    }
}
void orc_gen_f64_unit(uint64_t seed, uint64_t col, int64_t row0, int64_t n, double* out) {
    for (int64_t i = 0; i < n; ++i) out[i] = (double)(gen_u64(seed, col, (uint64_t)(row0 + i)) >> 11) * 0x1.0p-53;
}
void orc_gen_perm_pow2(int64_t row0, int64_t n, int bits, int64_t* out) {
    for (int64_t i = 0; i < n; ++i) out[i] = (int64_t)perm_pow2((uint64_t)(row0 + i), bits);
}
uint64_t orc_perm_pow2(uint64_t x, int bits) { return perm_pow2(x, bits); }
uint64_t orc_perm_pow2_inv(uint64_t x, int bits) { return perm_pow2_inv(x, bits); }

// ----------------------------------------------------------------------------
// Timed CPU baselines for bench.py (the reference's algorithm on host cores).
// Each returns seconds spent inside the operator loop, data generated outside.
// ----------------------------------------------------------------------------
}  // extern "C"

#include <chrono>
static double now_s() { return std::chrono::duration<double>(std::chrono::steady_clock::now().time_since_epoch()).count(); }

extern "C" {

// C2 filter+project over `nrows` generated rows, reference streaming shape:
// 1024-row batches, Filter then Projection, every expression node materialised.
// Returns seconds; *out_rows receives the surviving row count.
double orc_bench_filter_project(uint64_t seed, int64_t nrows, int32_t threshold, int64_t* out_rows) {
    std::vector<int32_t> a(nrows), b(nrows); std::vector<double> x(nrows);
    orc_gen_i32_uniform(seed, 0, 0, nrows, 1000, a.data());
    orc_gen_i32_uniform(seed, 1, 0, nrows, 1ull << 31, b.data());
    orc_gen_f64_unit(seed, 2, 0, nrows, x.data());
    orc_col cols[3] = {{IGL_T_I32, 0, nrows, a.data(), nullptr, nullptr, "a"},
                       {IGL_T_I32, 0, nrows, b.data(), nullptr, nullptr, "b"},
                       {IGL_T_F64, 0, nrows, x.data(), nullptr, nullptr, "x"}};
    orc_table* scan = orc_scan(cols, 3, nrows, 1024);
    // a < threshold
    uint8_t pred[16]; size_t pl = 0;
    pred[pl++] = IGL_OP_COL; pred[pl++] = 0; pred[pl++] = 0;
    pred[pl++] = IGL_OP_LIT_I32; memcpy(pred + pl, &threshold, 4); pl += 4;
    pred[pl++] = IGL_OP_LT;
    // b + a ; x * 1.5 + 0.25
    uint8_t e0[8] = {IGL_OP_COL, 1, 0, IGL_OP_COL, 0, 0, IGL_OP_ADD};
    uint8_t e1[32]; size_t l1 = 0; double c1 = 1.5, c2 = 0.25;
    e1[l1++] = IGL_OP_COL; e1[l1++] = 2; e1[l1++] = 0;
    e1[l1++] = IGL_OP_LIT_F64; memcpy(e1 + l1, &c1, 8); l1 += 8; e1[l1++] = IGL_OP_MUL;
    e1[l1++] = IGL_OP_LIT_F64; memcpy(e1 + l1, &c2, 8); l1 += 8; e1[l1++] = IGL_OP_ADD;
    const uint8_t* ex[2] = {e0, e1}; size_t ln[2] = {7, l1}; const char* nm[2] = {"s", "y"};
    double t0 = now_s();
    orc_table* f = orc_filter(scan, pred, pl);
    orc_table* p = f ? orc_project(f, ex, ln, nm, 2) : nullptr;
    double t1 = now_s();
    if (out_rows) *out_rows = p ? p->nrows : -1;
    orc_table_free(scan); if (f) orc_table_free(f); if (p) orc_table_free(p);
    return t1 - t0;
}

// C3-shaped inner join at reduced scale: build 2^build_bits unique keys (k, v),
// probe `nprobe` uniform keys (k, w).  mode 0 = faithful, 1 = fast.
double orc_bench_join(uint64_t seed, int build_bits, int64_t nprobe, int mode, int threads, int64_t* out_rows) {
    const int64_t nb = 1ll << build_bits;
    std::vector<int64_t> rk(nb), rv(nb), lk(nprobe), lw(nprobe);
    orc_gen_perm_pow2(0, nb, build_bits, rk.data());
    orc_gen_i64_uniform(seed, 1, 0, nb, 0, rv.data());
    orc_gen_i64_uniform(seed, 2, 0, nprobe, (uint64_t)nb, lk.data());
    orc_gen_i64_uniform(seed, 3, 0, nprobe, 0, lw.data());
    orc_col rc[2] = {{IGL_T_I64, 0, nb, rk.data(), nullptr, nullptr, "k"}, {IGL_T_I64, 0, nb, rv.data(), nullptr, nullptr, "v"}};
    orc_col lc[2] = {{IGL_T_I64, 0, nprobe, lk.data(), nullptr, nullptr, "k"}, {IGL_T_I64, 0, nprobe, lw.data(), nullptr, nullptr, "w"}};
    orc_table* R = orc_scan(rc, 2, nb, 1024);
    orc_table* L = orc_scan(lc, 2, nprobe, 1024);
    uint8_t key[3] = {IGL_OP_COL, 0, 0};
    const uint8_t* ks[1] = {key}; size_t kl[1] = {3};
    double t0 = now_s();
    orc_table* J = orc_hash_join(L, R, ks, kl, ks, kl, 1, JOIN_INNER, mode, threads);
    double t1 = now_s();
    if (out_rows) *out_rows = J ? J->nrows : -1;
    orc_table_free(R); orc_table_free(L); if (J) orc_table_free(J);
    return t1 - t0;
}

}  // extern "C"
