// expr_compile.cpp — expression bytecode -> typed tree -> three-address IR.
//
// Stands in for DataFusion's `create_physical_expr` + `PhysicalExpr::data_type`
// / `::nullable` as the reference's planner uses them
// (crates/engine/src/physical_planner.rs:60-71, :86-90, :107-116).  Type rules
// are DataFusion's after coercion: both operands of a binary operator have one
// type; comparisons and IS [NOT] NULL yield Boolean; AND / OR / NOT take
// Boolean; arithmetic keeps the operand type.
#include "expr_ir.h"

#include <cstring>

using namespace igl;

namespace {

bool is_numeric(int t) { return t == IGL_T_I32 || t == IGL_T_I64 || t == IGL_T_F64; }

}  // namespace

extern "C" int32_t igl_expr_compile(igl_ctx* ctx, const uint8_t* bc, size_t len, const int32_t* col_types,
                                    const uint8_t* col_nullable, int32_t ncols, igl_expr** out) {
    (void)ctx;
    if (!out) return fail(IGL_ERR_INVALID_ARG, "igl_expr_compile: out is NULL");
    *out = nullptr;
    if (!bc || len == 0) return fail(IGL_ERR_INVALID_ARG, "igl_expr_compile: empty bytecode");
    if (ncols < 0 || (ncols > 0 && !col_types)) return fail(IGL_ERR_INVALID_ARG, "igl_expr_compile: bad schema");
    // the IR addresses columns with 8 bits (Instr::a_idx / b_idx): wider schemas are refused, not truncated
    if (ncols > 256) return fail(IGL_ERR_UNSUPPORTED, "schemas of more than 256 columns are not supported (got %d)", ncols);
    auto e = std::make_unique<igl_expr>();
    e->ncols = ncols;
    e->col_types.assign(col_types, col_types + ncols);
    if (col_nullable) e->col_nullable.assign(col_nullable, col_nullable + ncols);
    else e->col_nullable.assign(ncols, 0);
    std::vector<int> st;
    size_t p = 0;
    auto& N = e->nodes;
#define NEED(k) if (p + (k) > len) return fail(IGL_ERR_INVALID_ARG, "truncated expression bytecode")
#define POP(var) if (st.empty()) return fail(IGL_ERR_INVALID_ARG, "expression stack underflow"); int var = st.back(); st.pop_back()
    while (p < len) {
        ExprNode n;
        n.op = bc[p++];
        switch (n.op) {
            case IGL_OP_COL: {
                NEED(2);
                uint16_t c; memcpy(&c, bc + p, 2); p += 2;
                if (c >= ncols) return fail(IGL_ERR_INVALID_ARG, "column index %u out of range (%d columns)", c, ncols);
                n.col = c; n.type = e->col_types[c]; n.nullable = e->col_nullable[c] != 0;
                if (n.type < IGL_T_BOOL || n.type > IGL_T_UTF8) return fail(IGL_ERR_UNSUPPORTED, "unsupported column type %d", n.type);
                break;
            }
            case IGL_OP_LIT_BOOL: NEED(1); n.type = IGL_T_BOOL; n.lit_bits = bc[p++] ? 1 : 0; break;
            case IGL_OP_LIT_I32: { NEED(4); int32_t v; memcpy(&v, bc + p, 4); p += 4; n.type = IGL_T_I32; n.lit_bits = (uint64_t)(uint32_t)v; break; }
            case IGL_OP_LIT_I64: NEED(8); memcpy(&n.lit_bits, bc + p, 8); p += 8; n.type = IGL_T_I64; break;
            case IGL_OP_LIT_F64: NEED(8); memcpy(&n.lit_bits, bc + p, 8); p += 8; n.type = IGL_T_F64; break;
            case IGL_OP_LIT_NULL: {
                NEED(1); n.type = bc[p++]; n.lit_null = true; n.nullable = true;
                if (n.type < IGL_T_BOOL || n.type > IGL_T_F64) return fail(IGL_ERR_UNSUPPORTED, "unsupported NULL literal type %d", n.type);
                break;
            }
            case IGL_OP_EQ: case IGL_OP_NE: case IGL_OP_LT: case IGL_OP_LE: case IGL_OP_GT: case IGL_OP_GE: {
                POP(b); POP(a);
                if (N[a].type != N[b].type) return fail(IGL_ERR_INVALID_ARG, "comparison operands have different types (%d vs %d); coerce in the planner", N[a].type, N[b].type);
                if (N[a].type == IGL_T_UTF8) return fail(IGL_ERR_UNSUPPORTED, "Utf8 columns are pass-through only");
                n.a = a; n.b = b; n.type = IGL_T_BOOL; n.nullable = N[a].nullable || N[b].nullable;
                break;
            }
            case IGL_OP_ADD: case IGL_OP_SUB: case IGL_OP_MUL: case IGL_OP_DIV: case IGL_OP_MOD: {
                POP(b); POP(a);
                if (N[a].type != N[b].type) return fail(IGL_ERR_INVALID_ARG, "arithmetic operands have different types (%d vs %d); coerce in the planner", N[a].type, N[b].type);
                if (!is_numeric(N[a].type)) return fail(IGL_ERR_INVALID_ARG, "arithmetic needs numeric operands");
                n.a = a; n.b = b; n.type = N[a].type; n.nullable = N[a].nullable || N[b].nullable;
                break;
            }
            case IGL_OP_AND: case IGL_OP_OR: {
                POP(b); POP(a);
                if (N[a].type != IGL_T_BOOL || N[b].type != IGL_T_BOOL) return fail(IGL_ERR_INVALID_ARG, "AND/OR need boolean operands");
                n.a = a; n.b = b; n.type = IGL_T_BOOL; n.nullable = N[a].nullable || N[b].nullable;
                break;
            }
            case IGL_OP_NOT: {
                POP(a);
                if (N[a].type != IGL_T_BOOL) return fail(IGL_ERR_INVALID_ARG, "NOT needs a boolean operand");
                n.a = a; n.type = IGL_T_BOOL; n.nullable = N[a].nullable;
                break;
            }
            case IGL_OP_NEG: {
                POP(a);
                if (!is_numeric(N[a].type)) return fail(IGL_ERR_INVALID_ARG, "unary minus needs a numeric operand");
                n.a = a; n.type = N[a].type; n.nullable = N[a].nullable;
                break;
            }
            case IGL_OP_IS_NULL: case IGL_OP_IS_NOT_NULL: {
                POP(a);
                if (N[a].type == IGL_T_UTF8) return fail(IGL_ERR_UNSUPPORTED, "Utf8 columns are pass-through only");
                n.a = a; n.type = IGL_T_BOOL; n.nullable = false;
                break;
            }
            case IGL_OP_CAST: {
                NEED(1); n.cast_to = bc[p++];
                POP(a);
                if (!is_numeric(N[a].type) || !is_numeric(n.cast_to)) return fail(IGL_ERR_UNSUPPORTED, "unsupported CAST %d -> %d", N[a].type, n.cast_to);
                n.a = a; n.type = n.cast_to; n.nullable = N[a].nullable;
                break;
            }
            default:
                return fail(IGL_ERR_INVALID_ARG, "unknown opcode 0x%02x in expression bytecode", n.op);
        }
        N.push_back(n);
        // lowering recurses once per node: bound the tree here, long before the host stack is at risk
        if (N.size() > (size_t)(4 * kMaxInstr)) return fail(IGL_ERR_UNSUPPORTED, "expression has more than %d nodes", 4 * kMaxInstr);
        st.push_back((int)N.size() - 1);
    }
#undef NEED
#undef POP
    if (st.size() != 1) return fail(IGL_ERR_INVALID_ARG, "expression must leave exactly one value (left %zu)", st.size());
    e->type = N.back().type;
    e->nullable = N.back().nullable;
    *out = e.release();
    return IGL_OK;
}

extern "C" int32_t igl_expr_type(const igl_expr* e, int32_t* type, int32_t* nullable) {
    if (!e) return fail(IGL_ERR_INVALID_ARG, "igl_expr_type: NULL expression");
    if (type) *type = e->type;
    if (nullable) *nullable = e->nullable ? 1 : 0;
    return IGL_OK;
}

extern "C" void igl_expr_destroy(igl_expr* e) { delete e; }

namespace igl {

namespace {
struct Operand { uint8_t kind; uint8_t idx; uint64_t lit; bool lit_null; int producer; };

struct Lowerer {
    const igl_expr& e;
    std::vector<Instr>& code;
    bool tmp_used[kMaxTmp] = {};
    int max_tmp = 0;
    int32_t err = IGL_OK;

    bool is_leaf(int ni) const {
        const int op = e.nodes[ni].op;
        return op == IGL_OP_COL || (op >= IGL_OP_LIT_BOOL && op <= IGL_OP_LIT_NULL);
    }
    int alloc_tmp() {
        for (int i = 0; i < kMaxTmp; ++i)
            if (!tmp_used[i]) { tmp_used[i] = true; if (i + 1 > max_tmp) max_tmp = i + 1; return i; }
        return -1;
    }
    void release(const Operand& o) { if (o.kind == OPK_TMP) tmp_used[o.idx] = false; }

    Operand lower(int ni) {
        const ExprNode& n = e.nodes[ni];
        if (n.op == IGL_OP_COL) return {OPK_COL, (uint8_t)n.col, 0, false, -1};
        if (n.op >= IGL_OP_LIT_BOOL && n.op <= IGL_OP_LIT_NULL) return {OPK_LIT, 0, n.lit_bits, n.lit_null, -1};
        Operand a{OPK_NONE, 0, 0, false, -1}, b{OPK_NONE, 0, 0, false, -1};
        if (n.a >= 0) { a = lower(n.a); if (err < 0) return a; }
        if (n.b >= 0) {
            if (a.kind == OPK_ACC && !is_leaf(n.b)) {
                // the right subtree will overwrite the accumulator: park the left
                // result in a temporary by retargeting the instruction that made it
                const int t = alloc_tmp();
                if (t < 0) { err = fail(IGL_ERR_UNSUPPORTED, "expression needs more than %d parked temporaries", kMaxTmp); return a; }
                code[a.producer].dst_kind = OPK_TMP;
                code[a.producer].dst_idx = (uint8_t)t;
                a.kind = OPK_TMP; a.idx = (uint8_t)t;
            }
            b = lower(n.b);
            if (err < 0) return b;
        }
        release(a); release(b);
        Instr in{};
        in.op = (uint8_t)n.op;
        in.type = (uint8_t)(n.a >= 0 ? e.nodes[n.a].type : n.type);  // operand type
        in.a_kind = a.kind; in.a_idx = a.idx; in.a_lit = a.lit;
        in.b_kind = b.kind; in.b_idx = b.idx; in.b_lit = b.lit;
        in.dst_kind = OPK_ACC;
        in.aux = (n.op == IGL_OP_CAST) ? (uint8_t)n.cast_to : (uint8_t)((a.lit_null ? AUX_A_NULL : 0) | (b.lit_null ? AUX_B_NULL : 0));
        if (n.op == IGL_OP_CAST && a.lit_null) {  // CAST(NULL) stays NULL: lower as a null literal move
            in.op = IOP_MOV; in.type = (uint8_t)n.type; in.aux = AUX_A_NULL;
        }
        code.push_back(in);
        return {OPK_ACC, 0, 0, false, (int)code.size() - 1};
    }
};
}  // namespace

int32_t lower_expr(const igl_expr& e, std::vector<Instr>& code, LoweredExpr& out) {
    Lowerer L{e, code};
    out.first = (int)code.size();
    Operand r = L.lower((int)e.nodes.size() - 1);
    if (L.err < 0) return L.err;
    if (r.kind != OPK_ACC) {  // a bare column or literal: one MOV so the result sits in the accumulator
        Instr in{};
        in.op = IOP_MOV; in.type = (uint8_t)e.type;
        in.a_kind = r.kind; in.a_idx = r.idx; in.a_lit = r.lit;
        in.dst_kind = OPK_ACC;
        in.aux = r.lit_null ? AUX_A_NULL : 0;
        code.push_back(in);
    }
    out.count = (int)code.size() - out.first;
    out.n_tmp = L.max_tmp;
    out.type = e.type;
    out.nullable = e.nullable;
    if ((int)code.size() > kMaxInstr) return fail(IGL_ERR_UNSUPPORTED, "fused operator exceeds %d instructions", kMaxInstr);
    return IGL_OK;
}

void referenced_columns(const igl_expr& e, std::vector<bool>& used) {
    for (auto& n : e.nodes) if (n.op == IGL_OP_COL) used[n.col] = true;
}

}  // namespace igl
