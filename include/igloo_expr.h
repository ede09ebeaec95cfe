/*
 * igloo_expr.h — wire format of the expression bytecode accepted by
 * igl_expr_compile() (include/igloo_cuda.h).
 *
 * Reference interface replaced: the `Arc<dyn PhysicalExpr>` values that
 * PhysicalPlanner::create_physical_plan lowers from DataFusion `Expr`s and
 * hands to the operators (crates/engine/src/physical_planner.rs:60-64,
 * :86-90, :107-116) and that the operators evaluate per batch
 * (filter.rs:47, projection.rs:61, hash_join.rs:110,147).  A trait object
 * cannot cross a C ABI, so the Rust shim serialises the (already type-coerced)
 * PhysicalExpr tree into this postfix byte string; see INTEGRATION.md.
 *
 * Encoding: a sequence of tokens in postfix (RPN) order.  Each token is one
 * opcode byte followed by its immediate (little endian).  A valid program
 * leaves exactly one value on the stack.  Types must already agree the way
 * DataFusion's type coercion leaves them (both sides of a binary operator have
 * the same type); igl_expr_compile() rejects anything else with
 * IGL_ERR_INVALID_ARG instead of guessing.
 */
#ifndef IGLOO_EXPR_H
#define IGLOO_EXPR_H

#ifdef __cplusplus
extern "C" {
#endif

/* Column / value types (subset of Arrow the configs use). */
enum igl_type {
    IGL_T_INVALID = 0,
    IGL_T_BOOL = 1, /* Arrow Boolean (bit-packed in Arrow buffers)        */
    IGL_T_I32 = 2,  /* Arrow Int32                                        */
    IGL_T_I64 = 3,  /* Arrow Int64                                        */
    IGL_T_F64 = 4,  /* Arrow Float64                                      */
    IGL_T_UTF8 = 5  /* Arrow Utf8 (i32 offsets); pass-through columns only */
};

enum igl_opcode {
    /* leaves */
    IGL_OP_COL = 0x01,      /* + u16 column index into the input schema    */
    IGL_OP_LIT_BOOL = 0x02, /* + u8                                         */
    IGL_OP_LIT_I32 = 0x03,  /* + i32                                        */
    IGL_OP_LIT_I64 = 0x04,  /* + i64                                        */
    IGL_OP_LIT_F64 = 0x05,  /* + f64 (IEEE-754 bits)                        */
    IGL_OP_LIT_NULL = 0x06, /* + u8 igl_type                                */
    /* comparisons -> BOOL (floats: IEEE-754 totalOrder, as arrow-ord cmp)  */
    IGL_OP_EQ = 0x10,
    IGL_OP_NE = 0x11,
    IGL_OP_LT = 0x12,
    IGL_OP_LE = 0x13,
    IGL_OP_GT = 0x14,
    IGL_OP_GE = 0x15,
    /* arithmetic: ints wrap; int / and % by zero raise; f64 is IEEE        */
    IGL_OP_ADD = 0x20,
    IGL_OP_SUB = 0x21,
    IGL_OP_MUL = 0x22,
    IGL_OP_DIV = 0x23,
    IGL_OP_MOD = 0x24,
    IGL_OP_NEG = 0x25,
    /* boolean (Kleene three-valued AND / OR)                               */
    IGL_OP_AND = 0x30,
    IGL_OP_OR = 0x31,
    IGL_OP_NOT = 0x32,
    IGL_OP_IS_NULL = 0x33,
    IGL_OP_IS_NOT_NULL = 0x34,
    /* CAST among I32 / I64 / F64: + u8 target igl_type; overflow raises    */
    IGL_OP_CAST = 0x40
};

#ifdef __cplusplus
}
#endif
#endif /* IGLOO_EXPR_H */
