// expr_isa.h — the fused-operator instruction set shared by the host lowering
// (expr_compile.cpp), the interpreter kernels and the run-time-specialised kernels
// (fp_jit.cpp compiles fp_device.cuh with NVRTC: this header must stay free of host-only
// includes).
#pragma once

#ifdef __CUDACC_RTC__
#include "igloo_cuda.h"
#include "igloo_expr.h"
#else
#include <stdint.h>
#include "../../include/igloo_cuda.h"
#include "../../include/igloo_expr.h"
#endif

namespace igl {

// ---- expression IR -----------------------------------------------------------
// Accumulator machine with operand modes.  A leaf (column or literal) is an
// operand of its consumer, never an instruction of its own, so `a < 250` is one
// instruction and `x * 1.5 + 0.25` is two.  An inner result lives in the
// accumulator (registers); when the right subtree is itself an expression the
// left result is parked in a per-warp shared-memory temporary instead.
enum : uint8_t { OPK_NONE = 0, OPK_ACC = 1, OPK_COL = 2, OPK_LIT = 3, OPK_TMP = 4 };
enum : uint8_t { IOP_MOV = 0x50, IOP_ROWID = 0x51 };  // besides the igl_opcode values

constexpr int kMaxTmp = 4;      // shared-memory temporaries per warp
constexpr int kMaxInstr = 48;   // per fused operator (predicate + all outputs)
constexpr int kMaxExprs = 18;   // predicate + 16 outputs + row id
constexpr int kMaxCols = 16;    // staged input columns per operator

struct Instr {            // 32 bytes
    uint8_t op;           // igl_opcode, IOP_MOV or IOP_ROWID
    uint8_t type;         // operand type (igl_type) the operation is carried out in
    uint8_t a_kind, b_kind;
    uint8_t a_idx, b_idx;     // COL: staged slot; TMP: temporary index
    uint8_t dst_kind, dst_idx;  // OPK_ACC or OPK_TMP k
    uint8_t aux;          // CAST: target type; else AUX_*_NULL flags of literal operands
    uint8_t fast;         // 0 = generic step; else 1 + ((pair * 4 + class) * 3 + type): a fully specialised handler
    uint8_t pad[6];
    uint64_t a_lit, b_lit;  // literal bits when *_kind == OPK_LIT
};
static_assert(sizeof(Instr) == 32, "Instr layout");

// Fast handlers (filter_project.cu): operand pair x operator class x type, results in the accumulator.
enum : uint8_t { FP_PAIR_COL_LIT = 0, FP_PAIR_COL_COL = 1, FP_PAIR_ACC_LIT = 2, FP_PAIR_ACC_COL = 3 };
enum : uint8_t { FP_CLS_CMP = 0, FP_CLS_ADD = 1, FP_CLS_SUB = 2, FP_CLS_MUL = 3 };
enum : uint8_t { FP_TY_I32 = 0, FP_TY_I64 = 1, FP_TY_F64 = 2 };
constexpr uint8_t kFastMov32 = 49, kFastMov64 = 50;   // bare column pass-through, by width

// literal flags carried in `aux` for non-CAST ops
enum : uint8_t { AUX_A_NULL = 1, AUX_B_NULL = 2 };

}  // namespace igl
