// expr_ir.h — lowering of compiled expressions to the interpreter's IR.
#pragma once

#include <memory>
#include <vector>

#include "internal.h"

namespace igl {

struct LoweredExpr {
    int first = 0;       // index of the first instruction in the fused program
    int count = 0;       // number of instructions
    int n_tmp = 0;       // shared-memory temporaries the expression needs
    int type = 0;
    bool nullable = false;
};

// Appends the instructions of `e` to `code`.
int32_t lower_expr(const igl_expr& e, std::vector<Instr>& code, LoweredExpr& out);
// Marks the input columns `e` reads.
void referenced_columns(const igl_expr& e, std::vector<bool>& used);

}  // namespace igl
