// fp_jit.h — run-time specialisation of the fused Filter / Projection kernel.
//
// The library's own fp_kernel interprets the operator's program (≈ 2 warp-instructions
// per row for `a < 500`, ≈ 5 per row for two projected expressions: the kernels are
// issue-bound, not HBM-bound).  For a plan that will see enough rows, fp_jit_build
// generates one translation unit that fixes the program, the staged-column layout and
// the output list as constants, compiles fp_device.cuh around it with NVRTC for sm_100a
// and loads the cubin; the interpreter folds into straight-line code for that operator.
// Same source, same arithmetic, same results — only the dispatch disappears.
//
// NVRTC is loaded with dlopen on first use; when it is not installed (or compilation
// fails) fp_jit_build returns a negative status and the caller keeps the interpreter.
#pragma once

#include <stdint.h>

#include <string>

#include "expr_isa.h"

namespace igl {

struct FpJitSpec {
    const Instr* code = nullptr;
    int n_instr = 0;
    const uint8_t* expr_first = nullptr;
    const uint8_t* expr_count = nullptr;
    int n_exprs = 0;
    int n_out = 0;
    const uint8_t* out_shift = nullptr;
    const uint32_t* smem_off = nullptr;
    const uint32_t* valid_off = nullptr;
    const uint8_t* shift = nullptr;
    int n_cols = 0;
    int rows_per_lane = 4;
    int mode = 0;          // FpMode
    bool has_nulls = false;
};

// The generated translation unit (exposed for tests and for reading).
std::string fp_jit_source(const FpJitSpec& spec);

// Compiles `spec` for sm_100a.  On success *cubin holds the image (needs no GPU).
int32_t fp_jit_compile(const FpJitSpec& spec, std::string* cubin, std::string* log);

// Compiles (or finds in the per-process cache) and loads the kernel on the current
// device.  *kernel is a cudaKernel_t usable with cudaLaunchKernel.
int32_t fp_jit_build(const FpJitSpec& spec, void** kernel);

}  // namespace igl
