// kernels.h — launchers shared between translation units.
#pragma once

#include <cuda_runtime.h>
#include <stdint.h>

#include "../../include/igloo_cuda.h"

namespace igl {

// byte-per-row {0,1} -> Arrow bitmap (LSB first).  The row count is read from
// *n_ptr on the device when n_ptr != NULL (a count produced by the same stream),
// else n_fixed.  `cap_rows` bounds the grid.
cudaError_t launch_pack_bits(const uint8_t* bytes, uint32_t* bits, const long long* n_ptr, long long n_fixed,
                             long long cap_rows, int sm_count, cudaStream_t st);

// igl_fp_create with two internal extras used by the streaming operators: `skip`
// hides input columns from a pass-through filter, `with_rowid` appends an Int32
// output holding the input row index of each surviving row.
int32_t fp_create_ex(igl_ctx* ctx, const int32_t* in_types, const uint8_t* in_nullable, int32_t n_in,
                     const igl_expr* pred, const igl_expr* const* proj, int32_t nproj, const uint8_t* skip,
                     bool with_rowid, igl_fp** out);

}  // namespace igl
