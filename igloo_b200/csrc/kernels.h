// kernels.h — launchers shared between translation units.
#pragma once

#include <cuda_runtime.h>
#include <stdint.h>

namespace igl {

// byte-per-row {0,1} -> Arrow bitmap (LSB first).  The row count is read from
// *n_ptr on the device when n_ptr != NULL (a count produced by the same stream),
// else n_fixed.  `cap_rows` bounds the grid.
cudaError_t launch_pack_bits(const uint8_t* bytes, uint32_t* bits, const long long* n_ptr, long long n_fixed,
                             long long cap_rows, int sm_count, cudaStream_t st);

}  // namespace igl
