// dense_registry.hpp -- dense bit-stream kernels with the layout compiled in (dense_static.cu).
#pragma once

#include "pint/cuda_kernels_dense.cuh"
#include "registry.hpp"

namespace pint_cuda_impl {

using pint::cuda::kernels::dense::dense_tuning;
using dense_binary_launch_t = cudaError_t (*)(const void* a, const void* b, void* out, size_t n, cudaStream_t,
                                              const launch_geom&, const dense_tuning&);

using dense_pair_launch_t = cudaError_t (*)(const void* a, const void* b, void* out0, void* out1, size_t n, cudaStream_t,
                                            const launch_geom&, const dense_tuning&);

struct dense_layout_kernels {
    uint32_t nfields;
    uint8_t bits[pint::cuda::kMaxFields];
    dense_binary_launch_t binary[pint::cuda::OP_NUM_BINARY];
    dense_pair_launch_t pair[pint::cuda::kernels::kNumOpPairs];  // one-pass kernels for kOpPairs
};

// the element side's Integer does not matter for a dense stream: lookup is by Bits... only
const dense_layout_kernels* find_dense_static(const layout_desc& d);

}  // namespace pint_cuda_impl
