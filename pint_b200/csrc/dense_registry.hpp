// dense_registry.hpp -- dense bit-stream kernels with the layout compiled in (dense_static.cu).
#pragma once

#include "pint/cuda_kernels_dense_convert.cuh"
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

// whole-chunk kernels with run-time masks, chunk length N = 3, 5, ..., 19 words (dense_generic_mw.cu,
// one translation unit per three values of N); false when the part does not cover N
#define PINT_MW_PART_DECL(P)                                                                                          \
    bool dense_mw_binary_part##P(int N, int op, const layout_desc& d, uint32_t S, const void* a, const void* b,       \
                                 void* out, size_t n, cudaStream_t s, const launch_geom& g, const dense_tuning& tune, \
                                 cudaError_t* rc);                                                                    \
    bool dense_mw_pair_part##P(int N, int pair, const layout_desc& d, uint32_t S, const void* a, const void* b,       \
                               void* out0, void* out1, size_t n, cudaStream_t s, const launch_geom& g,                \
                               const dense_tuning& tune, cudaError_t* rc);                                            \
    bool dense_mw_unary_part##P(int N, uint64_t elem_pre, uint32_t amount, bool left, uint32_t S, const void* in,     \
                                void* out, size_t n, cudaStream_t s, const launch_geom& g, const dense_tuning& tune,  \
                                cudaError_t* rc);
PINT_MW_PART_DECL(0)
PINT_MW_PART_DECL(1)
PINT_MW_PART_DECL(2)
#undef PINT_MW_PART_DECL

// conversions into a dense stream with S compiled in (dense_fixed.cu, part p covers S in [16 p, 16 p + 16));
// elem_bytes != 0: padded elements of that Integer, else planes of plane_bytes each.  false: no kernel for
// this (S, Integer).  Converts whole tiles only: *done elements, the caller finishes the rest.
// dense_extract_fixed_part: dense -> padded for uint8_t / uint16_t / uint64_t elements (elem_bytes != 0), dense -> planes
// of 1 or 2 bytes (elem_bytes == 0).
#define PINT_FIXED_PART_DECL(P)                                                                                       \
    bool dense_insert_fixed_part##P(uint32_t S, uint32_t elem_bytes, int plane_bytes, const void* other, void* dn,    \
                                    size_t n, const pint::cuda::kernels::dense::dense_fields& f, cudaStream_t s,      \
                                    const launch_geom& g, size_t* done, cudaError_t* rc);                             \
    bool dense_extract_fixed_part##P(uint32_t S, uint32_t elem_bytes, int plane_bytes, bool is_signed, const void* dn, \
                                     void* other, size_t n, const pint::cuda::kernels::dense::dense_fields& f,        \
                                     cudaStream_t s, const launch_geom& g, size_t* done, cudaError_t* rc);
PINT_FIXED_PART_DECL(0)
PINT_FIXED_PART_DECL(1)
PINT_FIXED_PART_DECL(2)
PINT_FIXED_PART_DECL(3)
#undef PINT_FIXED_PART_DECL

}  // namespace pint_cuda_impl
