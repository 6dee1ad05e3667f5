// registry.hpp -- table of pre-instantiated (static-mask) kernels, keyed by layout.
#pragma once

#include <cstring>

#include "pint/cuda_kernels.cuh"

namespace pint_cuda_impl {

using pint::cuda::kernels::launch_geom;
using pint::cuda::kernels::plane_ptrs;
using pint::cuda::layout_desc;

using binary_launch_t = cudaError_t (*)(const void* a, const void* b, void* out, size_t n, cudaStream_t, const launch_geom&);
using pair_launch_t = cudaError_t (*)(const void* a, const void* b, void* out0, void* out1, size_t n, cudaStream_t, const launch_geom&);
using scalar_launch_t = cudaError_t (*)(const void* a, uint64_t scalar, void* out, size_t n, cudaStream_t, const launch_geom&);
using pairs_launch_t = cudaError_t (*)(const void* pairs, void* out, size_t n, cudaStream_t, const launch_geom&);
using unpack_launch_t = cudaError_t (*)(const void* packed, const plane_ptrs&, size_t n, cudaStream_t, const launch_geom&);
using repack_launch_t = cudaError_t (*)(const plane_ptrs&, void* packed, size_t n, cudaStream_t, const launch_geom&);

struct layout_kernels {
    layout_desc desc;
    binary_launch_t binary[pint::cuda::OP_NUM_BINARY];
    binary_launch_t fused[pint::cuda::OP_NUM_BINARY];
    pair_launch_t pair[pint::cuda::kernels::kNumOpPairs];  // one-pass kernels for kOpPairs
    scalar_launch_t scalar[2][pint::cuda::OP_NUM_BINARY];  // [scalar_first][op]: OP(a[i], k) / OP(k, a[i])
    pairs_launch_t pairs[pint::cuda::OP_NUM_BINARY];       // interleaved operands -> results
    unpack_launch_t unpack[2][4];  // [is_signed][log2(plane_bytes)]
    repack_launch_t repack[4];     // [log2(plane_bytes)]
};

inline int plane_index(int plane_bytes) { return plane_bytes == 1 ? 0 : plane_bytes == 2 ? 1 : plane_bytes == 4 ? 2 : 3; }

inline bool same_layout(const layout_desc& a, const layout_desc& b)
{
    return a.elem_bytes == b.elem_bytes && a.nfields == b.nfields && std::memcmp(a.bits, b.bits, a.nfields) == 0;
}

// one per registry part (registry_part.cu)
const layout_kernels* registry_part0(int* count);
const layout_kernels* registry_part1(int* count);
const layout_kernels* registry_part2(int* count);
const layout_kernels* registry_part3(int* count);

// runtime-mask kernels (generic.cu)
cudaError_t generic_binary(int op, bool fused, const layout_desc& d, const void* a, const void* b, void* out,
                           size_t n, cudaStream_t s, const launch_geom& g);
cudaError_t generic_binary_pair(int pair, const layout_desc& d, const void* a, const void* b, void* out0, void* out1,
                                size_t n, cudaStream_t s, const launch_geom& g);
cudaError_t generic_unpack(const layout_desc& d, bool is_signed, int plane_bytes, const void* packed,
                           const plane_ptrs& planes, size_t n, cudaStream_t s, const launch_geom& g);
cudaError_t generic_repack(const layout_desc& d, int plane_bytes, const plane_ptrs& planes, void* packed,
                           size_t n, cudaStream_t s, const launch_geom& g);

}  // namespace pint_cuda_impl
