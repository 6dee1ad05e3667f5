// api_callers.cu -- C ABI of the two extra caller shapes (kernels: pint/cuda_kernels_callers.cuh):
// pint_cuda_binary_scalar (one operand the same packed_int for every element) and
// pint_cuda_binary_pairs (interleaved std::pair operands, tests/pint_bench.cpp:22).  Layouts of
// registry_layouts.def run functors with compiled-in masks, every other layout the run-time-mask
// functors (masks in the constant bank); results are identical.
#include "api_common.hpp"

namespace pint_cuda_impl {
namespace {

using namespace pint::cuda;
using namespace pint::cuda::kernels;

template <class Elem, int Op, bool ScalarFirst>
cudaError_t scalar_one(const layout_desc& d, const void* a, uint64_t scalar, void* out, size_t n, cudaStream_t s,
                       const launch_geom& g)
{
    using W = word_of<Elem>;
    using Bin = swar_fn<Op, runtime_masks<W>>;
    const scalar_fn<Bin, ScalarFirst> fn{Bin{runtime_masks<W>(make_word_masks<W>(d))}, replicate_scalar<Elem>(scalar)};
    return launch_map1<Elem>(a, out, n, fn, s, g);
}

template <class Elem, int... Ops>
cudaError_t scalar_ops(int op, bool first, const layout_desc& d, const void* a, uint64_t scalar, void* out, size_t n,
                       cudaStream_t s, const launch_geom& g, std::integer_sequence<int, Ops...>)
{
    cudaError_t rc = cudaErrorInvalidValue;
    if (first)
        ((op == Ops ? (rc = scalar_one<Elem, Ops, true>(d, a, scalar, out, n, s, g), 0) : 0), ...);
    else
        ((op == Ops ? (rc = scalar_one<Elem, Ops, false>(d, a, scalar, out, n, s, g), 0) : 0), ...);
    return rc;
}

template <class Elem, int Op>
cudaError_t pairs_one(const layout_desc& d, const void* pairs, void* out, size_t n, cudaStream_t s, const launch_geom& g)
{
    using W = word_of<Elem>;
    using Fn = swar_fn<Op, runtime_masks<W>>;
    return launch_pairs_map<Elem>(pairs, out, n, Fn{runtime_masks<W>(make_word_masks<W>(d))}, s, g);
}

template <class Elem, int... Ops>
cudaError_t pairs_map_ops(int op, const layout_desc& d, const void* pairs, void* out, size_t n, cudaStream_t s,
                          const launch_geom& g, std::integer_sequence<int, Ops...>)
{
    cudaError_t rc = cudaErrorInvalidValue;
    ((op == Ops ? (rc = pairs_one<Elem, Ops>(d, pairs, out, n, s, g), 0) : 0), ...);
    return rc;
}

}  // namespace
}  // namespace pint_cuda_impl

using namespace pint_cuda_impl;

extern "C" {

int pint_cuda_binary_scalar(int op, const pint_layout* layout, const void* a, uint64_t scalar, int scalar_first,
                            void* out, size_t n, pint_stream_t stream)
{
    layout_desc d;
    if (int rc = check_layout(layout, &d)) return rc;
    if (op < 0 || op >= PINT_NUM_BINARY_OPS) return PINT_E_OP;
    // the scalar is a packed_int::value(): it has to be representable in the Integer (pint.hpp:768)
    if (d.elem_bytes < 8 && (scalar >> (8 * d.elem_bytes))) return PINT_E_ARG;
    if (n == 0) return PINT_OK;
    if (!a || !out) return PINT_E_NULL;
    const uintptr_t i = (uintptr_t)a, o = (uintptr_t)out;
    const size_t bytes = n * d.elem_bytes;
    if (i != o && i < o + bytes && o < i + bytes) return PINT_E_ALIAS;
    if ((i | o) % d.elem_bytes) return PINT_E_ARG;
    const launch_geom g = geom();
    cudaStream_t s = (cudaStream_t)stream;
    const bool first = scalar_first != 0;
    if (const layout_kernels* k = find_static(d)) return (int)k->scalar[first][op](a, scalar, out, n, s, g);
    auto ops = std::make_integer_sequence<int, pint::cuda::OP_NUM_BINARY>();
    switch (d.elem_bytes) {
    case 1: return (int)scalar_ops<uint8_t>(op, first, d, a, scalar, out, n, s, g, ops);
    case 2: return (int)scalar_ops<uint16_t>(op, first, d, a, scalar, out, n, s, g, ops);
    case 4: return (int)scalar_ops<uint32_t>(op, first, d, a, scalar, out, n, s, g, ops);
    default: return (int)scalar_ops<uint64_t>(op, first, d, a, scalar, out, n, s, g, ops);
    }
}

int pint_cuda_binary_pairs(int op, const pint_layout* layout, const void* pairs, void* out, size_t n,
                           pint_stream_t stream)
{
    layout_desc d;
    if (int rc = check_layout(layout, &d)) return rc;
    if (op < 0 || op >= PINT_NUM_BINARY_OPS) return PINT_E_OP;
    if (n == 0) return PINT_OK;
    if (!pairs || !out) return PINT_E_NULL;
    const uintptr_t i = (uintptr_t)pairs, o = (uintptr_t)out;
    const size_t bytes = n * d.elem_bytes;
    // results are half the size of the operands: writing them over the pairs would overtake the reads
    if (i < o + bytes && o < i + 2 * bytes) return PINT_E_ALIAS;
    if ((i | o) % d.elem_bytes) return PINT_E_ARG;
    const launch_geom g = geom();
    cudaStream_t s = (cudaStream_t)stream;
    if (const layout_kernels* k = find_static(d)) return (int)k->pairs[op](pairs, out, n, s, g);
    auto ops = std::make_integer_sequence<int, pint::cuda::OP_NUM_BINARY>();
    switch (d.elem_bytes) {
    case 1: return (int)pairs_map_ops<uint8_t>(op, d, pairs, out, n, s, g, ops);
    case 2: return (int)pairs_map_ops<uint16_t>(op, d, pairs, out, n, s, g, ops);
    case 4: return (int)pairs_map_ops<uint32_t>(op, d, pairs, out, n, s, g, ops);
    default: return (int)pairs_map_ops<uint64_t>(op, d, pairs, out, n, s, g, ops);
    }
}

}  // extern "C"
