// generic.cu -- runtime-mask kernels: any valid layout, masks computed on the
// host by the same generator (pint/cuda_layout.hpp) and passed to the kernel as
// a __grid_constant__ parameter (constant bank).  Same kernel templates as the
// static-mask instantiations, different mask policy.
#include "registry.hpp"

namespace pint_cuda_impl {
namespace {

using namespace pint::cuda;
using namespace pint::cuda::kernels;

template <class Elem, int Op>
cudaError_t binary_one(bool fused, const layout_desc& d, const void* a, const void* b, void* out, size_t n,
                       cudaStream_t s, const launch_geom& g)
{
    using W = word_of<Elem>;
    if (fused) {
        using Fn = fused_fn<Op, runtime_fields<W>>;
        Fn fn{{make_field_table<W>(d), make_word_masks<W>(d).A}};
        return launch_binary<Elem, Fn>(a, b, out, n, fn, s, g);
    }
    const auto m = make_word_masks<W>(d);
    // ops that build a field-wide mask run only as many smear steps as the widest field needs
    // (2 steps: fields up to 4 bits, 4 steps: up to 16 bits; profiles/: <7,9,3> add_signed_saturate)
    constexpr bool fills = !(Op == OP_ADD_WRAP || Op == OP_SUB_WRAP || Op == OP_OR || Op == OP_AND || Op == OP_XOR);
    if constexpr (fills) {
        if (m.nsh > 1 && m.nsmear <= 2) {
            using Fn = swar_fn<Op, runtime_masks<W, 2>>;
            return launch_binary<Elem, Fn>(a, b, out, n, Fn{runtime_masks<W, 2>(m)}, s, g);
        }
        if (m.nsh > 1 && m.nsmear <= 4) {
            using Fn = swar_fn<Op, runtime_masks<W, 4>>;
            return launch_binary<Elem, Fn>(a, b, out, n, Fn{runtime_masks<W, 4>(m)}, s, g);
        }
    }
    using Fn = swar_fn<Op, runtime_masks<W>>;
    Fn fn{runtime_masks<W>(m)};
    return launch_binary<Elem, Fn>(a, b, out, n, fn, s, g);
}

template <class Elem, int... Ops>
cudaError_t binary_ops(int op, bool fused, const layout_desc& d, const void* a, const void* b, void* out, size_t n,
                       cudaStream_t s, const launch_geom& g, std::integer_sequence<int, Ops...>)
{
    cudaError_t rc = cudaErrorInvalidValue;
    ((op == Ops ? (rc = binary_one<Elem, Ops>(fused, d, a, b, out, n, s, g), 0) : 0), ...);
    return rc;
}

template <class Elem, int... Pairs>
cudaError_t pair_ops(int pair, const layout_desc& d, const void* a, const void* b, void* out0, void* out1, size_t n,
                     cudaStream_t s, const launch_geom& g, std::integer_sequence<int, Pairs...>)
{
    using W = word_of<Elem>;
    const runtime_masks<W> p(make_word_masks<W>(d));
    cudaError_t rc = cudaErrorInvalidValue;
    ((pair == Pairs ? (rc = launch_binary_pair<Elem, kOpPairs[Pairs][0], kOpPairs[Pairs][1]>(a, b, out0, out1, n, p, s, g), 0) : 0), ...);
    return rc;
}

template <class Elem>
cudaError_t unpack_elem(const layout_desc& d, bool is_signed, int pb, const void* packed, const plane_ptrs& planes,
                        size_t n, cudaStream_t s, const launch_geom& g)
{
    using W = word_of<Elem>;
    runtime_fields<W> fp{make_field_table<W>(d), make_word_masks<W>(d).A};
    // the value of get / get_signed is narrowed or widened to the plane type as a C cast would
    if (is_signed) {
        switch (pb) {
        case 1: return launch_unpack<Elem, 1, true>(packed, planes, d.nfields, n, fp, s, g);
        case 2: return launch_unpack<Elem, 2, true>(packed, planes, d.nfields, n, fp, s, g);
        case 4: return launch_unpack<Elem, 4, true>(packed, planes, d.nfields, n, fp, s, g);
        default: return launch_unpack<Elem, 8, true>(packed, planes, d.nfields, n, fp, s, g);
        }
    }
    switch (pb) {
    case 1: return launch_unpack<Elem, 1, false>(packed, planes, d.nfields, n, fp, s, g);
    case 2: return launch_unpack<Elem, 2, false>(packed, planes, d.nfields, n, fp, s, g);
    case 4: return launch_unpack<Elem, 4, false>(packed, planes, d.nfields, n, fp, s, g);
    default: return launch_unpack<Elem, 8, false>(packed, planes, d.nfields, n, fp, s, g);
    }
}

template <class Elem>
cudaError_t repack_elem(const layout_desc& d, int pb, const plane_ptrs& planes, void* packed, size_t n,
                        cudaStream_t s, const launch_geom& g)
{
    using W = word_of<Elem>;
    runtime_fields<W> fp{make_field_table<W>(d), make_word_masks<W>(d).A};
    switch (pb) {
    case 1: return launch_repack<Elem, 1>(planes, d.nfields, packed, n, fp, s, g);
    case 2: return launch_repack<Elem, 2>(planes, d.nfields, packed, n, fp, s, g);
    case 4: return launch_repack<Elem, 4>(planes, d.nfields, packed, n, fp, s, g);
    default: return launch_repack<Elem, 8>(planes, d.nfields, packed, n, fp, s, g);
    }
}

}  // namespace

cudaError_t generic_binary(int op, bool fused, const layout_desc& d, const void* a, const void* b, void* out,
                           size_t n, cudaStream_t s, const launch_geom& g)
{
    auto ops = std::make_integer_sequence<int, OP_NUM_BINARY>();
    switch (d.elem_bytes) {
    case 1: return binary_ops<uint8_t>(op, fused, d, a, b, out, n, s, g, ops);
    case 2: return binary_ops<uint16_t>(op, fused, d, a, b, out, n, s, g, ops);
    case 4: return binary_ops<uint32_t>(op, fused, d, a, b, out, n, s, g, ops);
    default: return binary_ops<uint64_t>(op, fused, d, a, b, out, n, s, g, ops);
    }
}

cudaError_t generic_binary_pair(int pair, const layout_desc& d, const void* a, const void* b, void* out0, void* out1,
                                size_t n, cudaStream_t s, const launch_geom& g)
{
    auto pairs = std::make_integer_sequence<int, kNumOpPairs>();
    switch (d.elem_bytes) {
    case 1: return pair_ops<uint8_t>(pair, d, a, b, out0, out1, n, s, g, pairs);
    case 2: return pair_ops<uint16_t>(pair, d, a, b, out0, out1, n, s, g, pairs);
    case 4: return pair_ops<uint32_t>(pair, d, a, b, out0, out1, n, s, g, pairs);
    default: return pair_ops<uint64_t>(pair, d, a, b, out0, out1, n, s, g, pairs);
    }
}

cudaError_t generic_unpack(const layout_desc& d, bool is_signed, int plane_bytes, const void* packed,
                           const plane_ptrs& planes, size_t n, cudaStream_t s, const launch_geom& g)
{
    switch (d.elem_bytes) {
    case 1: return unpack_elem<uint8_t>(d, is_signed, plane_bytes, packed, planes, n, s, g);
    case 2: return unpack_elem<uint16_t>(d, is_signed, plane_bytes, packed, planes, n, s, g);
    case 4: return unpack_elem<uint32_t>(d, is_signed, plane_bytes, packed, planes, n, s, g);
    default: return unpack_elem<uint64_t>(d, is_signed, plane_bytes, packed, planes, n, s, g);
    }
}

cudaError_t generic_repack(const layout_desc& d, int plane_bytes, const plane_ptrs& planes, void* packed,
                           size_t n, cudaStream_t s, const launch_geom& g)
{
    switch (d.elem_bytes) {
    case 1: return repack_elem<uint8_t>(d, plane_bytes, planes, packed, n, s, g);
    case 2: return repack_elem<uint16_t>(d, plane_bytes, planes, packed, n, s, g);
    case 4: return repack_elem<uint32_t>(d, plane_bytes, planes, packed, n, s, g);
    default: return repack_elem<uint64_t>(d, plane_bytes, planes, packed, n, s, g);
    }
}

}  // namespace pint_cuda_impl
