// registry_part.cu -- instantiates the static-mask kernels for the layouts of
// one part of registry_layouts.def.  Compile with -DPINT_REG_PART=<0..3>.
#include "registry.hpp"

#ifndef PINT_REG_PART
#error "compile with -DPINT_REG_PART=n"
#endif

namespace pint_cuda_impl {
namespace {

using namespace pint::cuda;
using namespace pint::cuda::kernels;

template <class SL, int Op>
cudaError_t binary_entry(const void* a, const void* b, void* out, size_t n, cudaStream_t s, const launch_geom& g)
{
    using Fn = swar_fn<Op, static_masks<SL>>;
    return launch_binary<typename SL::integer, Fn>(a, b, out, n, Fn{}, s, g);
}

template <class SL, int Op>
cudaError_t fused_entry(const void* a, const void* b, void* out, size_t n, cudaStream_t s, const launch_geom& g)
{
    using Fn = fused_fn<Op, static_fields<SL>>;
    return launch_binary<typename SL::integer, Fn>(a, b, out, n, Fn{}, s, g);
}

template <class SL, int Pair>
cudaError_t pair_entry(const void* a, const void* b, void* out0, void* out1, size_t n, cudaStream_t s, const launch_geom& g)
{
    return launch_binary_pair<typename SL::integer, kOpPairs[Pair][0], kOpPairs[Pair][1]>(a, b, out0, out1, n, static_masks<SL>{}, s, g);
}

// The functor of the extra caller shapes: at 2W bytes per element the ALU side weighs 1.5 x as much,
// so 64-bit layouts with several distinct widths take the per-field functor (C4: 7.0 against 6.6
// TB/s as a binary op), everything else the SWAR one.
template <class SL, int Op>
using caller_fn = typename std::conditional<sizeof(typename static_masks<SL>::word) == 8 && swar_fn<Op, static_masks<SL>>::heavy,
                                            fused_fn<Op, static_fields<SL>>, swar_fn<Op, static_masks<SL>>>::type;

template <class SL, int Op, bool First>
cudaError_t scalar_entry(const void* a, uint64_t scalar, void* out, size_t n, cudaStream_t s, const launch_geom& g)
{
    using E = typename SL::integer;
    const scalar_fn<caller_fn<SL, Op>, First> fn{caller_fn<SL, Op>{}, replicate_scalar<E>(scalar)};
    return launch_map1<E>(a, out, n, fn, s, g);
}

template <class SL, int Op>
cudaError_t pairs_entry(const void* pairs, void* out, size_t n, cudaStream_t s, const launch_geom& g)
{
    return launch_pairs_map<typename SL::integer>(pairs, out, n, caller_fn<SL, Op>{}, s, g);
}

template <class SL, int... Pairs>
void add_pairs(layout_kernels& k, std::integer_sequence<int, Pairs...>)
{
    ((k.pair[Pairs] = &pair_entry<SL, Pairs>), ...);
}

template <class SL, bool Signed, int PB>
cudaError_t unpack_entry(const void* packed, const plane_ptrs& planes, size_t n, cudaStream_t s, const launch_geom& g)
{
    return launch_unpack<typename SL::integer, PB, Signed>(packed, planes, SL::desc.nfields, n, static_fields<SL>{}, s, g);
}

template <class SL, int PB>
cudaError_t repack_entry(const plane_ptrs& planes, void* packed, size_t n, cudaStream_t s, const launch_geom& g)
{
    return launch_repack<typename SL::integer, PB>(planes, SL::desc.nfields, packed, n, static_fields<SL>{}, s, g);
}

template <class SL, int... Ops>
layout_kernels make_kernels(std::integer_sequence<int, Ops...>)
{
    layout_kernels k{};
    k.desc = SL::desc;
    ((k.binary[Ops] = &binary_entry<SL, Ops>), ...);
    ((k.fused[Ops] = &fused_entry<SL, Ops>), ...);
    add_pairs<SL>(k, std::make_integer_sequence<int, kNumOpPairs>());
    ((k.scalar[0][Ops] = &scalar_entry<SL, Ops, false>), ...);
    ((k.scalar[1][Ops] = &scalar_entry<SL, Ops, true>), ...);
    ((k.pairs[Ops] = &pairs_entry<SL, Ops>), ...);
    k.unpack[0][0] = &unpack_entry<SL, false, 1>;
    k.unpack[0][1] = &unpack_entry<SL, false, 2>;
    k.unpack[0][2] = &unpack_entry<SL, false, 4>;
    k.unpack[0][3] = &unpack_entry<SL, false, 8>;
    k.unpack[1][0] = &unpack_entry<SL, true, 1>;
    k.unpack[1][1] = &unpack_entry<SL, true, 2>;
    k.unpack[1][2] = &unpack_entry<SL, true, 4>;
    k.unpack[1][3] = &unpack_entry<SL, true, 8>;
    k.repack[0] = &repack_entry<SL, 1>;
    k.repack[1] = &repack_entry<SL, 2>;
    k.repack[2] = &repack_entry<SL, 4>;
    k.repack[3] = &repack_entry<SL, 8>;
    return k;
}

template <int Part, int P, class I, std::size_t... B>
void add(layout_kernels* tab, int& n)
{
    if constexpr (Part == P)
        tab[n++] = make_kernels<static_layout<I, B...>>(std::make_integer_sequence<int, OP_NUM_BINARY>());
}

}  // namespace

#define PINT_CAT2(a, b) a##b
#define PINT_CAT(a, b) PINT_CAT2(a, b)

const layout_kernels* PINT_CAT(registry_part, PINT_REG_PART)(int* count)
{
    static layout_kernels tab[16];
    static int n = [] {
        int k = 0;
#define X(part, I, ...) add<PINT_REG_PART, part, I, __VA_ARGS__>(tab, k);
#include "registry_layouts.def"
#undef X
        return k;
    }();
    *count = n;
    return tab;
}

}  // namespace pint_cuda_impl
