// dense_static.cu -- instantiates the dense bit-stream kernels for the layouts of dense_layouts.def.
#include "dense_registry.hpp"

namespace pint_cuda_impl {
namespace {

using namespace pint::cuda;
using namespace pint::cuda::kernels;

// CL = dense::chunk_layout<Bits...>: layouts whose chunk fits in registers compute whole chunks
// (cuda_kernels_dense_mw.cuh); the others slide windows over the chunk element by element.
template <class SL, class CL, int Op>
cudaError_t dense_entry(const void* a, const void* b, void* out, size_t n, cudaStream_t s, const launch_geom& g,
                        const dense_tuning& tune)
{
    if constexpr (CL::kN <= (int)dense::kMaxMwWords) {
        using Fn = swar_fn<Op, static_masks<CL>>;
        using Geo = dense::mw_geom<CL::kN>;
        return dense::launch_dense_binary(a, b, out, n, Fn{}, Geo{CL::kS, 1}, s, g, tune);
    } else {
        using W = typename SL::word;
        using Fn = swar_fn<Op, static_masks<SL>>;
        using Geo = dense::static_geom<W, SL::masks.total_bits, 8 * sizeof(typename SL::integer)>;
        return dense::launch_dense_binary(a, b, out, n, Fn{}, Geo{1}, s, g, tune);
    }
}

template <class SL, class CL, int Pair>
cudaError_t dense_pair_entry(const void* a, const void* b, void* out0, void* out1, size_t n, cudaStream_t s,
                             const launch_geom& g, const dense_tuning& tune)
{
    if constexpr (CL::kN <= (int)dense::kMaxMwWords) {
        using Fn = pair_fn<kOpPairs[Pair][0], kOpPairs[Pair][1], static_masks<CL>>;
        using Geo = dense::mw_geom<CL::kN>;
        return dense::launch_dense_binary<Fn, Geo, true>(a, b, out0, n, Fn{}, Geo{CL::kS, 1}, s, g, tune, out1);
    } else {
        return cudaErrorNotSupported;  // never installed (add_dense_pairs): window kernels run pairs as two launches
    }
}

template <class SL, class CL, int... Pairs>
void add_dense_pairs(dense_layout_kernels& k, std::integer_sequence<int, Pairs...>)
{
    if (CL::kN <= (int)dense::kMaxMwWords) ((k.pair[Pairs] = &dense_pair_entry<SL, CL, Pairs>), ...);
}

template <class SL, class CL, int... Ops>
dense_layout_kernels make_dense(std::integer_sequence<int, Ops...>)
{
    dense_layout_kernels k{};
    k.nfields = SL::desc.nfields;
    for (uint32_t i = 0; i < k.nfields; ++i) k.bits[i] = SL::desc.bits[i];
    ((k.binary[Ops] = &dense_entry<SL, CL, Ops>), ...);
    add_dense_pairs<SL, CL>(k, std::make_integer_sequence<int, kNumOpPairs>());
    return k;
}

}  // namespace

const dense_layout_kernels* find_dense_static(const layout_desc& d)
{
    static const dense_layout_kernels tab[] = {
#define X(W, ...) make_dense<static_layout<W, __VA_ARGS__>, dense::chunk_layout<__VA_ARGS__>>(std::make_integer_sequence<int, OP_NUM_BINARY>()),
#include "dense_layouts.def"
#undef X
    };
    for (const dense_layout_kernels& k : tab)
        if (k.nfields == d.nfields && std::memcmp(k.bits, d.bits, d.nfields) == 0) return &k;
    return nullptr;
}

}  // namespace pint_cuda_impl
