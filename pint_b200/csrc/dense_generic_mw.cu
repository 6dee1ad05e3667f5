// dense_generic_mw.cu -- whole-chunk dense kernels with RUN-TIME masks (cuda_kernels_dense_mw.cuh),
// instantiated per chunk length N (32-bit words).  Compile with -DPINT_MW_PART=<0..2>; each part
// covers three values of N.
#include "dense_registry.hpp"

#ifndef PINT_MW_PART
#error "compile with -DPINT_MW_PART=n"
#endif

namespace pint_cuda_impl {
namespace {

using namespace pint::cuda;
using namespace pint::cuda::kernels;

template <int N, int... Ops>
cudaError_t binary_n(int op, const layout_desc& d, uint32_t S, const void* a, const void* b, void* out, size_t n,
                     cudaStream_t s, const launch_geom& g, const dense_tuning& tune, std::integer_sequence<int, Ops...>)
{
    const auto m = dense::make_chunk_masks<N>(d);
    cudaError_t rc = cudaErrorInvalidValue;
    auto run = [&](auto steps_tag, auto op_tag) {
        constexpr int St = decltype(steps_tag)::value, Op = decltype(op_tag)::value;
        using P = dense::runtime_chunk_masks<N, St>;
        rc = dense::launch_dense_binary(a, b, out, n, swar_fn<Op, P>{P(m)}, dense::mw_geom<N>{S, 1}, s, g, tune);
    };
    // ops that build a field-wide mask run only as many smear steps as the widest field needs
    const int steps = m.nsh > 1 && m.nsmear <= 2 ? 2 : (m.nsh > 1 && m.nsmear <= 4 ? 4 : 6);
    auto pick = [&](auto op_tag) {
        constexpr int Op = decltype(op_tag)::value;
        constexpr bool fills = !(Op == OP_ADD_WRAP || Op == OP_SUB_WRAP || Op == OP_OR || Op == OP_AND || Op == OP_XOR);
        if constexpr (fills) {
            if (steps == 2) return run(std::integral_constant<int, 2>{}, op_tag);
            if (steps == 4) return run(std::integral_constant<int, 4>{}, op_tag);
        }
        run(std::integral_constant<int, 6>{}, op_tag);
    };
    ((op == Ops ? (pick(std::integral_constant<int, Ops>{}), 0) : 0), ...);
    return rc;
}

template <int N, int... Pairs>
cudaError_t pair_n(int pair, const layout_desc& d, uint32_t S, const void* a, const void* b, void* out0, void* out1,
                   size_t n, cudaStream_t s, const launch_geom& g, const dense_tuning& tune,
                   std::integer_sequence<int, Pairs...>)
{
    const auto m = dense::make_chunk_masks<N>(d);
    cudaError_t rc = cudaErrorInvalidValue;
    auto run = [&](auto steps_tag, auto pair_tag) {
        constexpr int St = decltype(steps_tag)::value, Pi = decltype(pair_tag)::value;
        using P = dense::runtime_chunk_masks<N, St>;
        using Fn = pair_fn<kOpPairs[Pi][0], kOpPairs[Pi][1], P>;
        rc = dense::launch_dense_binary<Fn, dense::mw_geom<N>, true>(a, b, out0, n, Fn{P(m)}, dense::mw_geom<N>{S, 1}, s, g, tune, out1);
    };
    const int steps = m.nsh > 1 && m.nsmear <= 2 ? 2 : (m.nsh > 1 && m.nsmear <= 4 ? 4 : 6);
    auto pick = [&](auto pair_tag) {
        constexpr int Pi = decltype(pair_tag)::value;
        constexpr bool fills = !(kOpPairs[Pi][0] == OP_ADD_WRAP && kOpPairs[Pi][1] == OP_SUB_WRAP);
        if constexpr (fills) {
            if (steps == 2) return run(std::integral_constant<int, 2>{}, pair_tag);
            if (steps == 4) return run(std::integral_constant<int, 4>{}, pair_tag);
        }
        run(std::integral_constant<int, 6>{}, pair_tag);
    };
    ((pair == Pairs ? (pick(std::integral_constant<int, Pairs>{}), 0) : 0), ...);
    return rc;
}

template <int N>
cudaError_t unary_n(uint64_t elem_pre, uint32_t amount, bool left, uint32_t S, const void* in, void* out, size_t n,
                    cudaStream_t s, const launch_geom& g, const dense_tuning& tune)
{
    dense::mw_unary_fn<N> fn{dense::replicate_element_mask<N>(elem_pre, S), amount, left ? 1u : 0u};
    return dense::launch_dense_binary<dense::mw_unary_fn<N>, dense::mw_geom<N>, 2>(in, nullptr, out, n, fn, dense::mw_geom<N>{S, 1},
                                                                                    s, g, tune);
}

constexpr int kFirst = 3 + 6 * PINT_MW_PART;  // N = kFirst, kFirst + 2, kFirst + 4

}  // namespace

#define PINT_CAT2(a, b) a##b
#define PINT_CAT(a, b) PINT_CAT2(a, b)

// false when this part does not cover N
bool PINT_CAT(dense_mw_binary_part, PINT_MW_PART)(int N, int op, const layout_desc& d, uint32_t S, const void* a,
                                                   const void* b, void* out, size_t n, cudaStream_t s,
                                                   const launch_geom& g, const dense_tuning& tune, cudaError_t* rc)
{
    auto ops = std::make_integer_sequence<int, OP_NUM_BINARY>();
    if (N == kFirst) *rc = binary_n<kFirst>(op, d, S, a, b, out, n, s, g, tune, ops);
    else if (N == kFirst + 2) *rc = binary_n<kFirst + 2>(op, d, S, a, b, out, n, s, g, tune, ops);
    else if (N == kFirst + 4) *rc = binary_n<kFirst + 4>(op, d, S, a, b, out, n, s, g, tune, ops);
    else return false;
    return true;
}

bool PINT_CAT(dense_mw_pair_part, PINT_MW_PART)(int N, int pair, const layout_desc& d, uint32_t S, const void* a,
                                                 const void* b, void* out0, void* out1, size_t n, cudaStream_t s,
                                                 const launch_geom& g, const dense_tuning& tune, cudaError_t* rc)
{
    auto pairs = std::make_integer_sequence<int, kNumOpPairs>();
    if (N == kFirst) *rc = pair_n<kFirst>(pair, d, S, a, b, out0, out1, n, s, g, tune, pairs);
    else if (N == kFirst + 2) *rc = pair_n<kFirst + 2>(pair, d, S, a, b, out0, out1, n, s, g, tune, pairs);
    else if (N == kFirst + 4) *rc = pair_n<kFirst + 4>(pair, d, S, a, b, out0, out1, n, s, g, tune, pairs);
    else return false;
    return true;
}

bool PINT_CAT(dense_mw_unary_part, PINT_MW_PART)(int N, uint64_t elem_pre, uint32_t amount, bool left, uint32_t S,
                                                  const void* in, void* out, size_t n, cudaStream_t s,
                                                  const launch_geom& g, const dense_tuning& tune, cudaError_t* rc)
{
    if (N == kFirst) *rc = unary_n<kFirst>(elem_pre, amount, left, S, in, out, n, s, g, tune);
    else if (N == kFirst + 2) *rc = unary_n<kFirst + 2>(elem_pre, amount, left, S, in, out, n, s, g, tune);
    else if (N == kFirst + 4) *rc = unary_n<kFirst + 4>(elem_pre, amount, left, S, in, out, n, s, g, tune);
    else return false;
    return true;
}

}  // namespace pint_cuda_impl
