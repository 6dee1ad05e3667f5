// pint/cuda.hpp -- the pint::cuda batch layer: host C++ templates over device arrays.
//
// The scalar API of pint (/root/reference/include/pint/pint.hpp, "pint.hpp:N"
// below) is unchanged; this header adds, beside it, the same operations applied
// elementwise to arrays of packed_int<Integer, Bits...> in GPU memory:
//
//     using rgba = pint::packed_int<uint32_t, 8, 8, 8, 8>;
//     rgba *a, *b, *out;                       // device pointers, n elements
//     pint::cuda::add_signed_saturate(a, b, out, n, stream);
//     // out[i] == pint::add_signed_saturate(a[i], b[i])      (pint.hpp:857-866)
//
// Names, argument order and per-element meaning are the scalar API's.  Each
// template turns the Bits... pack into a pint_layout descriptor at compile time
// and calls the C ABI of libpint_cuda.so (include/pint_cuda.h), where
// hand-written sm_100a kernels do the work; layouts listed in the library's
// registry run kernels with the field masks compiled in, all others run the
// same kernels with the masks as kernel parameters.  No CPU fallback.
//
// C++11, header-only, needs neither nvcc nor pint.hpp: packed_int is only
// forward-declared (exactly as pint.hpp:25 declares it), pointers to it are
// passed through untouched.  Include pint.hpp as well to use the scalar API.
//
// Errors: the conditions pint rejects at compile time are rejected here at
// compile time too (static_asserts below mirror pint.hpp:761-764, 748, 802,
// 813); runtime failures (CUDA errors, null or overlapping arrays) throw
// pint::cuda::error carrying the ABI's return code.  Every call is asynchronous
// on `stream`.
#pragma once

#include <cstddef>
#include <cstdint>
#include <stdexcept>
#include <string>
#include <type_traits>
#include <utility>

#include "../pint_cuda.h"

// Compiled by nvcc (C++17): the packed -> packed operations, unpack and repack instantiate the kernel
// templates for static_layout<Integer, Bits...> in the CALLER's translation unit, so that EVERY
// layout gets its mask constants at compile time exactly as in the reference (pint.hpp:324-337,
// 443-456), not only the ones built into libpint_cuda.so.  Same kernels, same results; heads,
// tails and misaligned operands are handled by the same launchers the library uses.  Define
// PINT_CUDA_NO_INLINE_KERNELS to keep nvcc translation units on the C ABI like any other compiler.
#if defined(__CUDACC__) && !defined(PINT_CUDA_NO_INLINE_KERNELS) && __cplusplus >= 201703L
#define PINT_CUDA_INLINE_KERNELS 1
#include "cuda_kernels.cuh"
#else
#define PINT_CUDA_INLINE_KERNELS 0
#endif

namespace pint {

using std::size_t;
template <class Integer, size_t Bits0, size_t... Bits> class packed_int;  // pint.hpp:25

namespace cuda {

typedef pint_stream_t stream_t;  // a cudaStream_t

class error : public std::runtime_error {
public:
    error(int rc, const char* what_call)
        : std::runtime_error(std::string(what_call) + ": " + pint_cuda_error_string(rc)), code(rc) {}
    int code;
};

inline void check(int rc, const char* what_call)
{
    if (rc != PINT_OK) throw error(rc, what_call);
}

namespace detail {

template <size_t... Bits> struct sum_bits;
template <> struct sum_bits<> { static const size_t value = 0; };
template <size_t B0, size_t... Bits> struct sum_bits<B0, Bits...> { static const size_t value = B0 + sum_bits<Bits...>::value; };

template <size_t... Bits> struct all_positive;
template <> struct all_positive<> { static const bool value = true; };
template <size_t B0, size_t... Bits> struct all_positive<B0, Bits...> { static const bool value = (B0 > 0) && all_positive<Bits...>::value; };

// sum of the first N values of the pack (slice offsets, pint.hpp:927-928)
template <size_t N, size_t... Bits> struct sum_front;
template <size_t... Bits> struct sum_front<0, Bits...> { static const size_t value = 0; };
template <size_t N, size_t B0, size_t... Bits> struct sum_front<N, B0, Bits...> { static const size_t value = B0 + sum_front<N - 1, Bits...>::value; };

// compile-time checks of packed_int (pint.hpp:761-764) + descriptor
template <class Integer, size_t... Bits>
struct layout_of {
    static_assert(std::is_integral<Integer>::value && std::is_unsigned<Integer>::value,
                  "Integer must be unsigned integer");
    static_assert(sizeof(Integer) == 1 || sizeof(Integer) == 2 || sizeof(Integer) == 4 || sizeof(Integer) == 8,
                  "Integer must be uint8_t, uint16_t, uint32_t or uint64_t");
    static_assert(sizeof(Integer) * 8 >= sum_bits<Bits...>::value, "Integer won't fit given number of bits");
    static_assert(all_positive<Bits...>::value, "fields must have at least 1 bit");
    static_assert(sizeof...(Bits) >= 1 && sizeof...(Bits) <= PINT_MAX_FIELDS, "1..64 fields");

    static pint_layout get()
    {
        pint_layout l = {(uint32_t)sizeof(Integer), (uint32_t)sizeof...(Bits), {(uint8_t)Bits...}};
        return l;
    }
};

// slice<Start, End> result type (sliced_int, pint.hpp:746-754), without pint.hpp's type lists:
// skip `Skip` widths, take `Take` widths, wrap them in packed_int.
template <size_t... Bits> struct bits_pack {};
template <int State, size_t Skip, size_t Take, class Integer, class Taken, size_t... Rest> struct slice_step;
template <size_t Skip, size_t Take, class Integer, class Taken, size_t... Rest>
struct slice_next {
    typedef typename slice_step<(Skip > 0 ? 0 : (Take > 0 ? 1 : 2)), Skip, Take, Integer, Taken, Rest...>::type type;
};
template <size_t Skip, size_t Take, class Integer, size_t B0, size_t... Rest>
struct slice_step<0, Skip, Take, Integer, bits_pack<>, B0, Rest...> {
    typedef typename slice_next<Skip - 1, Take, Integer, bits_pack<>, Rest...>::type type;
};
template <size_t Take, class Integer, size_t... Taken, size_t B0, size_t... Rest>
struct slice_step<1, 0, Take, Integer, bits_pack<Taken...>, B0, Rest...> {
    typedef typename slice_next<0, Take - 1, Integer, bits_pack<Taken..., B0>, Rest...>::type type;
};
template <class Integer, size_t T0, size_t... Taken, size_t... Rest>
struct slice_step<2, 0, 0, Integer, bits_pack<T0, Taken...>, Rest...> {
    typedef packed_int<Integer, T0, Taken...> type;
};

}  // namespace detail

// packed_int<Integer, Bits...> as the ABI's descriptor
namespace detail {
// bytes of one array element: a packed_int is its Integer (pint.hpp:25; the class may be incomplete here)
template <class T> struct elem_size { static const size_t value = sizeof(T); };
template <class Integer, size_t Bits0, size_t... Bits>
struct elem_size<packed_int<Integer, Bits0, Bits...> > { static const size_t value = sizeof(Integer); };
}  // namespace detail

template <class Integer, size_t Bits0, size_t... Bits>
inline pint_layout layout(const packed_int<Integer, Bits0, Bits...>* = nullptr)
{
    return detail::layout_of<Integer, Bits0, Bits...>::get();
}

template <size_t Start, size_t End, class Integer, size_t Bits0, size_t... Bits>
struct sliced_int {
    static_assert(Start < End && End <= sizeof...(Bits) + 1, "Incorrect slice bounds");  // pint.hpp:748
    typedef typename detail::slice_next<Start, End - Start, Integer, detail::bits_pack<>, Bits0, Bits...>::type type;
};

#if PINT_CUDA_INLINE_KERNELS
namespace detail {

// the argument checks of the C ABI (pint_b200/csrc/api.cu binary_impl), for launches made from this header
inline int check_binary_arrays(const void* a, const void* b, const void* out, size_t n, size_t elem_bytes)
{
    if (n == 0) return PINT_OK;
    if (!a || !b || !out) return PINT_E_NULL;
    const uintptr_t o = (uintptr_t)out, bytes = n * elem_bytes;
    const uintptr_t in[2] = {(uintptr_t)a, (uintptr_t)b};
    for (int i = 0; i < 2; ++i)
        if (in[i] != o && in[i] < o + bytes && o < in[i] + bytes) return PINT_E_ALIAS;
    if (((uintptr_t)a | (uintptr_t)b | o) % elem_bytes) return PINT_E_ARG;
    return PINT_OK;
}

template <int Op, bool Fused, class Integer, size_t... Bits>
inline int inline_binary(const void* a, const void* b, void* out, size_t n, stream_t stream)
{
    layout_of<Integer, Bits...>::get();  // the reference's static_asserts
    if (int rc = check_binary_arrays(a, b, out, n, sizeof(Integer))) return rc;
    using SL = static_layout<Integer, Bits...>;
    const kernels::launch_geom g;  // one CTA per tile
    if (Fused) {
        using Fn = kernels::fused_fn<Op, kernels::static_fields<SL>>;
        return (int)kernels::launch_binary<Integer, Fn>(a, b, out, n, Fn{}, (cudaStream_t)stream, g);
    }
    using Fn = kernels::swar_fn<Op, kernels::static_masks<SL>>;
    return (int)kernels::launch_binary<Integer, Fn>(a, b, out, n, Fn{}, (cudaStream_t)stream, g);
}

}  // namespace detail
#endif

// ---- the 13 binary operations: out[i] = OP(a[i], b[i]); out may be a or b ----
#if PINT_CUDA_INLINE_KERNELS
#define PINT_CUDA_BINARY(NAME, CODE)                                                                    \
    template <class Integer, size_t Bits0, size_t... Bits>                                              \
    inline void NAME(const packed_int<Integer, Bits0, Bits...>* a, const packed_int<Integer, Bits0, Bits...>* b, \
                     packed_int<Integer, Bits0, Bits...>* out, size_t n, stream_t stream = nullptr)     \
    {                                                                                                   \
        check(detail::inline_binary<CODE, false, Integer, Bits0, Bits...>(a, b, out, n, stream), "pint::cuda::" #NAME); \
    }
#else
#define PINT_CUDA_BINARY(NAME, CODE)                                                                    \
    template <class Integer, size_t Bits0, size_t... Bits>                                              \
    inline void NAME(const packed_int<Integer, Bits0, Bits...>* a, const packed_int<Integer, Bits0, Bits...>* b, \
                     packed_int<Integer, Bits0, Bits...>* out, size_t n, stream_t stream = nullptr)     \
    {                                                                                                   \
        const pint_layout l = detail::layout_of<Integer, Bits0, Bits...>::get();                        \
        check(pint_cuda_binary(CODE, &l, a, b, out, n, stream), "pint::cuda::" #NAME);                  \
    }
#endif

PINT_CUDA_BINARY(add_wrap, PINT_ADD_WRAP)                            // pint.hpp:826-838
PINT_CUDA_BINARY(add_unsigned_saturate, PINT_ADD_UNSIGNED_SATURATE)  // pint.hpp:840-855
PINT_CUDA_BINARY(add_signed_saturate, PINT_ADD_SIGNED_SATURATE)      // pint.hpp:857-866
PINT_CUDA_BINARY(sub_wrap, PINT_SUB_WRAP)                            // pint.hpp:870-884
PINT_CUDA_BINARY(sub_unsigned_saturate, PINT_SUB_UNSIGNED_SATURATE)  // pint.hpp:886-908
PINT_CUDA_BINARY(sub_signed_saturate, PINT_SUB_SIGNED_SATURATE)      // pint.hpp:910-919
PINT_CUDA_BINARY(bit_or, PINT_OR)                                    // pint.hpp:781
PINT_CUDA_BINARY(bit_and, PINT_AND)                                  // pint.hpp:782
PINT_CUDA_BINARY(bit_xor, PINT_XOR)                                  // pint.hpp:783
PINT_CUDA_BINARY(min_unsigned, PINT_MIN_UNSIGNED)                    // pint.hpp:936-950
PINT_CUDA_BINARY(max_unsigned, PINT_MAX_UNSIGNED)                    // pint.hpp:952-966
PINT_CUDA_BINARY(min_signed, PINT_MIN_SIGNED)                        // pint.hpp:968-985
PINT_CUDA_BINARY(max_signed, PINT_MAX_SIGNED)                        // pint.hpp:987-1004
#undef PINT_CUDA_BINARY

// Two (up to PINT_MAX_MULTI_OPS) operations on the SAME operands: pairs such as
// add_unsigned_saturate + add_signed_saturate run as ONE pass that reads a and b once
// (pint_cuda_binary_multi); outs[j][i] = ops[j](a[i], b[i]).  Outputs may not alias inputs.
template <size_t NOps, class Integer, size_t Bits0, size_t... Bits>
inline void binary_multi(const pint_op (&ops)[NOps], const packed_int<Integer, Bits0, Bits...>* a,
                         const packed_int<Integer, Bits0, Bits...>* b,
                         packed_int<Integer, Bits0, Bits...>* const (&outs)[NOps], size_t n, stream_t stream = nullptr)
{
    static_assert(NOps >= 1 && NOps <= PINT_MAX_MULTI_OPS, "1..PINT_MAX_MULTI_OPS operations");
    const pint_layout l = detail::layout_of<Integer, Bits0, Bits...>::get();
    int codes[NOps];
    void* o[NOps];
    for (size_t j = 0; j < NOps; ++j) { codes[j] = ops[j]; o[j] = outs[j]; }
    check(pint_cuda_binary_multi((int)NOps, codes, &l, a, b, o, n, stream), "pint::cuda::binary_multi");
}

// One operand the same packed_int for every element: out[i] = OP(a[i], k) (binary_scalar) or
// OP(k, a[i]) (scalar_binary); 2 * sizeof(Integer) bytes per element.  out may be a.  k is given
// as a packed_int (needs pint.hpp) or as its value() (pint.hpp:776).
template <class Integer, size_t Bits0, size_t... Bits>
inline void binary_scalar(pint_op op, const packed_int<Integer, Bits0, Bits...>* a, Integer k_value,
                          packed_int<Integer, Bits0, Bits...>* out, size_t n, stream_t stream = nullptr)
{
    const pint_layout l = detail::layout_of<Integer, Bits0, Bits...>::get();
    check(pint_cuda_binary_scalar(op, &l, a, k_value, 0, out, n, stream), "pint::cuda::binary_scalar");
}
template <class Integer, size_t Bits0, size_t... Bits>
inline void scalar_binary(pint_op op, Integer k_value, const packed_int<Integer, Bits0, Bits...>* a,
                          packed_int<Integer, Bits0, Bits...>* out, size_t n, stream_t stream = nullptr)
{
    const pint_layout l = detail::layout_of<Integer, Bits0, Bits...>::get();
    check(pint_cuda_binary_scalar(op, &l, a, k_value, 1, out, n, stream), "pint::cuda::scalar_binary");
}
template <class Integer, size_t Bits0, size_t... Bits>
inline void binary_scalar(pint_op op, const packed_int<Integer, Bits0, Bits...>* a, packed_int<Integer, Bits0, Bits...> k,
                          packed_int<Integer, Bits0, Bits...>* out, size_t n, stream_t stream = nullptr)
{
    binary_scalar(op, a, k.value(), out, n, stream);
}
template <class Integer, size_t Bits0, size_t... Bits>
inline void scalar_binary(pint_op op, packed_int<Integer, Bits0, Bits...> k, const packed_int<Integer, Bits0, Bits...>* a,
                          packed_int<Integer, Bits0, Bits...>* out, size_t n, stream_t stream = nullptr)
{
    scalar_binary(op, k.value(), a, out, n, stream);
}

// Interleaved operands, as the reference benchmark stores them (std::pair<Integer, Integer>,
// tests/pint_bench.cpp:22): out[i] = OP(packed_int(pairs[i].first), packed_int(pairs[i].second)).
template <class Integer, size_t Bits0, size_t... Bits>
inline void binary_pairs(pint_op op, const std::pair<Integer, Integer>* pairs, packed_int<Integer, Bits0, Bits...>* out,
                         size_t n, stream_t stream = nullptr)
{
    static_assert(sizeof(std::pair<Integer, Integer>) == 2 * sizeof(Integer), "pairs are two Integers back to back");
    const pint_layout l = detail::layout_of<Integer, Bits0, Bits...>::get();
    check(pint_cuda_binary_pairs(op, &l, pairs, out, n, stream), "pint::cuda::binary_pairs");
}

// packed_int::operator== over arrays (pint.hpp:778-779): equal_out[i] = (a[i] == b[i]) as bytes and / or
// the number of unequal elements in *not_equal_count (device uint64_t); either may be null.
template <class Integer, size_t Bits0, size_t... Bits>
inline void equal(const packed_int<Integer, Bits0, Bits...>* a, const packed_int<Integer, Bits0, Bits...>* b, size_t n,
                  uint8_t* equal_out, uint64_t* not_equal_count, stream_t stream = nullptr)
{
    const pint_layout l = detail::layout_of<Integer, Bits0, Bits...>::get();
    check(pint_cuda_equal(&l, a, b, n, equal_out, not_equal_count, stream), "pint::cuda::equal");
}

#if PINT_CUDA_INLINE_KERNELS
// The fused pipeline with the operation as a template argument, instantiated for this layout in the
// caller's translation unit (nvcc only; the run-time `op` form below goes through the library).
template <pint_op Op, class Integer, size_t Bits0, size_t... Bits>
inline void fused_unpack_arith_repack(const packed_int<Integer, Bits0, Bits...>* a, const packed_int<Integer, Bits0, Bits...>* b,
                                      packed_int<Integer, Bits0, Bits...>* out, size_t n, stream_t stream = nullptr)
{
    check(detail::inline_binary<(int)Op, true, Integer, Bits0, Bits...>(a, b, out, n, stream), "pint::cuda::fused_unpack_arith_repack");
}
#endif

// Any of the above, computed as unpack -> per-field arithmetic -> repack in one kernel.
template <class Integer, size_t Bits0, size_t... Bits>
inline void fused_unpack_arith_repack(pint_op op, const packed_int<Integer, Bits0, Bits...>* a,
                                      const packed_int<Integer, Bits0, Bits...>* b,
                                      packed_int<Integer, Bits0, Bits...>* out, size_t n, stream_t stream = nullptr)
{
    const pint_layout l = detail::layout_of<Integer, Bits0, Bits...>::get();
    check(pint_cuda_fused_unpack_arith_repack(op, &l, a, b, out, n, stream), "pint::cuda::fused_unpack_arith_repack");
}

// shift_left / shift_right_unsigned with a uniform amount (pint.hpp:1006-1029)
template <class Integer, size_t Bits0, size_t... Bits>
inline void shift_left(const packed_int<Integer, Bits0, Bits...>* in, size_t shift_amount,
                       packed_int<Integer, Bits0, Bits...>* out, size_t n, stream_t stream = nullptr)
{
    const pint_layout l = detail::layout_of<Integer, Bits0, Bits...>::get();
    check(pint_cuda_shift(1, &l, shift_amount, in, out, n, stream), "pint::cuda::shift_left");
}

template <class Integer, size_t Bits0, size_t... Bits>
inline void shift_right_unsigned(const packed_int<Integer, Bits0, Bits...>* in, size_t shift_amount,
                                 packed_int<Integer, Bits0, Bits...>* out, size_t n, stream_t stream = nullptr)
{
    const pint_layout l = detail::layout_of<Integer, Bits0, Bits...>::get();
    check(pint_cuda_shift(0, &l, shift_amount, in, out, n, stream), "pint::cuda::shift_right_unsigned");
}

// slice<Start, End> (pint.hpp:923-932): out is an array of the narrower layout, same Integer
template <size_t Start, size_t End, class Integer, size_t Bits0, size_t... Bits>
inline void slice(const packed_int<Integer, Bits0, Bits...>* in,
                  typename sliced_int<Start, End, Integer, Bits0, Bits...>::type* out, size_t n,
                  stream_t stream = nullptr)
{
    const pint_layout l = detail::layout_of<Integer, Bits0, Bits...>::get();
    check(pint_cuda_slice(&l, (uint32_t)Start, (uint32_t)End, in, out, n, stream), "pint::cuda::slice");
}

// get<I> / get_signed<I> over the whole array (pint.hpp:799-822): planes[f][i] = field f of
// packed[i].  Plane is int32_t/uint32_t or int64_t/uint64_t; a null plane skips the field.
template <class Plane, class Integer, size_t Bits0, size_t... Bits>
inline void unpack(const packed_int<Integer, Bits0, Bits...>* packed, Plane* const (&planes)[sizeof...(Bits) + 1],
                   size_t n, stream_t stream = nullptr)
{
    static_assert(std::is_integral<Plane>::value && (sizeof(Plane) == 1 || sizeof(Plane) == 2 || sizeof(Plane) == 4 || sizeof(Plane) == 8),
                  "planes are 8-, 16-, 32- or 64-bit integers");
    const pint_layout l = detail::layout_of<Integer, Bits0, Bits...>::get();
#if PINT_CUDA_INLINE_KERNELS
    if (n == 0) return;
    kernels::plane_ptrs pp{};
    for (size_t i = 0; i < sizeof...(Bits) + 1; ++i) {
        pp.p[i] = planes[i];
        if ((uintptr_t)planes[i] % sizeof(Plane)) check(PINT_E_ARG, "pint::cuda::unpack");
    }
    if (!packed) check(PINT_E_NULL, "pint::cuda::unpack");
    if ((uintptr_t)packed % sizeof(Integer)) check(PINT_E_ARG, "pint::cuda::unpack");
    using SL = static_layout<Integer, Bits0, Bits...>;
    check((int)kernels::launch_unpack<Integer, (int)sizeof(Plane), std::is_signed<Plane>::value>(
              packed, pp, l.nfields, n, kernels::static_fields<SL>{}, (cudaStream_t)stream, kernels::launch_geom{}),
          "pint::cuda::unpack");
#else
    void* p[sizeof...(Bits) + 1];
    for (size_t i = 0; i < sizeof...(Bits) + 1; ++i) p[i] = planes[i];
    check(pint_cuda_unpack(&l, std::is_signed<Plane>::value ? 1 : 0, packed, p, (int)sizeof(Plane), n, stream),
          "pint::cuda::unpack");
#endif
}

// the truncating field constructor over arrays (pint.hpp:770-774)
template <class Plane, class Integer, size_t Bits0, size_t... Bits>
inline void repack(const Plane* const (&planes)[sizeof...(Bits) + 1], packed_int<Integer, Bits0, Bits...>* packed,
                   size_t n, stream_t stream = nullptr)
{
    static_assert(std::is_integral<Plane>::value && (sizeof(Plane) == 1 || sizeof(Plane) == 2 || sizeof(Plane) == 4 || sizeof(Plane) == 8),
                  "planes are 8-, 16-, 32- or 64-bit integers");
    const pint_layout l = detail::layout_of<Integer, Bits0, Bits...>::get();
#if PINT_CUDA_INLINE_KERNELS
    if (n == 0) return;
    kernels::plane_ptrs pp{};
    for (size_t i = 0; i < sizeof...(Bits) + 1; ++i) {
        if (!planes[i]) check(PINT_E_NULL, "pint::cuda::repack");
        if ((uintptr_t)planes[i] % sizeof(Plane)) check(PINT_E_ARG, "pint::cuda::repack");
        pp.p[i] = const_cast<Plane*>(planes[i]);
    }
    if (!packed) check(PINT_E_NULL, "pint::cuda::repack");
    if ((uintptr_t)packed % sizeof(Integer)) check(PINT_E_ARG, "pint::cuda::repack");
    using SL = static_layout<Integer, Bits0, Bits...>;
    check((int)kernels::launch_repack<Integer, (int)sizeof(Plane)>(pp, l.nfields, packed, n, kernels::static_fields<SL>{},
                                                                   (cudaStream_t)stream, kernels::launch_geom{}),
          "pint::cuda::repack");
#else
    const void* p[sizeof...(Bits) + 1];
    for (size_t i = 0; i < sizeof...(Bits) + 1; ++i) p[i] = planes[i];
    check(pint_cuda_repack(&l, p, (int)sizeof(Plane), packed, n, stream), "pint::cuda::repack");
#endif
}

// Dense bit-stream storage (include/pint_cuda.h, "dense" section): element i of
// packed_int<Integer, Bits...> in bits [i*S, (i+1)*S) of a byte stream, S = sum(Bits).
// The layout is a template argument (a stream has no element type): e.g.
//     pint::cuda::dense::add_signed_saturate<rgb>(a, b, out, n, stream);
namespace dense {

typedef unsigned char stream_byte;

template <class P> struct layout_from;
template <class Integer, size_t Bits0, size_t... Bits>
struct layout_from<packed_int<Integer, Bits0, Bits...> > {
    static pint_layout get() { return detail::layout_of<Integer, Bits0, Bits...>::get(); }
    typedef Integer integer;
    static const size_t nfields = sizeof...(Bits) + 1;
};

// bytes of a dense stream of n elements (16-byte granular)
template <class P>
inline size_t bytes(size_t n)
{
    const pint_layout l = layout_from<P>::get();
    return pint_cuda_dense_bytes(&l, n);
}

template <class P>
inline void binary(pint_op op, const stream_byte* a, const stream_byte* b, stream_byte* out, size_t n,
                   stream_t stream = nullptr)
{
    const pint_layout l = layout_from<P>::get();
    check(pint_cuda_dense_binary(op, &l, a, b, out, n, stream), "pint::cuda::dense::binary");
}

#define PINT_CUDA_DENSE_BINARY(NAME, CODE)                                                                   \
    template <class P>                                                                                       \
    inline void NAME(const stream_byte* a, const stream_byte* b, stream_byte* out, size_t n, stream_t stream = nullptr) \
    {                                                                                                        \
        binary<P>(CODE, a, b, out, n, stream);                                                               \
    }
PINT_CUDA_DENSE_BINARY(add_wrap, PINT_ADD_WRAP)
PINT_CUDA_DENSE_BINARY(add_unsigned_saturate, PINT_ADD_UNSIGNED_SATURATE)
PINT_CUDA_DENSE_BINARY(add_signed_saturate, PINT_ADD_SIGNED_SATURATE)
PINT_CUDA_DENSE_BINARY(sub_wrap, PINT_SUB_WRAP)
PINT_CUDA_DENSE_BINARY(sub_unsigned_saturate, PINT_SUB_UNSIGNED_SATURATE)
PINT_CUDA_DENSE_BINARY(sub_signed_saturate, PINT_SUB_SIGNED_SATURATE)
PINT_CUDA_DENSE_BINARY(bit_or, PINT_OR)
PINT_CUDA_DENSE_BINARY(bit_and, PINT_AND)
PINT_CUDA_DENSE_BINARY(bit_xor, PINT_XOR)
PINT_CUDA_DENSE_BINARY(min_unsigned, PINT_MIN_UNSIGNED)
PINT_CUDA_DENSE_BINARY(max_unsigned, PINT_MAX_UNSIGNED)
PINT_CUDA_DENSE_BINARY(min_signed, PINT_MIN_SIGNED)
PINT_CUDA_DENSE_BINARY(max_signed, PINT_MAX_SIGNED)
#undef PINT_CUDA_DENSE_BINARY

// several ops on the same dense operands; pairs such as add_unsigned_saturate + add_signed_saturate in one pass
template <class P, size_t NOps>
inline void binary_multi(const pint_op (&ops)[NOps], const stream_byte* a, const stream_byte* b,
                         stream_byte* const (&outs)[NOps], size_t n, stream_t stream = nullptr)
{
    static_assert(NOps >= 1 && NOps <= PINT_MAX_MULTI_OPS, "1..PINT_MAX_MULTI_OPS operations");
    const pint_layout l = layout_from<P>::get();
    int codes[NOps];
    void* o[NOps];
    for (size_t j = 0; j < NOps; ++j) { codes[j] = ops[j]; o[j] = outs[j]; }
    check(pint_cuda_dense_binary_multi((int)NOps, codes, &l, a, b, o, n, stream), "pint::cuda::dense::binary_multi");
}

// shift_left / shift_right_unsigned of every field by one uniform amount (pint.hpp:1006-1029)
template <class P>
inline void shift_left(const stream_byte* in, stream_byte* out, size_t n, size_t amount, stream_t stream = nullptr)
{
    const pint_layout l = layout_from<P>::get();
    check(pint_cuda_dense_shift(1, &l, amount, in, out, n, stream), "pint::cuda::dense::shift_left");
}

template <class P>
inline void shift_right_unsigned(const stream_byte* in, stream_byte* out, size_t n, size_t amount, stream_t stream = nullptr)
{
    const pint_layout l = layout_from<P>::get();
    check(pint_cuda_dense_shift(0, &l, amount, in, out, n, stream), "pint::cuda::dense::shift_right_unsigned");
}

// padded packed_int array <-> dense stream
template <class Integer, size_t Bits0, size_t... Bits>
inline void pack(const packed_int<Integer, Bits0, Bits...>* padded, stream_byte* dense_out, size_t n, stream_t stream = nullptr)
{
    const pint_layout l = detail::layout_of<Integer, Bits0, Bits...>::get();
    check(pint_cuda_dense_pack(&l, padded, dense_out, n, stream), "pint::cuda::dense::pack");
}

template <class Integer, size_t Bits0, size_t... Bits>
inline void unpack(const stream_byte* dense_in, packed_int<Integer, Bits0, Bits...>* padded, size_t n, stream_t stream = nullptr)
{
    const pint_layout l = detail::layout_of<Integer, Bits0, Bits...>::get();
    check(pint_cuda_dense_unpack(&l, dense_in, padded, n, stream), "pint::cuda::dense::unpack");
}

// get<I> / get_signed<I> of every field straight from a dense stream, and the truncating
// field constructor straight into one
template <class P, class Plane>
inline void to_planes(const stream_byte* dense_in, Plane* const (&planes)[layout_from<P>::nfields], size_t n,
                      stream_t stream = nullptr)
{
    static_assert(std::is_integral<Plane>::value && (sizeof(Plane) == 1 || sizeof(Plane) == 2 || sizeof(Plane) == 4 || sizeof(Plane) == 8),
                  "planes are 8-, 16-, 32- or 64-bit integers");
    const pint_layout l = layout_from<P>::get();
    void* p[layout_from<P>::nfields];
    for (size_t i = 0; i < layout_from<P>::nfields; ++i) p[i] = planes[i];
    check(pint_cuda_dense_to_planes(&l, std::is_signed<Plane>::value ? 1 : 0, dense_in, p, (int)sizeof(Plane), n, stream),
          "pint::cuda::dense::to_planes");
}

template <class P, class Plane>
inline void from_planes(const Plane* const (&planes)[layout_from<P>::nfields], stream_byte* dense_out, size_t n,
                        stream_t stream = nullptr)
{
    static_assert(std::is_integral<Plane>::value && (sizeof(Plane) == 1 || sizeof(Plane) == 2 || sizeof(Plane) == 4 || sizeof(Plane) == 8),
                  "planes are 8-, 16-, 32- or 64-bit integers");
    const pint_layout l = layout_from<P>::get();
    const void* p[layout_from<P>::nfields];
    for (size_t i = 0; i < layout_from<P>::nfields; ++i) p[i] = planes[i];
    check(pint_cuda_dense_from_planes(&l, p, (int)sizeof(Plane), dense_out, n, stream), "pint::cuda::dense::from_planes");
}

}  // namespace dense

// multi-GPU: independent contiguous shards, one per device, no collective
namespace sharded {

inline void plan(size_t n, size_t elem_bytes, int nshards, size_t* begin, size_t* count)
{
    check(pint_cuda_plan_shards(n, (uint32_t)elem_bytes, nshards, begin, count), "pint::cuda::sharded::plan");
}

template <class Integer, size_t Bits0, size_t... Bits>
inline void binary(pint_op op, int ngpu, const int* devices, const packed_int<Integer, Bits0, Bits...>* const* a,
                   const packed_int<Integer, Bits0, Bits...>* const* b, packed_int<Integer, Bits0, Bits...>* const* out,
                   const size_t* counts, const stream_t* streams = nullptr)
{
    const pint_layout l = detail::layout_of<Integer, Bits0, Bits...>::get();
    check(pint_cuda_sharded_binary(op, &l, ngpu, devices, reinterpret_cast<const void* const*>(a),
                                   reinterpret_cast<const void* const*>(b), reinterpret_cast<void* const*>(out), counts,
                                   streams),
          "pint::cuda::sharded::binary");
}

// config C4's shape: the fused unpack -> arithmetic -> repack kernel on every device's shard
template <class Integer, size_t Bits0, size_t... Bits>
inline void fused_unpack_arith_repack(pint_op op, int ngpu, const int* devices,
                                      const packed_int<Integer, Bits0, Bits...>* const* a,
                                      const packed_int<Integer, Bits0, Bits...>* const* b,
                                      packed_int<Integer, Bits0, Bits...>* const* out, const size_t* counts,
                                      const stream_t* streams = nullptr)
{
    const pint_layout l = detail::layout_of<Integer, Bits0, Bits...>::get();
    check(pint_cuda_sharded_fused_unpack_arith_repack(op, &l, ngpu, devices, reinterpret_cast<const void* const*>(a),
                                                      reinterpret_cast<const void* const*>(b),
                                                      reinterpret_cast<void* const*>(out), counts, streams),
          "pint::cuda::sharded::fused_unpack_arith_repack");
}

// dense streams: shard starts are multiples of 128 elements (16-byte boundaries of the stream)
inline void dense_plan(size_t n, int nshards, size_t* begin, size_t* count)
{
    check(pint_cuda_dense_plan_shards(n, nshards, begin, count), "pint::cuda::sharded::dense_plan");
}

template <class P>
inline void dense_binary(pint_op op, int ngpu, const int* devices, const unsigned char* const* a,
                         const unsigned char* const* b, unsigned char* const* out, const size_t* counts,
                         const stream_t* streams = nullptr)
{
    const pint_layout l = dense::layout_from<P>::get();
    check(pint_cuda_sharded_dense_binary(op, &l, ngpu, devices, reinterpret_cast<const void* const*>(a),
                                         reinterpret_cast<const void* const*>(b), reinterpret_cast<void* const*>(out),
                                         counts, streams),
          "pint::cuda::sharded::dense_binary");
}

// one array on src_device -> the shards of plan() on devices[] (peer copies, one stream each) and back:
// the partitioner for an array that is already resident on ONE GPU
template <class T>
inline void scatter(const T* src, int src_device, size_t n, int ngpu, const int* devices, T* const* shards,
                    const stream_t* streams = nullptr)
{
    check(pint_cuda_scatter_shards(src, src_device, (uint32_t)detail::elem_size<T>::value, n, ngpu, devices,
                                   reinterpret_cast<void* const*>(shards), streams),
          "pint::cuda::sharded::scatter");
}

template <class T>
inline void gather(T* dst, int dst_device, size_t n, int ngpu, const int* devices, const T* const* shards,
                   const stream_t* streams = nullptr)
{
    check(pint_cuda_gather_shards(dst, dst_device, (uint32_t)detail::elem_size<T>::value, n, ngpu, devices,
                                  reinterpret_cast<const void* const*>(shards), streams),
          "pint::cuda::sharded::gather");
}

inline void enable_peer_access(int ngpu, const int* devices)
{
    check(pint_cuda_enable_peer_access(ngpu, devices), "pint::cuda::sharded::enable_peer_access");
}

}  // namespace sharded

// The reference benchmark's loop as one call (tests/pint_bench.cpp:90-94): *sum_dev (device) = sum over i of
// OP(packed_int(pairs[i].first), packed_int(pairs[i].second)).value() mod 2^64; P names the layout:
//     pint::cuda::pairs_reduce<rgba>(PINT_ADD_SIGNED_SATURATE, pairs, n, sum_dev, stream);
template <class P>
inline void pairs_reduce(pint_op op,
                         const std::pair<typename dense::layout_from<P>::integer, typename dense::layout_from<P>::integer>* pairs,
                         size_t n, uint64_t* sum_dev, stream_t stream = nullptr)
{
    const pint_layout l = dense::layout_from<P>::get();
    check(pint_cuda_pairs_reduce(op, &l, pairs, n, sum_dev, stream), "pint::cuda::pairs_reduce");
}

// The same for callers that may not allocate (no counterpart in the reference; pint_cuda.h:
// pint_cuda_pairs_reduce_static): n <= max_pairs_no_alloc, calls that overlap on one GPU must be serialised.
static const unsigned long long max_pairs_no_alloc = PINT_STATIC_MAX_PAIRS;
template <class P>
inline void pairs_reduce_no_alloc(pint_op op,
                                  const std::pair<typename dense::layout_from<P>::integer, typename dense::layout_from<P>::integer>* pairs,
                                  size_t n, uint64_t* sum_dev, stream_t stream = nullptr)
{
    const pint_layout l = dense::layout_from<P>::get();
    check(pint_cuda_pairs_reduce_static(op, &l, pairs, n, sum_dev, stream), "pint::cuda::pairs_reduce_no_alloc");
}

// Mask constants for ANY layout without nvcc in the caller's build: compiles the kernels for
// static_layout<Integer, Bits...> with NVRTC on the current device ahead of time (about a second per family, cached
// on disk) and switches the "jit" option on so later calls use them.  0 = run-time masks, 1 = registered at
// build time, 2 = specialised at run time.
template <class P>
inline int specialization()
{
    const pint_layout l = dense::layout_from<P>::get();
    return pint_cuda_layout_is_specialized(&l);
}

template <class P>
inline void specialize()
{
    const pint_layout l = dense::layout_from<P>::get();
    if (pint_cuda_get_option("jit") == 0) check(pint_cuda_set_option("jit", 1), "pint::cuda::specialize");
    const int rc = pint_cuda_jit_compile(&l);
    if (rc != PINT_OK) throw error(rc, (std::string("pint::cuda::specialize: ") + pint_cuda_jit_last_error()).c_str());
}

// HOST arrays (the reference's workload is a host vector, tests/pint_bench.cpp:22,75): copied through per-device
// staging rings so that H2D, kernel and D2H overlap; synchronous.  Pinned buffers (host::alloc) run at the
// link's rate, pageable ones at the driver's staging rate.
namespace host {

template <class T>
inline T* alloc(size_t count, unsigned flags = 0)
{
    void* p = nullptr;
    check(pint_cuda_host_alloc_flags(&p, count * detail::elem_size<T>::value, flags), "pint::cuda::host::alloc");
    return static_cast<T*>(p);
}

inline void dealloc(void* p) { check(pint_cuda_host_free(p), "pint::cuda::host::dealloc"); }
inline void release_staging() { check(pint_cuda_host_release(), "pint::cuda::host::release_staging"); }

template <class Integer, size_t Bits0, size_t... Bits>
inline void binary(pint_op op, const packed_int<Integer, Bits0, Bits...>* a, const packed_int<Integer, Bits0, Bits...>* b,
                   packed_int<Integer, Bits0, Bits...>* out, size_t n)
{
    const pint_layout l = detail::layout_of<Integer, Bits0, Bits...>::get();
    check(pint_cuda_binary_host(op, &l, a, b, out, n), "pint::cuda::host::binary");
}

// several ops over ONE upload of a and b; ngpu > 0 splits the arrays into plan()'s shards over devices[]
template <size_t NOps, class Integer, size_t Bits0, size_t... Bits>
inline void binary_multi(const pint_op (&ops)[NOps], const packed_int<Integer, Bits0, Bits...>* a,
                         const packed_int<Integer, Bits0, Bits...>* b, packed_int<Integer, Bits0, Bits...>* const (&outs)[NOps],
                         size_t n, int ngpu = 0, const int* devices = nullptr)
{
    static_assert(NOps >= 1 && NOps <= PINT_MAX_HOST_OPS, "1 .. PINT_MAX_HOST_OPS operations per pass");
    const pint_layout l = detail::layout_of<Integer, Bits0, Bits...>::get();
    int codes[NOps];
    void* o[NOps];
    for (size_t j = 0; j < NOps; ++j) { codes[j] = ops[j]; o[j] = outs[j]; }
    if (ngpu > 0)
        check(pint_cuda_sharded_binary_multi_host((int)NOps, codes, &l, ngpu, devices, a, b, o, n), "pint::cuda::host::binary_multi (sharded)");
    else
        check(pint_cuda_binary_multi_host((int)NOps, codes, &l, a, b, o, n), "pint::cuda::host::binary_multi");
}

}  // namespace host

}  // namespace cuda
}  // namespace pint
