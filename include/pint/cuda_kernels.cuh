// pint/cuda_kernels.cuh -- hand-written sm_100a kernels of the pint::cuda batch layer.
//
// Device code only (include from a .cu compiled with
//   nvcc -std=c++17 -gencode arch=compute_100a,code=sm_100a).
// libpint_cuda.so instantiates these templates for the layouts in
// pint_b200/csrc/registry_layouts.def (mask constants folded into the SASS)
// and once per Word for arbitrary runtime layouts (masks in a __grid_constant__
// kernel parameter).  A user with nvcc can instantiate them for any
// packed_int<Integer, Bits...> through static_layout<Integer, Bits...>.
//
// Every kernel is an elementwise map over a packed array -- the batched form of
// the loop in /root/reference/tests/pint_bench.cpp:90-94 -- and is bound by HBM
// bandwidth, not by arithmetic: 16-byte coalesced accesses, several independent
// loads in flight per thread, streaming cache policy (nothing is reused), grid
// sized in multiples of the SM count.  No tensor cores: nothing is a contraction.
//
// Per-element semantics are those of /root/reference/include/pint/pint.hpp
// ("pint.hpp:N" below); the formulas are our own where a cheaper one computes
// the identical function (each such place says why it is identical).
#pragma once
#ifndef PINT_CUDA_KERNELS_CUH_INCLUDED
#define PINT_CUDA_KERNELS_CUH_INCLUDED

#include "cuda_std.hpp"

#if !defined(__CUDACC_RTC__)
#include <cuda_runtime.h>
#endif

#include "cuda_layout.hpp"

namespace pint {
namespace cuda {

// Operation codes; identical to enum pint_op in include/pint_cuda.h.
enum op_code : int {
    OP_ADD_WRAP = 0,
    OP_ADD_UNSIGNED_SATURATE = 1,
    OP_ADD_SIGNED_SATURATE = 2,
    OP_SUB_WRAP = 3,
    OP_SUB_UNSIGNED_SATURATE = 4,
    OP_SUB_SIGNED_SATURATE = 5,
    OP_OR = 6,
    OP_AND = 7,
    OP_XOR = 8,
    OP_MIN_UNSIGNED = 9,
    OP_MAX_UNSIGNED = 10,
    OP_MIN_SIGNED = 11,
    OP_MAX_SIGNED = 12,
    OP_NUM_BINARY = 13
};

namespace kernels {

// ---------------------------------------------------------------------------
// Mask policies.  Both expose the same interface; the static one folds to
// immediates, the runtime one reads the constant bank.

template <class SL>
struct static_masks {
    using word = typename SL::word;
    // many distinct widths make the saturating ops ALU-heavy: their kernels run with a deeper unroll
    static constexpr bool heavy_fill = SL::masks.nsh >= 3;
    static constexpr bool compiled_in = true;
    __device__ __forceinline__ static constexpr word H() { return SL::masks.H; }
    __device__ __forceinline__ static constexpr word L() { return SL::masks.L; }
    __device__ __forceinline__ static constexpr word A() { return SL::masks.A; }
    __device__ __forceinline__ static constexpr word M1() { return SL::masks.M1; }
    __device__ __forceinline__ static constexpr word HL() { return SL::masks.HL; }
    __device__ __forceinline__ static constexpr word LM1() { return SL::masks.LM1; }

    template <std::size_t K>
    __device__ __forceinline__ static constexpr word step(word c)
    {
        constexpr uint32_t s = SL::masks.sh[K];
        constexpr word m = SL::masks.mk[K];
        constexpr bool masked = SL::masks.per_shift_mask != 0;
        return masked ? word((c >> s) & m) : word(c >> s);
    }
    template <std::size_t... Ks>
    __device__ __forceinline__ static constexpr word lsb_impl(word c, std::index_sequence<Ks...>)
    {
        word r = (step<Ks>(c) | ...);
        constexpr bool need_common = SL::masks.per_shift_mask == 0 && SL::masks.nsh > 1;
        return need_common ? word(r & SL::masks.L) : r;
    }
    // move each field's MSB flag to that field's LSB (c must be a subset of H)
    __device__ __forceinline__ static constexpr word lsb(word c)
    {
        return lsb_impl(c, std::make_index_sequence<SL::masks.nsh>());
    }
    // (a local constexpr copy: a class-type mask -- the multi-word chunks of the dense kernels --
    // must not be odr-used from device code)
    template <std::size_t K>
    __device__ __forceinline__ static constexpr word keep_k()
    {
        constexpr word m = SL::masks.keep[K];
        return m;
    }
    template <std::size_t... Ks>
    __device__ __forceinline__ static constexpr word smear_impl(word x, std::index_sequence<Ks...>)
    {
        ((x |= word((x >> (1u << Ks)) & keep_k<Ks>())), ...);
        return x;
    }
    // all bits of every field whose MSB flag is set in c (c subset of H): the
    // value of make_unsigned_saturation_mask, pint.hpp:544-551 ((c << 1) - lsb(c)
    // there; both are the union of the flagged fields' bit ranges)
    __device__ __forceinline__ static constexpr word fill(word c)
    {
        if constexpr (SL::masks.use_smear != 0) return smear_impl(c, std::make_index_sequence<SL::masks.nsmear>());
        else return c | (c - lsb(c));
    }
    // the bits under the MSB of every flagged field: make_signed_saturation_mask, pint.hpp:563-567
    __device__ __forceinline__ static constexpr word below(word c)
    {
        if constexpr (SL::masks.use_smear != 0) return fill(c) ^ c;
        else return c - lsb(c);
    }
};

// Runtime masks: a fixed-structure copy of the plan so that every constant sits
// at a fixed offset of the kernel parameter (constant bank operands cost
// nothing; a loop over a table with a dynamic index does -- measured 2x slower).
// Single-width layouts move the flag with one shift; all others smear.
// Steps: how many smear steps the fill runs.  The default (all of them: 5 for 32-bit Words, 6 for
// 64-bit ones) serves any layout; the run-time-mask launcher picks a 2- or 4-step instantiation
// when the widest field is at most 4 / 16 bits (generic.cu), so that narrow-field layouts do not
// pay for steps whose keep mask is zero.
template <class Word, int Steps = (sizeof(Word) == 8 ? 6 : 5)>
struct runtime_masks {
    using word = Word;
    static constexpr bool heavy_fill = true;
    static constexpr bool compiled_in = false;
    static constexpr int steps = Steps;
    Word mH, mL, mA, mM1, mHL, mLM1;
    Word keep[6];
    uint32_t shift0;   // w - 1 when all fields have the same width
    uint32_t nsmear;   // 0: single-width layout (shift plan), else number of smear steps

    runtime_masks() = default;
    explicit runtime_masks(const word_masks<Word>& m)
        : mH(m.H), mL(m.L), mA(m.A), mM1(m.M1), mHL(m.HL), mLM1(m.LM1), keep{m.keep[0], m.keep[1], m.keep[2], m.keep[3], m.keep[4], m.keep[5]},
          shift0(m.sh[0]), nsmear(m.nsh == 1 ? 0 : m.nsmear) {}

    __device__ __forceinline__ word H() const { return mH; }
    __device__ __forceinline__ word L() const { return mL; }
    __device__ __forceinline__ word A() const { return mA; }
    __device__ __forceinline__ word M1() const { return mM1; }
    __device__ __forceinline__ word HL() const { return mHL; }
    __device__ __forceinline__ word LM1() const { return mLM1; }
    __device__ __forceinline__ word fill(word c) const
    {
        if (nsmear == 0) return c | (c - (c >> shift0));
        // all Steps steps, unconditionally: keep[k] is 0 for the steps a layout does not need.  Skipping
        // them with warp-uniform branches on nsmear was measured and is slower where it matters
        // (padded <7,9,3> add_signed_saturate 6226 -> 5181 GB/s, profiles/r01q_extra.jsonl).
        word x = c;
        if constexpr (Steps >= 1) x |= (x >> 1) & keep[0];
        if constexpr (Steps >= 2) x |= (x >> 2) & keep[1];
        if constexpr (Steps >= 3) x |= (x >> 4) & keep[2];
        if constexpr (Steps >= 4) x |= (x >> 8) & keep[3];
        if constexpr (Steps >= 5) x |= (x >> 16) & keep[4];
        if constexpr (Steps >= 6 && sizeof(Word) == 8) x |= (x >> 32) & keep[5];
        return x;
    }
    __device__ __forceinline__ word below(word c) const
    {
        if (nsmear == 0) return c - (c >> shift0);
        return fill(c) ^ c;
    }
};

// ---------------------------------------------------------------------------
// SWAR arithmetic on one Word (several elements when the Integer is narrow).
// Bits above sum(Bits) of every element: zero after add_*/sub_* (M1 and H
// exclude them, pint.hpp:832-833, 877-878), passed through by | & ^
// (pint.hpp:781-783), taken from b by min/max (pint.hpp:603-606).

template <class P>
__device__ __forceinline__ typename P::word field_fill(typename P::word c, const P& p)
{
    return p.fill(c);
}

// add_wrap, pint.hpp:826-838
template <class P>
__device__ __forceinline__ typename P::word add_wrap(typename P::word a, typename P::word b, const P& p)
{
    return ((a & p.M1()) + (b & p.M1())) ^ ((a ^ b) & p.H());
}

// sub_wrap, pint.hpp:870-884: a + ~b + 1 per field with the MSB handled apart.
template <class P>
__device__ __forceinline__ typename P::word sub_wrap(typename P::word a, typename P::word b, const P& p)
{
    return ((a & p.M1()) + (~b & p.M1()) + p.LM1()) ^ (~(a ^ b) & p.H()) ^ p.HL();
}

// apply_signed_saturation, pint.hpp:569-574
template <class P>
__device__ __forceinline__ typename P::word signed_clamp(typename P::word r, typename P::word ov, const P& p)
{
    using W = typename P::word;
    const W below = p.below(ov);             // bits under the MSB of overflowed fields
    const W neg = ov & ~r;                   // overflowed towards -inf (wrapped result looks positive)
    const W below_neg = p.below(neg);
    return ((r ^ ov) | below) ^ below_neg;   // 0111..1 for +overflow, 1000..0 for -overflow
}

template <int Op, class P>
__device__ __forceinline__ typename P::word apply_op(typename P::word a, typename P::word b, const P& p)
{
    using W = typename P::word;
    if constexpr (Op == OP_ADD_WRAP) {
        return add_wrap(a, b, p);
    } else if constexpr (Op == OP_ADD_UNSIGNED_SATURATE) {
        // pint.hpp:840-855.  Carry out of each field from the wrapped per-field
        // sum instead of the full-width a + b (pint.hpp:375-378): exact per
        // field, and the reference's only deviation (a spurious flag on an
        // all-ones field that received a carry-in) saturates to the same value.
        const W s = add_wrap(a, b, p);
        const W c = ((a & b) | ((a | b) & ~s)) & p.H();
        return s | field_fill(c, p);
    } else if constexpr (Op == OP_ADD_SIGNED_SATURATE) {
        // pint.hpp:857-866, 576-582: overflow iff both operands' signs differ from the sum's
        const W s = add_wrap(a, b, p);
        return signed_clamp(s, W((s ^ a) & (s ^ b) & p.H()), p);
    } else if constexpr (Op == OP_SUB_WRAP) {
        return sub_wrap(a, b, p);
    } else if constexpr (Op == OP_SUB_UNSIGNED_SATURATE) {
        // pint.hpp:886-908 computes max(a - b, 0) per field as
        // add_wrap(sat(a + ~b), L); clearing the fields that borrowed gives
        // the same function with one SWAR add instead of three.
        const W d = sub_wrap(a, b, p);
        const W bo = ((~a & b) | (~(a ^ b) & d)) & p.H();
        return d & ~field_fill(bo, p);
    } else if constexpr (Op == OP_SUB_SIGNED_SATURATE) {
        // pint.hpp:910-919, 385-388, 584-590: overflow iff sign(a) != sign(b) and sign(d) != sign(a)
        const W d = sub_wrap(a, b, p);
        return signed_clamp(d, W((a ^ b) & (a ^ d) & p.H()), p);
    } else if constexpr (Op == OP_OR) {
        return a | b;
    } else if constexpr (Op == OP_AND) {
        return a & b;
    } else if constexpr (Op == OP_XOR) {
        return a ^ b;
    } else {
        // min/max, pint.hpp:936-1004: borrow flags of the full-width x - y
        // (carry_sub_vector, pint.hpp:380-383) select a or b per field; a borrow
        // that leaks in from below can only flip the choice between equal fields.
        constexpr bool is_signed = (Op == OP_MIN_SIGNED || Op == OP_MAX_SIGNED);
        constexpr bool is_max = (Op == OP_MAX_UNSIGNED || Op == OP_MAX_SIGNED);
        const W x0 = is_max ? b : a, y0 = is_max ? a : b;
        const W x = is_signed ? W(x0 ^ p.H()) : x0, y = is_signed ? W(y0 ^ p.H()) : y0;
        const W bo = ((~x & y) | (~(x ^ y) & (x - y))) & p.H();
        const W m = field_fill(bo, p);
        return (a & m) | (b & ~m);
    }
}

// unary "mask, shift, mask": covers slice (pint.hpp:923-932), shift_left
// (:1006-1017, 654-668) and shift_right_unsigned (:1019-1029, 692-705); the
// masks depend only on the layout and the uniform amount, so the host computes
// them once per call (unary_params_* in cuda_host.hpp).
template <class Word>
struct unary_params {
    Word pre;        // applied before the shift
    Word post;       // applied after the shift
    uint32_t amount; // < 8 * sizeof(Word)
    uint32_t left;   // 1: shift left, 0: shift right
};

template <class Word>
__device__ __forceinline__ Word apply_unary(Word v, const unary_params<Word>& u)
{
    const Word t = v & u.pre;
    return (u.left ? Word(t << u.amount) : Word(t >> u.amount)) & u.post;
}

}  // namespace kernels
}  // namespace cuda
}  // namespace pint

#include "cuda_kernels_stream.cuh"
#include "cuda_kernels_multi.cuh"
#include "cuda_kernels_planes.cuh"
#include "cuda_kernels_callers.cuh"

#endif  // PINT_CUDA_KERNELS_CUH_INCLUDED
