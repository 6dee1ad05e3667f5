// pint/cuda_layout.hpp -- the mask generator of the pint::cuda batch layer.
//
// One constexpr function, make_word_masks<Word>(layout_desc), turns a
// Bits... layout into every constant the sm_100a kernels need.  It is used
//   * at COMPILE time: static_layout<Integer, Bits...>::masks is a constexpr
//     object, so kernels instantiated for a layout see H, L, M1 and the
//     flag-move shifts as immediates ("kernel template constants"), and
//   * at RUN time on the host, for layouts that have no pre-instantiated
//     kernel: the same struct is passed as a __grid_constant__ kernel parameter.
//
// What it replaces in the reference (/root/reference/include/pint/pint.hpp):
//   mask_hiorder :324-329, mask_loorder :332-337, all_ones :340-348,
//   mask_offsets_vector :288-292, find_max :277-285,
//   detect_saturation_mask_type :424-456 and the three flag-move code shapes
//   :459-542, unsigned_saturation_mask_type_2 :503-515.
// tests/cpp/static_masks_check.cpp static_asserts equality with those
// metafunctions for every layout in oracle/ref_layouts.def.
//
// GPU registers are 32 bit wide, so packs narrower than 32 bit are processed
// several at a time: an array of uint16_t<5,6,5> IS an array of
// uint32_t<5,6,5,5,6,5> because no carry ever crosses a field (the reference
// does the same by hand in tests/pint_bench.cpp:468-485).  "Word" below is that
// processing unit (uint32_t for 1/2/4-byte Integers, uint64_t for 8-byte ones)
// and every mask is replicated across the elements a Word holds.
//
// Requires C++17.  Host and device.
#pragma once
#ifndef PINT_CUDA_LAYOUT_HPP_INCLUDED
#define PINT_CUDA_LAYOUT_HPP_INCLUDED

#include "cuda_std.hpp"

#if defined(__CUDACC__)
#define PINT_HD __host__ __device__
#else
#define PINT_HD
#endif

namespace pint {
namespace cuda {

constexpr int kMaxFields = 64;
// distinct field widths in <= 64 bits: 1+2+...+10 = 55 <= 64 < 66, so at most 10
constexpr int kMaxUnique = 10;

// Runtime description of packed_int<Integer, Bits...>; layout-compatible with
// the C ABI's pint_layout (include/pint_cuda.h).
struct layout_desc {
    uint32_t elem_bytes;       // sizeof(Integer): 1, 2, 4, 8
    uint32_t nfields;          // sizeof...(Bits) + 1
    uint8_t bits[kMaxFields];  // field 0 is least significant
};

// 0 when `d` describes a type the reference header accepts, else a negative
// code (the header's static_asserts: pint.hpp:761-764, 727-728).
constexpr int validate(const layout_desc& d)
{
    if (d.elem_bytes != 1 && d.elem_bytes != 2 && d.elem_bytes != 4 && d.elem_bytes != 8) return -2;
    if (d.nfields < 1 || d.nfields > (uint32_t)kMaxFields) return -3;
    uint32_t total = 0;
    for (uint32_t i = 0; i < d.nfields; ++i) {
        if (d.bits[i] == 0) return -4;
        total += d.bits[i];
    }
    if (total > 8 * d.elem_bytes) return -5;
    return 0;
}

// make_packed_int's storage choice (pint.hpp:789-795, clp2 :710-722); 0 if unsupported.
constexpr uint32_t storage_bytes(uint32_t total_bits)
{
    if (total_bits == 0 || total_bits > 64) return 0;
    uint32_t p = 8;
    while (p < total_bits) p *= 2;
    return p / 8;
}

template <class Word>
struct word_masks {
    Word H;    // MSB of every field, replicated over the elements in a Word
    Word L;    // LSB of every field
    Word A;    // all field bits (bits below sum(Bits) of every element)
    Word M1;   // A & ~H
    Word HL;   // H & L: the 1-bit fields (sub_wrap's extra term, pint.hpp:882)
    Word LM1;  // L & M1
    // Flag move "MSB of each field -> LSB of that field":
    //   lsb(c) = OR_k ((c >> sh[k]) & mk[k])         when per_shift_mask
    //          = (OR_k (c >> sh[k])) & L             otherwise
    uint32_t nsh;
    uint32_t per_shift_mask;
    uint32_t sh[kMaxUnique];
    Word mk[kMaxUnique];
    // Alternative "smear" plan for layouts with many distinct widths: all bits of
    // every flagged field directly, by log-step propagation of the MSB flag down
    // the field:  x = c; for k < nsmear: x |= (x >> (1 << k)) & keep[k];
    // keep[k] has bit i set iff bits i and i + 2^k belong to the same field.
    // use_smear is the generator's choice (cheaper of the two plans).
    uint32_t nsmear;
    uint32_t use_smear;
    Word keep[6];
    // Per unique width: the width and the LSBs of the fields of that width
    // (shift_left / shift_right_unsigned masks, pint.hpp:630-705).
    uint32_t nuniq;
    uint32_t uw[kMaxUnique];
    Word ulo[kMaxUnique];
    uint32_t elem_bits;   // 8 * sizeof(Integer)
    uint32_t total_bits;  // sum(Bits)
    uint32_t max_bits;    // max(Bits)
    uint32_t ref_sat_type;  // what detect_saturation_mask_type yields for the element (0/1/2)
};

namespace detail {

template <class Word>
constexpr Word ones(uint32_t w)
{
    return w >= 8 * sizeof(Word) ? ~Word(0) : Word((Word(1) << w) - 1);
}

constexpr int popcount64(uint64_t x)
{
    int c = 0;
    while (x) { x &= x - 1; ++c; }
    return c;
}

}  // namespace detail

// The mask generator.  `d` must be valid (validate(d) == 0) and
// 8 * d.elem_bytes must divide the Word width.
template <class Word>
constexpr word_masks<Word> make_word_masks(const layout_desc& d)
{
    word_masks<Word> m{};
    const uint32_t wbits = 8 * sizeof(Word);
    const uint32_t ebits = 8 * d.elem_bytes;
    const uint32_t reps = wbits / ebits;
    m.elem_bits = ebits;

    // element-level constants
    uint64_t eH = 0, eL = 0;
    uint32_t off = 0, maxb = 0;
    bool all_same = true;
    uint32_t nuniq = 0;
    uint32_t uw[kMaxFields] = {};
    uint64_t ulo[kMaxFields] = {};
    for (uint32_t i = 0; i < d.nfields; ++i) {
        const uint32_t w = d.bits[i];
        eL |= uint64_t(1) << off;
        eH |= uint64_t(1) << (off + w - 1);
        if (w > maxb) maxb = w;
        if (w != d.bits[0]) all_same = false;
        uint32_t k = 0;
        while (k < nuniq && uw[k] != w) ++k;
        if (k == nuniq) uw[nuniq++] = w;
        ulo[k] |= uint64_t(1) << off;
        off += w;
    }
    const uint64_t eA = detail::ones<uint64_t>(off);
    m.total_bits = off;
    m.max_bits = maxb;

    // the reference's own classification of the element (reported, and used
    // by tests to show all three code shapes are exercised)
    if (all_same) {
        m.ref_sat_type = 0;
    } else {
        int cnt = 0;
        for (uint32_t k = 0; k < nuniq; ++k) cnt += detail::popcount64((eH >> (uw[k] - 1)) & eL);
        m.ref_sat_type = (cnt == (int)d.nfields) ? 1 : 2;
    }

    // replicate across the Word
    uint64_t H = 0, L = 0, A = 0;
    uint64_t rlo[kMaxFields] = {};
    for (uint32_t r = 0; r < reps; ++r) {
        const uint32_t s = r * ebits;
        H |= eH << s;
        L |= eL << s;
        A |= eA << s;
        for (uint32_t k = 0; k < nuniq; ++k) rlo[k] |= ulo[k] << s;
    }
    m.H = Word(H);
    m.L = Word(L);
    m.A = Word(A);
    m.M1 = Word(A & ~H);
    m.HL = Word(H & L);
    m.LM1 = Word(L & A & ~H);
    m.nuniq = nuniq;
    for (uint32_t k = 0; k < nuniq && k < (uint32_t)kMaxUnique; ++k) {
        m.uw[k] = uw[k];
        m.ulo[k] = Word(rlo[k]);
    }

    // Flag-move plan, decided on the REPLICATED word (the reference decides on
    // the element, pint.hpp:443-456; all shapes compute the same function, so
    // the choice is free as long as it is alias-free here):
    //  - one unique width: a single shift lands every flag on its LSB, no mask;
    //  - alias-free OR of shifts: shifting H by (w-1) for a unique width w and
    //    masking with L must hit exactly the LSBs of the width-w fields;
    //  - otherwise one mask per shift.
    m.nsh = nuniq;
    for (uint32_t k = 0; k < nuniq && k < (uint32_t)kMaxUnique; ++k) m.sh[k] = uw[k] - 1;
    bool alias_free = true;
    for (uint32_t k = 0; k < nuniq; ++k)
        if (((H >> (uw[k] - 1)) & L) != rlo[k]) alias_free = false;
    if (nuniq == 1) {
        m.per_shift_mask = 0;
        m.mk[0] = Word(L);
    } else if (alias_free) {
        m.per_shift_mask = 0;
        for (uint32_t k = 0; k < nuniq && k < (uint32_t)kMaxUnique; ++k) m.mk[k] = Word(L);
    } else {
        m.per_shift_mask = 1;
        for (uint32_t k = 0; k < nuniq && k < (uint32_t)kMaxUnique; ++k) m.mk[k] = Word(rlo[k]);
    }
    // Smear plan and the choice between the two (cost in 32-bit ALU ops per
    // Word-half: a shift or a 3-input logic op each count 1).
    {
        uint32_t k = 0;
        while ((uint32_t(1) << k) < maxb) ++k;   // 2^k - 1 >= maxb - 1
        m.nsmear = k;
        // field id of every bit of the replicated word (0 = not in a field)
        uint32_t fid[64] = {};
        uint32_t id = 0;
        for (uint32_t r = 0; r < reps; ++r) {
            uint32_t o = r * ebits;
            for (uint32_t i = 0; i < d.nfields; ++i) {
                ++id;
                for (uint32_t b = 0; b < d.bits[i]; ++b) fid[o + b] = id;
                o += d.bits[i];
            }
        }
        for (uint32_t j = 0; j < k && j < 6; ++j) {
            const uint32_t s = uint32_t(1) << j;
            uint64_t keep = 0;
            for (uint32_t i = 0; i + s < wbits; ++i)
                if (fid[i] && fid[i] == fid[i + s]) keep |= uint64_t(1) << i;
            m.keep[j] = Word(keep);
        }
        const uint32_t cost_shift = m.nsh * (1 + m.per_shift_mask) + (m.nsh - 1) / 2 + 2 +
                                    ((m.nsh > 1 && !m.per_shift_mask) ? 1 : 0);
        const uint32_t cost_smear = 2 * k;
        m.use_smear = (m.nsh > 1 && cost_smear < cost_shift) ? 1 : 0;
    }
    return m;
}

// Per-field offsets and widths of one Word (the element's fields replicated
// over the elements a Word holds): what the per-field kernels (unpack, repack,
// fused unpack->arith->repack) iterate over.  mask_offsets_vector,
// pint.hpp:288-292; take_offset_and_mask, pint.hpp:367-373.
struct field_table {
    uint32_t nfields;            // fields per element
    uint32_t nword_fields;       // fields per Word (= nfields * elements per Word)
    uint8_t off[kMaxFields];     // offsets within the Word
    uint8_t bits[kMaxFields];
};

template <class Word>
constexpr field_table make_field_table(const layout_desc& d)
{
    field_table t{};
    const uint32_t ebits = 8 * d.elem_bytes;
    const uint32_t reps = 8 * sizeof(Word) / ebits;
    t.nfields = d.nfields;
    uint32_t k = 0;
    for (uint32_t r = 0; r < reps; ++r) {
        uint32_t off = r * ebits;
        for (uint32_t i = 0; i < d.nfields && k < (uint32_t)kMaxFields; ++i, ++k) {
            t.off[k] = uint8_t(off);
            t.bits[k] = d.bits[i];
            off += d.bits[i];
        }
    }
    t.nword_fields = k;
    return t;
}

// Processing word for an Integer.
template <class Integer>
using word_of = typename std::conditional<sizeof(Integer) == 8, uint64_t, uint32_t>::type;

// Compile-time layout: everything a kernel instantiated for
// packed_int<Integer, Bits...> needs, as constants.
template <class Integer, std::size_t... Bits>
struct static_layout {
    static_assert(std::is_integral<Integer>::value && std::is_unsigned<Integer>::value,
                  "Integer must be unsigned integer");  // pint.hpp:761-762
    using integer = Integer;
    using word = word_of<Integer>;
    static constexpr layout_desc desc{(uint32_t)sizeof(Integer), (uint32_t)sizeof...(Bits), {uint8_t(Bits)...}};
    static_assert(validate(desc) == 0, "Integer won't fit given number of bits");  // pint.hpp:763-764
    static constexpr word_masks<word> masks = make_word_masks<word>(desc);
    static constexpr field_table fields = make_field_table<word>(desc);
};

}  // namespace cuda
}  // namespace pint

#endif  // PINT_CUDA_LAYOUT_HPP_INCLUDED
