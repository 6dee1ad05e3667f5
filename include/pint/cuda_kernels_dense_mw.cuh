// pint/cuda_kernels_dense_mw.cuh -- "the chunk is one big packed_int": whole-chunk SWAR for
// dense bit streams (see cuda_kernels_dense.cuh for the format and the chunk mapping).
//
// A chunk of N = S / gcd(S, 32) 32-bit words holds G = 32 / gcd(S, 32) whole elements, i.e. it
// IS a packed_int whose fields are the layout's fields repeated G times over 32*N bits -- only
// wider than any Integer.  mw<N> below is that N-word unsigned integer with the operators the
// SWAR formulas use (+ - & | ^ ~ >>), and make_chunk_masks<N>() is the mask generator of
// cuda_layout.hpp run over the whole chunk (fields that straddle a word are simply fields whose
// bits live in two words).  With those two pieces apply_op<Op>() of cuda_kernels.cuh computes a
// whole chunk as it stands: the carry of the masked add cannot leave a field (M1 clears every
// field's MSB), so the only carries that cross a word are those inside a straddling field, and
// operator+ propagates exactly those; flag moves are multi-word shifts.  No per-element
// extract / deposit at all: 12-bit samples cost ~6 ALU ops per element instead of ~15 with the
// sliding windows, which turns the narrow layouts from ALU-bound into HBM-bound.
// Used for chunks of N <= kMaxMwWords words (the chunk must live in registers), with compiled-in
// masks (static_masks<chunk_layout<Bits...>>) or run-time masks (runtime_chunk_masks<N>).
// Included by cuda_kernels_dense.cuh (after its helpers, before its kernels).
#pragma once
#ifndef PINT_CUDA_KERNELS_DENSE_MW_CUH_INCLUDED
#define PINT_CUDA_KERNELS_DENSE_MW_CUH_INCLUDED

namespace pint {
namespace cuda {
namespace kernels {
namespace dense {

constexpr uint32_t kMaxMwWords = 19;

// N-word little-endian unsigned integer; every loop has constant bounds so that w[] stays in registers.
template <int N>
struct mw {
    uint32_t w[N];

    PINT_HD constexpr mw() : w{} {}

    PINT_HD friend constexpr mw operator&(const mw& x, const mw& y) { mw r; for (int j = 0; j < N; ++j) r.w[j] = x.w[j] & y.w[j]; return r; }
    PINT_HD friend constexpr mw operator|(const mw& x, const mw& y) { mw r; for (int j = 0; j < N; ++j) r.w[j] = x.w[j] | y.w[j]; return r; }
    PINT_HD friend constexpr mw operator^(const mw& x, const mw& y) { mw r; for (int j = 0; j < N; ++j) r.w[j] = x.w[j] ^ y.w[j]; return r; }
    PINT_HD friend constexpr mw operator~(const mw& x) { mw r; for (int j = 0; j < N; ++j) r.w[j] = ~x.w[j]; return r; }
    PINT_HD constexpr mw& operator|=(const mw& y) { for (int j = 0; j < N; ++j) w[j] |= y.w[j]; return *this; }
    PINT_HD constexpr mw& operator&=(const mw& y) { for (int j = 0; j < N; ++j) w[j] &= y.w[j]; return *this; }

    // carry chain over the words (IADD3 / IADD3.X in SASS)
    PINT_HD friend constexpr mw operator+(const mw& x, const mw& y)
    {
        mw r;
        uint32_t c = 0;
        for (int j = 0; j < N; ++j) {
            const uint64_t t = uint64_t(x.w[j]) + y.w[j] + c;
            r.w[j] = uint32_t(t);
            c = uint32_t(t >> 32);
        }
        return r;
    }
    PINT_HD friend constexpr mw operator-(const mw& x, const mw& y)
    {
        mw r;
        uint32_t c = 1;  // x + ~y + 1
        for (int j = 0; j < N; ++j) {
            const uint64_t t = uint64_t(x.w[j]) + uint32_t(~y.w[j]) + c;
            r.w[j] = uint32_t(t);
            c = uint32_t(t >> 32);
        }
        return r;
    }
    // logical shift right by s < 32 * N bits (s is a compile-time constant wherever the kernels use it)
    PINT_HD friend constexpr mw operator>>(const mw& x, uint32_t s)
    {
        mw r;
        const int ws = int(s / 32);
        const uint32_t bs = s % 32;
        for (int j = 0; j < N; ++j) {
            const uint32_t lo = j + ws < N ? x.w[j + ws] : 0u;
            const uint32_t hi = j + ws + 1 < N ? x.w[j + ws + 1] : 0u;
            r.w[j] = bs ? ((lo >> bs) | (hi << (32 - bs))) : lo;
        }
        return r;
    }
    // logical shift left by s < 32 * N bits
    PINT_HD friend constexpr mw operator<<(const mw& x, uint32_t s)
    {
        mw r;
        const int ws = int(s / 32);
        const uint32_t bs = s % 32;
        for (int j = 0; j < N; ++j) {
            const uint32_t hi = j - ws >= 0 ? x.w[j - ws] : 0u;
            const uint32_t lo = j - ws - 1 >= 0 ? x.w[j - ws - 1] : 0u;
            r.w[j] = bs ? ((hi << bs) | (lo >> (32 - bs))) : hi;
        }
        return r;
    }
    PINT_HD friend constexpr bool operator==(const mw& x, const mw& y)
    {
        for (int j = 0; j < N; ++j)
            if (x.w[j] != y.w[j]) return false;
        return true;
    }
    PINT_HD constexpr void set_bit(uint32_t i) { w[i / 32] |= 1u << (i % 32); }
};

// word_masks (cuda_layout.hpp) over a whole chunk: same members, same meaning, mw<N> instead of a Word.
template <int N>
struct chunk_masks {
    mw<N> H, L, A, M1, HL, LM1;
    uint32_t nsh;
    uint32_t per_shift_mask;
    uint32_t sh[kMaxUnique];
    mw<N> mk[kMaxUnique];
    uint32_t nsmear;
    uint32_t use_smear;
    mw<N> keep[6];
    uint32_t total_bits;
    uint32_t max_bits;
};

// The mask generator of cuda_layout.hpp (make_word_masks) over the 32*N bits of a chunk: the
// layout's fields repeated for every element of the chunk, wherever the word boundaries fall.
template <int N>
constexpr chunk_masks<N> make_chunk_masks(const layout_desc& d)
{
    chunk_masks<N> m{};
    constexpr uint32_t NB = 32u * N;
    uint32_t S = 0, maxb = 0;
    for (uint32_t i = 0; i < d.nfields; ++i) {
        S += d.bits[i];
        if (d.bits[i] > maxb) maxb = d.bits[i];
    }
    m.total_bits = S;
    m.max_bits = maxb;
    const uint32_t G = NB / S;  // whole elements per chunk (NB is a multiple of S by construction)

    uint32_t nuniq = 0;
    uint32_t uw[kMaxFields] = {};
    mw<N> ulo[kMaxUnique] = {};
    uint16_t fid[NB] = {};  // field id of every bit of the chunk
    uint32_t id = 0;
    for (uint32_t e = 0; e < G; ++e) {
        uint32_t off = e * S;
        for (uint32_t i = 0; i < d.nfields; ++i) {
            const uint32_t w = d.bits[i];
            m.L.set_bit(off);
            m.H.set_bit(off + w - 1);
            uint32_t k = 0;
            while (k < nuniq && uw[k] != w) ++k;
            if (k == nuniq) uw[nuniq++] = w;
            if (k < (uint32_t)kMaxUnique) ulo[k].set_bit(off);
            ++id;
            for (uint32_t b = 0; b < w; ++b) fid[off + b] = uint16_t(id);
            off += w;
        }
    }
    for (uint32_t i = 0; i < G * S; ++i) m.A.set_bit(i);
    m.M1 = m.A & ~m.H;
    m.HL = m.H & m.L;
    m.LM1 = m.L & m.M1;

    // flag-move plan, decided on the whole chunk exactly as make_word_masks decides on a Word
    m.nsh = nuniq;
    for (uint32_t k = 0; k < nuniq && k < (uint32_t)kMaxUnique; ++k) m.sh[k] = uw[k] - 1;
    bool alias_free = true;
    for (uint32_t k = 0; k < nuniq && k < (uint32_t)kMaxUnique; ++k)
        if (!(((m.H >> (uw[k] - 1)) & m.L) == ulo[k])) alias_free = false;
    m.per_shift_mask = (nuniq == 1 || alias_free) ? 0 : 1;
    for (uint32_t k = 0; k < nuniq && k < (uint32_t)kMaxUnique; ++k) m.mk[k] = m.per_shift_mask ? ulo[k] : m.L;

    uint32_t ns = 0;
    while ((uint32_t(1) << ns) < maxb) ++ns;
    m.nsmear = ns;
    for (uint32_t j = 0; j < ns && j < 6; ++j) {
        const uint32_t s = uint32_t(1) << j;
        for (uint32_t i = 0; i + s < NB; ++i)
            if (fid[i] && fid[i] == fid[i + s]) m.keep[j].set_bit(i);
    }
    const uint32_t cost_shift = m.nsh * (1 + m.per_shift_mask) + (m.nsh - 1) / 2 + 2 + ((m.nsh > 1 && !m.per_shift_mask) ? 1 : 0);
    // a multi-word subtraction (the shift plan's c - lsb(c)) costs a carry chain: count it as 2 more
    m.use_smear = (m.nsh > 1 && 2 * ns < cost_shift + 2) ? 1 : 0;
    return m;
}

// packed_int<_, Bits...> seen as the layout of one chunk of a dense stream: what static_masks<>
// (cuda_kernels.cuh) needs to compute the chunk with apply_op.
template <std::size_t... Bits>
struct chunk_layout {
    static constexpr uint32_t kS = uint32_t((Bits + ... + 0));
    static constexpr int kN = int(kS / gcd_u32(kS, 32));
    using word = mw<kN>;
    static constexpr layout_desc desc{8u, (uint32_t)sizeof...(Bits), {uint8_t(Bits)...}};
    static constexpr chunk_masks<kN> masks = make_chunk_masks<kN>(desc);
};

// <12>: field MSBs at bits 11, 23, 35, ... -- three words, eight fields, two of them straddling
static_assert(chunk_layout<12>::kN == 3 && chunk_layout<12>::masks.H.w[0] == 0x00800800u &&
                  chunk_layout<12>::masks.H.w[1] == 0x08008008u && chunk_layout<12>::masks.H.w[2] == 0x80080080u &&
                  chunk_layout<12>::masks.L.w[1] == 0x10010010u,
              "chunk mask generator");

// Run-time masks over a chunk: runtime_masks (cuda_kernels.cuh) with mw<N> words -- a fixed-structure
// copy of the plan in the kernel parameter, so every constant is a constant-bank operand.  Per
// 32-bit word of stream this costs what the padded run-time kernels cost per word, and moves the
// same 12 bytes per word -- so any layout with N <= kMaxMwWords runs at the padded kernels' speed
// without being compiled in (the element-by-element windows were ALU-bound at 2.8-4.8 TB/s).
// Steps: smear steps the fill runs (6 = any layout; the launcher picks 2 / 4 when the widest field has at
// most 4 / 16 bits -- a step on mw<N> is N funnel shifts and 2 N logic ops, so skipped steps count here).
template <int N, int Steps = 6>
struct runtime_chunk_masks {
    using word = mw<N>;
    static constexpr bool heavy_fill = true;
    mw<N> mH, mL, mA, mM1, mHL, mLM1;
    mw<N> keep[6];
    uint32_t shift0;  // w - 1 when all fields have the same width
    uint32_t nsmear;  // 0: single-width layout (shift plan), else number of smear steps

    runtime_chunk_masks() = default;
    explicit runtime_chunk_masks(const chunk_masks<N>& m)
        : mH(m.H), mL(m.L), mA(m.A), mM1(m.M1), mHL(m.HL), mLM1(m.LM1),
          keep{m.keep[0], m.keep[1], m.keep[2], m.keep[3], m.keep[4], m.keep[5]},
          shift0(m.sh[0]), nsmear(m.nsh == 1 ? 0 : m.nsmear) {}

    __device__ __forceinline__ word H() const { return mH; }
    __device__ __forceinline__ word L() const { return mL; }
    __device__ __forceinline__ word A() const { return mA; }
    __device__ __forceinline__ word M1() const { return mM1; }
    __device__ __forceinline__ word HL() const { return mHL; }
    __device__ __forceinline__ word LM1() const { return mLM1; }
    __device__ __forceinline__ word fill(const word& c) const
    {
        if (nsmear == 0) return c | (c - (c >> shift0));
        word x = c;
        if constexpr (Steps >= 1) x |= (x >> 1) & keep[0];
        if constexpr (Steps >= 2) x |= (x >> 2) & keep[1];
        if constexpr (Steps >= 3) x |= (x >> 4) & keep[2];
        if constexpr (Steps >= 4) x |= (x >> 8) & keep[3];
        if constexpr (Steps >= 5) x |= (x >> 16) & keep[4];
        if constexpr (Steps >= 6)
            if (nsmear > 5) x |= (x >> 32) & keep[5];  // fields wider than 32 bits only
        return x;
    }
    __device__ __forceinline__ word below(const word& c) const
    {
        if (nsmear == 0) return c - (c >> shift0);
        return fill(c) ^ c;
    }
};

// geometry of the whole-chunk kernels: the tile is handled in 32-bit words, N per chunk
template <int N>
struct mw_geom {
    static constexpr bool is_static = true;
    static constexpr bool whole_chunk = true;
    static constexpr bool wide64 = false;
    using unit = uint32_t;
    static constexpr uint32_t kCw = N;
    uint32_t s;  // sum(Bits)
    uint32_t chunks_per_thread;
    __host__ __device__ uint32_t S() const { return s; }
    __host__ __device__ static constexpr uint32_t Cw() { return kCw; }
};

template <class Fn, int N>
__device__ __forceinline__ void process_chunk_mw(uint32_t* A, const uint32_t* B, const Fn& fn)
{
    mw<N> a, b;
#pragma unroll
    for (int j = 0; j < N; ++j) { a.w[j] = A[j]; b.w[j] = B[j]; }
    const mw<N> r = fn(a, b);
#pragma unroll
    for (int j = 0; j < N; ++j) A[j] = r.w[j];
}

// shift_left / shift_right_unsigned with a uniform amount (pint.hpp:1006-1029) on a whole chunk:
// "mask, shift" with the per-field keep mask replicated over the chunk's elements (the shift
// cannot move a kept bit out of its field, so multi-word shifts are exact).
template <int N>
struct mw_unary_fn {
    using word = mw<N>;
    mw<N> pre;
    uint32_t amount;  // < max(Bits)
    uint32_t left;
    __device__ __forceinline__ mw<N> operator()(const mw<N>& v) const
    {
        const mw<N> t = v & pre;
        return left ? (t << amount) : (t >> amount);
    }
};

// element mask (S bits) replicated over the G elements of a chunk
template <int N>
inline mw<N> replicate_element_mask(uint64_t elem_mask, uint32_t S)
{
    mw<N> r;
    const uint32_t G = 32u * N / S;
    for (uint32_t e = 0; e < G; ++e)
        for (uint32_t b = 0; b < S; ++b)
            if ((elem_mask >> b) & 1) r.set_bit(e * S + b);
    return r;
}

template <class Fn, int N>
__device__ __forceinline__ void process_chunk_mw_unary(uint32_t* A, const Fn& fn)
{
    mw<N> a;
#pragma unroll
    for (int j = 0; j < N; ++j) a.w[j] = A[j];
    const mw<N> r = fn(a);
#pragma unroll
    for (int j = 0; j < N; ++j) A[j] = r.w[j];
}

template <class PFn, int N>
__device__ __forceinline__ void process_chunk_mw_pair(uint32_t* A, uint32_t* B, const PFn& fn)
{
    mw<N> a, b;
#pragma unroll
    for (int j = 0; j < N; ++j) { a.w[j] = A[j]; b.w[j] = B[j]; }
    const mw<N> r0 = fn.first(a, b), r1 = fn.second(a, b);
#pragma unroll
    for (int j = 0; j < N; ++j) { A[j] = r0.w[j]; B[j] = r1.w[j]; }
}

}  // namespace dense
}  // namespace kernels
}  // namespace cuda
}  // namespace pint

#endif  // PINT_CUDA_KERNELS_DENSE_MW_CUH_INCLUDED
