// pint/cuda_kernels_planes.cuh -- per-field kernels: unpack (get / get_signed),
// repack (field constructor / make_truncate) and the fused
// unpack -> arithmetic -> repack pipeline.  Included by cuda_kernels.cuh.
//
// Bytes per element (W = sizeof(Integer), F planes of P bytes):
//   unpack  W + P*F      repack  P*F + W      fused  3W (planes stay in registers)
#pragma once
#ifndef PINT_CUDA_KERNELS_PLANES_CUH_INCLUDED
#define PINT_CUDA_KERNELS_PLANES_CUH_INCLUDED

namespace pint {
namespace cuda {
namespace kernels {

// ---------------------------------------------------------------------------
// Field iteration policies: compile-time (offsets and widths are immediates,
// the loop is fully unrolled) and run-time (table in the constant bank).

template <class SL>
struct static_fields {
    using word = typename SL::word;
    template <class F, std::size_t... Is>
    __device__ __forceinline__ static void each_impl(F&& f, std::index_sequence<Is...>)
    {
        (f(uint32_t(Is), uint32_t(SL::fields.off[Is]), uint32_t(SL::fields.bits[Is])), ...);
    }
    // fields of one element: f(index, offset, width)
    template <class F>
    __device__ __forceinline__ static void each_elem_field(F&& f)
    {
        each_impl(static_cast<F&&>(f), std::make_index_sequence<SL::fields.nfields>());
    }
    // fields of one Word (element fields replicated)
    template <class F>
    __device__ __forceinline__ static void each_word_field(F&& f)
    {
        each_impl(static_cast<F&&>(f), std::make_index_sequence<SL::fields.nword_fields>());
    }
    __device__ __forceinline__ static constexpr word A() { return SL::masks.A; }
};

template <class Word>
struct runtime_fields {
    using word = Word;
    field_table t;
    Word a_mask;
    template <class F>
    __device__ __forceinline__ void each_elem_field(F&& f) const
    {
        for (uint32_t i = 0; i < t.nfields; ++i) f(i, uint32_t(t.off[i]), uint32_t(t.bits[i]));
    }
    template <class F>
    __device__ __forceinline__ void each_word_field(F&& f) const
    {
        for (uint32_t i = 0; i < t.nword_fields; ++i) f(i, uint32_t(t.off[i]), uint32_t(t.bits[i]));
    }
    __device__ __forceinline__ Word A() const { return a_mask; }
};

struct plane_ptrs {
    void* p[kMaxFields];  // device pointers, one per field; null = field not wanted (unpack only)
};

// ---------------------------------------------------------------------------
// Field extraction on a register holding one element zero-extended to a Word.

// get<I>, pint.hpp:799-807
template <class Word>
__device__ __forceinline__ Word field_get(Word e, uint32_t off, uint32_t w)
{
    constexpr uint32_t WB = 8 * sizeof(Word);
    return (e << (WB - off - w)) >> (WB - w);
}

// get_signed<I>, pint.hpp:809-822: field MSB to the register MSB, arithmetic
// shift back.  On a 32-bit register holding a narrower Integer this yields the
// reference's make_signed<Integer> value sign-extended to 32 bits.
template <class Word>
__device__ __forceinline__ typename std::make_signed<Word>::type field_get_signed(Word e, uint32_t off, uint32_t w)
{
    using S = typename std::make_signed<Word>::type;
    constexpr uint32_t WB = 8 * sizeof(Word);
    return S(e << (WB - off - w)) >> (WB - w);
}

// ---------------------------------------------------------------------------
// Four consecutive elements per thread ("quad"): plane accesses are 16-byte
// vectors (int32 planes), 2 x 16 bytes (int64 planes) or one 8- / 4-byte word
// (int16 / int8 planes); a warp covers 128 ... 1024 contiguous bytes of every plane.

template <class Elem, class Word>
__device__ __forceinline__ void load_quad(const Elem* p, Word (&e)[4])
{
    if constexpr (sizeof(Elem) == 8) {
        const uint4 v0 = ld16<CACHE_STREAMING>(reinterpret_cast<const uint4*>(p));
        const uint4 v1 = ld16<CACHE_STREAMING>(reinterpret_cast<const uint4*>(p) + 1);
        e[0] = (uint64_t(v0.y) << 32) | v0.x; e[1] = (uint64_t(v0.w) << 32) | v0.z;
        e[2] = (uint64_t(v1.y) << 32) | v1.x; e[3] = (uint64_t(v1.w) << 32) | v1.z;
    } else if constexpr (sizeof(Elem) == 4) {
        const uint4 v = ld16<CACHE_STREAMING>(reinterpret_cast<const uint4*>(p));
        e[0] = v.x; e[1] = v.y; e[2] = v.z; e[3] = v.w;
    } else if constexpr (sizeof(Elem) == 2) {
        const uint2 v = __ldcs(reinterpret_cast<const uint2*>(p));
        e[0] = v.x & 0xFFFFu; e[1] = v.x >> 16; e[2] = v.y & 0xFFFFu; e[3] = v.y >> 16;
    } else {
        const uint32_t v = __ldcs(reinterpret_cast<const uint32_t*>(p));
        e[0] = v & 0xFFu; e[1] = (v >> 8) & 0xFFu; e[2] = (v >> 16) & 0xFFu; e[3] = v >> 24;
    }
}

template <class Elem, class Word>
__device__ __forceinline__ void store_quad(Elem* p, const Word (&e)[4])
{
    if constexpr (sizeof(Elem) == 8) {
        // (a 32-bit Word under a 64-bit Integer happens in the dense kernels: sum(Bits) <= 32)
        const uint64_t w0 = e[0], w1 = e[1], w2 = e[2], w3 = e[3];
        st16<CACHE_STREAMING>(reinterpret_cast<uint4*>(p),
                              make_uint4(uint32_t(w0), uint32_t(w0 >> 32), uint32_t(w1), uint32_t(w1 >> 32)));
        st16<CACHE_STREAMING>(reinterpret_cast<uint4*>(p) + 1,
                              make_uint4(uint32_t(w2), uint32_t(w2 >> 32), uint32_t(w3), uint32_t(w3 >> 32)));
    } else if constexpr (sizeof(Elem) == 4) {
        st16<CACHE_STREAMING>(reinterpret_cast<uint4*>(p), make_uint4(e[0], e[1], e[2], e[3]));
    } else if constexpr (sizeof(Elem) == 2) {
        __stcs(reinterpret_cast<uint2*>(p),
               make_uint2((e[0] & 0xFFFFu) | (e[1] << 16), (e[2] & 0xFFFFu) | (e[3] << 16)));
    } else {
        __stcs(reinterpret_cast<uint32_t*>(p),
               (e[0] & 0xFFu) | ((e[1] & 0xFFu) << 8) | ((e[2] & 0xFFu) << 16) | (e[3] << 24));
    }
}

// Plane element type: int32_t/uint32_t/int64_t/uint64_t handled as 4- or 8-byte words.
template <int PlaneBytes, int St = CACHE_STREAMING>
__device__ __forceinline__ void store_plane_quad(void* plane, size_t q, const uint64_t (&v)[4])
{
    constexpr int CACHE_ST = St;
    if constexpr (PlaneBytes == 1) {
        // narrow planes (the plane type is the Integer itself for 8 / 16-bit packs): one 4- or 8-byte store
        __stcs(reinterpret_cast<uint32_t*>(plane) + q, (uint32_t(v[0]) & 0xFFu) | ((uint32_t(v[1]) & 0xFFu) << 8) |
                                                          ((uint32_t(v[2]) & 0xFFu) << 16) | (uint32_t(v[3]) << 24));
    } else if constexpr (PlaneBytes == 2) {
        __stcs(reinterpret_cast<uint2*>(plane) + q, make_uint2((uint32_t(v[0]) & 0xFFFFu) | (uint32_t(v[1]) << 16),
                                                               (uint32_t(v[2]) & 0xFFFFu) | (uint32_t(v[3]) << 16)));
    } else if constexpr (PlaneBytes == 4) {
        st16<CACHE_ST>(reinterpret_cast<uint4*>(plane) + q,
                              make_uint4(uint32_t(v[0]), uint32_t(v[1]), uint32_t(v[2]), uint32_t(v[3])));
    } else {
        uint4* d = reinterpret_cast<uint4*>(plane) + 2 * q;
        st16<CACHE_ST>(d, make_uint4(uint32_t(v[0]), uint32_t(v[0] >> 32), uint32_t(v[1]), uint32_t(v[1] >> 32)));
        st16<CACHE_ST>(d + 1, make_uint4(uint32_t(v[2]), uint32_t(v[2] >> 32), uint32_t(v[3]), uint32_t(v[3] >> 32)));
    }
}

template <int PlaneBytes>
__device__ __forceinline__ void load_plane_quad(const void* plane, size_t q, uint64_t (&v)[4])
{
    if constexpr (PlaneBytes == 1) {
        const uint32_t x = __ldcs(reinterpret_cast<const uint32_t*>(plane) + q);
        v[0] = uint64_t(int64_t(int8_t(x))); v[1] = uint64_t(int64_t(int8_t(x >> 8)));
        v[2] = uint64_t(int64_t(int8_t(x >> 16))); v[3] = uint64_t(int64_t(int8_t(x >> 24)));
    } else if constexpr (PlaneBytes == 2) {
        const uint2 x = __ldcs(reinterpret_cast<const uint2*>(plane) + q);
        v[0] = uint64_t(int64_t(int16_t(x.x))); v[1] = uint64_t(int64_t(int16_t(x.x >> 16)));
        v[2] = uint64_t(int64_t(int16_t(x.y))); v[3] = uint64_t(int64_t(int16_t(x.y >> 16)));
    } else if constexpr (PlaneBytes == 4) {
        const uint4 x = ld16<CACHE_STREAMING>(reinterpret_cast<const uint4*>(plane) + q);
        // int32 plane values are converted to the Integer as the field
        // constructor's value_type parameters are (pint.hpp:771-772): sign-extend
        v[0] = uint64_t(int64_t(int32_t(x.x))); v[1] = uint64_t(int64_t(int32_t(x.y)));
        v[2] = uint64_t(int64_t(int32_t(x.z))); v[3] = uint64_t(int64_t(int32_t(x.w)));
    } else {
        const uint4* s = reinterpret_cast<const uint4*>(plane) + 2 * q;
        const uint4 x = ld16<CACHE_STREAMING>(s), y = ld16<CACHE_STREAMING>(s + 1);
        v[0] = (uint64_t(x.y) << 32) | x.x; v[1] = (uint64_t(x.w) << 32) | x.z;
        v[2] = (uint64_t(y.y) << 32) | y.x; v[3] = (uint64_t(y.w) << 32) | y.z;
    }
}

// packed -> planes.  get<I> (pint.hpp:799-807) or get_signed<I> (:809-822) for
// every wanted field of Q x 4 consecutive elements per thread (Q quads, strided by
// the block so that every warp access stays contiguous).
template <class Elem, int PlaneBytes, bool Signed, class FP, int Q = 1, int St = CACHE_STREAMING>
__global__ void __launch_bounds__(256)
unpack_quad_kernel(const Elem* packed, const __grid_constant__ plane_ptrs planes, size_t nquads,
                   const __grid_constant__ FP fp)
{
    pdl_prologue();
    using W = typename FP::word;
    const size_t tile = size_t(Q) * blockDim.x;
    for (size_t t = blockIdx.x; t * tile < nquads; t += gridDim.x) {
        W e[Q][4];
        size_t q[Q];
#pragma unroll
        for (int k = 0; k < Q; ++k) {
            q[k] = t * tile + size_t(k) * blockDim.x + threadIdx.x;
            if (q[k] < nquads) load_quad<Elem, W>(packed + 4 * q[k], e[k]);
        }
        fp.each_elem_field([&](uint32_t i, uint32_t off, uint32_t w) {
            void* dst = planes.p[i];
            if (!dst) return;
#pragma unroll
            for (int k = 0; k < Q; ++k) {
                if (q[k] >= nquads) continue;
                uint64_t v[4];
#pragma unroll
                for (int j = 0; j < 4; ++j)
                    v[j] = Signed ? uint64_t(int64_t(field_get_signed<W>(e[k][j], off, w))) : uint64_t(field_get<W>(e[k][j], off, w));
                store_plane_quad<PlaneBytes, St>(dst, q[k], v);
            }
        });
    }
}

template <class Elem, int PlaneBytes, bool Signed, class FP>
__global__ void __launch_bounds__(256)
unpack_elem_kernel(const Elem* packed, const __grid_constant__ plane_ptrs planes, size_t first, size_t n,
                   const __grid_constant__ FP fp)
{
    using W = typename FP::word;
    for (size_t i = first + size_t(blockIdx.x) * blockDim.x + threadIdx.x; i < first + n;
         i += size_t(gridDim.x) * blockDim.x) {
        const W e = W(packed[i]);
        fp.each_elem_field([&](uint32_t f, uint32_t off, uint32_t w) {
            void* dst = planes.p[f];
            if (!dst) return;
            const uint64_t v = Signed ? uint64_t(int64_t(field_get_signed<W>(e, off, w))) : uint64_t(field_get<W>(e, off, w));
            if constexpr (PlaneBytes == 1) static_cast<uint8_t*>(dst)[i] = uint8_t(v);
            else if constexpr (PlaneBytes == 2) static_cast<uint16_t*>(dst)[i] = uint16_t(v);
            else if constexpr (PlaneBytes == 4) static_cast<uint32_t*>(dst)[i] = uint32_t(v);
            else static_cast<uint64_t*>(dst)[i] = v;
        });
    }
}

// planes -> packed.  The field constructor (pint.hpp:770-774) = make_truncate
// (:592-601) = OR_i ((v_i & ones(w_i)) << off_i) (:390-407).
template <class Elem, int PlaneBytes, class FP>
__global__ void __launch_bounds__(256)
repack_quad_kernel(const __grid_constant__ plane_ptrs planes, Elem* packed, size_t nquads,
                   const __grid_constant__ FP fp)
{
    pdl_prologue();
    using W = typename FP::word;
    constexpr uint32_t WB = 8 * sizeof(W);
    for (size_t q = size_t(blockIdx.x) * blockDim.x + threadIdx.x; q < nquads; q += size_t(gridDim.x) * blockDim.x) {
        W e[4] = {0, 0, 0, 0};
        fp.each_elem_field([&](uint32_t i, uint32_t off, uint32_t w) {
            uint64_t v[4];
            load_plane_quad<PlaneBytes>(planes.p[i], q, v);
#pragma unroll
            for (int j = 0; j < 4; ++j) e[j] |= (W(v[j]) << (WB - w)) >> (WB - w - off);
        });
        store_quad<Elem, W>(packed + 4 * q, e);
    }
}

// ---------------------------------------------------------------------------
// Narrow planes (1 or 2 bytes per value): a quad would be a 4- or 8-byte plane access and 20-32 bytes
// of work per thread (issue-bound: RGB565 -> three uint8 planes ran at 4.9 TB/s).  Here a thread takes
// G = 16 / PlaneBytes consecutive elements, so every plane access is a 16-byte vector and the packed
// side one or more 256-bit accesses (32 bytes per lane: whole sectors at any lane stride).
template <class Elem, int G>
struct group_words {
    static constexpr int BYTES = G * int(sizeof(Elem));
    static constexpr int NW = BYTES / 4;  // 32-bit words per group (>= 2)
};

template <class Elem, int G>
__device__ __forceinline__ void load_group_words(const Elem* p, uint32_t (&w)[group_words<Elem, G>::NW])
{
    constexpr int BYTES = group_words<Elem, G>::BYTES;
    if constexpr (BYTES >= 32) {
#pragma unroll
        for (int k = 0; k < BYTES / 32; ++k) {
            const vec32 x = ld32(reinterpret_cast<const vec32*>(p) + k);
            w[8 * k + 0] = x.lo.x; w[8 * k + 1] = x.lo.y; w[8 * k + 2] = x.lo.z; w[8 * k + 3] = x.lo.w;
            w[8 * k + 4] = x.hi.x; w[8 * k + 5] = x.hi.y; w[8 * k + 6] = x.hi.z; w[8 * k + 7] = x.hi.w;
        }
    } else if constexpr (BYTES == 16) {
        const uint4 x = ld16<CACHE_STREAMING>(reinterpret_cast<const uint4*>(p));
        w[0] = x.x; w[1] = x.y; w[2] = x.z; w[3] = x.w;
    } else {
        const uint2 x = __ldcs(reinterpret_cast<const uint2*>(p));
        w[0] = x.x; w[1] = x.y;
    }
}

template <class Elem, int G>
__device__ __forceinline__ void store_group_words(Elem* p, const uint32_t (&w)[group_words<Elem, G>::NW])
{
    constexpr int BYTES = group_words<Elem, G>::BYTES;
    if constexpr (BYTES >= 32) {
#pragma unroll
        for (int k = 0; k < BYTES / 32; ++k) {
            vec32 x;
            x.lo = make_uint4(w[8 * k + 0], w[8 * k + 1], w[8 * k + 2], w[8 * k + 3]);
            x.hi = make_uint4(w[8 * k + 4], w[8 * k + 5], w[8 * k + 6], w[8 * k + 7]);
            st32(reinterpret_cast<vec32*>(p) + k, x);
        }
    } else if constexpr (BYTES == 16) {
        st16<CACHE_STREAMING>(reinterpret_cast<uint4*>(p), make_uint4(w[0], w[1], w[2], w[3]));
    } else {
        __stcs(reinterpret_cast<uint2*>(p), make_uint2(w[0], w[1]));
    }
}

// element j of a group held as 32-bit words, zero-extended to the processing Word
template <class Elem, class Word, int NW>
__device__ __forceinline__ Word group_elem(const uint32_t (&w)[NW], int j)
{
    if constexpr (sizeof(Elem) == 8) return Word((uint64_t(w[2 * j + 1]) << 32) | w[2 * j]);
    else if constexpr (sizeof(Elem) == 4) return Word(w[j]);
    else if constexpr (sizeof(Elem) == 2) return Word((w[j >> 1] >> (16 * (j & 1))) & 0xFFFFu);
    else return Word((w[j >> 2] >> (8 * (j & 3))) & 0xFFu);
}

template <class Elem, int PlaneBytes, bool Signed, class FP>
__global__ void __launch_bounds__(256)
unpack_group_kernel(const Elem* packed, const __grid_constant__ plane_ptrs planes, size_t ngroups,
                    const __grid_constant__ FP fp)
{
    pdl_prologue();
    using W = typename FP::word;
    constexpr int G = 16 / PlaneBytes, NW = group_words<Elem, G>::NW;
    for (size_t g = size_t(blockIdx.x) * blockDim.x + threadIdx.x; g < ngroups; g += size_t(gridDim.x) * blockDim.x) {
        uint32_t w[NW];
        load_group_words<Elem, G>(packed + size_t(G) * g, w);
        fp.each_elem_field([&](uint32_t i, uint32_t off, uint32_t wd) {
            void* dst = planes.p[i];
            if (!dst) return;
            uint32_t o[4] = {0, 0, 0, 0};
#pragma unroll
            for (int j = 0; j < G; ++j) {
                const W e = group_elem<Elem, W, NW>(w, j);
                const uint32_t v = Signed ? uint32_t(field_get_signed<W>(e, off, wd)) : uint32_t(field_get<W>(e, off, wd));
                if constexpr (PlaneBytes == 1) o[j >> 2] |= (v & 0xFFu) << (8 * (j & 3));
                else o[j >> 1] |= (v & 0xFFFFu) << (16 * (j & 1));
            }
            st16<CACHE_STREAMING>(reinterpret_cast<uint4*>(dst) + g, make_uint4(o[0], o[1], o[2], o[3]));
        });
    }
}

// int32 planes, eight consecutive elements per thread: every plane store is one 256-bit vector
// (a warp writes 1 KB of each plane per instruction), the packed side 256-bit loads.
template <class Elem, bool Signed, class FP, int K = 1>
__global__ void __launch_bounds__(256)
unpack_oct_kernel(const Elem* packed, const __grid_constant__ plane_ptrs planes, size_t nocts, const __grid_constant__ FP fp)
{
    pdl_prologue();
    using W = typename FP::word;
    constexpr int G = 8, NW = group_words<Elem, G>::NW;
    // K octs per thread, a block apart (every access of a warp stays contiguous), all loads first
    for (size_t t = blockIdx.x; t * (size_t(K) * blockDim.x) < nocts; t += gridDim.x) {
        uint32_t w[K][NW];
        size_t g[K];
#pragma unroll
        for (int k = 0; k < K; ++k) {
            g[k] = (t * K + k) * blockDim.x + threadIdx.x;
            if (g[k] < nocts) load_group_words<Elem, G>(packed + size_t(G) * g[k], w[k]);
        }
        fp.each_elem_field([&](uint32_t i, uint32_t off, uint32_t wd) {
            void* dst = planes.p[i];
            if (!dst) return;
#pragma unroll
            for (int k = 0; k < K; ++k) {
                if (g[k] >= nocts) continue;
                uint32_t v[G];
#pragma unroll
                for (int j = 0; j < G; ++j) {
                    const W e = group_elem<Elem, W, NW>(w[k], j);
                    v[j] = Signed ? uint32_t(field_get_signed<W>(e, off, wd)) : uint32_t(field_get<W>(e, off, wd));
                }
                vec32 r;
                r.lo = make_uint4(v[0], v[1], v[2], v[3]);
                r.hi = make_uint4(v[4], v[5], v[6], v[7]);
                st32(reinterpret_cast<vec32*>(dst) + g[k], r);
            }
        });
    }
}

template <class Elem, int PlaneBytes, class FP>
__global__ void __launch_bounds__(256)
repack_group_kernel(const __grid_constant__ plane_ptrs planes, Elem* packed, size_t ngroups, const __grid_constant__ FP fp)
{
    pdl_prologue();
    using W = typename FP::word;
    constexpr uint32_t WB = 8 * sizeof(W);
    constexpr int G = 16 / PlaneBytes, NW = group_words<Elem, G>::NW;
    for (size_t g = size_t(blockIdx.x) * blockDim.x + threadIdx.x; g < ngroups; g += size_t(gridDim.x) * blockDim.x) {
        W e[G];
#pragma unroll
        for (int j = 0; j < G; ++j) e[j] = 0;
        fp.each_elem_field([&](uint32_t i, uint32_t off, uint32_t wd) {
            const uint4 x = ld16<CACHE_STREAMING>(reinterpret_cast<const uint4*>(planes.p[i]) + g);
            const uint32_t o[4] = {x.x, x.y, x.z, x.w};
#pragma unroll
            for (int j = 0; j < G; ++j) {
                // plane values are signed integers converted to the Integer (pint.hpp:771-772), then truncated
                const int32_t v = PlaneBytes == 1 ? int32_t(int8_t(o[j >> 2] >> (8 * (j & 3)))) : int32_t(int16_t(o[j >> 1] >> (16 * (j & 1))));
                e[j] |= (W(int64_t(v)) << (WB - wd)) >> (WB - wd - off);
            }
        });
        uint32_t w[NW];
#pragma unroll
        for (int k = 0; k < NW; ++k) w[k] = 0;
#pragma unroll
        for (int j = 0; j < G; ++j) {
            if constexpr (sizeof(Elem) == 8) { w[2 * j] = uint32_t(e[j]); w[2 * j + 1] = uint32_t(uint64_t(e[j]) >> 32); }
            else if constexpr (sizeof(Elem) == 4) w[j] = uint32_t(e[j]);
            else if constexpr (sizeof(Elem) == 2) w[j >> 1] |= (uint32_t(e[j]) & 0xFFFFu) << (16 * (j & 1));
            else w[j >> 2] |= (uint32_t(e[j]) & 0xFFu) << (8 * (j & 3));
        }
        store_group_words<Elem, G>(packed + size_t(G) * g, w);
    }
}

template <class Elem, int PlaneBytes, class FP>
__global__ void __launch_bounds__(256)
repack_elem_kernel(const __grid_constant__ plane_ptrs planes, Elem* packed, size_t first, size_t n,
                   const __grid_constant__ FP fp)
{
    using W = typename FP::word;
    constexpr uint32_t WB = 8 * sizeof(W);
    for (size_t i = first + size_t(blockIdx.x) * blockDim.x + threadIdx.x; i < first + n;
         i += size_t(gridDim.x) * blockDim.x) {
        W e = 0;
        fp.each_elem_field([&](uint32_t f, uint32_t off, uint32_t w) {
            uint64_t v;
            if constexpr (PlaneBytes == 1) v = uint64_t(int64_t(static_cast<const int8_t*>(planes.p[f])[i]));
            else if constexpr (PlaneBytes == 2) v = uint64_t(int64_t(static_cast<const int16_t*>(planes.p[f])[i]));
            else if constexpr (PlaneBytes == 4) v = uint64_t(int64_t(static_cast<const int32_t*>(planes.p[f])[i]));
            else v = static_cast<const uint64_t*>(planes.p[f])[i];
            e |= (W(v) << (WB - w)) >> (WB - w - off);
        });
        packed[i] = Elem(e);
    }
}

}  // namespace kernels
}  // namespace cuda
}  // namespace pint

#include "cuda_kernels_planes_bulk.cuh"
#include "cuda_kernels_fused.cuh"

#endif  // PINT_CUDA_KERNELS_PLANES_CUH_INCLUDED
