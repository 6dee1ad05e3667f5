// pint/cuda_kernels_fused.cuh -- fused unpack -> arithmetic -> repack, and the
// launchers of the per-field kernels.  Included by cuda_kernels_planes.cuh.
#pragma once

namespace pint {
namespace cuda {
namespace kernels {

// One Word in, one Word out; every field is unpacked into a register, the
// arithmetic is done on the unpacked value, and the result is repacked -- the
// planes of the unfused pipeline (unpack kernel, arithmetic kernel per plane,
// repack kernel: 168 bytes/element for the 6-field u64 layout) never exist in
// memory, so the traffic is the 3W bytes/element of a plain binary op.
//
// Fields are unpacked LEFT-ALIGNED (the first half of get_signed,
// pint.hpp:820: value << (W - hi_order_bit_no)): the field's MSB sits on the
// register's MSB, so the register's own carry flag / sign bit are the field's,
// for every width from 1 to the full Word, and wrap-around needs no masking.
// The result must equal the SWAR op of swar_fn bit for bit; tests check both
// against the oracle and against each other.
template <int Op, class FP>
struct fused_fn {
    using word = typename FP::word;
    FP fp;

    __device__ __forceinline__ word operator()(word a, word b) const
    {
        using W = word;
        using S = typename std::make_signed<W>::type;
        constexpr uint32_t WB = 8 * sizeof(W);
        constexpr W SIGN = W(1) << (WB - 1);
        if constexpr (Op == OP_OR) return a | b;
        if constexpr (Op == OP_AND) return a & b;
        if constexpr (Op == OP_XOR) return a ^ b;
        W r = 0;
        fp.each_word_field([&](uint32_t, uint32_t off, uint32_t w) {
            const uint32_t sl = WB - off - w;
            const W top = ~W(0) << (WB - w);  // the field's bits once left-aligned
            const W xa = (a << sl) & top, xb = (b << sl) & top;
            W t;
            if constexpr (Op == OP_ADD_WRAP) {
                t = xa + xb;
            } else if constexpr (Op == OP_ADD_UNSIGNED_SATURATE) {
                t = xa + xb;
                t = t < xa ? top : t;
            } else if constexpr (Op == OP_ADD_SIGNED_SATURATE) {
                t = xa + xb;
                const W sat = (xa & SIGN) ? SIGN : W(~SIGN & top);
                t = ((xa ^ t) & (xb ^ t) & SIGN) ? sat : t;
            } else if constexpr (Op == OP_SUB_WRAP) {
                t = xa - xb;
            } else if constexpr (Op == OP_SUB_UNSIGNED_SATURATE) {
                t = xa < xb ? W(0) : W(xa - xb);
            } else if constexpr (Op == OP_SUB_SIGNED_SATURATE) {
                t = xa - xb;
                const W sat = (xa & SIGN) ? SIGN : W(~SIGN & top);
                t = ((xa ^ xb) & (xa ^ t) & SIGN) ? sat : t;
            } else if constexpr (Op == OP_MIN_UNSIGNED) {
                t = xa < xb ? xa : xb;
            } else if constexpr (Op == OP_MAX_UNSIGNED) {
                t = xa > xb ? xa : xb;
            } else if constexpr (Op == OP_MIN_SIGNED) {
                t = S(xa) < S(xb) ? xa : xb;
            } else {
                t = S(xa) > S(xb) ? xa : xb;
            }
            r |= (t >> (WB - w)) << off;  // repack, pint.hpp:394
        });
        if constexpr (Op >= OP_MIN_UNSIGNED) r |= b & ~fp.A();  // min/max keep b's unused bits
        return r;
    }
};

// ---------------------------------------------------------------------------
// Launchers for unpack / repack.

inline bool aligned16(const void* p) { return (reinterpret_cast<uintptr_t>(p) & 15u) == 0; }

template <class Elem, int PlaneBytes, bool Signed, class FP>
cudaError_t launch_unpack(const void* packed, const plane_ptrs& planes, uint32_t nfields, size_t n, const FP& fp,
                          cudaStream_t stream, const launch_geom& g)
{
    if (n == 0) return cudaSuccess;
    const Elem* src = static_cast<const Elem*>(packed);
    constexpr size_t quad_align = sizeof(Elem) * 4 < 16 ? sizeof(Elem) * 4 : 16;
    bool ok = (reinterpret_cast<uintptr_t>(packed) % quad_align) == 0;
    for (uint32_t f = 0; f < nfields; ++f)
        if (planes.p[f] && !aligned16(planes.p[f])) ok = false;
    const size_t nquads = ok ? n / 4 : 0;
    if (nquads) {
        const unsigned grid = grid_for((nquads + 255) / 256, g);
        unpack_quad_kernel<Elem, PlaneBytes, Signed, FP><<<grid, 256, 0, stream>>>(src, planes, nquads, fp);
        count_launch(g);
    }
    const size_t rest = n - 4 * nquads;
    if (rest) {
        const unsigned grid = grid_for((rest + 255) / 256, g);
        unpack_elem_kernel<Elem, PlaneBytes, Signed, FP><<<grid, 256, 0, stream>>>(src, planes, 4 * nquads, rest, fp);
        count_launch(g);
    }
    return cudaGetLastError();
}

template <class Elem, int PlaneBytes, class FP>
cudaError_t launch_repack(const plane_ptrs& planes, uint32_t nfields, void* packed, size_t n, const FP& fp,
                          cudaStream_t stream, const launch_geom& g)
{
    if (n == 0) return cudaSuccess;
    Elem* dst = static_cast<Elem*>(packed);
    constexpr size_t quad_align = sizeof(Elem) * 4 < 16 ? sizeof(Elem) * 4 : 16;
    bool ok = (reinterpret_cast<uintptr_t>(packed) % quad_align) == 0;
    for (uint32_t f = 0; f < nfields; ++f)
        if (!aligned16(planes.p[f])) ok = false;
    const size_t nquads = ok ? n / 4 : 0;
    if (nquads) {
        const unsigned grid = grid_for((nquads + 255) / 256, g);
        repack_quad_kernel<Elem, PlaneBytes, FP><<<grid, 256, 0, stream>>>(planes, dst, nquads, fp);
        count_launch(g);
    }
    const size_t rest = n - 4 * nquads;
    if (rest) {
        const unsigned grid = grid_for((rest + 255) / 256, g);
        repack_elem_kernel<Elem, PlaneBytes, FP><<<grid, 256, 0, stream>>>(planes, dst, 4 * nquads, rest, fp);
        count_launch(g);
    }
    return cudaGetLastError();
}

}  // namespace kernels
}  // namespace cuda
}  // namespace pint
