// pint/cuda_kernels_fused.cuh -- fused unpack -> arithmetic -> repack, and the
// launchers of the per-field kernels.  Included by cuda_kernels_planes.cuh.
#pragma once
#ifndef PINT_CUDA_KERNELS_FUSED_CUH_INCLUDED
#define PINT_CUDA_KERNELS_FUSED_CUH_INCLUDED

namespace pint {
namespace cuda {
namespace kernels {

// One Word in, one Word out; every field is unpacked into a register, the
// arithmetic is done on the unpacked value, and the result is repacked -- the
// planes of the unfused pipeline (unpack kernel, arithmetic kernel per plane,
// repack kernel: 168 bytes/element for the 6-field u64 layout) never exist in
// memory, so the traffic is the 3W bytes/element of a plain binary op.
// The result must equal the SWAR op of swar_fn bit for bit; tests check both
// against the oracle and against each other.
//
// Two unpacked forms, chosen per field (at compile time for static layouts, by a
// warp-uniform branch for runtime ones):
//  * fields of up to 31 bits are unpacked RIGHT-ALIGNED into a 32-bit register
//    exactly as get (pint.hpp:805-806) / get_signed (pint.hpp:819-821) do, the
//    arithmetic is plain int32/uint32 with min/max for the saturation (the
//    sum of two w-bit values fits when w <= 31), and the result is truncated and
//    shifted back (make_truncated_int, pint.hpp:394).  A field of a 64-bit
//    element that lies in one 32-bit half is handled in that half; one that
//    straddles the halves is fetched with a funnel shift.
//  * wider fields (32..64 bits) are unpacked LEFT-ALIGNED (the first half of
//    get_signed, value << (W - hi_order_bit_no)): the register's own carry and
//    sign bits are then the field's, for every width up to the full Word.

// right-aligned 32-bit field arithmetic: xa/xb hold the field at bit offset o
// (o + w <= 32, 1 <= w <= 31, bits above the field may be anything); returns the
// result in the low w bits, higher bits unspecified (the caller masks on insert)
template <int Op>
__device__ __forceinline__ uint32_t field_op32(uint32_t xa, uint32_t xb, uint32_t o, uint32_t w)
{
    constexpr bool is_signed = (Op == OP_ADD_SIGNED_SATURATE || Op == OP_SUB_SIGNED_SATURATE ||
                                Op == OP_MIN_SIGNED || Op == OP_MAX_SIGNED);
    const uint32_t fm = 0xFFFFFFFFu >> (32 - w);
    if constexpr (is_signed) {
        const int sa = int(xa << (32 - o - w)) >> (32 - w);  // get_signed, pint.hpp:819-821
        const int sb = int(xb << (32 - o - w)) >> (32 - w);
        const int smax = int(fm >> 1), smin = ~smax;
        int t;
        if constexpr (Op == OP_ADD_SIGNED_SATURATE) t = max(min(sa + sb, smax), smin);
        else if constexpr (Op == OP_SUB_SIGNED_SATURATE) t = max(min(sa - sb, smax), smin);
        else if constexpr (Op == OP_MIN_SIGNED) t = min(sa, sb);
        else t = max(sa, sb);
        return uint32_t(t);
    } else {
        const uint32_t ua = (xa >> o) & fm, ub = (xb >> o) & fm;  // get, pint.hpp:805-806
        if constexpr (Op == OP_ADD_WRAP) return ua + ub;
        else if constexpr (Op == OP_ADD_UNSIGNED_SATURATE) return min(ua + ub, fm);
        else if constexpr (Op == OP_SUB_WRAP) return ua - ub;
        else if constexpr (Op == OP_SUB_UNSIGNED_SATURATE) return ua - min(ua, ub);
        else if constexpr (Op == OP_MIN_UNSIGNED) return min(ua, ub);
        else return max(ua, ub);
    }
}

// left-aligned arithmetic on a full Word (fields of 32..64 bits, or the full 32-bit field)
template <int Op, class W>
__device__ __forceinline__ W field_op_wide(W a, W b, uint32_t off, uint32_t w)
{
    using S = typename std::make_signed<W>::type;
    constexpr uint32_t WB = 8 * sizeof(W);
    constexpr W SIGN = W(1) << (WB - 1);
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
    return (t >> (WB - w)) << off;  // repack, pint.hpp:394
}

template <int Op, class FP>
struct fused_fn {
    using word = typename FP::word;
    static constexpr bool heavy = !(Op == OP_OR || Op == OP_AND || Op == OP_XOR);
    static constexpr bool vec256 = true;  // 32-byte accesses measured faster (profiles/r01z_ab_vec256.log)
    FP fp;

    __device__ __forceinline__ word operator()(word a, word b) const
    {
        using W = word;
        if constexpr (Op == OP_OR) return a | b;
        if constexpr (Op == OP_AND) return a & b;
        if constexpr (Op == OP_XOR) return a ^ b;
        W r = 0;
        if constexpr (sizeof(W) == 4) {
            fp.each_word_field([&](uint32_t, uint32_t off, uint32_t w) {
                if (w <= 31) {
                    const uint32_t fm = 0xFFFFFFFFu >> (32 - w);
                    r |= (field_op32<Op>(a, b, off, w) << off) & (fm << off);
                } else {
                    r |= field_op_wide<Op, W>(a, b, off, w);
                }
            });
        } else {
            const uint32_t alo = uint32_t(a), ahi = uint32_t(a >> 32), blo = uint32_t(b), bhi = uint32_t(b >> 32);
            uint32_t rlo = 0, rhi = 0;
            fp.each_word_field([&](uint32_t, uint32_t off, uint32_t w) {
                if (w > 31) {
                    r |= field_op_wide<Op, W>(a, b, off, w);
                } else {
                    const uint32_t fm = 0xFFFFFFFFu >> (32 - w);
                    if (off + w <= 32) {          // in the low half
                        rlo |= (field_op32<Op>(alo, blo, off, w) << off) & (fm << off);
                    } else if (off >= 32) {       // in the high half
                        const uint32_t o = off - 32;
                        rhi |= (field_op32<Op>(ahi, bhi, o, w) << o) & (fm << o);
                    } else {                      // straddles: funnel-shift it down to bit 0
                        const uint32_t xa = __funnelshift_r(alo, ahi, off), xb = __funnelshift_r(blo, bhi, off);
                        const uint32_t t = field_op32<Op>(xa, xb, 0, w) & fm;
                        rlo |= t << off;
                        rhi |= t >> (32 - off);
                    }
                }
            });
            r |= (uint64_t(rhi) << 32) | rlo;
        }
        if constexpr (Op >= OP_MIN_UNSIGNED) r |= b & ~fp.A();  // min/max keep b's unused bits
        return r;
    }
};

// ---------------------------------------------------------------------------
// Launchers for unpack / repack.

inline bool aligned16(const void* p) { return (reinterpret_cast<uintptr_t>(p) & 15u) == 0; }

#if !defined(__CUDACC_RTC__)  // host-side launchers: not part of a run-time (NVRTC) specialisation
template <class Elem, int PlaneBytes, bool Signed, class FP>
cudaError_t launch_unpack(const void* packed, const plane_ptrs& planes, uint32_t nfields, size_t n, const FP& fp,
                          cudaStream_t stream, const launch_geom& g)
{
    if (n == 0) return cudaSuccess;
    const Elem* src = static_cast<const Elem*>(packed);
    constexpr size_t quad_align = sizeof(Elem) * 4 < 16 ? sizeof(Elem) * 4 : 16;
    constexpr size_t plane_quad_align = PlaneBytes * 4 < 16 ? PlaneBytes * 4 : 16;  // one quad of a plane is one access
    bool ok = (reinterpret_cast<uintptr_t>(packed) % quad_align) == 0;
    for (uint32_t f = 0; f < nfields; ++f)
        if (planes.p[f] && (reinterpret_cast<uintptr_t>(planes.p[f]) % plane_quad_align) != 0) ok = false;
    size_t done = 0;
    if constexpr (PlaneBytes <= 2) {
        // narrow planes: groups of 16 / PlaneBytes elements per thread when the packed array reaches a
        // group boundary (at most 32 bytes) and the planes 16-byte boundaries
        constexpr size_t G = 16 / PlaneBytes, gbytes = G * sizeof(Elem) < 32 ? G * sizeof(Elem) : 32;
        bool gok = (reinterpret_cast<uintptr_t>(packed) % gbytes) == 0;
        for (uint32_t f = 0; f < nfields; ++f)
            if (planes.p[f] && !aligned16(planes.p[f])) gok = false;
        const size_t ngroups = gok ? n / G : 0;
        if (ngroups) {
            if (cudaError_t e = launch_stream_kernel(unpack_group_kernel<Elem, PlaneBytes, Signed, FP>, grid_for((ngroups + 255) / 256, g), 256, stream, g.pdl != 0, src, planes, ngroups, fp)) return e;
            count_launch(g);
            done = ngroups * G;
        }
    }
    if constexpr (PlaneBytes == 4) {
        // plane tiles staged in shared memory and written by bulk (TMA) stores (cuda_kernels_planes_bulk.cuh):
        // +0.8 ... +0.9 % over the oct kernel on the same box (C3 6760 -> 6825, 6670 -> 6725 GB/s,
        // profiles/r12_ab_unpack_bulk.log, r12_tune_unpack_bulk.log; flat in tile and CTA size), the default
        // up to 12 fields (at most 48 KB of staging per CTA); option "unpack_bulk" = 0 keeps the register kernels.
        // Whole tiles only, the remainder goes on below.
        if (g.unpack_bulk > 0 && nfields <= 12 && (reinterpret_cast<uintptr_t>(packed) & 31u) == 0) {
            bool bok = true;
            for (uint32_t f = 0; f < nfields; ++f)
                if (planes.p[f] && !aligned16(planes.p[f])) bok = false;
            const size_t tile = nfields <= 6 ? 2048 : 1024;
            const size_t ntiles = bok ? n / tile : 0;
            if (ntiles) {
                const size_t smem = size_t(nfields) * tile * 4;
                if (tile == 2048)
                    { if (cudaError_t e = launch_stream_kernel_smem(unpack_bulk_kernel<Elem, Signed, FP, 2048, 256>, grid_for(ntiles, g), 256, smem, stream, g.pdl != 0, src, planes, ntiles, nfields, fp)) return e; }
                else
                    { if (cudaError_t e = launch_stream_kernel_smem(unpack_bulk_kernel<Elem, Signed, FP, 1024, 256>, grid_for(ntiles, g), 256, smem, stream, g.pdl != 0, src, planes, ntiles, nfields, fp)) return e; }
                count_launch(g);
                done = ntiles * tile;
                if (done == n) return cudaGetLastError();
                // the tail: the same launcher on the remaining elements (planes advanced)
                plane_ptrs rest{};
                for (uint32_t f = 0; f < nfields; ++f)
                    rest.p[f] = planes.p[f] ? static_cast<char*>(planes.p[f]) + done * 4 : nullptr;
                launch_geom g2 = g;
                g2.unpack_bulk = 0;
                return launch_unpack<Elem, PlaneBytes, Signed, FP>(src + done, rest, nfields, n - done, fp, stream, g2);
            }
        }
        // 256-bit plane stores, eight elements per thread: +0.4 ... +2.3 % over the quad kernel on the same
        // box (C3 6515 -> 6665 GB/s, profiles/r07d_ab_unpack_oct.log); option vec256 = 0 keeps the quads
        constexpr size_t obytes = 8 * sizeof(Elem) < 32 ? 8 * sizeof(Elem) : 32;
        bool ook = g.vec256 != 0 && (reinterpret_cast<uintptr_t>(packed) % obytes) == 0;
        for (uint32_t f = 0; f < nfields; ++f)
            if (planes.p[f] && (reinterpret_cast<uintptr_t>(planes.p[f]) & 31u)) ook = false;
        const size_t nocts = ook ? n / 8 : 0;
        if (nocts) {
            if (g.vec256 == 2)
                { if (cudaError_t e = launch_stream_kernel(unpack_oct_kernel<Elem, Signed, FP, 2>, grid_for((nocts + 511) / 512, g), 256, stream, g.pdl != 0, src, planes, nocts, fp)) return e; }
            else
                { if (cudaError_t e = launch_stream_kernel(unpack_oct_kernel<Elem, Signed, FP, 1>, grid_for((nocts + 255) / 256, g), 256, stream, g.pdl != 0, src, planes, nocts, fp)) return e; }
            count_launch(g);
            done = nocts * 8;
        }
    }
    const size_t nquads = (ok && !done) ? n / 4 : 0;
    if (nquads) {
        const unsigned grid = grid_for((nquads + 255) / 256, g);
        if (cudaError_t e = launch_stream_kernel(unpack_quad_kernel<Elem, PlaneBytes, Signed, FP, 1, CACHE_STREAMING>, grid, 256, stream, g.pdl != 0, src, planes, nquads, fp)) return e;
        count_launch(g);
        done = 4 * nquads;
    }
    const size_t rest = n - done;
    if (rest) {
        const unsigned grid = grid_for((rest + 255) / 256, g);
        unpack_elem_kernel<Elem, PlaneBytes, Signed, FP><<<grid, 256, 0, stream>>>(src, planes, done, rest, fp);
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
    constexpr size_t plane_quad_align = PlaneBytes * 4 < 16 ? PlaneBytes * 4 : 16;  // one quad of a plane is one access
    bool ok = (reinterpret_cast<uintptr_t>(packed) % quad_align) == 0;
    for (uint32_t f = 0; f < nfields; ++f)
        if ((reinterpret_cast<uintptr_t>(planes.p[f]) % plane_quad_align) != 0) ok = false;
    size_t done = 0;
    if constexpr (PlaneBytes <= 2) {
        constexpr size_t G = 16 / PlaneBytes, gbytes = G * sizeof(Elem) < 32 ? G * sizeof(Elem) : 32;
        bool gok = (reinterpret_cast<uintptr_t>(packed) % gbytes) == 0;
        for (uint32_t f = 0; f < nfields; ++f)
            if (!aligned16(planes.p[f])) gok = false;
        const size_t ngroups = gok ? n / G : 0;
        if (ngroups) {
            if (cudaError_t e = launch_stream_kernel(repack_group_kernel<Elem, PlaneBytes, FP>, grid_for((ngroups + 255) / 256, g), 256, stream, g.pdl != 0, planes, dst, ngroups, fp)) return e;
            count_launch(g);
            done = ngroups * G;
        }
    }
    const size_t nquads = (ok && !done) ? n / 4 : 0;
    if (nquads) {
        const unsigned grid = grid_for((nquads + 255) / 256, g);
        if (cudaError_t e = launch_stream_kernel(repack_quad_kernel<Elem, PlaneBytes, FP>, grid, 256, stream, g.pdl != 0, planes, dst, nquads, fp)) return e;
        count_launch(g);
        done = 4 * nquads;
    }
    const size_t rest = n - done;
    if (rest) {
        const unsigned grid = grid_for((rest + 255) / 256, g);
        repack_elem_kernel<Elem, PlaneBytes, FP><<<grid, 256, 0, stream>>>(planes, dst, done, rest, fp);
        count_launch(g);
    }
    return cudaGetLastError();
}

#endif  // !__CUDACC_RTC__

}  // namespace kernels
}  // namespace cuda
}  // namespace pint

#endif  // PINT_CUDA_KERNELS_FUSED_CUH_INCLUDED
