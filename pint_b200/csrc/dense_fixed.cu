// dense_fixed.cu -- padded / planes -> dense conversions with the element width S compiled in
// (cuda_kernels_dense_fixed.cuh), one kernel per S.  Compile with -DPINT_FIXED_PART=<0..3>; part p
// covers S in [16 p, 16 p + 16).  The padded side is instantiated for the natural Integer of S
// (make_packed_int's choice, pint.hpp:724-734, 789-795); other Integers use the run-time-geometry kernel.
#include "dense_registry.hpp"
#include "pint/cuda_kernels_dense_fixed.cuh"

#ifndef PINT_FIXED_PART
#error "compile with -DPINT_FIXED_PART=n"
#endif

namespace pint_cuda_impl {
namespace {

using namespace pint::cuda;
using namespace pint::cuda::kernels;

template <uint32_t S>
using natural_int = typename std::conditional<(S <= 8), uint8_t,
                    typename std::conditional<(S <= 16), uint16_t,
                    typename std::conditional<(S <= 32), uint32_t, uint64_t>::type>::type>::type;

template <uint32_t S>
bool insert_s(uint32_t elem_bytes, int plane_bytes, const void* other, void* dn, size_t n, const dense::dense_fields& f,
              cudaStream_t s, const launch_geom& g, size_t* done, cudaError_t* rc)
{
    using E = natural_int<S>;
    using W = typename dense::fixed_geom<S, 4>::W;
    if (elem_bytes) {
        if (elem_bytes != sizeof(E)) return false;
        *rc = dense::launch_dense_insert_fixed<E, S, dense::SIDE_PADDED, 4>(other, dn, n, f, s, g, done);
    } else if (plane_bytes == 1) {
        *rc = dense::launch_dense_insert_fixed<W, S, dense::SIDE_PLANES, 1>(nullptr, dn, n, f, s, g, done);
    } else if (plane_bytes == 2) {
        *rc = dense::launch_dense_insert_fixed<W, S, dense::SIDE_PLANES, 2>(nullptr, dn, n, f, s, g, done);
    } else if (plane_bytes == 4) {
        *rc = dense::launch_dense_insert_fixed<W, S, dense::SIDE_PLANES, 4>(nullptr, dn, n, f, s, g, done);
    } else {
        *rc = dense::launch_dense_insert_fixed<W, S, dense::SIDE_PLANES, 8>(nullptr, dn, n, f, s, g, done);
    }
    return true;
}

// dense -> padded (not for uint32_t) and dense -> planes of 1 or 2 bytes: same-box A/B (profiles/r03m_ab_dense_convert.log) has the
// group-owner kernel ahead for uint8_t / uint16_t / uint64_t elements (5454 vs 3662, 6035 vs 5669,
// 6785 vs 5277 GB/s) and 2 % behind the run-time-geometry kernel for uint32_t and for planes.
template <uint32_t S>
bool extract_s(uint32_t elem_bytes, int plane_bytes, bool is_signed, const void* dn, void* other, size_t n,
               const dense::dense_fields& f, cudaStream_t s, const launch_geom& g, size_t* done, cudaError_t* rc)
{
    using E = natural_int<S>;
    if (!elem_bytes) {
        // dense -> planes of 1 or 2 bytes: 16 / plane_bytes elements per thread, one 16-byte store per plane
        using W = typename dense::fixed_geom<S, 4>::W;
        if (plane_bytes == 1) *rc = dense::launch_dense_extract_fixed<W, S, dense::SIDE_PLANES, 1, false>(dn, nullptr, n, f, s, g, done, is_signed ? 1u : 0u);
        else if (plane_bytes == 2) *rc = dense::launch_dense_extract_fixed<W, S, dense::SIDE_PLANES, 2, false>(dn, nullptr, n, f, s, g, done, is_signed ? 1u : 0u);
        else return false;
        return true;
    }
    if constexpr (sizeof(E) == 4) {
        return false;
    } else {
        if (elem_bytes != sizeof(E)) return false;
        if (sizeof(E) == 8 && reinterpret_cast<uintptr_t>(other) % 32) return false;  // 32-byte stores
        *rc = dense::launch_dense_extract_fixed<E, S, dense::SIDE_PADDED, 4, false>(dn, other, n, f, s, g, done);
        return true;
    }
}

template <uint32_t Lo, uint32_t Hi>
bool extract_range(uint32_t S, uint32_t elem_bytes, int plane_bytes, bool is_signed, const void* dn, void* other, size_t n,
                   const dense::dense_fields& f, cudaStream_t s, const launch_geom& g, size_t* done, cudaError_t* rc)
{
    if constexpr (Lo >= Hi) {
        return false;
    } else {
        if constexpr (Lo != 0 && Lo != 8 && Lo != 16 && Lo != 32)
            if (S == Lo) return extract_s<Lo>(elem_bytes, plane_bytes, is_signed, dn, other, n, f, s, g, done, rc);
        return extract_range<Lo + 1, Hi>(S, elem_bytes, plane_bytes, is_signed, dn, other, n, f, s, g, done, rc);
    }
}

template <uint32_t Lo, uint32_t Hi>
bool insert_range(uint32_t S, uint32_t elem_bytes, int plane_bytes, const void* other, void* dn, size_t n,
                  const dense::dense_fields& f, cudaStream_t s, const launch_geom& g, size_t* done, cudaError_t* rc)
{
    if constexpr (Lo >= Hi) {
        return false;
    } else {
        // S = 8, 16, 32, 64 is the padded format itself (api_dense.cu)
        if constexpr (Lo != 0 && Lo != 8 && Lo != 16 && Lo != 32)
            if (S == Lo) return insert_s<Lo>(elem_bytes, plane_bytes, other, dn, n, f, s, g, done, rc);
        return insert_range<Lo + 1, Hi>(S, elem_bytes, plane_bytes, other, dn, n, f, s, g, done, rc);
    }
}

}  // namespace

#define PINT_CAT2(a, b) a##b
#define PINT_CAT(a, b) PINT_CAT2(a, b)

bool PINT_CAT(dense_insert_fixed_part, PINT_FIXED_PART)(uint32_t S, uint32_t elem_bytes, int plane_bytes, const void* other,
                                                        void* dn, size_t n, const dense::dense_fields& f, cudaStream_t s,
                                                        const launch_geom& g, size_t* done, cudaError_t* rc)
{
    return insert_range<16 * PINT_FIXED_PART, 16 * PINT_FIXED_PART + 16>(S, elem_bytes, plane_bytes, other, dn, n, f, s, g,
                                                                          done, rc);
}

bool PINT_CAT(dense_extract_fixed_part, PINT_FIXED_PART)(uint32_t S, uint32_t elem_bytes, int plane_bytes, bool is_signed,
                                                         const void* dn, void* other, size_t n, const dense::dense_fields& f,
                                                         cudaStream_t s, const launch_geom& g, size_t* done, cudaError_t* rc)
{
    return extract_range<16 * PINT_FIXED_PART, 16 * PINT_FIXED_PART + 16>(S, elem_bytes, plane_bytes, is_signed, dn, other, n, f, s,
                                                                           g, done, rc);
}

}  // namespace pint_cuda_impl
