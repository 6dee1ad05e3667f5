// pint/cuda_kernels_planes_bulk.cuh -- packed -> int32 planes with the plane tiles staged in shared
// memory and written by the TMA engine (cp.async.bulk.global.shared::cta, SASS UBLKCP.G.S).
//
// unpack is the write-heavy map of the path (C3: 8 bytes read, 24 written per element, six output
// streams).  The register-to-global kernels (unpack_quad / unpack_oct) leave every plane as 32-byte
// stores per lane, 1 KB per warp instruction; here a CTA assembles TILE consecutive values of every
// plane in shared memory (conflict-free 16-byte stores) and hands each plane tile (TILE * 4 bytes,
// contiguous) to the TMA engine as ONE bulk store, so the L2 sees whole multi-KB writes per plane.
// One tile per CTA; several CTAs per SM overlap load, extract and store.  Which of the two families
// the library launches is a measured decision (tools/tune.cu mode 11, profiles/).
#pragma once
#ifndef PINT_CUDA_KERNELS_PLANES_BULK_CUH_INCLUDED
#define PINT_CUDA_KERNELS_PLANES_BULK_CUH_INCLUDED

#include "cuda_kernels.cuh"
#include "cuda_kernels_bulk.cuh"

namespace pint {
namespace cuda {
namespace kernels {

// four consecutive elements as ONE access where the Integer allows (32 bytes for uint64_t)
template <class Elem, class Word>
__device__ __forceinline__ void load_quad_wide(const Elem* p, Word (&e)[4])
{
    if constexpr (sizeof(Elem) == 8) {
        const vec32 v = ld32(reinterpret_cast<const vec32*>(p));
        e[0] = (uint64_t(v.lo.y) << 32) | v.lo.x; e[1] = (uint64_t(v.lo.w) << 32) | v.lo.z;
        e[2] = (uint64_t(v.hi.y) << 32) | v.hi.x; e[3] = (uint64_t(v.hi.w) << 32) | v.hi.z;
    } else {
        load_quad<Elem, Word>(p, e);
    }
}

// get<I> / get_signed<I> (pint.hpp:799-822) of every wanted field of TILE consecutive elements per CTA.
// Thread t takes the quads t, t + BLOCK, ... of the tile: global loads and shared stores of a warp are
// contiguous.  Dynamic shared memory: nfields * TILE * 4 bytes.
template <class Elem, bool Signed, class FP, int TILE, int BLOCK>
__global__ void __launch_bounds__(BLOCK)
unpack_bulk_kernel(const Elem* packed, const __grid_constant__ plane_ptrs planes, size_t ntiles, uint32_t nfields,
                   const __grid_constant__ FP fp)
{
    pdl_prologue();
    using W = typename FP::word;
    static_assert(TILE % (4 * BLOCK) == 0, "a tile is a whole number of quads per thread");
    constexpr int QPT = TILE / 4 / BLOCK;
    extern __shared__ __align__(128) unsigned char unpack_bulk_smem[];
    uint4* stage = reinterpret_cast<uint4*>(unpack_bulk_smem);  // [field][TILE / 4]
    const uint64_t pol = bulk::evict_first_policy();
    const bool issuer = threadIdx.x < nfields && planes.p[threadIdx.x] != nullptr;
    bool first = true;
    for (size_t t = blockIdx.x; t < ntiles; t += gridDim.x) {
        W e[QPT][4];
#pragma unroll
        for (int k = 0; k < QPT; ++k)
            load_quad_wide<Elem, W>(packed + t * TILE + 4 * (size_t(k) * BLOCK + threadIdx.x), e[k]);
        if (!first) {  // (persistent grids only) the previous tile's bulk stores must have read the stage
            if (issuer) bulk::bulk_wait_read<0>();
            __syncthreads();
        }
        first = false;
        fp.each_elem_field([&](uint32_t i, uint32_t off, uint32_t w) {
            if (!planes.p[i]) return;
#pragma unroll
            for (int k = 0; k < QPT; ++k) {
                uint32_t v[4];
#pragma unroll
                for (int j = 0; j < 4; ++j)
                    v[j] = Signed ? uint32_t(field_get_signed<W>(e[k][j], off, w)) : uint32_t(field_get<W>(e[k][j], off, w));
                stage[size_t(i) * (TILE / 4) + k * BLOCK + threadIdx.x] = make_uint4(v[0], v[1], v[2], v[3]);
            }
        });
        asm volatile("fence.proxy.async.shared::cta;" ::: "memory");  // generic-proxy writes -> async proxy
        __syncthreads();
        if (issuer)
            bulk::bulk_store(static_cast<char*>(planes.p[threadIdx.x]) + (t * TILE) * 4, stage + size_t(threadIdx.x) * (TILE / 4),
                             uint32_t(TILE) * 4u, pol);
    }
    if (issuer) bulk::bulk_wait_read<0>();  // the stage must outlive the engine's reads
}

}  // namespace kernels
}  // namespace cuda
}  // namespace pint

#endif  // PINT_CUDA_KERNELS_PLANES_BULK_CUH_INCLUDED
