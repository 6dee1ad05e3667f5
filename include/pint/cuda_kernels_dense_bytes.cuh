// pint/cuda_kernels_dense_bytes.cuh -- dense stream <-> planes when every field is exactly one plane
// value wide (all Bits == 8 * PlaneBytes): interleaved <-> planar, e.g. dense RGB24 <-> three uint8_t
// channels, dense RGB48 <-> three uint16_t channels.
//
// In that case get / get_signed (pint.hpp:799-822) narrowed to the plane type, and the truncating field
// constructor (pint.hpp:770-774) of plane values, move whole bytes and nothing else: element i of the
// stream is the record (plane_0[i], ..., plane_{F-1}[i]).  The general narrow-plane kernels spend ~20
// instructions per element on shifts, masks and byte packing (5999 / 4362 GB/s for RGB24 -> / <- uint8
// planes, profiles/r14_configs_extra.jsonl); here a thread owns a BLOCK of 16 / P records = 16 F bytes
// of stream = one 16-byte vector of every plane, and turns one into the other with byte permutes (PRMT)
// whose selectors are compile-time constants: 1-3 instructions per 32-bit word.
//   * plane side: one 16-byte access per plane per thread, consecutive lanes consecutive vectors;
//   * stream side: the CTA's tile (256 blocks) passes through shared memory, filled / drained with
//     coalesced 16-byte accesses; a thread's block is F vectors at a stride of 4 F words (conflict-free
//     for F = 3, 5, 7; two-way for F = 6).
// Whole tiles only; the launcher's caller finishes the remainder with the general kernels.
#pragma once

#include "cuda_kernels.cuh"

namespace pint {
namespace cuda {
namespace kernels {
namespace dense {

// byte j of the result word comes from flat byte index B_j of a 32-bit word array `w`
template <int B0, int B1, int B2, int B3, int NW>
__device__ __forceinline__ uint32_t gather_bytes(const uint32_t (&w)[NW])
{
    constexpr int W[4] = {B0 / 4, B1 / 4, B2 / 4, B3 / 4};
    constexpr int Y[4] = {B0 % 4, B1 % 4, B2 % 4, B3 % 4};
    // distinct source words in order of first use
    constexpr int u0 = W[0];
    constexpr int u1 = W[1] != u0 ? W[1] : (W[2] != u0 ? W[2] : (W[3] != u0 ? W[3] : u0));
    constexpr bool in01_2 = W[2] == u0 || W[2] == u1, in01_3 = W[3] == u0 || W[3] == u1;
    constexpr int u2 = !in01_2 ? W[2] : (!in01_3 ? W[3] : -1);
    constexpr int u3 = (u2 >= 0 && !in01_3 && W[3] != u2) ? W[3] : -1;
    constexpr uint32_t nib0 = uint32_t(W[0] == u0 ? Y[0] : 4 + Y[0]);
    constexpr uint32_t nib1 = uint32_t(W[1] == u0 ? Y[1] : (W[1] == u1 ? 4 + Y[1] : 0));
    constexpr uint32_t nib2 = uint32_t(W[2] == u0 ? Y[2] : (W[2] == u1 ? 4 + Y[2] : 0));
    constexpr uint32_t nib3 = uint32_t(W[3] == u0 ? Y[3] : (W[3] == u1 ? 4 + Y[3] : 0));
    uint32_t r = __byte_perm(w[u0], w[u1], nib0 | (nib1 << 4) | (nib2 << 8) | (nib3 << 12));
    if constexpr (u2 >= 0) {
        constexpr uint32_t s = uint32_t(W[0] == u2 ? 4 + Y[0] : 0) | (uint32_t(W[1] == u2 ? 4 + Y[1] : 1) << 4) |
                               (uint32_t(W[2] == u2 ? 4 + Y[2] : 2) << 8) | (uint32_t(W[3] == u2 ? 4 + Y[3] : 3) << 12);
        r = __byte_perm(r, w[u2], s);
    }
    if constexpr (u3 >= 0) {
        constexpr uint32_t s = uint32_t(W[0] == u3 ? 4 + Y[0] : 0) | (uint32_t(W[1] == u3 ? 4 + Y[1] : 1) << 4) |
                               (uint32_t(W[2] == u3 ? 4 + Y[2] : 2) << 8) | (uint32_t(W[3] == u3 ? 4 + Y[3] : 3) << 12);
        r = __byte_perm(r, w[u3], s);
    }
    return r;
}

// A block: R = 16 / P records of F fields of P bytes.  Stream byte of (record r, field f, byte b) =
// (r * F + f) * P + b; plane f byte of the same = r * P + b within that plane's 16-byte vector, i.e.
// flat byte 16 * f + r * P + b of the F plane vectors laid end to end.
template <int F, int P>
struct byte_block {
    static constexpr int R = 16 / P;
    static constexpr int NW = 4 * F;  // 32-bit words on either side
    // plane side byte index -> stream side byte index
    __host__ __device__ static constexpr int stream_of_plane(int pb) { return (((pb % 16) / P) * F + pb / 16) * P + (pb % 16) % P; }
    __host__ __device__ static constexpr int plane_of_stream(int sb) { return 16 * ((sb / P) % F) + ((sb / P) / F) * P + sb % P; }

    template <int I>
    __device__ __forceinline__ static uint32_t plane_word(const uint32_t (&rec)[NW])
    {
        return gather_bytes<stream_of_plane(4 * I), stream_of_plane(4 * I + 1), stream_of_plane(4 * I + 2), stream_of_plane(4 * I + 3)>(rec);
    }
    template <int I>
    __device__ __forceinline__ static uint32_t stream_word(const uint32_t (&pl)[NW])
    {
        return gather_bytes<plane_of_stream(4 * I), plane_of_stream(4 * I + 1), plane_of_stream(4 * I + 2), plane_of_stream(4 * I + 3)>(pl);
    }
    template <std::size_t... Is>
    __device__ __forceinline__ static void to_planes(const uint32_t (&rec)[NW], uint32_t (&pl)[NW], std::index_sequence<Is...>)
    {
        ((pl[Is] = plane_word<int(Is)>(rec)), ...);
    }
    template <std::size_t... Is>
    __device__ __forceinline__ static void to_stream(const uint32_t (&pl)[NW], uint32_t (&rec)[NW], std::index_sequence<Is...>)
    {
        ((rec[Is] = stream_word<int(Is)>(pl)), ...);
    }
};

constexpr int kByteTileBlocks = 256;  // blocks per CTA tile = threads per CTA

// dense -> planes: ntiles whole tiles of 256 blocks; a null plane is skipped
template <int F, int P>
__global__ void __launch_bounds__(kByteTileBlocks)
dense_bytes_to_planes_kernel(const uint4* stream, const __grid_constant__ plane_ptrs planes, size_t ntiles)
{
    pdl_prologue();
    using BB = byte_block<F, P>;
    __shared__ __align__(16) uint4 tile[kByteTileBlocks * F];
    for (size_t t = blockIdx.x; t < ntiles; t += gridDim.x) {
        const uint4* src = stream + t * (size_t(kByteTileBlocks) * F);
        uint4 v[F];
#pragma unroll
        for (int k = 0; k < F; ++k) v[k] = ld16<CACHE_STREAMING>(src + k * kByteTileBlocks + threadIdx.x);
        __syncthreads();  // the previous tile has been read
#pragma unroll
        for (int k = 0; k < F; ++k) tile[k * kByteTileBlocks + threadIdx.x] = v[k];
        __syncthreads();
        uint32_t rec[BB::NW], pl[BB::NW];
#pragma unroll
        for (int k = 0; k < F; ++k) {
            const uint4 q = tile[threadIdx.x * F + k];
            rec[4 * k] = q.x; rec[4 * k + 1] = q.y; rec[4 * k + 2] = q.z; rec[4 * k + 3] = q.w;
        }
        BB::to_planes(rec, pl, std::make_index_sequence<BB::NW>());
        const size_t block = t * kByteTileBlocks + threadIdx.x;
#pragma unroll
        for (int f = 0; f < F; ++f)
            if (planes.p[f])
                st16<CACHE_STREAMING>(static_cast<uint4*>(planes.p[f]) + block, make_uint4(pl[4 * f], pl[4 * f + 1], pl[4 * f + 2], pl[4 * f + 3]));
    }
}

// planes -> dense
template <int F, int P>
__global__ void __launch_bounds__(kByteTileBlocks)
dense_bytes_from_planes_kernel(const __grid_constant__ plane_ptrs planes, uint4* stream, size_t ntiles)
{
    pdl_prologue();
    using BB = byte_block<F, P>;
    __shared__ __align__(16) uint4 tile[kByteTileBlocks * F];
    for (size_t t = blockIdx.x; t < ntiles; t += gridDim.x) {
        const size_t block = t * kByteTileBlocks + threadIdx.x;
        uint32_t pl[BB::NW], rec[BB::NW];
#pragma unroll
        for (int f = 0; f < F; ++f) {
            const uint4 q = ld16<CACHE_STREAMING>(static_cast<const uint4*>(planes.p[f]) + block);
            pl[4 * f] = q.x; pl[4 * f + 1] = q.y; pl[4 * f + 2] = q.z; pl[4 * f + 3] = q.w;
        }
        BB::to_stream(pl, rec, std::make_index_sequence<BB::NW>());
        __syncthreads();  // the previous tile has been drained
#pragma unroll
        for (int k = 0; k < F; ++k) tile[threadIdx.x * F + k] = make_uint4(rec[4 * k], rec[4 * k + 1], rec[4 * k + 2], rec[4 * k + 3]);
        __syncthreads();
        uint4* dst = stream + t * (size_t(kByteTileBlocks) * F);
#pragma unroll
        for (int k = 0; k < F; ++k) st16<CACHE_STREAMING>(dst + k * kByteTileBlocks + threadIdx.x, tile[k * kByteTileBlocks + threadIdx.x]);
    }
}

#if !defined(__CUDACC_RTC__)
// *done = elements converted (whole tiles of 256 * 16 / P elements); false when (F, P) has no kernel.
// Planes and stream must be 16-byte aligned (the caller checks); to_stream needs every plane.
template <int F, int P>
cudaError_t launch_dense_bytes(bool to_stream, const plane_ptrs& planes, void* stream, size_t n, cudaStream_t s, const launch_geom& g,
                               size_t* done)
{
    const size_t tile_elems = size_t(kByteTileBlocks) * (16 / P);
    const size_t ntiles = n / tile_elems;
    *done = ntiles * tile_elems;
    if (!ntiles) return cudaSuccess;
    if (cudaError_t e = to_stream ? launch_stream_kernel(dense_bytes_from_planes_kernel<F, P>, grid_for(ntiles, g), kByteTileBlocks, s, g.pdl != 0, planes,
                                                         static_cast<uint4*>(stream), ntiles)
                                  : launch_stream_kernel(dense_bytes_to_planes_kernel<F, P>, grid_for(ntiles, g), kByteTileBlocks, s, g.pdl != 0,
                                                         static_cast<const uint4*>(stream), planes, ntiles))
        return e;
    count_launch(g);
    return cudaGetLastError();
}

inline bool dense_bytes_dispatch(uint32_t nfields, int plane_bytes, bool to_stream, const plane_ptrs& planes, void* stream, size_t n,
                                 cudaStream_t s, const launch_geom& g, size_t* done, cudaError_t* rc)
{
    *done = 0;
#define PINT_BYTES_CASE(F, P)                                                                              \
    if (nfields == F && plane_bytes == P) { *rc = launch_dense_bytes<F, P>(to_stream, planes, stream, n, s, g, done); return true; }
    PINT_BYTES_CASE(3, 1) PINT_BYTES_CASE(5, 1) PINT_BYTES_CASE(6, 1) PINT_BYTES_CASE(7, 1) PINT_BYTES_CASE(3, 2)
#undef PINT_BYTES_CASE
    return false;
}
#endif

}  // namespace dense
}  // namespace kernels
}  // namespace cuda
}  // namespace pint
