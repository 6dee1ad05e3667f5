// pint/cuda_kernels_dense_convert.cuh -- dense bit stream <-> padded packed_int
// arrays and <-> per-field planes ("element" mapping, see cuda_kernels_dense.cuh).
//
//   dense_extract_kernel: the dense tile is staged in shared memory (TMA for whole
//     tiles), thread e takes element e with a funnel shift at its own bit offset and
//     writes the padded element (or its fields: get / get_signed, pint.hpp:799-822)
//     with coalesced stores.
//   dense_insert_kernel: thread e reads the padded element (or builds it from the
//     planes: truncating field constructor, pint.hpp:770-774, 592-601) with
//     coalesced loads and ORs it into a zeroed shared tile with shared-memory
//     atomics (neighbouring elements share words); the tile leaves by TMA.
// Bytes per element: S/8 on the dense side + sizeof(Integer) (or P*F) on the other.
#pragma once

#include "cuda_kernels_dense.cuh"

namespace pint {
namespace cuda {
namespace kernels {
namespace dense {

// what the element side of a conversion is
enum : int { SIDE_PADDED = 0, SIDE_PLANES = 1 };

struct dense_fields {
    field_table t;      // fields of ONE element (offsets within the element)
    plane_ptrs planes;  // SIDE_PLANES only
};

// tile_elems is a multiple of 128, so a tile is a whole number of 16-byte vectors.
template <class W, class Elem, int Side, int PlaneBytes, bool Signed>
__global__ void dense_extract_kernel(const uint4* dense, Elem* padded, size_t n, uint32_t S, uint32_t tile_elems,
                                     size_t total_vec, const __grid_constant__ dense_fields f)
{
    constexpr uint32_t WB = 8 * sizeof(W);
    extern __shared__ __align__(128) unsigned char dense_smem[];
    __shared__ __align__(8) uint64_t full;
    W* D = reinterpret_cast<W*>(dense_smem);
    const uint32_t T = blockDim.x;
    const uint32_t tile_vecs = uint32_t(uint64_t(tile_elems) * S / 128);
    const uint32_t tile_bytes = tile_vecs * 16;
    const uint32_t tile_words = tile_bytes / sizeof(W);
    const size_t t = blockIdx.x;
    const size_t first = t * tile_elems;
    const uint32_t cnt = uint32_t(n - first < tile_elems ? n - first : tile_elems);
    const bool whole = (t + 1) * size_t(tile_vecs) <= total_vec;
    if (threadIdx.x == 0) D[tile_words] = 0;  // the word a last element's window may peek at
    if (whole) {
        if (threadIdx.x == 0) {
            const uint64_t pol = evict_first_policy();
            mbar_init(&full, 1);
            asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
            asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
            mbar_expect_tx(&full, tile_bytes);
            bulk_load(D, dense + t * tile_vecs, tile_bytes, &full, pol);
        }
        __syncthreads();
        mbar_wait(&full, 0);
    } else {
        for (uint32_t v = threadIdx.x; v < tile_vecs; v += T) {
            const size_t gv = t * tile_vecs + v;
            reinterpret_cast<uint4*>(D)[v] = gv < total_vec ? ld16<CACHE_STREAMING>(dense + gv) : make_uint4(0, 0, 0, 0);
        }
        __syncthreads();
    }
    const W SM = ones_w<W>(S);
    for (uint32_t e = threadIdx.x; e < cnt; e += T) {
        const uint64_t p = uint64_t(e) * S;
        const uint32_t w = uint32_t(p / WB), sh = uint32_t(p % WB);
        const W x = funnel_r(D[w], D[w + 1], sh) & SM;
        if constexpr (Side == SIDE_PADDED) {
            padded[first + e] = Elem(x);
        } else {
            using SW = typename std::make_signed<W>::type;
            using PT = typename std::conditional<PlaneBytes == 8, uint64_t, uint32_t>::type;
            for (uint32_t i = 0; i < f.t.nfields; ++i) {
                PT* plane = static_cast<PT*>(f.planes.p[i]);
                if (!plane) continue;
                const uint32_t off = f.t.off[i], wd = f.t.bits[i];
                // get_signed sign-extends in make_signed<Integer> (pint.hpp:819-821); the value is
                // the same in any wider signed type, which is what the plane cast produces
                const W v = Signed ? W(SW(x << (WB - off - wd)) >> (WB - wd)) : W((x << (WB - off - wd)) >> (WB - wd));
                plane[first + e] = Signed ? PT(int64_t(SW(v))) : PT(v);
            }
        }
    }
}

template <class W>
__device__ __forceinline__ void smem_or(W* p, W v)
{
    if constexpr (sizeof(W) == 4) atomicOr(reinterpret_cast<unsigned int*>(p), (unsigned int)v);
    else atomicOr(reinterpret_cast<unsigned long long*>(p), (unsigned long long)v);
}

template <class W, class Elem, int Side, int PlaneBytes>
__global__ void dense_insert_kernel(const Elem* padded, uint4* dense, size_t n, uint32_t S, uint32_t tile_elems,
                                    size_t total_vec, const __grid_constant__ dense_fields f)
{
    constexpr uint32_t WB = 8 * sizeof(W);
    extern __shared__ __align__(128) unsigned char dense_smem[];
    W* D = reinterpret_cast<W*>(dense_smem);
    const uint32_t T = blockDim.x;
    const uint32_t tile_vecs = uint32_t(uint64_t(tile_elems) * S / 128);
    const uint32_t tile_bytes = tile_vecs * 16;
    const size_t t = blockIdx.x;
    const size_t first = t * tile_elems;
    const uint32_t cnt = uint32_t(n - first < tile_elems ? n - first : tile_elems);
    const bool whole = (t + 1) * size_t(tile_vecs) <= total_vec;
    for (uint32_t v = threadIdx.x; v <= tile_vecs; v += T)  // one vector more: spill target of the last element
        reinterpret_cast<uint4*>(D)[v] = make_uint4(0, 0, 0, 0);
    __syncthreads();
    const W SM = ones_w<W>(S);
    for (uint32_t e = threadIdx.x; e < cnt; e += T) {
        W x;
        if constexpr (Side == SIDE_PADDED) {
            x = W(padded[first + e]) & SM;
        } else {
            x = 0;
            for (uint32_t i = 0; i < f.t.nfields; ++i) {
                // 4-byte planes are int32_t values converted to Integer (the field constructor's
                // parameters are value_type, pint.hpp:771-772): sign-extend before truncating
                using ST = typename std::conditional<PlaneBytes == 8, int64_t, int32_t>::type;
                const ST v = static_cast<const ST*>(f.planes.p[i])[first + e];
                x |= (W(int64_t(v)) & ones_w<W>(f.t.bits[i])) << f.t.off[i];  // make_truncated_int, pint.hpp:390-407
            }
        }
        const uint64_t p = uint64_t(e) * S;
        const uint32_t w = uint32_t(p / WB), sh = uint32_t(p % WB);
        smem_or(D + w, W(x << sh));
        if (sh + S > WB) smem_or(D + w + 1, spill_l(x, sh));
    }
    if (whole) {
        asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
        __syncthreads();
        if (threadIdx.x == 0) {
            bulk_store(dense + t * tile_vecs, D, tile_bytes, evict_first_policy());
            bulk_wait_read<0>();
        }
    } else {
        __syncthreads();
        for (uint32_t v = threadIdx.x; v < tile_vecs; v += T) {
            const size_t gv = t * tile_vecs + v;
            if (gv < total_vec) st16<CACHE_STREAMING>(dense + gv, reinterpret_cast<const uint4*>(D)[v]);
        }
    }
}

inline uint32_t convert_tile_elems(uint32_t S)
{
    uint32_t k = 1024 / S;  // 128 * k elements * S bits ~ 16 KB
    if (k < 1) k = 1;
    return 128 * k;
}

// Side == SIDE_PADDED: `other` is the Elem array; SIDE_PLANES: f.planes holds the planes.
template <class W, class Elem, int Side, int PlaneBytes, bool Signed>
cudaError_t launch_dense_extract(const void* dense, void* other, size_t n, uint32_t S, const dense_fields& f,
                                 cudaStream_t stream, const launch_geom& g)
{
    if (n == 0) return cudaSuccess;
    const uint32_t te = convert_tile_elems(S);
    const size_t total_vec = size_t((uint64_t(n) * S + 127) / 128);
    const size_t ntiles = (n + te - 1) / te;
    const size_t smem = size_t(te) * S / 8 + 16;
    if (ntiles > 0x7FFFFFFFull) return cudaErrorInvalidValue;
    dense_extract_kernel<W, Elem, Side, PlaneBytes, Signed><<<unsigned(ntiles), 256, smem, stream>>>(
        static_cast<const uint4*>(dense), static_cast<Elem*>(other), n, S, te, total_vec, f);
    count_launch(g);
    return cudaGetLastError();
}

template <class W, class Elem, int Side, int PlaneBytes>
cudaError_t launch_dense_insert(const void* other, void* dense, size_t n, uint32_t S, const dense_fields& f,
                                cudaStream_t stream, const launch_geom& g)
{
    if (n == 0) return cudaSuccess;
    const uint32_t te = convert_tile_elems(S);
    const size_t total_vec = size_t((uint64_t(n) * S + 127) / 128);
    const size_t ntiles = (n + te - 1) / te;
    const size_t smem = size_t(te) * S / 8 + 16;
    if (ntiles > 0x7FFFFFFFull) return cudaErrorInvalidValue;
    dense_insert_kernel<W, Elem, Side, PlaneBytes><<<unsigned(ntiles), 256, smem, stream>>>(
        static_cast<const Elem*>(other), static_cast<uint4*>(dense), n, S, te, total_vec, f);
    count_launch(g);
    return cudaGetLastError();
}

}  // namespace dense
}  // namespace kernels
}  // namespace cuda
}  // namespace pint
