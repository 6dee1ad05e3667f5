// pint/cuda_kernels_dense_convert.cuh -- dense bit stream <-> padded packed_int
// arrays and <-> per-field planes ("element" mapping, see cuda_kernels_dense.cuh).
//
//   dense_extract_kernel: the dense tile is staged in shared memory (TMA for whole
//     tiles), thread e takes element e with a funnel shift at its own bit offset and
//     writes the padded element (or its fields: get / get_signed, pint.hpp:799-822)
//     with coalesced stores.
//   dense_insert_kernel: one thread per output word gathers the padded elements (or
//     builds them from the planes: truncating field constructor, pint.hpp:770-774,
//     592-601) that overlap its word and stores it; no shared memory, no atomics.
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

// Insert ("word-owner" mapping): the padded elements (or the F planes) of a tile are
// staged in shared memory with coalesced, independent loads; then one thread per
// OUTPUT word gathers the elements that overlap its word -- word j of the tile holds
// (parts of) elements floor(j*WB/S) .. floor(((j+1)*WB-1)/S) -- shifts each into place
// and stores the finished word (consecutive threads, consecutive words).  No atomics.
// History (profiles/r01e_dense_sweep.jsonl, r01g_dense.jsonl): OR-ing elements into a
// zeroed shared tile with atomicOr ran at 3.1-3.9 TB/s; gathering straight from global
// memory (dependent loads in a data-dependent loop, one word per thread) at 1.7-3.2.
template <class W, class Elem, int Side, int PlaneBytes>
__global__ void __launch_bounds__(256)
dense_insert_kernel(const Elem* padded, W* dense, size_t n, uint32_t S, uint32_t tile_elems, size_t total_words,
                    const __grid_constant__ dense_fields f)
{
    constexpr uint32_t WB = 8 * sizeof(W);
    using ST = typename std::conditional<PlaneBytes == 8, int64_t, int32_t>::type;
    using Staged = typename std::conditional<Side == SIDE_PADDED, Elem, ST>::type;
    extern __shared__ __align__(128) unsigned char dense_smem[];
    Staged* tile = reinterpret_cast<Staged*>(dense_smem);
    const uint32_t T = blockDim.x;
    const size_t t = blockIdx.x;
    const size_t first = t * tile_elems;
    const uint32_t cnt = uint32_t(n - first < tile_elems ? n - first : tile_elems);
    const uint32_t nsrc = Side == SIDE_PADDED ? 1u : f.t.nfields;
    for (uint32_t k = 0; k < nsrc; ++k) {
        const Staged* src = (Side == SIDE_PADDED ? reinterpret_cast<const Staged*>(padded)
                                                 : static_cast<const Staged*>(f.planes.p[k])) + first;
        Staged* dst = tile + size_t(k) * tile_elems;
        if (cnt == tile_elems && (reinterpret_cast<uintptr_t>(src) & 15u) == 0) {
            const uint32_t nv = tile_elems * uint32_t(sizeof(Staged)) / 16;  // tile_elems is a multiple of 128
#pragma unroll 4
            for (uint32_t v = threadIdx.x; v < nv; v += T)
                reinterpret_cast<uint4*>(dst)[v] = ld16<CACHE_STREAMING>(reinterpret_cast<const uint4*>(src) + v);
        } else {
#pragma unroll 4
            for (uint32_t i = threadIdx.x; i < tile_elems; i += T) dst[i] = i < cnt ? src[i] : Staged(0);
        }
    }
    __syncthreads();
    const W SM = ones_w<W>(S);
    const uint32_t tile_words = uint32_t(uint64_t(tile_elems) * S / WB);
    for (uint32_t j = threadIdx.x; j < tile_words; j += T) {
        const size_t gj = t * size_t(tile_words) + j;
        if (gj >= total_words) break;
        const uint32_t bit0 = j * WB;
        uint32_t e = bit0 / S;
        int32_t rel = int32_t(e * S) - int32_t(bit0);  // in (-S, 0]
        W acc = 0;
        while (rel < int32_t(WB) && e < tile_elems) {
            W x;
            if constexpr (Side == SIDE_PADDED) {
                x = W(tile[e]) & SM;
            } else {
                x = 0;
                for (uint32_t i = 0; i < f.t.nfields; ++i) {
                    // plane values are signed integers converted to Integer (the field constructor's
                    // parameters are value_type, pint.hpp:771-772): sign-extend, then truncate to the field
                    const ST v = tile[size_t(i) * tile_elems + e];
                    x |= (W(int64_t(v)) & ones_w<W>(f.t.bits[i])) << f.t.off[i];  // make_truncated_int, pint.hpp:390-407
                }
            }
            acc |= rel >= 0 ? W(x << rel) : W(x >> (-rel));
            ++e;
            rel += int32_t(S);
        }
        dense[gj] = acc;
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
    const size_t total_words = size_t((uint64_t(n) * S + 127) / 128) * (16 / sizeof(W));
    // staged bytes per element: the padded element, or one value per plane
    const uint32_t per_elem = Side == SIDE_PADDED ? uint32_t(sizeof(Elem)) : uint32_t(PlaneBytes) * f.t.nfields;
    uint32_t te = (32768u / per_elem) & ~127u;
    if (te < 128) te = 128;
    if (te > 8192) te = 8192;
    const size_t smem = size_t(te) * per_elem;
    const size_t ntiles = (n + te - 1) / te;
    if (ntiles > 0x7FFFFFFFull) return cudaErrorInvalidValue;
    auto kernel = dense_insert_kernel<W, Elem, Side, PlaneBytes>;
    if (smem > 48 * 1024) {
        cudaError_t rc = cudaFuncSetAttribute(kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, int(smem));
        if (rc != cudaSuccess) return rc;
    }
    kernel<<<unsigned(ntiles), 256, smem, stream>>>(static_cast<const Elem*>(other), static_cast<W*>(dense), n, S, te,
                                                    total_words, f);
    count_launch(g);
    return cudaGetLastError();
}

}  // namespace dense
}  // namespace kernels
}  // namespace cuda
}  // namespace pint
