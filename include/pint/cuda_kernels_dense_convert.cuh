// pint/cuda_kernels_dense_convert.cuh -- dense bit stream <-> padded packed_int
// arrays and <-> per-field planes ("element" mapping, see cuda_kernels_dense.cuh).
//
//   dense_extract_kernel: the dense tile is staged in shared memory (TMA for whole
//     tiles), thread e takes element e with a funnel shift at its own bit offset and
//     writes the padded element (or its fields: get / get_signed, pint.hpp:799-822)
//     with coalesced stores.
//   dense_insert_kernel: the padded elements (or the planes: truncating field
//     constructor, pint.hpp:770-774, 592-601) are staged in shared memory, one thread
//     per chunk deposits its G elements into Cw whole words; no atomics.
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
// `quads` (host-checked: the padded array / every plane is 16-byte aligned): a thread takes
// FOUR consecutive elements and writes them with one 16-byte store per array (two for 8-byte
// values), which cuts the per-element pointer arithmetic and store count by four -- the
// one-element-per-thread form was instruction-issue-bound (ncu: 87 % issue active, 5.6 TB/s
// for three int32 planes; profiles/r01i_ncu_full_dense_summary.txt).
template <class W, class Elem, int Side, int PlaneBytes, bool Signed>
__global__ void dense_extract_kernel(const uint4* dense, Elem* padded, size_t n, uint32_t S, uint32_t tile_elems,
                                     size_t total_vec, uint32_t quads, const __grid_constant__ dense_fields f)
{
    pdl_prologue();
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
    auto element = [&](uint32_t e) {
        const uint32_t p = e * S;  // tile_elems * S < 2^32
        return W(funnel_r(D[p / WB], D[p / WB + 1], p % WB) & SM);
    };
    auto field = [&](W x, uint32_t i) {
        // get<I> (pint.hpp:799-807) / get_signed<I> (:809-822); the signed value is the same in
        // any wider signed type, which is what the plane cast produces
        const uint32_t off = f.t.off[i], wd = f.t.bits[i];
        return Signed ? uint64_t(int64_t(field_get_signed<W>(x, off, wd))) : uint64_t(field_get<W>(x, off, wd));
    };
    const uint32_t nquad = quads ? cnt / 4 : 0;
    for (uint32_t q = threadIdx.x; q < nquad; q += T) {
        W e[4];
#pragma unroll
        for (int j = 0; j < 4; ++j) e[j] = element(4 * q + j);
        if constexpr (Side == SIDE_PADDED) {
            store_quad<Elem, W>(padded + first + 4 * size_t(q), e);
        } else {
            for (uint32_t i = 0; i < f.t.nfields; ++i) {
                void* plane = f.planes.p[i];
                if (!plane) continue;
                uint64_t v[4];
#pragma unroll
                for (int j = 0; j < 4; ++j) v[j] = field(e[j], i);
                store_plane_quad<PlaneBytes>(plane, first / 4 + q, v);
            }
        }
    }
    for (uint32_t e = 4 * nquad + threadIdx.x; e < cnt; e += T) {  // everything when !quads, else the last < 4
        const W x = element(e);
        if constexpr (Side == SIDE_PADDED) {
            padded[first + e] = Elem(x);
        } else {
            using PT = typename std::conditional<PlaneBytes == 8, uint64_t, typename std::conditional<PlaneBytes == 4, uint32_t,
                       typename std::conditional<PlaneBytes == 2, uint16_t, uint8_t>::type>::type>::type;
            for (uint32_t i = 0; i < f.t.nfields; ++i) {
                PT* plane = static_cast<PT*>(f.planes.p[i]);
                if (plane) plane[first + e] = PT(field(x, i));
            }
        }
    }
}

// Insert ("chunk" mapping, as in dense_binary_kernel).  The padded elements (or the F
// planes) of a tile are staged in shared memory with coalesced, independent loads, at
// index i + i / G so that a thread reading G consecutive values walks a row of odd stride
// G + 1 (conflict-free).  Then one thread per CHUNK (G elements = Cw whole output words)
// deposits its elements into a two-word window with shifts that are the same for every
// thread, and writes its Cw words into the output tile (stride Cw, odd).  The tile leaves
// by TMA.  No element and no output word is shared between threads: no atomics.
// History (profiles/r01e_dense_sweep.jsonl, r01g/r01h_dense.jsonl): atomicOr into a zeroed
// shared tile 3.1-3.9 TB/s; one thread per output word gathering from global 1.7-3.2, from
// a staged tile 2.7-4.5 (issue-bound on per-word division and data-dependent loops).
template <class W, class Elem, int Side, int PlaneBytes>
__global__ void dense_insert_kernel(const Elem* padded, uint4* dense, size_t n, uint32_t S, uint32_t log2G, uint32_t Cw,
                                    uint32_t chunks_per_tile, size_t total_vec, uint32_t quads,
                                    const __grid_constant__ dense_fields f)
{
    pdl_prologue();
    constexpr uint32_t WB = 8 * sizeof(W);
    using ST = typename std::conditional<PlaneBytes == 8, int64_t, typename std::conditional<PlaneBytes == 4, int32_t,
               typename std::conditional<PlaneBytes == 2, int16_t, int8_t>::type>::type>::type;
    // what is staged per element: the padded element, or the element assembled from its planes
    using Staged = typename std::conditional<Side == SIDE_PADDED, Elem, W>::type;
    extern __shared__ __align__(128) unsigned char dense_smem[];
    const uint32_t T = blockDim.x, G = 1u << log2G;
    const uint32_t tile_elems = chunks_per_tile << log2G;
    const uint32_t tile_words = chunks_per_tile * Cw;          // a multiple of 16 bytes (host)
    const uint32_t tile_vecs = tile_words * uint32_t(sizeof(W)) / 16;
    W* out = reinterpret_cast<W*>(dense_smem);
    Staged* in = reinterpret_cast<Staged*>(dense_smem + size_t(tile_vecs) * 16);
    const size_t t = blockIdx.x;
    const size_t first = t * size_t(tile_elems);
    const uint32_t cnt = uint32_t(n - first < tile_elems ? n - first : tile_elems);
    const W SM = ones_w<W>(S);
    if constexpr (Side == SIDE_PADDED) {
        const Elem* src = padded + first;
        if (cnt == tile_elems && (reinterpret_cast<uintptr_t>(src) & 15u) == 0) {
            // 16-byte loads (several in flight per thread), scattered to the padded rows value by value
            constexpr uint32_t PV = 16 / sizeof(Elem);
            const uint32_t nv = tile_elems / PV;  // tile_elems is a multiple of 128
#pragma unroll 4
            for (uint32_t v = threadIdx.x; v < nv; v += T) {
                union { uint4 q; Elem x[PV]; } u;
                u.q = ld16<CACHE_STREAMING>(reinterpret_cast<const uint4*>(src) + v);
#pragma unroll
                for (uint32_t k2 = 0; k2 < PV; ++k2) {
                    const uint32_t i = v * PV + k2;
                    in[i + (i >> log2G)] = u.x[k2];
                }
            }
        } else {
#pragma unroll 4
            for (uint32_t i = threadIdx.x; i < tile_elems; i += T) in[i + (i >> log2G)] = i < cnt ? src[i] : Elem(0);
        }
    } else {
        // the truncating field constructor (pint.hpp:770-774, 592-601, 390-407) on four elements at
        // a time: one 16-byte load per plane (two for 8-byte planes); plane values are signed
        // integers converted to Integer (value_type parameters, :771-772): sign-extend, then truncate
        const uint32_t nquad = quads && cnt == tile_elems ? tile_elems / 4 : 0;
        for (uint32_t q = threadIdx.x; q < nquad; q += T) {
            W e[4] = {0, 0, 0, 0};
            for (uint32_t i = 0; i < f.t.nfields; ++i) {
                uint64_t v[4];
                load_plane_quad<PlaneBytes>(f.planes.p[i], first / 4 + q, v);
                const W m = ones_w<W>(f.t.bits[i]);
                const uint32_t off = f.t.off[i];
#pragma unroll
                for (int jj = 0; jj < 4; ++jj) e[jj] |= (W(v[jj]) & m) << off;
            }
#pragma unroll
            for (uint32_t jj = 0; jj < 4; ++jj) {
                const uint32_t i = 4 * q + jj;
                in[i + (i >> log2G)] = e[jj];
            }
        }
        for (uint32_t i = 4 * nquad + threadIdx.x; i < tile_elems; i += T) {
            W x = 0;
            if (i < cnt)
                for (uint32_t k = 0; k < f.t.nfields; ++k) {
                    const ST v = static_cast<const ST*>(f.planes.p[k])[first + i];
                    x |= (W(int64_t(v)) & ones_w<W>(f.t.bits[k])) << f.t.off[k];
                }
            in[i + (i >> log2G)] = x;
        }
    }
    __syncthreads();
    for (uint32_t c = threadIdx.x; c < chunks_per_tile; c += T) {
        const Staged* row = in + size_t(c) * (G + 1);
        W* o = out + size_t(c) * Cw;
        W ocur = 0, onxt = 0;
        uint32_t sh = 0, w = 0;
        for (uint32_t e = 0; e < G; ++e) {
            const W x = W(row[e]) & SM;
            ocur |= x << sh;
            onxt |= spill_l(x, sh);
            sh += S;
            if (sh >= WB) {
                sh -= WB;
                o[w++] = ocur;
                ocur = onxt;
                onxt = 0;
            }
        }
    }
    const bool whole = (t + 1) * size_t(tile_vecs) <= total_vec;
    if (whole) {
        asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
        __syncthreads();
        if (threadIdx.x == 0) {
            bulk_store(dense + t * tile_vecs, out, tile_vecs * 16, evict_first_policy());
            bulk_wait_read<0>();
        }
    } else {
        __syncthreads();
        for (uint32_t v = threadIdx.x; v < tile_vecs; v += T) {
            const size_t gv = t * tile_vecs + v;
            if (gv < total_vec) st16<CACHE_STREAMING>(dense + gv, reinterpret_cast<const uint4*>(out)[v]);
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
    // 16-byte stores of four values need aligned arrays (tiles start on multiples of 128 elements)
    uintptr_t align = reinterpret_cast<uintptr_t>(other);
    if (Side == SIDE_PLANES) {
        align = 0;
        for (uint32_t i = 0; i < f.t.nfields; ++i) align |= reinterpret_cast<uintptr_t>(f.planes.p[i]);
    }
    const uint32_t quads = (align & 15u) == 0 ? 1u : 0u;
    if (cudaError_t e = launch_stream_kernel_smem(dense_extract_kernel<W, Elem, Side, PlaneBytes, Signed>, unsigned(ntiles), 256, smem, stream, g.pdl != 0,
                                                  static_cast<const uint4*>(dense), static_cast<Elem*>(other), n, S, te, total_vec, quads, f))
        return e;
    count_launch(g);
    return cudaGetLastError();
}

template <class W, class Elem, int Side, int PlaneBytes>
cudaError_t launch_dense_insert(const void* other, void* dense, size_t n, uint32_t S, const dense_fields& f,
                                cudaStream_t stream, const launch_geom& g)
{
    if (n == 0) return cudaSuccess;
    constexpr uint32_t WB = 8 * sizeof(W);
    const uint32_t gc = gcd_u32(S, WB), G = WB / gc, Cw = S / gc;
    uint32_t log2G = 0;
    while ((1u << log2G) < G) ++log2G;
    // staged bytes per element: the padded element, or the element assembled from its planes
    const uint32_t per_elem = Side == SIDE_PADDED ? uint32_t(sizeof(Elem)) : uint32_t(sizeof(W));
    const uint32_t unit = 128 / WB;  // chunks per tile must make the output tile a whole number of 16-byte vectors
    uint32_t cpt = (12288u / (per_elem * G)) / unit * unit;
    if (cpt < unit) cpt = unit;
    if (cpt > 2048) cpt = 2048;
    const size_t out_bytes = size_t(cpt) * Cw * sizeof(W);
    const size_t in_bytes = size_t(cpt) * (G + 1) * per_elem;
    const size_t smem = out_bytes + in_bytes;
    const size_t tile_elems = size_t(cpt) * G;
    const size_t total_vec = size_t((uint64_t(n) * S + 127) / 128);
    const size_t ntiles = (n + tile_elems - 1) / tile_elems;
    if (ntiles > 0x7FFFFFFFull) return cudaErrorInvalidValue;
    auto kernel = dense_insert_kernel<W, Elem, Side, PlaneBytes>;
    if (smem > 48 * 1024) {
        cudaError_t rc = cudaFuncSetAttribute(kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, int(smem));
        if (rc != cudaSuccess) return rc;
    }
    const unsigned threads = cpt >= 192 ? 256u : (cpt >= 96 ? 128u : 64u);
    uintptr_t align = 0;  // 16-byte plane loads of four values need aligned planes
    if (Side == SIDE_PLANES)
        for (uint32_t i = 0; i < f.t.nfields; ++i) align |= reinterpret_cast<uintptr_t>(f.planes.p[i]);
    const uint32_t quads = (align & 15u) == 0 ? 1u : 0u;
    if (cudaError_t e = launch_stream_kernel_smem(kernel, unsigned(ntiles), threads, smem, stream, g.pdl != 0, static_cast<const Elem*>(other),
                                                  static_cast<uint4*>(dense), n, S, log2G, Cw, cpt, total_vec, quads, f))
        return e;
    count_launch(g);
    return cudaGetLastError();
}

}  // namespace dense
}  // namespace kernels
}  // namespace cuda
}  // namespace pint
