// pint/cuda_kernels_dense_fixed.cuh -- padded elements / planes -> dense stream with the
// element width S known at COMPILE time.
//
// The run-time-geometry insert kernel (cuda_kernels_dense_convert.cuh) spends ~40
// instructions per element on staging, index arithmetic and data-dependent shifts and is
// issue-bound at 3.4-4.7 TB/s (profiles/r03b_ncu_insert_summary.txt).  Here everything that
// depends on S is a compile-time constant:
//   * a thread owns whole UNITS: GE consecutive elements = CW whole words of the stream
//     (GE = WB / gcd(S, WB) elements, times K when that is less than 16 bytes of input), so
//     no word is shared between threads and there is nothing to stage on the input side;
//   * it reads them with 16-byte loads straight from global memory (consecutive lanes read
//     consecutive units; several loads in flight), deposits each element with immediate
//     shifts and writes every finished word into the CTA's output tile in shared memory
//     (unit stride CW words, odd: conflict-free);
//   * the tile leaves by one TMA bulk store.
// Only whole tiles; the launcher hands the remainder to the run-time-geometry kernel.
#pragma once

#include "cuda_kernels_dense_convert.cuh"

namespace pint {
namespace cuda {
namespace kernels {
namespace dense {

constexpr uint32_t cgcd(uint32_t a, uint32_t b) { return b == 0 ? a : cgcd(b, a % b); }

// Src: the type read from memory (the padded Integer, or the plane value type)
template <uint32_t S, uint32_t SrcBytes>
struct fixed_geom {
    using W = typename std::conditional<(S <= 32), uint32_t, uint64_t>::type;
    static constexpr uint32_t WB = 8 * sizeof(W);
    static constexpr uint32_t gc = cgcd(S, WB);
    static constexpr uint32_t G0 = WB / gc, Cw0 = S / gc;                                  // one chunk
    static constexpr uint32_t K = G0 * SrcBytes >= 16 ? 1 : 16 / (G0 * SrcBytes);          // chunks per unit
    static constexpr uint32_t GE = K * G0, CW = K * Cw0;                                   // one unit
    static constexpr uint32_t unit_in_bytes = GE * SrcBytes;                               // a multiple of 16
    static constexpr uint32_t unit_out_bytes = CW * uint32_t(sizeof(W));
    // elements a thread has in flight: at least 16 and at least 64 bytes of input
    static constexpr uint32_t want = SrcBytes >= 4 ? 16 : 64 / SrcBytes;
    static constexpr uint32_t UB = GE >= want ? 1 : want / GE;                             // units per iteration
    static constexpr uint32_t BEL = GE >= want ? want : GE;                                // elements per load batch
};

__device__ __forceinline__ void st_stream(uint32_t* p, uint32_t v) { asm volatile("st.global.cs.u32 [%0], %1;" ::"l"(p), "r"(v) : "memory"); }
__device__ __forceinline__ void st_stream(uint64_t* p, uint64_t v) { asm volatile("st.global.cs.u64 [%0], %1;" ::"l"(p), "l"(v) : "memory"); }

// deposit of the E-th element of a unit; folds to immediates once the callers' loops are unrolled
template <class W, uint32_t S>
__device__ __forceinline__ void deposit(W x, uint32_t e, W& cur, W* o)
{
    constexpr uint32_t WB = 8 * sizeof(W);
    const uint32_t p = e * S, sh = p % WB, w = p / WB;
    x &= ones_w<W>(S);
    cur |= W(x << sh);
    if (sh + S >= WB) {
        o[w] = cur;
        cur = sh + S > WB ? W(x >> (WB - sh)) : W(0);
    }
}

// NQ consecutive quads of one plane (sign-extended to 64 bits as load_plane_quad does); V32: 256-bit
// loads where the run is a whole number of 32-byte sectors, so that a lane never leaves half a
// sector to a second instruction
template <int PlaneBytes, uint32_t NQ, bool V32>
__device__ __forceinline__ void load_plane_quads(const void* plane, size_t q0, uint64_t (&v)[NQ][4])
{
    if constexpr (V32 && PlaneBytes == 8) {
#pragma unroll
        for (uint32_t q = 0; q < NQ; ++q) {
            const vec32 x = ld32(reinterpret_cast<const vec32*>(plane) + q0 + q);
            v[q][0] = (uint64_t(x.lo.y) << 32) | x.lo.x; v[q][1] = (uint64_t(x.lo.w) << 32) | x.lo.z;
            v[q][2] = (uint64_t(x.hi.y) << 32) | x.hi.x; v[q][3] = (uint64_t(x.hi.w) << 32) | x.hi.z;
        }
    } else if constexpr (V32 && PlaneBytes == 4 && NQ % 2 == 0) {
#pragma unroll
        for (uint32_t q = 0; q < NQ; q += 2) {
            const vec32 x = ld32(reinterpret_cast<const vec32*>(static_cast<const uint4*>(plane) + q0 + q));
            v[q][0] = uint64_t(int64_t(int32_t(x.lo.x))); v[q][1] = uint64_t(int64_t(int32_t(x.lo.y)));
            v[q][2] = uint64_t(int64_t(int32_t(x.lo.z))); v[q][3] = uint64_t(int64_t(int32_t(x.lo.w)));
            v[q + 1][0] = uint64_t(int64_t(int32_t(x.hi.x))); v[q + 1][1] = uint64_t(int64_t(int32_t(x.hi.y)));
            v[q + 1][2] = uint64_t(int64_t(int32_t(x.hi.z))); v[q + 1][3] = uint64_t(int64_t(int32_t(x.hi.w)));
        }
    } else if constexpr (PlaneBytes == 1 && NQ % 4 == 0) {
        // narrow planes: 16 consecutive int8 values per 16-byte load (the caller's units start on 16-byte
        // boundaries of every plane: unit_in_bytes is a multiple of 16)
#pragma unroll
        for (uint32_t q = 0; q < NQ; q += 4) {
            const uint4 x = ld16<CACHE_STREAMING>(reinterpret_cast<const uint4*>(static_cast<const uint32_t*>(plane) + q0 + q));
            const uint32_t w[4] = {x.x, x.y, x.z, x.w};
#pragma unroll
            for (uint32_t k = 0; k < 4; ++k)
#pragma unroll
                for (uint32_t i = 0; i < 4; ++i) v[q + k][i] = uint64_t(int64_t(int8_t(w[k] >> (8 * i))));
        }
    } else if constexpr (PlaneBytes == 2 && NQ % 2 == 0) {
#pragma unroll
        for (uint32_t q = 0; q < NQ; q += 2) {
            const uint4 x = ld16<CACHE_STREAMING>(reinterpret_cast<const uint4*>(static_cast<const uint2*>(plane) + q0 + q));
            const uint32_t w[4] = {x.x, x.y, x.z, x.w};
#pragma unroll
            for (uint32_t k = 0; k < 2; ++k)
#pragma unroll
                for (uint32_t i = 0; i < 4; ++i) v[q + k][i] = uint64_t(int64_t(int16_t(w[2 * k + (i >> 1)] >> (16 * (i & 1)))));
        }
    } else {
#pragma unroll
        for (uint32_t q = 0; q < NQ; ++q) load_plane_quad<PlaneBytes>(plane, q0 + q, v[q]);
    }
}

template <class Elem, uint32_t S, int Side, int PlaneBytes, bool V32 = false>
__global__ void __launch_bounds__(256)
dense_insert_fixed_kernel(const Elem* padded, uint4* dense, uint32_t units_per_thread, const __grid_constant__ dense_fields f)
{
    pdl_prologue();
    using Geo = fixed_geom<S, (Side == SIDE_PADDED ? uint32_t(sizeof(Elem)) : uint32_t(PlaneBytes))>;
    using W = typename Geo::W;
    constexpr uint32_t GE = Geo::GE, CW = Geo::CW, UB = Geo::UB, BEL = Geo::BEL;
    extern __shared__ __align__(128) unsigned char dense_smem[];
    W* out = reinterpret_cast<W*>(dense_smem);
    const uint32_t T = blockDim.x;
    const uint32_t units_per_tile = T * units_per_thread;
    const size_t unit0 = size_t(blockIdx.x) * units_per_tile;
    for (uint32_t it = 0; it < units_per_thread; it += UB) {
        if constexpr (Side == SIDE_PADDED) {
            // 16-byte loads, or 32-byte ones (V32) when a batch is a whole number of sectors
            constexpr bool Wide = V32 && (BEL * sizeof(Elem)) % 32 == 0;
            constexpr uint32_t PV = (Wide ? 32 : 16) / sizeof(Elem);  // elements per load
            constexpr uint32_t NVB = BEL / PV;                        // loads per batch
            using Vec = typename std::conditional<Wide, vec32, uint4>::type;
            // UB > 1: one batch holds UB whole units; else the unit is cut into GE / BEL batches and the
            // word under construction (`cur`) carries over from one batch to the next
            constexpr uint32_t NU = UB;
            static_assert(UB == 1 || GE == BEL, "a multi-unit batch holds whole units");
            W cur[NU];
#pragma unroll
            for (uint32_t j = 0; j < NU; ++j) cur[j] = 0;
#pragma unroll
            for (uint32_t b0 = 0; b0 < GE; b0 += BEL) {
                union { Vec q; Elem x[PV]; } v[NU][NVB];
#pragma unroll
                for (uint32_t j = 0; j < NU; ++j) {
                    const uint32_t u = (it + j) * T + threadIdx.x;
                    const Vec* src = reinterpret_cast<const Vec*>(padded + (unit0 + u) * GE + b0);
#pragma unroll
                    for (uint32_t k = 0; k < NVB; ++k) {
                        if constexpr (Wide) v[j][k].q = ld32(src + k);
                        else v[j][k].q = ld16<CACHE_STREAMING>(src + k);
                    }
                }
#pragma unroll
                for (uint32_t j = 0; j < NU; ++j) {
                    const uint32_t u = (it + j) * T + threadIdx.x;
                    W* o = out + size_t(u) * CW;
#pragma unroll
                    for (uint32_t k = 0; k < NVB; ++k)
#pragma unroll
                        for (uint32_t i = 0; i < PV; ++i) deposit<W, S>(W(v[j][k].x[i]), b0 + k * PV + i, cur[j], o);
                }
            }
        } else {
            // the truncating field constructor (pint.hpp:770-774, 592-601, 390-407), four elements per
            // plane load; plane values are signed integers converted to Integer: sign-extend, then truncate
            constexpr uint32_t NQ = BEL / 4;
            constexpr uint32_t NU = UB;
            static_assert(UB == 1 || GE == BEL, "a multi-unit batch holds whole units");
            W cur[NU];
#pragma unroll
            for (uint32_t j = 0; j < NU; ++j) cur[j] = 0;
#pragma unroll
            for (uint32_t b0 = 0; b0 < GE; b0 += BEL) {
                W e[NU][BEL];
#pragma unroll
                for (uint32_t j = 0; j < NU; ++j)
#pragma unroll
                    for (uint32_t i = 0; i < BEL; ++i) e[j][i] = 0;
                for (uint32_t fi = 0; fi < f.t.nfields; ++fi) {
                    const W m = ones_w<W>(f.t.bits[fi]);
                    const uint32_t off = f.t.off[fi];
                    const void* plane = f.planes.p[fi];
                    uint64_t pv[NU][NQ][4];
#pragma unroll
                    for (uint32_t j = 0; j < NU; ++j) {
                        const uint32_t u = (it + j) * T + threadIdx.x;
                        const size_t q0 = ((unit0 + u) * GE + b0) / 4;
                        load_plane_quads<PlaneBytes, NQ, V32>(plane, q0, pv[j]);
                    }
#pragma unroll
                    for (uint32_t j = 0; j < NU; ++j)
#pragma unroll
                        for (uint32_t q = 0; q < NQ; ++q)
#pragma unroll
                            for (uint32_t i = 0; i < 4; ++i) e[j][4 * q + i] |= W(W(pv[j][q][i]) & m) << off;
                }
#pragma unroll
                for (uint32_t j = 0; j < NU; ++j) {
                    const uint32_t u = (it + j) * T + threadIdx.x;
                    W* o = out + size_t(u) * CW;
#pragma unroll
                    for (uint32_t i = 0; i < BEL; ++i) deposit<W, S>(e[j][i], b0 + i, cur[j], o);
                }
            }
        }
    }
    const uint32_t tile_bytes = units_per_tile * Geo::unit_out_bytes;  // a multiple of 128 (T is a multiple of 32)
    asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
    __syncthreads();
    if (threadIdx.x == 0) {
        bulk_store(reinterpret_cast<unsigned char*>(dense) + size_t(blockIdx.x) * tile_bytes, out, tile_bytes, evict_first_policy());
        bulk_wait_read<0>();
    }
}

// Units of more than 64 bytes of input (odd S: 32 or 64 elements per unit) make the direct form's
// loads one whole 128-byte line per lane and instruction (measured 3.5-5.1 TB/s, r03c_dense.jsonl).
// For those the tile's elements are first staged in shared memory with coalesced 16-byte loads
// (assembled from the planes on the way for SIDE_PLANES), element i at index i + i / GE so that
// the thread owning a unit walks a row of odd stride; the deposit is the compile-time one above.
template <class Elem, uint32_t S, int Side, int PlaneBytes>
__global__ void __launch_bounds__(256)
dense_insert_fixed_staged_kernel(const Elem* padded, uint4* dense, uint32_t units_per_tile, const __grid_constant__ dense_fields f)
{
    pdl_prologue();
    using Geo = fixed_geom<S, (Side == SIDE_PADDED ? uint32_t(sizeof(Elem)) : uint32_t(PlaneBytes))>;
    using W = typename Geo::W;
    using Staged = typename std::conditional<Side == SIDE_PADDED, Elem, W>::type;
    constexpr uint32_t GE = Geo::GE, CW = Geo::CW;
    // 64-bit words: a unit is up to 64 elements and one thread per unit leaves most of the CTA idle
    // (3.5 TB/s); there every thread assembles whole output WORDS instead (word w of the tile starts
    // at element 64 w / S, a multiply-high), and the staged elements need no row padding
    constexpr bool WordOwner = sizeof(W) == 8;
    constexpr uint32_t PAD = WordOwner ? 0 : 1;
    extern __shared__ __align__(128) unsigned char dense_smem[];
    const uint32_t T = blockDim.x;
    const uint32_t tile_elems = units_per_tile * GE;
    const uint32_t tile_bytes = units_per_tile * Geo::unit_out_bytes;  // a multiple of 128 (units_per_tile is one of 32)
    W* out = reinterpret_cast<W*>(dense_smem);  // unit owners only; word owners store straight to global memory
    Staged* in = reinterpret_cast<Staged*>(dense_smem + (WordOwner ? 0 : tile_bytes));
    const size_t first = size_t(blockIdx.x) * tile_elems;
    // four independent 16-byte loads (per plane) in flight per thread; the launcher makes the tile a
    // multiple of 4 * T vectors / quads
    if constexpr (Side == SIDE_PADDED) {
        constexpr uint32_t PV = 16 / sizeof(Elem);
        const uint4* src = reinterpret_cast<const uint4*>(padded + first);
        const uint32_t nv = tile_elems / PV;
        for (uint32_t v0 = threadIdx.x; v0 < nv; v0 += 4 * T) {
            union { uint4 q; Elem x[PV]; } u[4];
#pragma unroll
            for (uint32_t j = 0; j < 4; ++j)
                if (v0 + j * T < nv) u[j].q = ld16<CACHE_STREAMING>(src + v0 + j * T);
#pragma unroll
            for (uint32_t j = 0; j < 4; ++j) {
                if (v0 + j * T >= nv) break;
                const uint32_t i = (v0 + j * T) * PV;
                Staged* dst = in + i + PAD * (i / GE);  // PV <= GE: the vector stays in one row
#pragma unroll
                for (uint32_t k = 0; k < PV; ++k) dst[k] = u[j].x[k];
            }
        }
    } else {
        // the truncating field constructor (pint.hpp:770-774, 592-601, 390-407) on four elements at a time
        // quads in flight per thread: 64-bit words double the registers per value (89 registers with four)
        constexpr uint32_t LIF = sizeof(W) == 8 ? 2 : 4;
        const uint32_t nq = tile_elems / 4;
        for (uint32_t q0 = threadIdx.x; q0 < nq; q0 += LIF * T) {
            W e[LIF][4];
#pragma unroll
            for (uint32_t j = 0; j < LIF; ++j)
#pragma unroll
                for (uint32_t k = 0; k < 4; ++k) e[j][k] = 0;
            for (uint32_t fi = 0; fi < f.t.nfields; ++fi) {
                uint64_t v[LIF][4];
#pragma unroll
                for (uint32_t j = 0; j < LIF; ++j)
                    if (q0 + j * T < nq) load_plane_quad<PlaneBytes>(f.planes.p[fi], first / 4 + q0 + j * T, v[j]);
                const W m = ones_w<W>(f.t.bits[fi]);
                const uint32_t off = f.t.off[fi];
#pragma unroll
                for (uint32_t j = 0; j < LIF; ++j)
#pragma unroll
                    for (uint32_t k = 0; k < 4; ++k) e[j][k] |= W(W(v[j][k]) & m) << off;
            }
#pragma unroll
            for (uint32_t j = 0; j < LIF; ++j) {
                if (q0 + j * T >= nq) break;
                const uint32_t i = 4 * (q0 + j * T);
                Staged* dst = in + i + PAD * (i / GE);
#pragma unroll
                for (uint32_t k = 0; k < 4; ++k) dst[k] = e[j][k];
            }
        }
    }
    __syncthreads();
    if constexpr (WordOwner) {
        constexpr uint32_t WB = Geo::WB;
        constexpr uint32_t MAXE = (WB - 1 + S - 1) / S;  // elements after the first that can reach into one word
        const W SM = ones_w<W>(S);
        const uint32_t nwords = units_per_tile * CW;
        for (uint32_t w = threadIdx.x; w < nwords; w += T) {
            const uint32_t p = w * WB;  // tile_elems * S < 2^32
            const uint32_t i = p / S, off = p - i * S;
            W r = W(W(in[i]) & SM) >> off;
            uint32_t have = S - off;
#pragma unroll
            for (uint32_t k = 1; k <= MAXE; ++k)
                if (have < WB) {  // the tile ends on a word boundary: never reads past its last element
                    r |= W(W(in[i + k]) & SM) << have;
                    have += S;
                }
            st_stream(reinterpret_cast<W*>(dense) + size_t(blockIdx.x) * nwords + w, r);  // consecutive lanes, consecutive words
        }
        return;
    } else {
        for (uint32_t u = threadIdx.x; u < units_per_tile; u += T) {
            const Staged* row = in + size_t(u) * (GE + 1);
            W* o = out + size_t(u) * CW;
            W cur = 0;
#pragma unroll
            for (uint32_t e = 0; e < GE; ++e) deposit<W, S>(W(row[e]), e, cur, o);
        }
    }
    asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
    __syncthreads();
    if (threadIdx.x == 0) {
        bulk_store(reinterpret_cast<unsigned char*>(dense) + size_t(blockIdx.x) * tile_bytes, out, tile_bytes, evict_first_policy());
        bulk_wait_read<0>();
    }
}

// Whole tiles only: *done = the elements converted (a multiple of 128); the caller finishes
// [*done, n) with the run-time-geometry kernel.  Needs 16-byte aligned inputs (checked by the caller).
template <class Elem, uint32_t S, int Side, int PlaneBytes>
cudaError_t launch_dense_insert_fixed(const void* other, void* dense, size_t n, const dense_fields& f, cudaStream_t stream,
                                      const launch_geom& g, size_t* done)
{
    using Geo = fixed_geom<S, (Side == SIDE_PADDED ? uint32_t(sizeof(Elem)) : uint32_t(PlaneBytes))>;
    if constexpr (Geo::unit_in_bytes > 64) {
        using Staged = typename std::conditional<Side == SIDE_PADDED, Elem, typename Geo::W>::type;
        // units per tile: ~8 KB of output, at most ~16 KB of staged input, a whole number of load rounds
        // (4 vectors or quads per thread: 2048 elements cover every source width)
        const uint32_t by_out = 8192 / Geo::unit_out_bytes, by_in = 16384 / ((Geo::GE + 1) * uint32_t(sizeof(Staged)));
        const uint32_t round = 2048 / Geo::GE > 32 ? 2048 / Geo::GE : 32;
        uint32_t U = (by_out < by_in ? by_out : by_in) / round * round;
        if (U < round) U = round;
        const bool word_owner = sizeof(typename Geo::W) == 8;  // as in the kernel: no output tile, no row padding
        const size_t smem = word_owner ? size_t(U) * Geo::GE * sizeof(Staged)
                                       : size_t(U) * Geo::unit_out_bytes + size_t(U) * (Geo::GE + 1) * sizeof(Staged);
        const size_t tile_elems = size_t(U) * Geo::GE;
        const size_t ntiles = n / tile_elems;
        *done = ntiles * tile_elems;
        if (ntiles == 0) return cudaSuccess;
        if (ntiles > 0x7FFFFFFFull) return cudaErrorInvalidValue;
        if (cudaError_t e = launch_stream_kernel_smem(dense_insert_fixed_staged_kernel<Elem, S, Side, PlaneBytes>, unsigned(ntiles), word_owner ? 256u : 128u,
                                                      smem, stream, g.pdl != 0, static_cast<const Elem*>(other), static_cast<uint4*>(dense), U, f))
            return e;
        count_launch(g);
        return cudaGetLastError();
    } else {
    // CTA size: the largest whose one-unit-per-thread tile stays within 16 KB; then units per thread
    // to bring the output tile to ~16 KB
    unsigned T = 256;
    while (T > 32 && T * Geo::unit_out_bytes > 16384) T >>= 1;
    // a thread takes at least UB units per round (narrow planes: up to 64 elements in flight): keep that tile
    // within the 48 KB a kernel gets without opting in
    while (T > 32 && T * Geo::UB * Geo::unit_out_bytes > 32768) T >>= 1;
    uint32_t upt = 16384 / (T * Geo::unit_out_bytes) / Geo::UB * Geo::UB;
    if (upt < Geo::UB) upt = Geo::UB;
    const size_t smem = size_t(T) * upt * Geo::unit_out_bytes;
    const size_t tile_elems = size_t(T) * upt * Geo::GE;
    const size_t ntiles = n / tile_elems;
    *done = ntiles * tile_elems;
    if (ntiles == 0) return cudaSuccess;
    if (ntiles > 0x7FFFFFFFull) return cudaErrorInvalidValue;
    // 256-bit loads when every input is 32-byte aligned (and option vec256 != 0)
    uintptr_t align = reinterpret_cast<uintptr_t>(other);
    if (Side == SIDE_PLANES)
        for (uint32_t i = 0; i < f.t.nfields; ++i) align |= reinterpret_cast<uintptr_t>(f.planes.p[i]);
    constexpr uint32_t src_bytes = Side == SIDE_PADDED ? uint32_t(sizeof(Elem)) : uint32_t(PlaneBytes);
    // same-box A/B (profiles/r03r_ab_insert_vec256.log): +10-15 % for 64-bit words (48-bit pack 6198 -> 6818,
    // 40-bit from_planes 6105 -> 7014 GB/s), 0 to -3 % for 32-bit words: instantiated for the former only
    bool wide = false;
    if constexpr (sizeof(typename Geo::W) == 8 && (Geo::BEL * src_bytes) % 32 == 0) {
        if (g.vec256 != 0 && align % 32 == 0) {
            wide = true;
            if (cudaError_t e = launch_stream_kernel_smem(dense_insert_fixed_kernel<Elem, S, Side, PlaneBytes, true>, unsigned(ntiles), T, smem, stream,
                                                          g.pdl != 0, static_cast<const Elem*>(other), static_cast<uint4*>(dense), upt, f))
                return e;
        }
    }
    if (!wide)
        if (cudaError_t e = launch_stream_kernel_smem(dense_insert_fixed_kernel<Elem, S, Side, PlaneBytes, false>, unsigned(ntiles), T, smem, stream,
                                                      g.pdl != 0, static_cast<const Elem*>(other), static_cast<uint4*>(dense), upt, f))
            return e;
    count_launch(g);
    return cudaGetLastError();
    }
}

// The other direction.  Owning whole units does not pay here: a lane's 16-byte stores would be
// 32-128 bytes apart and partial-sector stores cost more than they save (measured: slower than the
// run-time-geometry kernel for every unit above 16 bytes, profiles/r03k_dense_extract_unit_owner_experiment.jsonl).
// Instead a thread owns a GROUP of GP consecutive elements -- 16 bytes of padded output (32 for
// uint64_t), 4 values per plane -- so that consecutive lanes store consecutive vectors.  The group
// starts at bit g * GP * S of the TMA-staged tile: the thread loads the NW words that cover it ONCE,
// re-aligns them to the group start with NW - 1 funnel shifts by its own offset, and every element
// is then cut out with IMMEDIATE shifts.  The run-time-geometry kernel loads two words per element
// and costs 14 (padded) to 22 (one plane) instructions per element, 66-75 % issue-active on 12-bit
// elements (profiles/r03j_ncu_convert_summary.txt).  Instantiated for dense -> padded with uint8_t /
// uint16_t / uint64_t elements, where it wins on the same box (profiles/r03m_ab_dense_convert.log:
// 7-bit 3662 -> 5454, 12-bit 5669 -> 6035, 63-bit 5277 -> 6785 GB/s; 32-byte stores for uint64_t);
// for uint32_t elements and for planes the run-time-geometry kernel stays 0-2 % ahead and is kept.
template <class Elem, uint32_t S, int Side, int PlaneBytes = 4>
struct group_geom {
    using W = typename std::conditional<(S <= 32), uint32_t, uint64_t>::type;
    static constexpr uint32_t WB = 8 * sizeof(W);
    // planes of 1 or 2 bytes: 16 / PlaneBytes elements per thread, so that every plane store is one 16-byte vector
    // (dense RGB24 -> three uint8 planes ran at 4.6 TB/s with quads: 4-byte stores, issue-bound)
    static constexpr uint32_t GP = Side == SIDE_PADDED ? (sizeof(Elem) >= 4 ? 4u : 16u / uint32_t(sizeof(Elem)))
                                                       : (PlaneBytes <= 2 ? 16u / uint32_t(PlaneBytes) : 4u);
    static constexpr uint32_t NA = (GP * S + WB - 1) / WB;  // aligned words that hold the group
    static constexpr uint32_t NW = NA + 1;                  // words loaded: the group may start anywhere in the first
};

template <class Elem, uint32_t S, int Side, int PlaneBytes, bool Signed>
__global__ void __launch_bounds__(256)
dense_extract_fixed_kernel(const uint4* dense, Elem* padded, uint32_t tile_elems, const __grid_constant__ dense_fields f,
                           uint32_t narrow_signed = 0)
{
    pdl_prologue();
    using Geo = group_geom<Elem, S, Side, PlaneBytes>;
    using W = typename Geo::W;
    constexpr uint32_t GP = Geo::GP, NA = Geo::NA, NW = Geo::NW, WB = Geo::WB;
    extern __shared__ __align__(128) unsigned char dense_smem[];
    __shared__ __align__(8) uint64_t full;
    W* D = reinterpret_cast<W*>(dense_smem);
    const uint32_t T = blockDim.x;
    const uint32_t tile_bytes = tile_elems / 8 * S;  // tile_elems is a multiple of 128
    const size_t first = size_t(blockIdx.x) * tile_elems;
    if (threadIdx.x == 0) {
        mbar_init(&full, 1);
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
        asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
        mbar_expect_tx(&full, tile_bytes);
        bulk_load(dense_smem, reinterpret_cast<const unsigned char*>(dense) + size_t(blockIdx.x) * tile_bytes, tile_bytes, &full,
                  evict_first_policy());
    }
    // the last group's extra word lies past the tile (its bits are never used): keep it defined
    if (threadIdx.x < 16 / sizeof(W)) D[tile_bytes / sizeof(W) + threadIdx.x] = 0;
    __syncthreads();
    mbar_wait(&full, 0);
    const W SM = ones_w<W>(S);
    const uint32_t ngroups = tile_elems / GP;
    for (uint32_t g = threadIdx.x; g < ngroups; g += T) {
        const uint32_t p0 = g * (GP * S);  // tile_elems * S < 2^32
        const uint32_t w0 = p0 / WB, off = p0 % WB;
        W wd[NW];
#pragma unroll
        for (uint32_t k = 0; k < NW; ++k) wd[k] = D[w0 + k];
        W al[NA];
#pragma unroll
        for (uint32_t k = 0; k < NA; ++k) al[k] = funnel_r(wd[k], wd[k + 1], off);
        W x[GP];
#pragma unroll
        for (uint32_t e = 0; e < GP; ++e) {
            const uint32_t p = e * S, sh = p % WB, w = p / WB;  // immediates once unrolled
            W v = al[w] >> sh;
            if (sh + S > WB) v |= W(al[w + 1] << (WB - sh));
            x[e] = v & SM;
        }
        const size_t e0 = first + size_t(g) * GP;
        if constexpr (Side == SIDE_PADDED) {
            if constexpr (sizeof(Elem) == 8) {
                const uint64_t y0 = x[0], y1 = x[1], y2 = x[2], y3 = x[3];
                vec32 v;
                v.lo = make_uint4(uint32_t(y0), uint32_t(y0 >> 32), uint32_t(y1), uint32_t(y1 >> 32));
                v.hi = make_uint4(uint32_t(y2), uint32_t(y2 >> 32), uint32_t(y3), uint32_t(y3 >> 32));
                st32(reinterpret_cast<vec32*>(padded + e0), v);
            } else {
                constexpr uint32_t PV = 16 / sizeof(Elem);
                static_assert(PV == GP, "one 16-byte store per group");
                union { uint4 q; Elem x[PV]; } v;
#pragma unroll
                for (uint32_t i = 0; i < PV; ++i) v.x[i] = Elem(x[i]);
                st16<CACHE_STREAMING>(reinterpret_cast<uint4*>(padded + e0), v.q);
            }
        } else if constexpr (PlaneBytes <= 2) {
            // GP values per plane as ONE 16-byte store; get / get_signed (pint.hpp:799-822) narrowed to the plane
            // type as a C cast does.  The sign is a run-time (uniform) flag here: (v ^ m) - m with m = the field's
            // sign bit extends, m = 0 leaves the unsigned value.
            for (uint32_t fi = 0; fi < f.t.nfields; ++fi) {
                void* plane = f.planes.p[fi];
                if (!plane) continue;
                const uint32_t foff = f.t.off[fi], wdt = f.t.bits[fi];
                const W fm = ones_w<W>(wdt);
                const W sm = narrow_signed ? W(W(1) << (wdt - 1)) : W(0);
                uint32_t o[4] = {0, 0, 0, 0};
#pragma unroll
                for (uint32_t i = 0; i < GP; ++i) {
                    const W u = W(x[i] >> foff) & fm;
                    const uint32_t v = uint32_t(W(u ^ sm) - sm);
                    if constexpr (PlaneBytes == 1) o[i >> 2] |= (v & 0xFFu) << (8 * (i & 3));
                    else o[i >> 1] |= (v & 0xFFFFu) << (16 * (i & 1));
                }
                st16<CACHE_STREAMING>(reinterpret_cast<uint4*>(static_cast<unsigned char*>(plane) + e0 * PlaneBytes), make_uint4(o[0], o[1], o[2], o[3]));
            }
        } else {
            for (uint32_t fi = 0; fi < f.t.nfields; ++fi) {
                void* plane = f.planes.p[fi];
                if (!plane) continue;
                const uint32_t foff = f.t.off[fi], wdt = f.t.bits[fi];
                uint64_t v[4];
#pragma unroll
                for (uint32_t i = 0; i < 4; ++i)
                    v[i] = Signed ? uint64_t(int64_t(field_get_signed<W>(x[i], foff, wdt))) : uint64_t(field_get<W>(x[i], foff, wdt));
                store_plane_quad<PlaneBytes>(plane, e0 / 4, v);
            }
        }
    }
}

// Whole tiles only, as for the insert kernels.  The caller checks alignment: 16 bytes (32 for a
// uint64_t padded array, which is stored with 32-byte vectors).
template <class Elem, uint32_t S, int Side, int PlaneBytes, bool Signed>
cudaError_t launch_dense_extract_fixed(const void* dense, void* other, size_t n, const dense_fields& f, cudaStream_t stream,
                                       const launch_geom& g, size_t* done, uint32_t narrow_signed = 0)
{
    using W = typename group_geom<Elem, S, Side, PlaneBytes>::W;
    const uint32_t tile_elems = convert_tile_elems(S);  // ~16 KB of stream, a multiple of 128 elements
    const size_t smem = size_t(tile_elems) / 8 * S + 16;
    const size_t ntiles = n / tile_elems;
    *done = ntiles * size_t(tile_elems);
    if (ntiles == 0) return cudaSuccess;
    if (ntiles > 0x7FFFFFFFull) return cudaErrorInvalidValue;
    (void)sizeof(W);
    if (cudaError_t e = launch_stream_kernel_smem(dense_extract_fixed_kernel<Elem, S, Side, PlaneBytes, Signed>, unsigned(ntiles), 256, smem, stream,
                                                  g.pdl != 0, static_cast<const uint4*>(dense), static_cast<Elem*>(other), tile_elems, f, narrow_signed))
        return e;
    count_launch(g);
    return cudaGetLastError();
}

}  // namespace dense
}  // namespace kernels
}  // namespace cuda
}  // namespace pint
