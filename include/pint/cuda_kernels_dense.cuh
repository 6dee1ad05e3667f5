// pint/cuda_kernels_dense.cuh -- kernels for the DENSE bit-stream storage format
// (SURVEY.md 8(f)4): element i of a layout with S = sum(Bits) bits occupies bits
// [i*S, (i+1)*S) of a little-endian bit stream, with no padding to
// find_appropriate_int (pint.hpp:724-734).  Elements -- and fields -- straddle
// word boundaries here, and a packed array costs S/8 bytes per element instead
// of sizeof(Integer): 3 instead of 4 bytes for <8,8,8>, 2.375 instead of 4 for
// <5,6,5,3>.  Per-element semantics are unchanged: a dense op is
// pack(op(unpack(a), unpack(b))) with op the SWAR functor of cuda_kernels.cuh.
//
// Data movement.  A straddling layout has no fixed relation between threads and
// 16-byte vectors, so tiles are staged through shared memory by the TMA engine
// (1-D bulk copies, cp.async.bulk + mbarrier complete_tx; the result tile goes
// back with cp.async.bulk.global.shared::cta).  Global traffic is therefore
// fully coalesced whatever S is, and costs no per-thread address arithmetic.
//
// Thread mapping for dense -> dense ops ("chunk" mapping).  The bit pattern of a
// layout repeats every lcm(S, WB) bits (WB = bits of the processing word): a
// CHUNK of Cw = S / gcd(S, WB) words holds exactly G = WB / gcd(S, WB) whole
// elements.  One thread owns a chunk, so no element is shared between threads:
// no shuffles, no atomics.  Cw is odd for every S that is not 8, 16, 32 or 64
// (those ARE the padded format), so the chunk stride is conflict-free in shared
// memory.  Two ways to compute a chunk:
//   * whole-chunk SWAR (cuda_kernels_dense_mw.cuh; chunks of up to 19 32-bit
//     words): the chunk is treated as ONE wide packed_int and the SWAR functor
//     runs on it directly with multi-word carries and shifts;
//   * sliding windows (below; longer chunks): the thread walks the chunk with a
//     two-word window, extracts each element with a funnel shift, applies the
//     functor, and deposits the result into a two-word output window flushed word
//     by word over the input tile; all shifts are the same for every thread
//     (immediates when the layout is compiled in).
//
// Thread mapping for format conversions (cuda_kernels_dense_convert.cuh): dense ->
// padded/planes takes one thread per four ELEMENTS (funnel shifts at per-lane offsets
// out of the staged tile, 16-byte stores); padded/planes -> dense stages the inputs in
// shared memory and uses the chunk mapping again, so both sides stay coalesced and
// nothing needs atomics.
#pragma once
#ifndef PINT_CUDA_KERNELS_DENSE_CUH_INCLUDED
#define PINT_CUDA_KERNELS_DENSE_CUH_INCLUDED

#include "cuda_kernels.cuh"
#include "cuda_kernels_bulk.cuh"

namespace pint {
namespace cuda {
namespace kernels {
namespace dense {

using bulk::bulk_load;
using bulk::bulk_store;
using bulk::bulk_wait_read;
using bulk::evict_first_policy;
using bulk::mbar_expect_tx;
using bulk::mbar_init;
using bulk::mbar_wait;

constexpr uint32_t gcd_u32(uint32_t a, uint32_t b) { return b == 0 ? a : gcd_u32(b, a % b); }

template <class W>
__host__ __device__ __forceinline__ constexpr W ones_w(uint32_t n)
{
    return n >= 8 * sizeof(W) ? ~W(0) : W((W(1) << n) - 1);
}

// low word of (hi:lo) >> sh, sh in [0, WB)
template <class W>
__device__ __forceinline__ W funnel_r(W lo, W hi, uint32_t sh)
{
    if constexpr (sizeof(W) == 4) return __funnelshift_r(lo, hi, sh);
    else return (lo >> sh) | ((hi << 1) << (63u - sh));
}

// the bits of (r << sh) that leave the word: r >> (WB - sh), 0 when sh == 0
template <class W>
__device__ __forceinline__ W spill_l(W r, uint32_t sh)
{
    if constexpr (sizeof(W) == 4) return __funnelshift_l(r, 0u, sh);
    else return (r >> 1) >> (63u - sh);
}

// ---------------------------------------------------------------------------
// Chunk geometry: compile-time (static layouts) or run-time (constant bank).
// EB = bits of one SLOT of the processing word: elements narrower than 17 (9) bits
// are gathered two (four) to a 32-bit word at slot boundaries, which is exactly the
// replicated-mask form the SWAR functors use for uint16_t (uint8_t) packs, so one
// functor call serves 2 (4) elements.

template <class W, uint32_t S_, uint32_t EB_ = 8 * sizeof(W)>
struct static_geom {
    static constexpr bool is_static = true;
    static constexpr bool whole_chunk = false;
    static constexpr bool wide64 = false;
    using unit = W;  // the tile is handled in words of this type
    static constexpr uint32_t WB = 8 * sizeof(W);
    static constexpr uint32_t kS = S_, kCw = S_ / gcd_u32(S_, WB), kG = WB / gcd_u32(S_, WB), kEB = EB_, kEPW = WB / EB_;
    static_assert(kG % kEPW == 0, "a chunk holds a whole number of gathered words");
    uint32_t chunks_per_thread;
    __host__ __device__ static constexpr uint32_t S() { return kS; }
    __host__ __device__ static constexpr uint32_t Cw() { return kCw; }
    __host__ __device__ static constexpr uint32_t G() { return kG; }
    __host__ __device__ static constexpr uint32_t EB() { return kEB; }
};

template <class W>
struct runtime_geom {
    static constexpr bool is_static = false;
    static constexpr bool whole_chunk = false;
    static constexpr bool wide64 = false;
    using unit = W;
    static constexpr uint32_t WB = 8 * sizeof(W);
    uint32_t s, cw, g, eb;
    uint32_t chunks_per_thread;
    runtime_geom() = default;
    // slot_bits: 8 * sizeof(Integer) of the layout the masks were generated for
    runtime_geom(uint32_t S_, uint32_t slot_bits)
        : s(S_), cw(S_ / gcd_u32(S_, WB)), g(WB / gcd_u32(S_, WB)), eb(slot_bits), chunks_per_thread(1) {}
    __host__ __device__ uint32_t S() const { return s; }
    __host__ __device__ uint32_t Cw() const { return cw; }
    __host__ __device__ uint32_t G() const { return g; }
    __host__ __device__ uint32_t EB() const { return eb; }
};

// 33..63-bit elements (64-bit functors) over a tile of 32-bit words: the bit pattern repeats every
// lcm(S, 32) bits, half the 64-bit-word chunk for odd S, so twice as many threads fit beside the
// staged tiles (S = 63: 252 instead of 504 bytes of each operand per thread).
struct wide_geom {
    static constexpr bool is_static = false;
    static constexpr bool whole_chunk = false;
    static constexpr bool wide64 = true;
    using unit = uint32_t;
    uint32_t s, cw, g;
    uint32_t chunks_per_thread;
    wide_geom() = default;
    explicit wide_geom(uint32_t S_) : s(S_), cw(S_ / gcd_u32(S_, 32)), g(32 / gcd_u32(S_, 32)), chunks_per_thread(1) {}
    __host__ __device__ uint32_t S() const { return s; }
    __host__ __device__ uint32_t Cw() const { return cw; }  // 32-bit words
    __host__ __device__ uint32_t G() const { return g; }
};

}  // namespace dense
}  // namespace kernels
}  // namespace cuda
}  // namespace pint

#include "cuda_kernels_dense_mw.cuh"  // whole-chunk SWAR for compiled-in layouts with small chunks

namespace pint {
namespace cuda {
namespace kernels {
namespace dense {

// One chunk, in place: A[0..Cw) <- op(A, B) element by element.  A and B are
// shared-memory pointers to the same chunk of the two operand tiles.  An input
// window (two words per operand) and an output window (two words) slide over the
// chunk; the output never runs ahead of the input, and a word is written only
// after every element that touches it has been read, so in place is safe.
template <class Fn, class Geo>
__device__ __forceinline__ void process_chunk(typename Fn::word* A, const typename Fn::word* B, const Fn& fn, const Geo& geo)
{
    using W = typename Fn::word;
    constexpr uint32_t WB = 8 * sizeof(W);
    const uint32_t S = geo.S(), Cw = geo.Cw(), EB = geo.EB();
    const W SM = ones_w<W>(S);
    W ca = A[0], cb = B[0];
    W na = Cw > 1 ? A[1] : W(0), nb = Cw > 1 ? B[1] : W(0);
    W ocur = 0, onxt = 0;
    uint32_t shi = 0, wi = 0, sho = 0, wo = 0;
    auto take = [&](W& xa, W& xb) {
        xa = funnel_r(ca, na, shi);
        xb = funnel_r(cb, nb, shi);
        shi += S;
        if (shi >= WB) {  // the window moves on by one word (S <= WB: never by more)
            shi -= WB;
            ++wi;
            ca = na;
            cb = nb;
            na = wi + 1 < Cw ? A[wi + 1] : W(0);
            nb = wi + 1 < Cw ? B[wi + 1] : W(0);
        }
    };
    auto put = [&](W r) {
        ocur |= r << sho;
        onxt |= spill_l(r, sho);
        sho += S;
        if (sho >= WB) {
            sho -= WB;
            A[wo] = ocur;
            ocur = onxt;
            onxt = 0;
            ++wo;
        }
    };
    auto group = [&]() {  // run-time slots: 2 or 4 elements per functor call
        W xa = 0, xb = 0;
        for (uint32_t k = 0; k < WB; k += EB) {
            W ta, tb;
            take(ta, tb);
            xa |= (ta & SM) << k;
            xb |= (tb & SM) << k;
        }
        const W r = fn(xa, xb);
        for (uint32_t k = 0; k < WB; k += EB) put((r >> k) & SM);
    };
    if constexpr (Geo::is_static) {
#pragma unroll
        for (uint32_t e = 0; e < Geo::kG; e += Geo::kEPW) {
            if constexpr (Geo::kEPW == 1) {
                // Bits above S of xa / xb belong to the next element.  The functors ignore them for
                // add/sub (M1 and H cover field bits only), pass them through for | & ^ and take them
                // from b for min/max, so masking the RESULT is enough.
                W xa, xb;
                take(xa, xb);
                put(fn(xa, xb) & SM);
            } else {
                W xa = 0, xb = 0;
#pragma unroll
                for (uint32_t k = 0; k < Geo::kEPW; ++k) {
                    W ta, tb;
                    take(ta, tb);
                    xa |= (ta & SM) << (k * Geo::kEB);
                    xb |= (tb & SM) << (k * Geo::kEB);
                }
                const W r = fn(xa, xb);
#pragma unroll
                for (uint32_t k = 0; k < Geo::kEPW; ++k) put((r >> (k * Geo::kEB)) & SM);
            }
        }
    } else if (EB == WB) {
        // one element per word: input and output windows move together -- one shift, one branch
        const uint32_t G = geo.G();
        uint32_t sh = 0, w = 0;
        for (uint32_t e = 0; e < G; ++e) {
            const W r = fn(funnel_r(ca, na, sh), funnel_r(cb, nb, sh)) & SM;
            ocur |= r << sh;
            onxt |= spill_l(r, sh);
            sh += S;
            if (sh >= WB) {
                sh -= WB;
                A[w] = ocur;
                ocur = onxt;
                onxt = 0;
                ++w;
                ca = na;
                cb = nb;
                na = w + 1 < Cw ? A[w + 1] : W(0);
                nb = w + 1 < Cw ? B[w + 1] : W(0);
            }
        }
    } else {
        const uint32_t G = geo.G(), step = WB / EB;
        for (uint32_t e = 0; e < G; e += step) group();
    }
    // G * S == Cw * WB: the last element ends exactly on the chunk's last bit, so the last
    // word was flushed by put() and both windows are back at shift 0.
}

// 33..63-bit elements, tile seen as 32-bit words (wide_geom).  64-bit funnel shifts by run-time
// amounts are 2-3 instructions each on the 32-bit datapath and made the 64-bit-word form of this
// path ALU-bound at 4.1-4.9 TB/s (profiles/r03u_ncu_windows_summary.txt: 83-88 % ALU pipe).  On
// 32-bit words an element is two native funnel shifts over three words and its result three shifts
// into a three-word output window; an element always completes the word it starts in (S > 32) and
// sometimes the next.  Word w + 2 of the chunk's last element is the word after the chunk: read
// but never used (the launcher pads the tiles by 16 bytes).  In place as process_chunk: a word is
// written only after every element that touches it has been read.
template <class Fn>
__device__ __forceinline__ void process_chunk_wide64(uint32_t* A32, const uint32_t* B32, const Fn& fn, const wide_geom& geo)
{
    static_assert(sizeof(typename Fn::word) == 8, "64-bit functor");
    const uint32_t S = geo.S(), G = geo.G();
    const uint64_t SM = ones_w<uint64_t>(S);
    uint32_t o0 = 0, o1 = 0, p = 0;
    for (uint32_t e = 0; e < G; ++e) {
        const uint32_t w = p >> 5, sh = p & 31u;
        const uint32_t a0 = A32[w], a1 = A32[w + 1], a2 = A32[w + 2];
        const uint32_t b0 = B32[w], b1 = B32[w + 1], b2 = B32[w + 2];
        const uint64_t xa = (uint64_t(__funnelshift_r(a1, a2, sh)) << 32) | __funnelshift_r(a0, a1, sh);
        const uint64_t xb = (uint64_t(__funnelshift_r(b1, b2, sh)) << 32) | __funnelshift_r(b0, b1, sh);
        const uint64_t r = fn(xa, xb) & SM;
        const uint32_t lo = uint32_t(r), hi = uint32_t(r >> 32);
        A32[w] = o0 | (lo << sh);
        const uint32_t m1 = o1 | __funnelshift_l(lo, hi, sh), m2 = __funnelshift_l(hi, 0u, sh);
        if (sh + S >= 64) {
            A32[w + 1] = m1;
            o0 = m2;
            o1 = 0;
        } else {
            o0 = m1;
            o1 = m2;
        }
        p += S;
    }
}

// Two results per element exist only for whole-chunk geometries (process_chunk_mw_pair): with the
// sliding windows a second output window costs more than the second pass over a and b saves
// (two launches 447 / 767 / 208 against one pass 411 / 660 / 179 Gelem-results/s for S = 31 / 21 / 63,
// profiles/r03p_ab_dense_pair.log), so the C ABI runs such layouts as two launches.

// One-in/one-out maps (shift_left / shift_right_unsigned) element by element: one input window,
// one output window, in place.
template <class W>
struct window_unary_fn {
    using word = W;
    unary_params<W> u;
    __device__ __forceinline__ W operator()(W v) const { return apply_unary(v, u); }
};

template <class Fn, class Geo>
__device__ __forceinline__ void process_chunk_unary(typename Fn::word* A, const Fn& fn, const Geo& geo)
{
    using W = typename Fn::word;
    constexpr uint32_t WB = 8 * sizeof(W);
    const uint32_t S = geo.S(), Cw = geo.Cw(), G = geo.G();
    const W SM = ones_w<W>(S);
    W ca = A[0], na = Cw > 1 ? A[1] : W(0);
    W ocur = 0, onxt = 0;
    uint32_t sh = 0, w = 0;
    for (uint32_t e = 0; e < G; ++e) {
        const W r = fn(funnel_r(ca, na, sh) & SM) & SM;
        ocur |= r << sh;
        onxt |= spill_l(r, sh);
        sh += S;
        if (sh >= WB) {
            sh -= WB;
            A[w] = ocur;
            ocur = onxt;
            onxt = 0;
            ++w;
            ca = na;
            na = w + 1 < Cw ? A[w + 1] : W(0);
        }
    }
}

// zero the bits of `word` (which starts at bit `start` of the tile) at or above bit `limit`
template <class W>
__device__ __forceinline__ W clip_word(W word, uint64_t start, uint64_t limit)
{
    constexpr uint32_t WB = 8 * sizeof(W);
    if (start >= limit) return 0;
    if (start + WB > limit) return word & ones_w<W>(uint32_t(limit - start));
    return word;
}

__device__ __forceinline__ uint4 clip_vec(uint4 v, uint64_t start_bit, uint64_t limit)
{
    v.x = clip_word<uint32_t>(v.x, start_bit, limit);
    v.y = clip_word<uint32_t>(v.y, start_bit + 32, limit);
    v.z = clip_word<uint32_t>(v.z, start_bit + 64, limit);
    v.w = clip_word<uint32_t>(v.w, start_bit + 96, limit);
    return v;
}

// ---------------------------------------------------------------------------
// out = op(a, b) on dense streams.  One tile per CTA (tiles are handed out in
// address order by the block scheduler, as in binary_stream_kernel); a tile is
// blockDim.x * chunks_per_thread chunks.  Tiles that lie entirely below bit
// n*S travel by TMA; the one tile that contains the end of the data is loaded
// and stored with guarded 16-byte vectors, with the bits at or above n*S
// forced to zero (op(0, 0) == 0 for every op, so the output tail is zero too).
// `out` may alias a or b: the tile is complete in shared memory before any
// store.  Dynamic shared memory: 2 * tile bytes.
// Mode 0: out = fn(a, b).  Mode 1 (pair): Fn is a pair_fn; the second result is produced in place over the
// b tile and stored to out1.  Mode 2 (unary): out = fn(a); b is not read and has no tile.
template <class Fn, class Geo, int Mode = 0>
__global__ void dense_binary_kernel(const uint4* a, const uint4* b, uint4* out, uint4* out1, size_t total_vec,
                                    uint64_t valid_bits, const __grid_constant__ Fn fn, const __grid_constant__ Geo geo)
{
    pdl_prologue();
    using W = typename Geo::unit;
    extern __shared__ __align__(128) unsigned char dense_smem[];
    __shared__ __align__(8) uint64_t full;
    const uint32_t T = blockDim.x, R = geo.chunks_per_thread, Cw = geo.Cw();
    const uint32_t tile_words = T * R * Cw;
    const uint32_t tile_bytes = tile_words * uint32_t(sizeof(W));
    const uint32_t tile_vecs = tile_bytes / 16;
    W* A = reinterpret_cast<W*>(dense_smem);
    W* B = reinterpret_cast<W*>(dense_smem + tile_bytes);
    const size_t t = blockIdx.x;
    const uint64_t tile_bits = uint64_t(tile_bytes) * 8;
    const bool whole = (t + 1) * tile_bits <= valid_bits;
    uint64_t pol = 0;
    constexpr bool Pair = Mode == 1, Unary = Mode == 2;

    if (whole) {
        if (threadIdx.x == 0) {
            pol = evict_first_policy();
            mbar_init(&full, 1);
            asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
            asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
            mbar_expect_tx(&full, (Unary ? 1 : 2) * tile_bytes);
            bulk_load(A, a + t * tile_vecs, tile_bytes, &full, pol);
            if constexpr (!Unary) bulk_load(B, b + t * tile_vecs, tile_bytes, &full, pol);
        }
        __syncthreads();  // the barrier is initialised before anyone polls it
        mbar_wait(&full, 0);
    } else {
        const uint64_t limit = valid_bits - t * tile_bits;  // first invalid bit, relative to the tile
        for (uint32_t v = threadIdx.x; v < tile_vecs; v += T) {
            const size_t gv = t * tile_vecs + v;
            uint4 x = make_uint4(0, 0, 0, 0), y = x;
            if (gv < total_vec) {
                x = clip_vec(ld16<CACHE_STREAMING>(a + gv), uint64_t(v) * 128, limit);
                if constexpr (!Unary) y = clip_vec(ld16<CACHE_STREAMING>(b + gv), uint64_t(v) * 128, limit);
            }
            reinterpret_cast<uint4*>(A)[v] = x;
            if constexpr (!Unary) reinterpret_cast<uint4*>(B)[v] = y;
        }
        __syncthreads();
    }

    for (uint32_t r = 0; r < R; ++r) {
        const uint32_t c = (r * T + threadIdx.x) * Cw;
        if constexpr (Unary) {
            static_assert(!Geo::wide64, "shifts use the 64-bit-word windows (HBM-bound as they are)");
            if constexpr (Geo::whole_chunk) process_chunk_mw_unary<Fn, int(Geo::kCw)>(A + c, fn);
            else process_chunk_unary(A + c, fn, geo);
        } else if constexpr (Geo::whole_chunk) {
            if constexpr (Pair) process_chunk_mw_pair<Fn, int(Geo::kCw)>(A + c, B + c, fn);
            else process_chunk_mw<Fn, int(Geo::kCw)>(A + c, B + c, fn);
        } else {
            static_assert(!Pair, "op pairs in one pass: whole-chunk geometries only");
            if constexpr (Geo::wide64) process_chunk_wide64(A + c, B + c, fn, geo);
            else process_chunk(A + c, B + c, fn, geo);
        }
    }

    if (whole) {
        asm volatile("fence.proxy.async.shared::cta;" ::: "memory");  // generic-proxy writes -> async proxy
        __syncthreads();
        if (threadIdx.x == 0) {
            bulk_store(out + t * tile_vecs, A, tile_bytes, pol);
            if constexpr (Pair) bulk_store(out1 + t * tile_vecs, B, tile_bytes, pol);
            bulk_wait_read<0>();  // shared memory must stay valid until the TMA engine has read it
        }
    } else {
        __syncthreads();
        for (uint32_t v = threadIdx.x; v < tile_vecs; v += T) {
            const size_t gv = t * tile_vecs + v;
            if (gv < total_vec) {
                st16<CACHE_STREAMING>(out + gv, reinterpret_cast<const uint4*>(A)[v]);
                if constexpr (Pair) st16<CACHE_STREAMING>(out1 + gv, reinterpret_cast<const uint4*>(B)[v]);
            }
        }
    }
}

struct dense_tuning {
    int threads = 0;     // 0 = choose from the chunk size
    int tile_bytes = 0;  // per operand; 0 = default target
};

template <class W>
inline void choose_tile(uint32_t cw, const dense_tuning& tune, uint32_t* threads, uint32_t* chunks_per_thread)
{
    const uint32_t chunk_bytes = cw * uint32_t(sizeof(W));
    uint32_t T = tune.threads > 0 ? uint32_t(tune.threads) : (chunk_bytes > 192 ? 64u : 128u);
    const uint32_t target = tune.tile_bytes > 0 ? uint32_t(tune.tile_bytes) : 16384u;
    uint32_t R = target / (T * chunk_bytes);
    if (R < 1) R = 1;
    // 2 operand tiles must fit in 227 KB
    while (R > 1 && 2ull * T * R * chunk_bytes > 200u * 1024u) --R;
    while (T > 32 && 2ull * T * R * chunk_bytes > 200u * 1024u) T -= 32;
    *threads = T;
    *chunks_per_thread = R;
}

#if !defined(__CUDACC_RTC__)  // host-side launcher: not part of a run-time (NVRTC) specialisation
// a, b, out: 16-byte aligned dense streams of n elements (ceil(n*S/128)*16 bytes each)
template <class Fn, class Geo, int Mode = 0>
cudaError_t launch_dense_binary(const void* a, const void* b, void* out, size_t n, const Fn& fn, Geo geo,
                                cudaStream_t stream, const launch_geom& g, const dense_tuning& tune = dense_tuning(),
                                void* out1 = nullptr)
{
    using W = typename Geo::unit;
    if (n == 0) return cudaSuccess;
    uint32_t T = 0, R = 0;
    choose_tile<W>(geo.Cw(), tune, &T, &R);
    geo.chunks_per_thread = R;
    const uint64_t valid_bits = uint64_t(n) * geo.S();
    const size_t total_vec = size_t((valid_bits + 127) / 128);
    const size_t tile_bytes = size_t(T) * R * geo.Cw() * sizeof(W);
    const size_t ntiles = (total_vec * 16 + tile_bytes - 1) / tile_bytes;
    const size_t smem = (Mode == 2 ? 1 : 2) * tile_bytes + 16;  // + 16: the 64-bit window form peeks one word past a chunk
    auto kernel = dense_binary_kernel<Fn, Geo, Mode>;
    if (smem > 48 * 1024) {
        cudaError_t rc = cudaFuncSetAttribute(kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, int(smem));
        if (rc != cudaSuccess) return rc;
    }
    if (ntiles > 0x7FFFFFFFull) return cudaErrorInvalidValue;
    // programmatic dependent launch like the padded stream kernels: frame-sized streams (RGB24 1080p is 6 MB) are launch-bound
    if (cudaError_t e = launch_stream_kernel_smem(kernel, unsigned(ntiles), T, smem, stream, g.pdl != 0, static_cast<const uint4*>(a),
                                                  static_cast<const uint4*>(b), static_cast<uint4*>(out), static_cast<uint4*>(out1),
                                                  total_vec, valid_bits, fn, geo))
        return e;
    count_launch(g);
    return cudaGetLastError();
}

#endif  // !__CUDACC_RTC__

}  // namespace dense
}  // namespace kernels
}  // namespace cuda
}  // namespace pint

#endif  // PINT_CUDA_KERNELS_DENSE_CUH_INCLUDED
