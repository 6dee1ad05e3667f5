// pint/cuda_kernels_dense_shift.cuh -- shift_left / shift_right_unsigned (pint.hpp:1006-1029) on a dense bit stream as a
// plain word stream: out = shift(stream & KEEP), KEEP being the per-field keep mask of one element repeated every
// sum(Bits) bits.  The bits KEEP clears are exactly the ones that would cross a field boundary, so after the mask the
// whole stream can be shifted as one long integer: a funnel shift of neighbouring 32-bit words.  No chunks, no shared
// memory staging: 16-byte vectors in and out, the neighbour word comes from the next (previous) lane and only
// the last (first) lane of a warp loads one extra word.
//
// KEEP has a period of N = S / gcd(S, 32) words; the host builds one period plus nine words of overlap
// (dense_shift_table) and every CTA copies it to shared memory.  The funnel shift of word t needs word t + 1 (right)
// or t - 1 (left) of the MASKED stream.
//
// Tiles are whole chunks (lcm(S, 128) bits), i.e. they start and end on element boundaries, where nothing crosses; a
// CTA performs all loads of its tile before its first store, so the same kernel serves in-place calls.  Only amounts
// below 32 are taken here: the caller (pint_cuda_dense_shift) keeps the whole-chunk kernels for wider shifts.
#pragma once
#ifndef PINT_CUDA_KERNELS_DENSE_SHIFT_CUH_INCLUDED
#define PINT_CUDA_KERNELS_DENSE_SHIFT_CUH_INCLUDED

#include "cuda_kernels.cuh"

namespace pint {
namespace cuda {
namespace kernels {
namespace dense {

constexpr int kShiftTableWords = 64 + 9;  // N <= 63 (S <= 64), + the nine words one thread looks ahead

struct dense_shift_table {
    uint32_t keep[kShiftTableWords];  // keep[x] = KEEP bits [32 x, 32 x + 32) of the periodic pattern, x < N + 9
    uint32_t nwords;                  // N: period of the pattern in words
    uint32_t amount;                  // < 32
    uint32_t left;                    // 1 = shift_left, 0 = shift_right_unsigned
};

// pre = keep mask of ONE element (bits above S are 0), S = sum(Bits)
inline dense_shift_table make_dense_shift_table(uint64_t pre, uint32_t S, uint32_t amount, bool left)
{
    dense_shift_table t{};
    uint32_t g = S, b = 32;
    while (b) { const uint32_t r = g % b; g = b; b = r; }
    t.nwords = S / g;
    t.amount = amount;
    t.left = left ? 1u : 0u;
    for (uint32_t x = 0; x < t.nwords + 9; ++x) {
        uint32_t w = 0;
        for (uint32_t bit = 0; bit < 32; ++bit) {
            const uint64_t pos = (uint64_t(x % t.nwords) * 32 + bit) % S;  // position inside the element
            w |= uint32_t((pre >> pos) & 1u) << bit;
        }
        t.keep[x] = w;
    }
    return t;
}

// in / out: the stream as 16-byte vectors (nvec of them, 4 words each); valid_bits = n * S: everything above is written as 0.
// A CTA's tile is tile_vecs vectors.  cross_tile = 1 (out of place): 1024-vector tiles, the edge word of a tile comes from
// the neighbouring tile of `in`.  cross_tile = 0 (in place): a whole number of chunks (lcm(S, 128) bits), so tiles begin
// and end on element boundaries where nothing crosses (the bits that would are masked) and start at phase 0 of the mask;
// every load of a tile is performed before its first store (the barrier below), which makes the kernel safe in place.
template <int U>
__global__ void __launch_bounds__(256)
dense_shift_stream_kernel(const uint4* in, uint4* out, size_t nvec, uint32_t tile_vecs, uint32_t cross_tile,
                          unsigned long long valid_bits, const __grid_constant__ dense_shift_table tab)
{
    static_assert(U * 256 == 1024, "the phase of a cross_tile tile assumes 1024 vectors per tile");
    pdl_prologue();
    __shared__ uint32_t keep[kShiftTableWords];
    const uint32_t N = tab.nwords, k = tab.amount;
    const bool left = tab.left != 0;
    const unsigned lane = threadIdx.x & 31u;
    const uint32_t* words = reinterpret_cast<const uint32_t*>(in);
    // position of a word in the mask period without a modulo per vector (a 64-bit one made the first version ALU-bound,
    // 5.6 TB/s): (4 * thread) mod N once per thread, then adds with a conditional subtract
    const uint32_t step_phase = 1024u % N;  // 256 threads x 4 words between a thread's consecutive vectors
    auto wrap = [N](uint32_t x) { return x >= N ? x - N : x; };
    uint32_t phase0 = (4u * threadIdx.x) % N;
    if (left) phase0 = wrap(phase0 + N - 1);  // left shifts look one word back
    bool first_tile = true;
    for (size_t base = size_t(blockIdx.x) * tile_vecs; base < nvec; base += size_t(gridDim.x) * tile_vecs) {
        // every load of the tile is in flight before anything waits: the vectors, and on the last (first) lane of each
        // warp the one word beyond the warp's 512 bytes that the funnel shift pulls bits from
        uint4 q[U];
        uint32_t far_edge[U];
#pragma unroll
        for (int u = 0; u < U; ++u) {
            const uint32_t local = uint32_t(u) * 256u + threadIdx.x;
            const size_t v = base + local;
            const bool have = local < tile_vecs && v < nvec;
            q[u] = have ? ld16<CACHE_STREAMING>(in + v) : make_uint4(0, 0, 0, 0);
            far_edge[u] = 0;
            // (plain loads: in place the word is written later by this CTA)
            if (!left) { if (lane == 31 && have && v + 1 < nvec && (cross_tile || local + 1 < tile_vecs)) far_edge[u] = words[4 * v + 4]; }
            else       { if (lane == 0 && have && v > 0 && (cross_tile || local > 0)) far_edge[u] = words[4 * v - 1]; }
        }
        if (first_tile)
            for (int x = threadIdx.x; x < kShiftTableWords; x += 256) keep[x] = tab.keep[x];
        first_tile = false;
        __syncthreads();  // the mask table is there; and (in place) every load of the tile precedes its first store
        // a chunk-aligned tile starts at phase 0; a 1024-vector tile t at (4096 t) mod N, in 32-bit arithmetic (a stream has
        // fewer than 2^32 tiles)
        uint32_t phase = phase0;
        if (cross_tile) phase = wrap(phase + ((4096u % N) * (uint32_t(base >> 10) % N)) % N);  // cross_tile tiles are U * 256 = 1024 vectors
#pragma unroll
        for (int u = 0; u < U; ++u) {
            const uint32_t local = uint32_t(u) * 256u + threadIdx.x;
            const size_t v = base + local;
            const bool have = local < tile_vecs && v < nvec;  // whole warps run on so that the shuffles stay convergent
            const size_t i0 = 4 * v;                            // first word of this thread's vector
            const uint32_t w[4] = {q[u].x, q[u].y, q[u].z, q[u].w};
            uint32_t edge = far_edge[u];
            if (!left) {
                const uint32_t nb = __shfl_down_sync(0xffffffffu, w[0], 1);
                if (lane != 31) edge = nb;
            } else {
                const uint32_t nb = __shfl_up_sync(0xffffffffu, w[3], 1);
                if (lane != 0) edge = nb;
            }
            // masks of words i0 .. i0 + 4 (right) or i0 - 1 .. i0 + 3 (left)
            const uint32_t j0 = phase;
            phase = wrap(phase + step_phase);
            if (!have) continue;
            uint32_t m[5];
            if (!left) {
#pragma unroll
                for (int t = 0; t < 4; ++t) m[t] = w[t] & keep[j0 + t];
                m[4] = edge & keep[j0 + 4];
            } else {
                m[0] = edge & keep[j0];
#pragma unroll
                for (int t = 0; t < 4; ++t) m[t + 1] = w[t] & keep[j0 + 1 + t];
            }
            uint32_t r[4];
#pragma unroll
            for (int t = 0; t < 4; ++t)
                r[t] = left ? __funnelshift_l(m[t], m[t + 1], k) : __funnelshift_r(m[t], m[t + 1], k);
            // padding above the last element stays 0 whatever the input holds there
            if ((unsigned long long)(i0 + 4) * 32ull > valid_bits) {
#pragma unroll
                for (int t = 0; t < 4; ++t) {
                    const unsigned long long lo = (unsigned long long)(i0 + t) * 32ull;
                    if (lo >= valid_bits) r[t] = 0;
                    else if (valid_bits - lo < 32ull) r[t] &= (1u << unsigned(valid_bits - lo)) - 1u;
                }
            }
            st16<CACHE_STREAMING>(out + v, make_uint4(r[0], r[1], r[2], r[3]));
        }
        __syncthreads();  // (grid-stride only) the next tile's table reads and loads start after this tile is done
    }
}

#if !defined(__CUDACC_RTC__)
// bytes: length of both streams (a multiple of 16); in and out 16-byte aligned, the same stream or disjoint; amount < 32
inline cudaError_t launch_dense_shift_stream(const void* in, void* out, size_t bytes, unsigned long long valid_bits, uint32_t S,
                                             const dense_shift_table& tab, cudaStream_t s, const launch_geom& g)
{
    constexpr int U = 4;
    const size_t nvec = bytes / 16;
    if (nvec) {
        // Out of place: tiles of 1024 vectors (16 KiB, every thread busy, 6.5-6.9 TB/s); the one word a tile needs from its
        // neighbour is read from `in`, which nobody writes.  In place that word could already be overwritten, so tiles are
        // whole chunks there -- lcm(S, 128) bits, nothing crosses their edges -- and whole 128-byte lines (tiles 16 bytes off
        // a line cost 6 %); up to 11 % of the threads idle in the last row, 6.0-6.6 TB/s (profiles/r36b, r37).
        uint32_t tile_vecs = uint32_t(U) * 256u, cross_tile = 1;
        if (in == out) {
            uint32_t a = S, b = 128;
            while (b) { const uint32_t r = a % b; a = b; b = r; }
            const uint32_t chunk_vecs = S / a;  // lcm(S, 128) bits in 16-byte vectors: at most 63
            uint32_t unit = chunk_vecs;
            while (unit % 8) unit += chunk_vecs;  // lcm(chunk_vecs, 8) <= 504
            tile_vecs = uint32_t(U) * 256u / unit * unit;
            cross_tile = 0;
        }
        const size_t ntiles = (nvec + tile_vecs - 1) / tile_vecs;
        if (cudaError_t e = launch_stream_kernel(dense_shift_stream_kernel<U>, grid_for(ntiles, g), 256, s, g.pdl != 0,
                                                 static_cast<const uint4*>(in), static_cast<uint4*>(out), nvec, tile_vecs, cross_tile, valid_bits, tab))
            return e;
        count_launch(g);
    }
    return cudaGetLastError();
}
#endif

}  // namespace dense
}  // namespace kernels
}  // namespace cuda
}  // namespace pint

#endif  // PINT_CUDA_KERNELS_DENSE_SHIFT_CUH_INCLUDED
