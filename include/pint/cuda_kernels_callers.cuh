// pint/cuda_kernels_callers.cuh -- two more caller shapes of the 13 binary operations:
//   * one operand the SAME packed_int for every element (out[i] = OP(a[i], k) or OP(k, a[i])):
//     a 1-in/1-out map, 2W bytes per element instead of 3W;
//   * operands interleaved as the reference benchmark holds them
//     (std::vector<std::pair<Integer, Integer>>, tests/pint_bench.cpp:22), results as a plain
//     array: 3W bytes per element, de-interleaved in registers.
// Both wrap any binary functor of the stream kernels (swar_fn / fused_fn, compiled-in or run-time
// masks).  Included by cuda_kernels.cuh.
#pragma once
#ifndef PINT_CUDA_KERNELS_CALLERS_CUH_INCLUDED
#define PINT_CUDA_KERNELS_CALLERS_CUH_INCLUDED

namespace pint {
namespace cuda {
namespace kernels {

// x -> OP(x, k) (or OP(k, x)): k is the scalar operand replicated over the processing Word, exactly
// as the masks are, so a Word of 2 / 4 narrow elements is one call.
template <class BinFn, bool ScalarFirst>
struct scalar_fn {
    using word = typename BinFn::word;
    static constexpr bool heavy = BinFn::heavy;
    BinFn fn;
    word k;
    __device__ __forceinline__ word operator()(word x) const { return ScalarFirst ? fn(k, x) : fn(x, k); }
};

template <class Elem>
inline word_of<Elem> replicate_scalar(uint64_t scalar)
{
    using W = word_of<Elem>;
    W k = W(Elem(scalar));
    if (sizeof(Elem) == 2) k = W(k * 0x00010001u);
    if (sizeof(Elem) == 1) k = W(k * 0x01010101u);
    return k;
}

// out[v] = map(in[v]) over 16-byte vectors; same structure as unary_stream_kernel (one CTA per
// tile, U independent loads per thread, streaming policy both ways).
template <class Fn, int U>
__global__ void __launch_bounds__(256)
map1_stream_kernel(const uint4* in, uint4* out, size_t nvec, const __grid_constant__ Fn fn)
{
    pdl_prologue();
    using W = typename Fn::word;
    constexpr int B = 256;
    constexpr size_t tile = size_t(U) * B;
    const size_t nfull = nvec / tile;
    auto f = [&](W x) { return fn(x); };
    for (size_t t = blockIdx.x; t < nfull; t += gridDim.x) {
        const size_t base = t * tile + threadIdx.x;
        uint4 v[U];
#pragma unroll
        for (int k = 0; k < U; ++k) v[k] = ld16<CACHE_STREAMING>(in + base + size_t(k) * B);
#pragma unroll
        for (int k = 0; k < U; ++k) st16<CACHE_STREAMING>(out + base + size_t(k) * B, map_words1<W>(v[k], f));
    }
    if (blockIdx.x == nfull % gridDim.x) {
        for (size_t i = nfull * tile + threadIdx.x; i < nvec; i += B)
            st16<CACHE_STREAMING>(out + i, map_words1<W>(ld16<CACHE_STREAMING>(in + i), f));
    }
}

template <class Elem, class Fn>
__global__ void __launch_bounds__(256)
map1_elem_kernel(const Elem* in, Elem* out, size_t n, const __grid_constant__ Fn fn)
{
    using W = typename Fn::word;
    for (size_t i = size_t(blockIdx.x) * blockDim.x + threadIdx.x; i < n; i += size_t(gridDim.x) * blockDim.x)
        out[i] = Elem(fn(W(in[i])));
}

#if !defined(__CUDACC_RTC__)  // host-side launchers: not part of a run-time (NVRTC) specialisation
// out[i] = fn(in[i]) for a one-operand functor (scalar_fn); `out` may be `in`.
template <class Elem, class Fn>
cudaError_t launch_map1(const void* in, void* out, size_t n, const Fn& fn, cudaStream_t s, const launch_geom& g)
{
    if (n == 0) return cudaSuccess;
    const Elem* ei = static_cast<const Elem*>(in);
    Elem* eo = static_cast<Elem*>(out);
    auto elems = [&](size_t first, size_t count) {
        if (!count) return;
        map1_elem_kernel<Elem, Fn><<<grid_for((count + 255) / 256, g), 256, 0, s>>>(ei + first, eo + first, count, fn);
        count_launch(g);
    };
    const void* ptrs[2] = {in, out};
    size_t head = 0, nvec = 0, tail = 0;
    if (!split_aligned(sizeof(Elem), n, ptrs, 2, &head, &nvec, &tail)) {
        elems(0, n);
        return cudaGetLastError();
    }
    elems(0, head);
    if (nvec) {
        // 64 bytes in flight per thread, as the two-operand kernels have at UNROLL 2 (U = 2 measured
        // 6516-6896 GB/s on RGBA8 depending on the op's ALU weight, U = 4 6780-6800 for all of them)
        constexpr int U = 4;
        const size_t ntiles = (nvec + size_t(U) * 256 - 1) / (size_t(U) * 256);
        if (cudaError_t e = launch_stream_kernel(map1_stream_kernel<Fn, U>, grid_for(ntiles, g), 256, s, g.pdl != 0,
                                                 reinterpret_cast<const uint4*>(ei + head), reinterpret_cast<uint4*>(eo + head), nvec, fn))
            return e;
        count_launch(g);
    }
    elems(n - tail, tail);
    return cudaGetLastError();
}

#endif  // !__CUDACC_RTC__

// ---------------------------------------------------------------------------
// Interleaved operands.  A thread turns 32 bytes of pairs into 16 bytes of results; the operands
// are separated in registers: 64-bit and 32-bit Integers by word selection, 16-bit and 8-bit ones
// with one byte permute per operand Word (PRMT), which also packs them 2 / 4 to the Word the
// replicated masks expect.
template <class Elem, class Fn>
__device__ __forceinline__ uint4 pairs_to_results(const uint4& lo, const uint4& hi, const Fn& fn)
{
    uint4 r;
    if constexpr (sizeof(Elem) == 8) {
        const uint64_t r0 = fn((uint64_t(lo.y) << 32) | lo.x, (uint64_t(lo.w) << 32) | lo.z);
        const uint64_t r1 = fn((uint64_t(hi.y) << 32) | hi.x, (uint64_t(hi.w) << 32) | hi.z);
        r = make_uint4(uint32_t(r0), uint32_t(r0 >> 32), uint32_t(r1), uint32_t(r1 >> 32));
    } else if constexpr (sizeof(Elem) == 4) {
        r = make_uint4(fn(lo.x, lo.y), fn(lo.z, lo.w), fn(hi.x, hi.y), fn(hi.z, hi.w));
    } else {
        // 16-bit: a word is (first | second << 16); 8-bit: (first0, second0, first1, second1)
        constexpr uint32_t SA = sizeof(Elem) == 2 ? 0x5410u : 0x6420u;
        constexpr uint32_t SB = sizeof(Elem) == 2 ? 0x7632u : 0x7531u;
        r.x = fn(__byte_perm(lo.x, lo.y, SA), __byte_perm(lo.x, lo.y, SB));
        r.y = fn(__byte_perm(lo.z, lo.w, SA), __byte_perm(lo.z, lo.w, SB));
        r.z = fn(__byte_perm(hi.x, hi.y, SA), __byte_perm(hi.x, hi.y, SB));
        r.w = fn(__byte_perm(hi.z, hi.w, SA), __byte_perm(hi.z, hi.w, SB));
    }
    return r;
}

// out[v] (16 bytes) = OP over pairs[v] (32 bytes, 256-bit loads); U vectors per thread.
template <class Elem, class Fn, int U>
__global__ void __launch_bounds__(256)
pairs_map_kernel(const vec32* pairs, uint4* out, size_t nvec, const __grid_constant__ Fn fn)
{
    pdl_prologue();
    constexpr int B = 256;
    constexpr size_t tile = size_t(U) * B;
    const size_t nfull = nvec / tile;
    for (size_t t = blockIdx.x; t < nfull; t += gridDim.x) {
        const size_t base = t * tile + threadIdx.x;
        vec32 v[U];
#pragma unroll
        for (int k = 0; k < U; ++k) v[k] = ld32(pairs + base + size_t(k) * B);
#pragma unroll
        for (int k = 0; k < U; ++k)
            st16<CACHE_STREAMING>(out + base + size_t(k) * B, pairs_to_results<Elem>(v[k].lo, v[k].hi, fn));
    }
    if (blockIdx.x == nfull % gridDim.x) {
        for (size_t i = nfull * tile + threadIdx.x; i < nvec; i += B) {
            const vec32 v = ld32(pairs + i);
            st16<CACHE_STREAMING>(out + i, pairs_to_results<Elem>(v.lo, v.hi, fn));
        }
    }
}

template <class Elem, class Fn>
__global__ void __launch_bounds__(256)
pairs_map_elem_kernel(const Elem* pairs, Elem* out, size_t n, const __grid_constant__ Fn fn)
{
    using W = typename Fn::word;
    for (size_t i = size_t(blockIdx.x) * blockDim.x + threadIdx.x; i < n; i += size_t(gridDim.x) * blockDim.x)
        out[i] = Elem(fn(W(pairs[2 * i]), W(pairs[2 * i + 1])));
}

// out[i] = fn(pairs[2i], pairs[2i+1]) for a binary functor; `out` must not overlap `pairs`.
#if !defined(__CUDACC_RTC__)  // host-side launchers: not part of a run-time (NVRTC) specialisation
template <class Elem, class Fn>
cudaError_t launch_pairs_map(const void* pairs, void* out, size_t n, const Fn& fn, cudaStream_t s, const launch_geom& g)
{
    if (n == 0) return cudaSuccess;
    const Elem* ep = static_cast<const Elem*>(pairs);
    Elem* eo = static_cast<Elem*>(out);
    auto elems = [&](size_t first, size_t count) {
        if (!count) return;
        pairs_map_elem_kernel<Elem, Fn><<<grid_for((count + 255) / 256, g), 256, 0, s>>>(ep + 2 * first, eo + first, count, fn);
        count_launch(g);
    };
    // head: elements up to the first 16-byte boundary of `out`; the vector body needs the pairs
    // to reach a 32-byte boundary at the same element
    constexpr size_t per = 16 / sizeof(Elem);
    const uintptr_t off = reinterpret_cast<uintptr_t>(out) & 15u;
    size_t head = off ? (16 - off) / sizeof(Elem) : 0;
    if (head > n) head = n;
    const bool vec_ok = (reinterpret_cast<uintptr_t>(ep + 2 * head) & 31u) == 0;
    const size_t nvec = vec_ok ? (n - head) / per : 0;
    if (!nvec) {
        elems(0, n);
        return cudaGetLastError();
    }
    elems(0, head);
    constexpr int U = 2;
    const size_t ntiles = (nvec + size_t(U) * 256 - 1) / (size_t(U) * 256);
    if (cudaError_t e = launch_stream_kernel(pairs_map_kernel<Elem, Fn, U>, grid_for(ntiles, g), 256, s, g.pdl != 0,
                                             reinterpret_cast<const vec32*>(ep + 2 * head), reinterpret_cast<uint4*>(eo + head), nvec, fn))
        return e;
    count_launch(g);
    elems(head + nvec * per, n - head - nvec * per);
    return cudaGetLastError();
}

#endif  // !__CUDACC_RTC__

}  // namespace kernels
}  // namespace cuda
}  // namespace pint

#endif  // PINT_CUDA_KERNELS_CALLERS_CUH_INCLUDED
