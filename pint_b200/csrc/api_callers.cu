// api_callers.cu -- two more caller shapes of the same 13 operations (the callers and
// data formats either side of the path):
//   * pint_cuda_binary_scalar: one operand is the SAME packed_int for every element
//     (out[i] = OP(a[i], k) or OP(k, a[i])): a 1-in/1-out map, 2W bytes per element instead of 3W;
//   * pint_cuda_binary_pairs: the operands arrive interleaved, as the reference benchmark holds
//     them (std::vector<std::pair<Integer, Integer>>, tests/pint_bench.cpp:22), and the results
//     leave as a plain array: 3W bytes per element, de-interleaved in registers.
// Both use the run-time-mask functors (masks in the constant bank): one instantiation per op and
// Word serves every layout, and these maps are HBM-bound with those functors.
#include "api_common.hpp"

namespace pint_cuda_impl {
namespace {

using namespace pint::cuda;
using namespace pint::cuda::kernels;

// x -> OP(x, k) (or OP(k, x)): k is the scalar operand replicated over the processing Word, exactly
// as the masks are, so a Word of 2 / 4 narrow elements is one call.
template <int Op, class W, bool ScalarFirst>
struct scalar_fn {
    using word = W;
    static constexpr bool heavy = swar_fn<Op, runtime_masks<W>>::heavy;
    runtime_masks<W> p;
    W k;
    __device__ __forceinline__ W operator()(W x) const
    {
        return ScalarFirst ? apply_op<Op>(k, x, p) : apply_op<Op>(x, k, p);
    }
};

// out[v] = map(in[v]) over 16-byte vectors; same structure as unary_stream_kernel (one CTA per
// tile, U independent loads per thread, streaming policy both ways).
template <class Fn, int U>
__global__ void __launch_bounds__(256)
map1_stream_kernel(const uint4* in, uint4* out, size_t nvec, const __grid_constant__ Fn fn)
{
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

template <class Elem, int Op, bool ScalarFirst>
cudaError_t scalar_one(const layout_desc& d, const void* a, uint64_t scalar, void* out, size_t n, cudaStream_t s,
                       const launch_geom& g)
{
    using W = word_of<Elem>;
    using Fn = scalar_fn<Op, W, ScalarFirst>;
    W k = W(Elem(scalar));
    if (sizeof(Elem) == 2) k = W(k * 0x00010001u);
    if (sizeof(Elem) == 1) k = W(k * 0x01010101u);
    const Fn fn{runtime_masks<W>(make_word_masks<W>(d)), k};
    const Elem* ei = static_cast<const Elem*>(a);
    Elem* eo = static_cast<Elem*>(out);
    auto elems = [&](size_t first, size_t count) {
        if (!count) return;
        map1_elem_kernel<Elem, Fn><<<grid_for((count + 255) / 256, g), 256, 0, s>>>(ei + first, eo + first, count, fn);
        count_launch(g);
    };
    const void* ptrs[2] = {a, out};
    size_t head = 0, nvec = 0, tail = 0;
    if (!split_aligned(sizeof(Elem), n, ptrs, 2, &head, &nvec, &tail)) {
        elems(0, n);
        return cudaGetLastError();
    }
    elems(0, head);
    if (nvec) {
        constexpr int U = Fn::heavy ? 4 : 2;
        const size_t ntiles = (nvec + size_t(U) * 256 - 1) / (size_t(U) * 256);
        map1_stream_kernel<Fn, U><<<grid_for(ntiles, g), 256, 0, s>>>(
            reinterpret_cast<const uint4*>(ei + head), reinterpret_cast<uint4*>(eo + head), nvec, fn);
        count_launch(g);
    }
    elems(n - tail, tail);
    return cudaGetLastError();
}

template <class Elem, int... Ops>
cudaError_t scalar_ops(int op, bool first, const layout_desc& d, const void* a, uint64_t scalar, void* out, size_t n,
                       cudaStream_t s, const launch_geom& g, std::integer_sequence<int, Ops...>)
{
    cudaError_t rc = cudaErrorInvalidValue;
    if (first)
        ((op == Ops ? (rc = scalar_one<Elem, Ops, true>(d, a, scalar, out, n, s, g), 0) : 0), ...);
    else
        ((op == Ops ? (rc = scalar_one<Elem, Ops, false>(d, a, scalar, out, n, s, g), 0) : 0), ...);
    return rc;
}

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

template <class Elem, int Op>
cudaError_t pairs_one(const layout_desc& d, const void* pairs, void* out, size_t n, cudaStream_t s, const launch_geom& g)
{
    using W = word_of<Elem>;
    using Fn = swar_fn<Op, runtime_masks<W>>;
    const Fn fn{runtime_masks<W>(make_word_masks<W>(d))};
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
    pairs_map_kernel<Elem, Fn, U><<<grid_for(ntiles, g), 256, 0, s>>>(
        reinterpret_cast<const vec32*>(ep + 2 * head), reinterpret_cast<uint4*>(eo + head), nvec, fn);
    count_launch(g);
    elems(head + nvec * per, n - head - nvec * per);
    return cudaGetLastError();
}

template <class Elem, int... Ops>
cudaError_t pairs_map_ops(int op, const layout_desc& d, const void* pairs, void* out, size_t n, cudaStream_t s,
                          const launch_geom& g, std::integer_sequence<int, Ops...>)
{
    cudaError_t rc = cudaErrorInvalidValue;
    ((op == Ops ? (rc = pairs_one<Elem, Ops>(d, pairs, out, n, s, g), 0) : 0), ...);
    return rc;
}

}  // namespace
}  // namespace pint_cuda_impl

using namespace pint_cuda_impl;

extern "C" {

int pint_cuda_binary_scalar(int op, const pint_layout* layout, const void* a, uint64_t scalar, int scalar_first,
                            void* out, size_t n, pint_stream_t stream)
{
    layout_desc d;
    if (int rc = check_layout(layout, &d)) return rc;
    if (op < 0 || op >= PINT_NUM_BINARY_OPS) return PINT_E_OP;
    // the scalar is a packed_int::value(): it has to be representable in the Integer (pint.hpp:768)
    if (d.elem_bytes < 8 && (scalar >> (8 * d.elem_bytes))) return PINT_E_ARG;
    if (n == 0) return PINT_OK;
    if (!a || !out) return PINT_E_NULL;
    const uintptr_t i = (uintptr_t)a, o = (uintptr_t)out;
    const size_t bytes = n * d.elem_bytes;
    if (i != o && i < o + bytes && o < i + bytes) return PINT_E_ALIAS;
    if ((i | o) % d.elem_bytes) return PINT_E_ARG;
    const launch_geom g = geom();
    cudaStream_t s = (cudaStream_t)stream;
    auto ops = std::make_integer_sequence<int, pint::cuda::OP_NUM_BINARY>();
    const bool first = scalar_first != 0;
    switch (d.elem_bytes) {
    case 1: return (int)scalar_ops<uint8_t>(op, first, d, a, scalar, out, n, s, g, ops);
    case 2: return (int)scalar_ops<uint16_t>(op, first, d, a, scalar, out, n, s, g, ops);
    case 4: return (int)scalar_ops<uint32_t>(op, first, d, a, scalar, out, n, s, g, ops);
    default: return (int)scalar_ops<uint64_t>(op, first, d, a, scalar, out, n, s, g, ops);
    }
}

int pint_cuda_binary_pairs(int op, const pint_layout* layout, const void* pairs, void* out, size_t n,
                           pint_stream_t stream)
{
    layout_desc d;
    if (int rc = check_layout(layout, &d)) return rc;
    if (op < 0 || op >= PINT_NUM_BINARY_OPS) return PINT_E_OP;
    if (n == 0) return PINT_OK;
    if (!pairs || !out) return PINT_E_NULL;
    const uintptr_t i = (uintptr_t)pairs, o = (uintptr_t)out;
    const size_t bytes = n * d.elem_bytes;
    // results are half the size of the operands: writing them over the pairs would overtake the reads
    if (i < o + bytes && o < i + 2 * bytes) return PINT_E_ALIAS;
    if ((i | o) % d.elem_bytes) return PINT_E_ARG;
    const launch_geom g = geom();
    cudaStream_t s = (cudaStream_t)stream;
    auto ops = std::make_integer_sequence<int, pint::cuda::OP_NUM_BINARY>();
    switch (d.elem_bytes) {
    case 1: return (int)pairs_map_ops<uint8_t>(op, d, pairs, out, n, s, g, ops);
    case 2: return (int)pairs_map_ops<uint16_t>(op, d, pairs, out, n, s, g, ops);
    case 4: return (int)pairs_map_ops<uint32_t>(op, d, pairs, out, n, s, g, ops);
    default: return (int)pairs_map_ops<uint64_t>(op, d, pairs, out, n, s, g, ops);
    }
}

}  // extern "C"
