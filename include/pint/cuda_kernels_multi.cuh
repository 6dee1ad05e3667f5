// pint/cuda_kernels_multi.cuh -- two binary ops on the SAME operands in one pass.
//
// out0[i] = Op0(a[i], b[i]), out1[i] = Op1(a[i], b[i]): a and b are read once and
// both results are written, so a step such as BASELINE.json's C1
// (add_unsigned_saturate + add_signed_saturate over the same arrays) moves
// 4W bytes per element instead of 2 * 3W -- 16 instead of 24 bytes on uint32 packs.
//
// Both op codes are TEMPLATE parameters.  The first version of this kernel took
// the op codes at run time (a warp-uniform switch per vector): ptxas kept all 13
// op bodies live, 115 registers per thread, 21 % of the warps resident, 3.2 TB/s
// (profiles/r01b_ncu_multi_runtime_switch.txt).  With compile-time ops the kernel
// is binary_stream_kernel with a second store.  libpint_cuda.so instantiates the
// pairs of kOpPairs below; pint_cuda_binary_multi decomposes any op list into
// those pairs plus single-op passes.  Included by cuda_kernels.cuh.
#pragma once
#ifndef PINT_CUDA_KERNELS_MULTI_CUH_INCLUDED
#define PINT_CUDA_KERNELS_MULTI_CUH_INCLUDED

namespace pint {
namespace cuda {
namespace kernels {

constexpr int kMaxMultiOps = 4;

// op pairs with a one-pass kernel (unordered; the dispatcher swaps the outputs)
constexpr int kNumOpPairs = 7;
constexpr int kOpPairs[kNumOpPairs][2] = {
    {OP_ADD_UNSIGNED_SATURATE, OP_ADD_SIGNED_SATURATE},  // BASELINE.json C1
    {OP_SUB_UNSIGNED_SATURATE, OP_SUB_SIGNED_SATURATE},
    {OP_ADD_SIGNED_SATURATE, OP_SUB_SIGNED_SATURATE},    // BASELINE.json C2
    {OP_ADD_UNSIGNED_SATURATE, OP_SUB_UNSIGNED_SATURATE},
    {OP_ADD_WRAP, OP_SUB_WRAP},
    {OP_MIN_UNSIGNED, OP_MAX_UNSIGNED},
    {OP_MIN_SIGNED, OP_MAX_SIGNED},
};

// index into kOpPairs and whether (x, y) is stored there as (y, x); -1 if the pair has no kernel
inline int find_op_pair(int x, int y, bool* swapped)
{
    for (int i = 0; i < kNumOpPairs; ++i) {
        if (kOpPairs[i][0] == x && kOpPairs[i][1] == y) { *swapped = false; return i; }
        if (kOpPairs[i][0] == y && kOpPairs[i][1] == x) { *swapped = true; return i; }
    }
    return -1;
}

template <int Op0, int Op1, class P>
struct pair_fn {
    using word = typename P::word;
    static constexpr bool heavy = swar_fn<Op0, P>::heavy || swar_fn<Op1, P>::heavy;
    P p;
    __device__ __forceinline__ word first(word a, word b) const { return apply_op<Op0>(a, b, p); }
    __device__ __forceinline__ word second(word a, word b) const { return apply_op<Op1>(a, b, p); }
};

// Same structure as binary_stream_kernel: one CTA per tile (grid stride when the
// grid is capped), all 2 * UNROLL loads issued before the first use, two
// streaming stores per input vector pair.
template <class Fn, class Cfg>
__global__ void __launch_bounds__(Cfg::BLOCK)
binary_pair_stream_kernel(const uint4* a, const uint4* b, uint4* out0, uint4* out1, size_t nvec,
                          const __grid_constant__ Fn fn)
{
    pdl_prologue();
    using W = typename Fn::word;
    constexpr int U = Cfg::UNROLL, B = Cfg::BLOCK;
    constexpr size_t tile = size_t(U) * B;
    const size_t nfull = nvec / tile;
    auto f0 = [&](W x, W y) { return fn.first(x, y); };
    auto f1 = [&](W x, W y) { return fn.second(x, y); };

    for (size_t t = blockIdx.x; t < nfull; t += gridDim.x) {
        const size_t base = t * tile + threadIdx.x;
        uint4 va[U], vb[U];
#pragma unroll
        for (int u = 0; u < U; ++u) va[u] = ld16<Cfg::LD>(a + base + size_t(u) * B);
#pragma unroll
        for (int u = 0; u < U; ++u) vb[u] = ld16<Cfg::LD>(b + base + size_t(u) * B);
#pragma unroll
        for (int u = 0; u < U; ++u) {
            st16<Cfg::ST>(out0 + base + size_t(u) * B, map_words2<W>(va[u], vb[u], f0));
            st16<Cfg::ST>(out1 + base + size_t(u) * B, map_words2<W>(va[u], vb[u], f1));
        }
    }
    if (blockIdx.x == nfull % gridDim.x) {
        for (size_t i = nfull * tile + threadIdx.x; i < nvec; i += B) {
            const uint4 x = ld16<Cfg::LD>(a + i), y = ld16<Cfg::LD>(b + i);
            st16<Cfg::ST>(out0 + i, map_words2<W>(x, y, f0));
            st16<Cfg::ST>(out1 + i, map_words2<W>(x, y, f1));
        }
    }
}

// 32-byte-vector form (see vec32 in cuda_kernels_stream.cuh)
template <class Fn, class Cfg>
__global__ void __launch_bounds__(Cfg::BLOCK)
binary_pair_stream256_kernel(const vec32* a, const vec32* b, vec32* out0, vec32* out1, size_t nvec,
                             const __grid_constant__ Fn fn)
{
    pdl_prologue();
    using W = typename Fn::word;
    constexpr int U = Cfg::UNROLL / 2 > 0 ? Cfg::UNROLL / 2 : 1, B = Cfg::BLOCK;
    constexpr size_t tile = size_t(U) * B;
    const size_t nfull = nvec / tile;
    auto f0 = [&](W x, W y) { return fn.first(x, y); };
    auto f1 = [&](W x, W y) { return fn.second(x, y); };
    auto emit = [&](size_t i, const vec32& x, const vec32& y) {
        vec32 r;
        r.lo = map_words2<W>(x.lo, y.lo, f0);
        r.hi = map_words2<W>(x.hi, y.hi, f0);
        st32(out0 + i, r);
        r.lo = map_words2<W>(x.lo, y.lo, f1);
        r.hi = map_words2<W>(x.hi, y.hi, f1);
        st32(out1 + i, r);
    };
    for (size_t t = blockIdx.x; t < nfull; t += gridDim.x) {
        const size_t base = t * tile + threadIdx.x;
        vec32 va[U], vb[U];
#pragma unroll
        for (int u = 0; u < U; ++u) va[u] = ld32(a + base + size_t(u) * B);
#pragma unroll
        for (int u = 0; u < U; ++u) vb[u] = ld32(b + base + size_t(u) * B);
#pragma unroll
        for (int u = 0; u < U; ++u) emit(base + size_t(u) * B, va[u], vb[u]);
    }
    if (blockIdx.x == nfull % gridDim.x) {
        for (size_t i = nfull * tile + threadIdx.x; i < nvec; i += B) emit(i, ld32(a + i), ld32(b + i));
    }
}

// The vector body runs the pair kernel; heads, tails and operands with different
// misalignments go through the single-op launcher (element-granular kernels that
// the library instantiates anyway).  Outputs must not overlap the inputs or each
// other (checked at the C ABI).
#if !defined(__CUDACC_RTC__)  // host-side launchers: not part of a run-time (NVRTC) specialisation
template <class Elem, int Op0, int Op1, class P, class Cfg = stream_cfg_for<pair_fn<Op0, Op1, P>>>
cudaError_t launch_binary_pair(const void* a, const void* b, void* out0, void* out1, size_t n, const P& p,
                               cudaStream_t stream, const launch_geom& g)
{
    if (n == 0) return cudaSuccess;
    using Fn = pair_fn<Op0, Op1, P>;
    const Elem* ea = static_cast<const Elem*>(a);
    const Elem* eb = static_cast<const Elem*>(b);
    Elem* e0 = static_cast<Elem*>(out0);
    Elem* e1 = static_cast<Elem*>(out1);
    auto single = [&](size_t first, size_t count) -> cudaError_t {
        if (!count) return cudaSuccess;
        cudaError_t rc = launch_binary<Elem, swar_fn<Op0, P>>(ea + first, eb + first, e0 + first, count, swar_fn<Op0, P>{p}, stream, g);
        if (rc != cudaSuccess) return rc;
        return launch_binary<Elem, swar_fn<Op1, P>>(ea + first, eb + first, e1 + first, count, swar_fn<Op1, P>{p}, stream, g);
    };
    const void* ptrs[4] = {a, b, out0, out1};
    size_t head = 0, nvec = 0, tail = 0;
    // 32-byte vectors for light functors on 32-bit words only (a 64-bit pair is 5-10 % slower with
    // them, RGBA8 pair 1 % faster: same-box A/B in profiles/r01z_ab_vec256.log)
    if ((g.vec256 == 2 || (g.vec256 == 1 && !Fn::heavy && sizeof(typename Fn::word) == 4)) &&
        Cfg::LD == CACHE_STREAMING && Cfg::ST == CACHE_STREAMING &&
        split_aligned(sizeof(Elem), n, ptrs, 4, &head, &nvec, &tail, 32) && nvec) {
        if (cudaError_t rc = single(0, head)) return rc;
        constexpr size_t tile = size_t(Cfg::UNROLL / 2 > 0 ? Cfg::UNROLL / 2 : 1) * Cfg::BLOCK;
        const size_t ntiles = (nvec + tile - 1) / tile;
        if (cudaError_t e = launch_stream_kernel(binary_pair_stream256_kernel<Fn, Cfg>, grid_for(ntiles, g), Cfg::BLOCK, stream, g.pdl != 0,
                                                 reinterpret_cast<const vec32*>(ea + head), reinterpret_cast<const vec32*>(eb + head),
                                                 reinterpret_cast<vec32*>(e0 + head), reinterpret_cast<vec32*>(e1 + head), nvec, Fn{p}))
            return e;
        count_launch(g);
        if (cudaError_t rc = single(n - tail, tail)) return rc;
        return cudaGetLastError();
    }
    if (!split_aligned(sizeof(Elem), n, ptrs, 4, &head, &nvec, &tail)) return single(0, n);
    if (cudaError_t rc = single(0, head)) return rc;
    if (nvec) {
        constexpr size_t tile = size_t(Cfg::UNROLL) * Cfg::BLOCK;
        const size_t ntiles = (nvec + tile - 1) / tile;
        if (cudaError_t e = launch_stream_kernel(binary_pair_stream_kernel<Fn, Cfg>, grid_for(ntiles, g), Cfg::BLOCK, stream, g.pdl != 0,
                                                 reinterpret_cast<const uint4*>(ea + head), reinterpret_cast<const uint4*>(eb + head),
                                                 reinterpret_cast<uint4*>(e0 + head), reinterpret_cast<uint4*>(e1 + head), nvec, Fn{p}))
            return e;
        count_launch(g);
    }
    if (cudaError_t rc = single(n - tail, tail)) return rc;
    return cudaGetLastError();
}

#endif  // !__CUDACC_RTC__

}  // namespace kernels
}  // namespace cuda
}  // namespace pint

#endif  // PINT_CUDA_KERNELS_MULTI_CUH_INCLUDED
