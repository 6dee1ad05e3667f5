// pint/cuda_kernels_stream.cuh -- packed -> packed streaming kernels
// (binary ops: 2 reads + 1 write = 3W bytes/element; unary ops: 2W).
// Included by cuda_kernels.cuh.
#pragma once
#ifndef PINT_CUDA_KERNELS_STREAM_CUH_INCLUDED
#define PINT_CUDA_KERNELS_STREAM_CUH_INCLUDED

namespace pint {
namespace cuda {
namespace kernels {

// ---------------------------------------------------------------------------
// 128-bit global accesses with an explicit cache policy.  Nothing in these
// kernels is read twice, so the default is "streaming" (evict-first) in both
// directions; the variants exist so the choice is a measured one (tools/tune.cu).
enum : int { CACHE_DEFAULT = 0, CACHE_NO_ALLOCATE = 1, CACHE_STREAMING = 2 };

template <int Policy>
__device__ __forceinline__ uint4 ld16(const uint4* p)
{
    uint4 r;
    if constexpr (Policy == CACHE_NO_ALLOCATE) {
        asm volatile("ld.global.L1::no_allocate.v4.u32 {%0,%1,%2,%3}, [%4];"
                     : "=r"(r.x), "=r"(r.y), "=r"(r.z), "=r"(r.w) : "l"(p));
    } else if constexpr (Policy == CACHE_STREAMING) {
        asm volatile("ld.global.cs.v4.u32 {%0,%1,%2,%3}, [%4];"
                     : "=r"(r.x), "=r"(r.y), "=r"(r.z), "=r"(r.w) : "l"(p));
    } else {
        asm volatile("ld.global.v4.u32 {%0,%1,%2,%3}, [%4];"
                     : "=r"(r.x), "=r"(r.y), "=r"(r.z), "=r"(r.w) : "l"(p));
    }
    return r;
}

template <int Policy>
__device__ __forceinline__ void st16(uint4* p, const uint4& v)
{
    if constexpr (Policy == CACHE_NO_ALLOCATE) {
        asm volatile("st.global.L1::no_allocate.v4.u32 [%0], {%1,%2,%3,%4};"
                     :: "l"(p), "r"(v.x), "r"(v.y), "r"(v.z), "r"(v.w) : "memory");
    } else if constexpr (Policy == CACHE_STREAMING) {
        asm volatile("st.global.cs.v4.u32 [%0], {%1,%2,%3,%4};"
                     :: "l"(p), "r"(v.x), "r"(v.y), "r"(v.z), "r"(v.w) : "memory");
    } else {
        asm volatile("st.global.v4.u32 [%0], {%1,%2,%3,%4};"
                     :: "l"(p), "r"(v.x), "r"(v.y), "r"(v.z), "r"(v.w) : "memory");
    }
}

// 256-bit global accesses (LDG.E.256 / STG.E.256, new with sm_100): half the memory instructions
// per byte.  Measured +0.3-0.9 % over 128-bit accesses on the bandwidth-bound kernels
// (profiles/r01v_tune256.log: 7.05 vs 7.00 TB/s single op, 6.93 vs 6.88 for an op pair), so the
// vector kernels use them whenever all operands are 32-byte congruent; streaming policy only.
struct vec32 {
    uint4 lo, hi;
};

__device__ __forceinline__ vec32 ld32(const vec32* p)
{
    vec32 r;
    asm volatile("ld.global.cs.v8.u32 {%0,%1,%2,%3,%4,%5,%6,%7}, [%8];"
                 : "=r"(r.lo.x), "=r"(r.lo.y), "=r"(r.lo.z), "=r"(r.lo.w), "=r"(r.hi.x), "=r"(r.hi.y), "=r"(r.hi.z), "=r"(r.hi.w)
                 : "l"(p));
    return r;
}

__device__ __forceinline__ void st32(vec32* p, const vec32& v)
{
    asm volatile("st.global.cs.v8.u32 [%0], {%1,%2,%3,%4,%5,%6,%7,%8};"
                 :: "l"(p), "r"(v.lo.x), "r"(v.lo.y), "r"(v.lo.z), "r"(v.lo.w), "r"(v.hi.x), "r"(v.hi.y), "r"(v.hi.z), "r"(v.hi.w)
                 : "memory");
}

template <int Unroll, int Block, int Ld = CACHE_STREAMING, int St = CACHE_STREAMING>
struct stream_cfg {
    static constexpr int UNROLL = Unroll;  // independent 16-byte loads in flight per operand per thread
    static constexpr int BLOCK = Block;    // threads per CTA
    static constexpr int LD = Ld;
    static constexpr int ST = St;
};
// Measured on B200 (profiles/r01_tune_sweep2.log): bandwidth-bound ops are best
// with 2 vectors per operand per thread (7.03-7.08 TB/s vs 6.93-6.95 with 4),
// ALU-heavy ones (many distinct widths, fused per-field arithmetic) with 4
// (irregular u64 add_signed_saturate 6.58 vs 5.80 TB/s).
using default_stream_cfg = stream_cfg<2, 256>;
using heavy_stream_cfg = stream_cfg<4, 256>;
template <class Fn>
using stream_cfg_for = typename std::conditional<Fn::heavy, heavy_stream_cfg, default_stream_cfg>::type;

// apply a Word function to the 16 bytes of a vector
template <class Word, class F>
__device__ __forceinline__ uint4 map_words2(const uint4& a, const uint4& b, F f)
{
    uint4 r;
    if constexpr (sizeof(Word) == 4) {
        r.x = f(a.x, b.x); r.y = f(a.y, b.y); r.z = f(a.z, b.z); r.w = f(a.w, b.w);
    } else {
        const uint64_t a0 = (uint64_t(a.y) << 32) | a.x, a1 = (uint64_t(a.w) << 32) | a.z;
        const uint64_t b0 = (uint64_t(b.y) << 32) | b.x, b1 = (uint64_t(b.w) << 32) | b.z;
        const uint64_t r0 = f(a0, b0), r1 = f(a1, b1);
        r.x = uint32_t(r0); r.y = uint32_t(r0 >> 32); r.z = uint32_t(r1); r.w = uint32_t(r1 >> 32);
    }
    return r;
}

template <class Word, class F>
__device__ __forceinline__ uint4 map_words1(const uint4& a, F f)
{
    uint4 r;
    if constexpr (sizeof(Word) == 4) {
        r.x = f(a.x); r.y = f(a.y); r.z = f(a.z); r.w = f(a.w);
    } else {
        const uint64_t a0 = (uint64_t(a.y) << 32) | a.x, a1 = (uint64_t(a.w) << 32) | a.z;
        const uint64_t r0 = f(a0), r1 = f(a1);
        r.x = uint32_t(r0); r.y = uint32_t(r0 >> 32); r.z = uint32_t(r1); r.w = uint32_t(r1 >> 32);
    }
    return r;
}

// The SWAR op as a function object over one Word.
template <int Op, class P>
struct swar_fn {
    using word = typename P::word;
    static constexpr bool heavy = P::heavy_fill && !(Op == OP_ADD_WRAP || Op == OP_SUB_WRAP || Op == OP_OR || Op == OP_AND || Op == OP_XOR);
    // 32-byte global accesses: faster for light functors and for heavy compiled-in ones on 32-bit
    // words, slower for heavy 64-bit ones (registers), neutral to slower for heavy run-time-mask
    // ones; same-box A/B in profiles/r01z_ab_vec256.log
    static constexpr bool vec256 = !heavy || (sizeof(word) == 4 && P::compiled_in);
    P p;
    __device__ __forceinline__ word operator()(word a, word b) const { return apply_op<Op>(a, b, p); }
};

// ---------------------------------------------------------------------------
// out[i] = op(a[i], b[i]) over 16-byte vectors.  A CTA walks tiles of
// BLOCK * UNROLL vectors with a grid stride; within a tile consecutive threads
// touch consecutive vectors (every warp access is 512 contiguous bytes) and
// all 2 * UNROLL loads are issued before the first use.  `out` may alias a or b:
// each vector is read and written by the same thread, read first.
//
// Fn is the per-Word function object (swar_fn below, or fused_fn in
// cuda_kernels_planes.cuh); it travels as a __grid_constant__ parameter, so a
// static-mask Fn is empty and a runtime-mask Fn lives in the constant bank.
// Programmatic dependent launch (option "pdl"): the next launch in the stream may bring its CTAs onto the
// SMs while this grid drains (launch_dependents), and nothing is read or written before the grids this
// one depends on have completed and flushed (wait).  Both are no-ops for an ordinary launch.
__device__ __forceinline__ void pdl_prologue()
{
    asm volatile("griddepcontrol.launch_dependents;" ::: "memory");
    asm volatile("griddepcontrol.wait;" ::: "memory");
}

template <class Fn, class Cfg>
__global__ void __launch_bounds__(Cfg::BLOCK)
binary_stream_kernel(const uint4* a, const uint4* b, uint4* out, size_t nvec, const __grid_constant__ Fn fn)
{
    pdl_prologue();
    using W = typename Fn::word;
    constexpr int U = Cfg::UNROLL, B = Cfg::BLOCK;
    constexpr size_t tile = size_t(U) * B;
    const size_t nfull = nvec / tile;
    auto f = [&](W x, W y) { return fn(x, y); };

    for (size_t t = blockIdx.x; t < nfull; t += gridDim.x) {
        const size_t base = t * tile + threadIdx.x;
        uint4 va[U], vb[U];
#pragma unroll
        for (int u = 0; u < U; ++u) va[u] = ld16<Cfg::LD>(a + base + size_t(u) * B);
#pragma unroll
        for (int u = 0; u < U; ++u) vb[u] = ld16<Cfg::LD>(b + base + size_t(u) * B);
#pragma unroll
        for (int u = 0; u < U; ++u) st16<Cfg::ST>(out + base + size_t(u) * B, map_words2<W>(va[u], vb[u], f));
    }
    // the partial last tile goes to the CTA whose turn it would be
    if (blockIdx.x == nfull % gridDim.x) {
        for (size_t i = nfull * tile + threadIdx.x; i < nvec; i += B)
            st16<Cfg::ST>(out + i, map_words2<W>(ld16<Cfg::LD>(a + i), ld16<Cfg::LD>(b + i), f));
    }
}

// The same map with 32-byte vectors: UNROLL / 2 (at least 1) vectors per operand per thread, i.e.
// the same bytes in flight per thread as the 16-byte kernel.
template <class Fn, class Cfg>
__global__ void __launch_bounds__(Cfg::BLOCK)
binary_stream256_kernel(const vec32* a, const vec32* b, vec32* out, size_t nvec, const __grid_constant__ Fn fn)
{
    pdl_prologue();
    using W = typename Fn::word;
    constexpr int U = Cfg::UNROLL / 2 > 0 ? Cfg::UNROLL / 2 : 1, B = Cfg::BLOCK;
    constexpr size_t tile = size_t(U) * B;
    const size_t nfull = nvec / tile;
    auto f = [&](W x, W y) { return fn(x, y); };
    auto apply = [&](const vec32& x, const vec32& y) {
        vec32 r;
        r.lo = map_words2<W>(x.lo, y.lo, f);
        r.hi = map_words2<W>(x.hi, y.hi, f);
        return r;
    };
    for (size_t t = blockIdx.x; t < nfull; t += gridDim.x) {
        const size_t base = t * tile + threadIdx.x;
        vec32 va[U], vb[U];
#pragma unroll
        for (int u = 0; u < U; ++u) va[u] = ld32(a + base + size_t(u) * B);
#pragma unroll
        for (int u = 0; u < U; ++u) vb[u] = ld32(b + base + size_t(u) * B);
#pragma unroll
        for (int u = 0; u < U; ++u) st32(out + base + size_t(u) * B, apply(va[u], vb[u]));
    }
    if (blockIdx.x == nfull % gridDim.x) {
        for (size_t i = nfull * tile + threadIdx.x; i < nvec; i += B) st32(out + i, apply(ld32(a + i), ld32(b + i)));
    }
}

struct tile_counter {
    unsigned long long next;  // next tile to hand out
    unsigned long long done;  // CTAs that have finished; the last one re-arms both fields
};
constexpr unsigned kTileCounterSlots = 4096;

// Ordered persistent variant for bandwidth-bound functors (OPTIONAL, off by default).
// num_sms * k CTAs stay resident and draw tiles in ADDRESS ORDER from a global
// counter, so the window of the arrays the chip is working on stays as compact as
// with one CTA per tile while the number of resident warps is set by the grid size
// rather than the register budget.  On B200 4 CTAs of 256 threads per SM reached
// 7.13-7.22 TB/s on the light ops in some runs (profiles/r01_tune_dyn_scheduler.log)
// against a rock-steady 7.02-7.05 for one CTA per tile, but the gain does not
// reproduce run to run and neighbouring grid sizes fall off a cliff (5.3-6.6 TB/s;
// profiles/r01_tune_ordered_vs_tile.log), so the library default stays one CTA per
// tile; pint_cuda_set_option("ordered_ctas_per_sm", k) enables this path.
// The last CTA to finish re-arms the counter, so a launch needs no memset; launches
// that may overlap use different slots of a counter ring.
template <class Fn, class Cfg>
__global__ void __launch_bounds__(Cfg::BLOCK)
binary_stream_ordered_kernel(const uint4* a, const uint4* b, uint4* out, size_t nvec, tile_counter* tc,
                             const __grid_constant__ Fn fn)
{
    using W = typename Fn::word;
    constexpr int U = Cfg::UNROLL, B = Cfg::BLOCK;
    constexpr size_t tile = size_t(U) * B;
    const size_t nfull = nvec / tile;
    auto f = [&](W x, W y) { return fn(x, y); };
    __shared__ unsigned long long drawn;
    for (;;) {
        if (threadIdx.x == 0) drawn = atomicAdd(&tc->next, 1ULL);
        __syncthreads();
        const size_t t = drawn;
        __syncthreads();
        if (t > nfull) break;
        if (t == nfull) {  // the partial last tile
            for (size_t i = nfull * tile + threadIdx.x; i < nvec; i += B)
                st16<Cfg::ST>(out + i, map_words2<W>(ld16<Cfg::LD>(a + i), ld16<Cfg::LD>(b + i), f));
            continue;
        }
        const size_t base = t * tile + threadIdx.x;
        uint4 va[U], vb[U];
#pragma unroll
        for (int u = 0; u < U; ++u) va[u] = ld16<Cfg::LD>(a + base + size_t(u) * B);
#pragma unroll
        for (int u = 0; u < U; ++u) vb[u] = ld16<Cfg::LD>(b + base + size_t(u) * B);
#pragma unroll
        for (int u = 0; u < U; ++u) st16<Cfg::ST>(out + base + size_t(u) * B, map_words2<W>(va[u], vb[u], f));
    }
    if (threadIdx.x == 0 && atomicAdd(&tc->done, 1ULL) == gridDim.x - 1) {
        tc->next = 0;  // every CTA has made its last draw: re-arm for the next launch
        tc->done = 0;
    }
}

template <class Word, class Cfg>
__global__ void __launch_bounds__(Cfg::BLOCK)
unary_stream_ordered_kernel(const uint4* in, uint4* out, size_t nvec, tile_counter* tc,
                            const __grid_constant__ unary_params<Word> u)
{
    constexpr int U = Cfg::UNROLL, B = Cfg::BLOCK;
    constexpr size_t tile = size_t(U) * B;
    const size_t nfull = nvec / tile;
    auto f = [&](Word x) { return apply_unary(x, u); };
    __shared__ unsigned long long drawn;
    for (;;) {
        if (threadIdx.x == 0) drawn = atomicAdd(&tc->next, 1ULL);
        __syncthreads();
        const size_t t = drawn;
        __syncthreads();
        if (t > nfull) break;
        if (t == nfull) {
            for (size_t i = nfull * tile + threadIdx.x; i < nvec; i += B)
                st16<Cfg::ST>(out + i, map_words1<Word>(ld16<Cfg::LD>(in + i), f));
            continue;
        }
        const size_t base = t * tile + threadIdx.x;
        uint4 v[U];
#pragma unroll
        for (int k = 0; k < U; ++k) v[k] = ld16<Cfg::LD>(in + base + size_t(k) * B);
#pragma unroll
        for (int k = 0; k < U; ++k) st16<Cfg::ST>(out + base + size_t(k) * B, map_words1<Word>(v[k], f));
    }
    if (threadIdx.x == 0 && atomicAdd(&tc->done, 1ULL) == gridDim.x - 1) {
        tc->next = 0;
        tc->done = 0;
    }
}

// Element-granular variant: array heads/tails shorter than a vector and
// operands whose addresses are not congruent modulo 16.  One element is
// zero-extended into a Word; the replicated masks act on zeros above it.
template <class Elem, class Fn>
__global__ void __launch_bounds__(256)
binary_elem_kernel(const Elem* a, const Elem* b, Elem* out, size_t n, const __grid_constant__ Fn fn)
{
    using W = typename Fn::word;
    for (size_t i = size_t(blockIdx.x) * blockDim.x + threadIdx.x; i < n; i += size_t(gridDim.x) * blockDim.x)
        out[i] = Elem(fn(W(a[i]), W(b[i])));
}

template <class Word, class Cfg>
__global__ void __launch_bounds__(Cfg::BLOCK)
unary_stream_kernel(const uint4* in, uint4* out, size_t nvec, const __grid_constant__ unary_params<Word> u)
{
    pdl_prologue();
    constexpr int U = Cfg::UNROLL, B = Cfg::BLOCK;
    constexpr size_t tile = size_t(U) * B;
    const size_t nfull = nvec / tile;
    auto f = [&](Word x) { return apply_unary(x, u); };
    for (size_t t = blockIdx.x; t < nfull; t += gridDim.x) {
        const size_t base = t * tile + threadIdx.x;
        uint4 v[U];
#pragma unroll
        for (int k = 0; k < U; ++k) v[k] = ld16<Cfg::LD>(in + base + size_t(k) * B);
#pragma unroll
        for (int k = 0; k < U; ++k) st16<Cfg::ST>(out + base + size_t(k) * B, map_words1<Word>(v[k], f));
    }
    if (blockIdx.x == nfull % gridDim.x) {
        for (size_t i = nfull * tile + threadIdx.x; i < nvec; i += B)
            st16<Cfg::ST>(out + i, map_words1<Word>(ld16<Cfg::LD>(in + i), f));
    }
}

template <class Word, class Cfg>
__global__ void __launch_bounds__(Cfg::BLOCK)
unary_stream256_kernel(const vec32* in, vec32* out, size_t nvec, const __grid_constant__ unary_params<Word> u)
{
    pdl_prologue();
    constexpr int U = Cfg::UNROLL, B = Cfg::BLOCK;  // twice the bytes per thread of the 16-byte kernel: one stream only
    constexpr size_t tile = size_t(U) * B;
    const size_t nfull = nvec / tile;
    auto f = [&](Word x) { return apply_unary(x, u); };
    auto apply = [&](const vec32& x) {
        vec32 r;
        r.lo = map_words1<Word>(x.lo, f);
        r.hi = map_words1<Word>(x.hi, f);
        return r;
    };
    for (size_t t = blockIdx.x; t < nfull; t += gridDim.x) {
        const size_t base = t * tile + threadIdx.x;
        vec32 v[U];
#pragma unroll
        for (int k = 0; k < U; ++k) v[k] = ld32(in + base + size_t(k) * B);
#pragma unroll
        for (int k = 0; k < U; ++k) st32(out + base + size_t(k) * B, apply(v[k]));
    }
    if (blockIdx.x == nfull % gridDim.x) {
        for (size_t i = nfull * tile + threadIdx.x; i < nvec; i += B) st32(out + i, apply(ld32(in + i)));
    }
}

template <class Elem, class Word>
__global__ void __launch_bounds__(256)
unary_elem_kernel(const Elem* in, Elem* out, size_t n, const __grid_constant__ unary_params<Word> u)
{
    for (size_t i = size_t(blockIdx.x) * blockDim.x + threadIdx.x; i < n; i += size_t(gridDim.x) * blockDim.x)
        out[i] = Elem(apply_unary(Word(in[i]), u));
}

#if !defined(__CUDACC_RTC__)  // host-side launchers: not part of a run-time (NVRTC) specialisation
// ---------------------------------------------------------------------------
// Launch geometry.  ctas_per_sm == 0 (default): one CTA per tile, handed out in
// address order by the hardware block scheduler -- measured 10-15 % faster on
// B200 than any persistent grid-stride grid (profiles/r01_tune_sweep1.log: 7.0-7.1
// vs 6.0-6.5 TB/s), because the SMs' working window over the arrays stays
// compact.  ctas_per_sm > 0: persistent grid of num_sms * ctas_per_sm CTAs
// walking the tiles with a grid stride (kept for comparison and for tuning).
struct launch_geom {
    int num_sms = 148;
    int ctas_per_sm = 0;
    unsigned long long* launch_counter = nullptr;  // bumped once per kernel launch
    // Ordered persistent mode for bandwidth-bound functors (see binary_stream_ordered_kernel):
    // a ring of zeroed device counters, a host cursor into it, and the grid size per SM.
    tile_counter* counters = nullptr;
    unsigned* counter_cursor = nullptr;
    int ordered_ctas_per_sm = 0;  // 0 = off
    int vec256 = 1;               // 32-byte vectors: 0 never, 1 where measured faster (light functors), 2 wherever aligned
    int pdl = 1;                  // the binary stream kernels are launched with programmatic stream serialization (0 = plain launches)
    int unpack_bulk = 1;          // unpack to int32 planes through shared memory + bulk stores (cuda_kernels_planes_bulk.cuh)
};

// The ordered persistent kernels take a counter slot chosen on the host at launch time: a graph
// capture would bake the slot in and concurrent replays would share it (tiles split between the
// replays, output left unwritten).  While the stream is capturing the plain one-CTA-per-tile kernels
// run instead.
inline bool stream_is_capturing(cudaStream_t s)
{
    cudaStreamCaptureStatus st = cudaStreamCaptureStatusNone;
    if (cudaStreamIsCapturing(s, &st) != cudaSuccess) { cudaGetLastError(); return true; }
    return st != cudaStreamCaptureStatusNone;
}

inline tile_counter* next_tile_counter(const launch_geom& g)
{
    const unsigned slot = __atomic_fetch_add(g.counter_cursor, 1u, __ATOMIC_RELAXED) % kTileCounterSlots;
    return g.counters + slot;
}

inline unsigned grid_for(size_t ntiles, const launch_geom& g)
{
    size_t n = ntiles;
    if (g.ctas_per_sm > 0) {
        const size_t cap = size_t(g.num_sms) * size_t(g.ctas_per_sm);
        if (n > cap) n = cap;
    }
    if (n > 0x7FFFFFFFull) n = 0x7FFFFFFFull;  // gridDim.x limit; the kernels grid-stride beyond it
    return unsigned(n ? n : 1);
}

// kern<<<grid, block, 0, stream>>>(args...), or the same launch with programmatic stream serialization
// allowed: the grid may start while the previous launch of the stream drains (the kernel's pdl_prologue
// holds its work until that launch has completed)
template <class... KArgs, class... Args>
inline cudaError_t launch_stream_kernel_smem(void (*kern)(KArgs...), unsigned grid, unsigned block, size_t smem, cudaStream_t stream, bool pdl,
                                            Args&&... args)
{
    if (!pdl) {
        kern<<<grid, block, smem, stream>>>(KArgs(args)...);
        return cudaSuccess;
    }
    cudaLaunchConfig_t cfg{};
    cfg.gridDim = dim3(grid);
    cfg.blockDim = dim3(block);
    cfg.dynamicSmemBytes = smem;
    cfg.stream = stream;
    cudaLaunchAttribute attr[1];
    attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
    attr[0].val.programmaticStreamSerializationAllowed = 1;
    cfg.attrs = attr;
    cfg.numAttrs = 1;
    return cudaLaunchKernelEx(&cfg, kern, KArgs(args)...);
}

template <class... KArgs, class... Args>
inline cudaError_t launch_stream_kernel(void (*kern)(KArgs...), unsigned grid, unsigned block, cudaStream_t stream, bool pdl, Args&&... args)
{
    return launch_stream_kernel_smem(kern, grid, block, 0, stream, pdl, static_cast<Args&&>(args)...);
}

inline void count_launch(const launch_geom& g)
{
    if (g.launch_counter) __atomic_fetch_add(g.launch_counter, 1ULL, __ATOMIC_RELAXED);
}

// Split [0, n) elements into head (element kernel), body (vector kernel) and
// tail (element kernel) given the operand addresses.  Returns false when the
// addresses are not congruent modulo 16 (everything goes element-wise).
inline bool split_aligned(size_t elem_bytes, size_t n, const void* const* ptrs, int nptrs,
                          size_t* head, size_t* nvec, size_t* tail, size_t vec_bytes = 16)
{
    const uintptr_t mask = uintptr_t(vec_bytes - 1);
    const uintptr_t off = reinterpret_cast<uintptr_t>(ptrs[0]) & mask;
    for (int i = 1; i < nptrs; ++i)
        if ((reinterpret_cast<uintptr_t>(ptrs[i]) & mask) != off) return false;
    if (off % elem_bytes) return false;
    size_t h = off ? (vec_bytes - off) / elem_bytes : 0;
    if (h > n) h = n;
    const size_t per_vec = vec_bytes / elem_bytes;
    *head = h;
    *nvec = (n - h) / per_vec;
    *tail = (n - h) - *nvec * per_vec;
    return true;
}

template <class Elem, class Fn, class Cfg = stream_cfg_for<Fn>>
cudaError_t launch_binary(const void* a, const void* b, void* out, size_t n, const Fn& fn,
                          cudaStream_t stream, const launch_geom& g)
{
    if (n == 0) return cudaSuccess;
    const Elem* ea = static_cast<const Elem*>(a);
    const Elem* eb = static_cast<const Elem*>(b);
    Elem* eo = static_cast<Elem*>(out);
    const void* ptrs[3] = {a, b, out};
    size_t head = 0, nvec = 0, tail = 0;
    auto elems = [&](size_t first, size_t count) {
        if (!count) return;
        const unsigned grid = grid_for((count + 255) / 256, g);
        binary_elem_kernel<Elem, Fn><<<grid, 256, 0, stream>>>(ea + first, eb + first, eo + first, count, fn);
        count_launch(g);
    };
    // 32-byte vectors when the operands allow it (and the ordered persistent mode is not asked for)
    if ((g.vec256 == 2 || (g.vec256 == 1 && Fn::vec256)) && Cfg::LD == CACHE_STREAMING && Cfg::ST == CACHE_STREAMING &&
        g.ordered_ctas_per_sm <= 0 &&
        split_aligned(sizeof(Elem), n, ptrs, 3, &head, &nvec, &tail, 32) && nvec) {
        elems(0, head);
        constexpr size_t tile = size_t(Cfg::UNROLL / 2 > 0 ? Cfg::UNROLL / 2 : 1) * Cfg::BLOCK;
        const size_t ntiles = (nvec + tile - 1) / tile;
        if (cudaError_t e = launch_stream_kernel(binary_stream256_kernel<Fn, Cfg>, grid_for(ntiles, g), Cfg::BLOCK, stream, g.pdl != 0,
                                                 reinterpret_cast<const vec32*>(ea + head), reinterpret_cast<const vec32*>(eb + head),
                                                 reinterpret_cast<vec32*>(eo + head), nvec, fn))
            return e;
        count_launch(g);
        elems(n - tail, tail);
        return cudaGetLastError();
    }
    const bool vec_ok = split_aligned(sizeof(Elem), n, ptrs, 3, &head, &nvec, &tail);
    if (!vec_ok) {
        elems(0, n);
        return cudaGetLastError();
    }
    elems(0, head);
    if (nvec) {
        constexpr size_t tile = size_t(Cfg::UNROLL) * Cfg::BLOCK;
        const size_t ntiles = (nvec + tile - 1) / tile;
        const size_t ordered_grid = size_t(g.num_sms) * size_t(g.ordered_ctas_per_sm > 0 ? g.ordered_ctas_per_sm : 0);
        if (!Fn::heavy && g.counters && g.ctas_per_sm == 0 && ordered_grid && ntiles > 2 * ordered_grid &&
            !stream_is_capturing(stream)) {
            binary_stream_ordered_kernel<Fn, Cfg><<<unsigned(ordered_grid), Cfg::BLOCK, 0, stream>>>(
                reinterpret_cast<const uint4*>(ea + head), reinterpret_cast<const uint4*>(eb + head),
                reinterpret_cast<uint4*>(eo + head), nvec, next_tile_counter(g), fn);
        } else {
            if (cudaError_t e = launch_stream_kernel(binary_stream_kernel<Fn, Cfg>, grid_for(ntiles, g), Cfg::BLOCK, stream, g.pdl != 0,
                                                     reinterpret_cast<const uint4*>(ea + head), reinterpret_cast<const uint4*>(eb + head),
                                                     reinterpret_cast<uint4*>(eo + head), nvec, fn))
                return e;
        }
        count_launch(g);
    }
    elems(n - tail, tail);
    return cudaGetLastError();
}

template <class Elem, class Word, class Cfg = default_stream_cfg>
cudaError_t launch_unary(const void* in, void* out, size_t n, const unary_params<Word>& u,
                         cudaStream_t stream, const launch_geom& g)
{
    if (n == 0) return cudaSuccess;
    const Elem* ei = static_cast<const Elem*>(in);
    Elem* eo = static_cast<Elem*>(out);
    const void* ptrs[2] = {in, out};
    size_t head = 0, nvec = 0, tail = 0;
    auto elems = [&](size_t first, size_t count) {
        if (!count) return;
        const unsigned grid = grid_for((count + 255) / 256, g);
        unary_elem_kernel<Elem, Word><<<grid, 256, 0, stream>>>(ei + first, eo + first, count, u);
        count_launch(g);
    };
    // 32-byte accesses measured 0.5 % SLOWER for the one-in/one-out maps (profiles/r02f_ab_vec256_unary.log):
    // only when forced (option "vec256" = 2, kept for parity tests and re-measurement)
    if (g.vec256 == 2 && Cfg::LD == CACHE_STREAMING && Cfg::ST == CACHE_STREAMING && g.ordered_ctas_per_sm <= 0 &&
        split_aligned(sizeof(Elem), n, ptrs, 2, &head, &nvec, &tail, 32) && nvec) {
        elems(0, head);
        constexpr size_t tile = size_t(Cfg::UNROLL) * Cfg::BLOCK;
        const size_t ntiles = (nvec + tile - 1) / tile;
        if (cudaError_t e = launch_stream_kernel(unary_stream256_kernel<Word, Cfg>, grid_for(ntiles, g), Cfg::BLOCK, stream, g.pdl != 0,
                                                 reinterpret_cast<const vec32*>(ei + head), reinterpret_cast<vec32*>(eo + head), nvec, u))
            return e;
        count_launch(g);
        elems(n - tail, tail);
        return cudaGetLastError();
    }
    const bool vec_ok = split_aligned(sizeof(Elem), n, ptrs, 2, &head, &nvec, &tail);
    if (!vec_ok) {
        elems(0, n);
        return cudaGetLastError();
    }
    elems(0, head);
    if (nvec) {
        constexpr size_t tile = size_t(Cfg::UNROLL) * Cfg::BLOCK;
        const size_t ntiles = (nvec + tile - 1) / tile;
        const size_t ordered_grid = size_t(g.num_sms) * size_t(g.ordered_ctas_per_sm > 0 ? g.ordered_ctas_per_sm : 0);
        if (g.counters && g.ctas_per_sm == 0 && ordered_grid && ntiles > 2 * ordered_grid && !stream_is_capturing(stream)) {
            unary_stream_ordered_kernel<Word, Cfg><<<unsigned(ordered_grid), Cfg::BLOCK, 0, stream>>>(
                reinterpret_cast<const uint4*>(ei + head), reinterpret_cast<uint4*>(eo + head), nvec,
                next_tile_counter(g), u);
        } else {
            if (cudaError_t e = launch_stream_kernel(unary_stream_kernel<Word, Cfg>, grid_for(ntiles, g), Cfg::BLOCK, stream, g.pdl != 0,
                                                     reinterpret_cast<const uint4*>(ei + head), reinterpret_cast<uint4*>(eo + head), nvec, u))
                return e;
        }
        count_launch(g);
    }
    elems(n - tail, tail);
    return cudaGetLastError();
}

#endif  // !__CUDACC_RTC__

}  // namespace kernels
}  // namespace cuda
}  // namespace pint

#endif  // PINT_CUDA_KERNELS_STREAM_CUH_INCLUDED
