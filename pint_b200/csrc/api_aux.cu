// api_aux.cu -- synthetic inputs, checksums, the benchmark-shaped pair
// reduction, the multi-GPU partitioner and the host-buffer (end-to-end) path.
#include <mutex>

#include "api_common.hpp"

namespace pint_cuda_impl {
namespace {

using namespace pint::cuda;
using namespace pint::cuda::kernels;

__device__ __forceinline__ uint64_t splitmix64(uint64_t x)
{
    x += 0x9E3779B97F4A7C15ULL;
    x = (x ^ (x >> 30)) * 0xBF58476D1CE4E5B9ULL;
    x = (x ^ (x >> 27)) * 0x94D049BB133111EBULL;
    return x ^ (x >> 31);
}

// out[i] = (Elem) splitmix64(seed + start + i)
template <class Elem>
__global__ void __launch_bounds__(256) fill_kernel(Elem* out, size_t n, uint64_t base)
{
    constexpr int PER = 16 / sizeof(Elem);
    const size_t nvec = (reinterpret_cast<uintptr_t>(out) & 15u) ? 0 : n / PER;
    for (size_t v = size_t(blockIdx.x) * blockDim.x + threadIdx.x; v < nvec; v += size_t(gridDim.x) * blockDim.x) {
        union { uint4 q; Elem e[PER]; } u;
#pragma unroll
        for (int j = 0; j < PER; ++j) u.e[j] = Elem(splitmix64(base + v * PER + j));
        st16<CACHE_STREAMING>(reinterpret_cast<uint4*>(out) + v, u.q);
    }
    for (size_t i = nvec * PER + size_t(blockIdx.x) * blockDim.x + threadIdx.x; i < n; i += size_t(gridDim.x) * blockDim.x)
        out[i] = Elem(splitmix64(base + i));
}

// One atomic per CTA into result[0] (sum) and, only when WithXor, into result[1] (xor): a sum-only
// caller (pairs_reduce: ONE device uint64_t, pint_cuda.h) never touches result[1].
template <bool WithXor>
__device__ __forceinline__ void block_reduce_sum_xor(uint64_t s, uint64_t x, uint64_t* result)
{
    __shared__ uint64_t ss[8], sx[8];
    for (int o = 16; o; o >>= 1) {
        s += __shfl_down_sync(0xffffffffu, s, o);
        if constexpr (WithXor) x ^= __shfl_down_sync(0xffffffffu, x, o);
    }
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    if (lane == 0) { ss[warp] = s; if constexpr (WithXor) sx[warp] = x; }
    __syncthreads();
    if (threadIdx.x == 0) {
        for (int w = 1; w < (int)(blockDim.x >> 5); ++w) { s += ss[w]; if constexpr (WithXor) x ^= sx[w]; }
        atomicAdd(reinterpret_cast<unsigned long long*>(result), (unsigned long long)s);
        if constexpr (WithXor) atomicXor(reinterpret_cast<unsigned long long*>(result + 1), (unsigned long long)x);
    }
}

// result[0] += sum of elements, result[1] ^= xor of elements (integer, so the
// order of the atomics does not matter: bit-exact and deterministic)
template <class Elem>
__global__ void __launch_bounds__(256) checksum_kernel(const Elem* x, size_t n, uint64_t* result)
{
    uint64_t s = 0, r = 0;
    for (size_t i = size_t(blockIdx.x) * blockDim.x + threadIdx.x; i < n; i += size_t(gridDim.x) * blockDim.x) {
        const uint64_t v = x[i];
        s += v;
        r ^= v;
    }
    block_reduce_sum_xor<true>(s, r, result);
}

template <class Elem>
__global__ void __launch_bounds__(256) checksum_vec_kernel(const uint4* x, size_t nvec, uint64_t* result)
{
    constexpr int PER = 16 / sizeof(Elem);
    uint64_t s = 0, r = 0;
    for (size_t i = size_t(blockIdx.x) * blockDim.x + threadIdx.x; i < nvec; i += size_t(gridDim.x) * blockDim.x) {
        union { uint4 q; Elem e[PER]; } u;
        u.q = ld16<CACHE_STREAMING>(x + i);
#pragma unroll
        for (int j = 0; j < PER; ++j) { s += u.e[j]; r ^= u.e[j]; }
    }
    block_reduce_sum_xor<true>(s, r, result);
}

// tests/pint_bench.cpp:90-94: sum += op(P(pair.first), P(pair.second)).value()
// P: the mask policy -- static_masks<SL> for the layouts of registry_layouts.def (constants folded into the code),
// runtime_masks<W, Steps> otherwise (Steps smear steps: 2 / 4 for layouts whose widest field has at most 4 / 16 bits)
__global__ void zero_sum_kernel(uint64_t* sum)
{
    pint::cuda::kernels::pdl_prologue();
    if (threadIdx.x == 0) *sum = 0;
}

// loads per thread and tile: 4 keeps the memory-bound ops at the read ceiling (7.2 TB/s at 10^8 pairs); the ops that are
// bound by the integer pipe (saturate / min / max on layouts with three or more field widths, e.g. the benchmark's
// <1,2,3,4,5,6,11>) gain from CTAs that live twice as long: 3.3-3.8 -> 5.1-5.2 TB/s at 2^29 pairs (profiles/r26_ab_reduce.jsonl)
template <int Op, class P>
__host__ __device__ constexpr int reduce_unroll()
{
    return (P::heavy_fill && !(Op == OP_ADD_WRAP || Op == OP_SUB_WRAP || Op == OP_OR || Op == OP_AND || Op == OP_XOR)) ? 8 : 4;
}

template <class Elem, int Op, bool V32, class P>
__global__ void __launch_bounds__(256)
pairs_reduce_kernel(const Elem* pairs, size_t n, uint64_t* sum, const __grid_constant__ P p)
{
    pint::cuda::kernels::pdl_prologue();
    using W = word_of<Elem>;
    constexpr int PER = 8 / sizeof(Elem);  // pairs per 16-byte vector
    uint64_t s = 0;
    const bool vec_ok = (reinterpret_cast<uintptr_t>(pairs) & 15u) == 0;
    const size_t nvec = vec_ok ? n / PER : 0;
    // tiles of U * blockDim vectors, U independent loads per thread issued before the first use
    constexpr int U = reduce_unroll<Op, P>();
    const uint4* vec = reinterpret_cast<const uint4*>(pairs);
    const size_t tile = size_t(U) * blockDim.x, nfull = nvec / tile;
    auto add_vec = [&](const uint4& q) {
        union { uint4 q; Elem e[2 * PER]; } u;
        u.q = q;
#pragma unroll
        for (int j = 0; j < PER; ++j) s += Elem(apply_op<Op>(W(u.e[2 * j]), W(u.e[2 * j + 1]), p));
    };
    if constexpr (V32) {
        // the array is 32-byte aligned (host-checked): U 256-bit loads per thread, tiles of 2 * U * blockDim vectors
        const pint::cuda::kernels::vec32* v32 = reinterpret_cast<const pint::cuda::kernels::vec32*>(pairs);
        const size_t nfull32 = nvec / (2 * tile);
        for (size_t t = blockIdx.x; t < nfull32; t += gridDim.x) {
            pint::cuda::kernels::vec32 q[U];
#pragma unroll
            for (int k = 0; k < U; ++k) q[k] = pint::cuda::kernels::ld32(v32 + t * tile + size_t(k) * blockDim.x + threadIdx.x);
#pragma unroll
            for (int k = 0; k < U; ++k) { add_vec(q[k].lo); add_vec(q[k].hi); }
        }
        for (size_t v = nfull32 * 2 * tile + size_t(blockIdx.x) * blockDim.x + threadIdx.x; v < nfull * tile; v += size_t(gridDim.x) * blockDim.x)
            add_vec(ld16<CACHE_STREAMING>(vec + v));
    } else {
    for (size_t t = blockIdx.x; t < nfull; t += gridDim.x) {
        uint4 q[U];
#pragma unroll
        for (int k = 0; k < U; ++k) q[k] = ld16<CACHE_STREAMING>(vec + t * tile + size_t(k) * blockDim.x + threadIdx.x);
#pragma unroll
        for (int k = 0; k < U; ++k) add_vec(q[k]);
    }
    }
    for (size_t v = nfull * tile + size_t(blockIdx.x) * blockDim.x + threadIdx.x; v < nvec; v += size_t(gridDim.x) * blockDim.x)
        add_vec(ld16<CACHE_STREAMING>(vec + v));
    for (size_t i = nvec * PER + size_t(blockIdx.x) * blockDim.x + threadIdx.x; i < n; i += size_t(gridDim.x) * blockDim.x)
        s += Elem(apply_op<Op>(W(pairs[2 * i]), W(pairs[2 * i + 1]), p));
    block_reduce_sum_xor<false>(s, 0, sum);
}

template <class Elem, class P, int... Ops>
cudaError_t pairs_ops(int op, const P& p, const void* pairs, size_t n, uint64_t* sum, cudaStream_t s,
                      const launch_geom& g, std::integer_sequence<int, Ops...>)
{
    // one CTA per tile of the main loop (U loads of 16 or 32 bytes per thread): a grid sized in PAIRS / 1024 launched
    // 2 ... 16 times more CTAs than tiles, and every empty one still paid its launch, reduction and atomic
    // (10^8 uint32_t pairs, add_wrap: 5131 -> 7224 GB/s, profiles/r26_ab_reduce.jsonl)
    // 256-bit loads, 128 bytes in flight per thread: 6221 -> 6950 GB/s (profiles/r03s_ab_pairs_reduce.log)
    const bool v32 = g.vec256 != 0 && (uintptr_t)pairs % 32 == 0;
    cudaError_t rc = cudaSuccess;
    auto launch = [&](auto opc) {
        constexpr int O = decltype(opc)::value;
        const size_t tile_pairs = size_t(v32 ? 2 : 1) * reduce_unroll<O, P>() * 256 * (8 / sizeof(Elem));
        const unsigned grid = grid_for((n + tile_pairs - 1) / tile_pairs, g);
        if (v32) rc = launch_stream_kernel(pairs_reduce_kernel<Elem, O, true, P>, grid, 256, s, g.pdl != 0, (const Elem*)pairs, n, sum, p);
        else rc = launch_stream_kernel(pairs_reduce_kernel<Elem, O, false, P>, grid, 256, s, g.pdl != 0, (const Elem*)pairs, n, sum, p);
        return 0;
    };
    ((op == Ops ? launch(std::integral_constant<int, Ops>{}) : 0), ...);
    count_launch(g);
    return rc != cudaSuccess ? rc : cudaGetLastError();
}

// run-time masks: as many smear steps as the widest field needs (the benchmark's own type-2 layout
// <1,2,3,4,5,6,11> with add_signed_saturate was ALU-bound at 3.3 TB/s with all five, profiles/r17_replay_pint_bench.jsonl)
template <class Elem>
cudaError_t pairs_runtime(int op, const layout_desc& d, const void* pairs, size_t n, uint64_t* sum, cudaStream_t s, const launch_geom& g)
{
    using W = word_of<Elem>;
    const auto m = make_word_masks<W>(d);
    auto ops = std::make_integer_sequence<int, pint::cuda::OP_NUM_BINARY>();
    if (m.nsh > 1 && m.nsmear <= 2) return pairs_ops<Elem>(op, runtime_masks<W, 2>(m), pairs, n, sum, s, g, ops);
    if (m.nsh > 1 && m.nsmear <= 4) return pairs_ops<Elem>(op, runtime_masks<W, 4>(m), pairs, n, sum, s, g, ops);
    return pairs_ops<Elem>(op, runtime_masks<W>(m), pairs, n, sum, s, g, ops);
}

// compiled-in masks for the layouts of registry_layouts.def (which include the five layouts of the reference's
// benchmark, tests/pint_bench.cpp:88,158,376,469,639)
struct static_pairs_entry {
    layout_desc desc;
    cudaError_t (*run)(int op, const void* pairs, size_t n, uint64_t* sum, cudaStream_t s, const launch_geom& g);
};

template <class SL>
cudaError_t pairs_static(int op, const void* pairs, size_t n, uint64_t* sum, cudaStream_t s, const launch_geom& g)
{
    return pairs_ops<typename SL::integer>(op, static_masks<SL>{}, pairs, n, sum, s, g,
                                           std::make_integer_sequence<int, pint::cuda::OP_NUM_BINARY>());
}

const static_pairs_entry kStaticPairs[] = {
#define X(part, I, ...) {static_layout<I, __VA_ARGS__>::desc, &pairs_static<static_layout<I, __VA_ARGS__>>},
#include "registry_layouts.def"
#undef X
};

}  // namespace
}  // namespace pint_cuda_impl

using namespace pint_cuda_impl;

extern "C" {

int pint_cuda_fill_splitmix(void* out, uint32_t elem_bytes, size_t n, uint64_t seed, uint64_t start, pint_stream_t stream)
{
    if (elem_bytes != 1 && elem_bytes != 2 && elem_bytes != 4 && elem_bytes != 8) return PINT_E_ELEM_BYTES;
    if (n == 0) return PINT_OK;
    if (!out) return PINT_E_NULL;
    if ((uintptr_t)out % elem_bytes) return PINT_E_ARG;
    const launch_geom g = geom();
    const unsigned grid = grid_for((n * elem_bytes / 16 + 255) / 256, g);
    cudaStream_t s = (cudaStream_t)stream;
    const uint64_t base = seed + start;
    switch (elem_bytes) {
    case 1: fill_kernel<uint8_t><<<grid, 256, 0, s>>>((uint8_t*)out, n, base); break;
    case 2: fill_kernel<uint16_t><<<grid, 256, 0, s>>>((uint16_t*)out, n, base); break;
    case 4: fill_kernel<uint32_t><<<grid, 256, 0, s>>>((uint32_t*)out, n, base); break;
    default: fill_kernel<uint64_t><<<grid, 256, 0, s>>>((uint64_t*)out, n, base); break;
    }
    count_launch(g);
    return (int)cudaGetLastError();
}

int pint_cuda_checksum(const void* x, uint32_t elem_bytes, size_t n, uint64_t* result_dev, pint_stream_t stream)
{
    if (elem_bytes != 1 && elem_bytes != 2 && elem_bytes != 4 && elem_bytes != 8) return PINT_E_ELEM_BYTES;
    if (!result_dev) return PINT_E_NULL;
    cudaStream_t s = (cudaStream_t)stream;
    if (cudaError_t e = cudaMemsetAsync(result_dev, 0, 2 * sizeof(uint64_t), s)) return (int)e;
    if (n == 0) return PINT_OK;
    if (!x) return PINT_E_NULL;
    if ((uintptr_t)x % elem_bytes) return PINT_E_ARG;
    launch_geom g = geom();
    g.ctas_per_sm = 8;  // reductions end in one atomic per CTA: keep the grid at 8 CTAs per SM
    const bool vec_ok = ((uintptr_t)x & 15u) == 0;
    const size_t per = 16 / elem_bytes;
    const size_t nvec = vec_ok ? n / per : 0, rest = n - nvec * per;
    const char* tail = (const char*)x + nvec * 16;
#define PINT_CK(T)                                                                                            \
    do {                                                                                                      \
        if (nvec) { checksum_vec_kernel<T><<<grid_for((nvec + 1023) / 1024, g), 256, 0, s>>>((const uint4*)x, nvec, result_dev); count_launch(g); } \
        if (rest) { checksum_kernel<T><<<grid_for((rest + 255) / 256, g), 256, 0, s>>>((const T*)tail, rest, result_dev); count_launch(g); }        \
    } while (0)
    switch (elem_bytes) {
    case 1: PINT_CK(uint8_t); break;
    case 2: PINT_CK(uint16_t); break;
    case 4: PINT_CK(uint32_t); break;
    default: PINT_CK(uint64_t); break;
    }
#undef PINT_CK
    return (int)cudaGetLastError();
}

int pint_cuda_pairs_reduce(int op, const pint_layout* layout, const void* pairs, size_t n, uint64_t* sum_dev,
                           pint_stream_t stream)
{
    layout_desc d;
    if (int rc = check_layout(layout, &d)) return rc;
    if (op < 0 || op >= PINT_NUM_BINARY_OPS) return PINT_E_OP;
    if (!sum_dev) return PINT_E_NULL;
    cudaStream_t s = (cudaStream_t)stream;
    if (n == 0) return (int)cudaMemsetAsync(sum_dev, 0, sizeof(uint64_t), s);
    if (!pairs) return PINT_E_NULL;
    if ((uintptr_t)pairs % d.elem_bytes) return PINT_E_ARG;
    // One CTA per tile and one atomic per CTA (24 414 atomics for the benchmark's 10^8 uint32_t pairs): measured
    // faster than 8 persistent CTAs per SM, 7224 against 6943 GB/s (profiles/r26_ab_reduce.jsonl).
    const launch_geom g = geom();
    // *sum_dev = 0 by a one-thread kernel instead of cudaMemsetAsync, and both launches with programmatic stream
    // serialization (option "pdl"): the reduction's CTAs are resident and waiting when the zeroing completes, and the
    // zeroing itself is resident when the previous pass of the stream drains.  Back-to-back passes over 10^8 pairs,
    // with the grid above: 132 -> 121 us (profiles/r25_ / r28_replay_pint_bench.jsonl; letting the reduction read before
    // its wait changed nothing, r28b).  Order is unchanged: every kernel waits for the launch before it.
    if (cudaError_t e = launch_stream_kernel(zero_sum_kernel, 1, 32, s, g.pdl != 0, sum_dev)) return (int)e;
    count_launch(g);
    if (!opts().force_generic)
        for (const static_pairs_entry& e : kStaticPairs)
            if (same_layout(e.desc, d)) return (int)e.run(op, pairs, n, sum_dev, s, g);
    switch (d.elem_bytes) {
    case 1: return (int)pairs_runtime<uint8_t>(op, d, pairs, n, sum_dev, s, g);
    case 2: return (int)pairs_runtime<uint16_t>(op, d, pairs, n, sum_dev, s, g);
    case 4: return (int)pairs_runtime<uint32_t>(op, d, pairs, n, sum_dev, s, g);
    default: return (int)pairs_runtime<uint64_t>(op, d, pairs, n, sum_dev, s, g);
    }
}

int pint_cuda_plan_shards(size_t n, uint32_t elem_bytes, int nshards, size_t* begin, size_t* count)
{
    if (elem_bytes != 1 && elem_bytes != 2 && elem_bytes != 4 && elem_bytes != 8) return PINT_E_ELEM_BYTES;
    if (nshards < 1 || !begin || !count) return PINT_E_ARG;
    const size_t per128 = 128 / elem_bytes;           // elements per 128 bytes
    size_t base = (n / (size_t)nshards) / per128 * per128;  // shard length rounded down to 128 B
    size_t pos = 0;
    for (int k = 0; k < nshards; ++k) {
        begin[k] = pos;
        count[k] = (k == nshards - 1) ? n - pos : base;
        pos += count[k];
    }
    return PINT_OK;
}

int pint_cuda_sharded_binary(int op, const pint_layout* layout, int ngpu, const int* devices, const void* const* a,
                             const void* const* b, void* const* out, const size_t* counts, const pint_stream_t* streams)
{
    if (ngpu < 1 || !devices || !a || !b || !out || !counts) return PINT_E_ARG;
    int prev = 0;
    if (cudaError_t e = cudaGetDevice(&prev)) return (int)e;
    int rc = PINT_OK;
    for (int k = 0; k < ngpu && rc == PINT_OK; ++k) {
        if (cudaError_t e = cudaSetDevice(devices[k])) { rc = (int)e; break; }
        rc = binary_impl(op, false, layout, a[k], b[k], out[k], counts[k], streams ? (cudaStream_t)streams[k] : nullptr);
    }
    cudaSetDevice(prev);
    return rc;
}

int pint_cuda_sharded_fused_unpack_arith_repack(int op, const pint_layout* layout, int ngpu, const int* devices,
                                                const void* const* a, const void* const* b, void* const* out,
                                                const size_t* counts, const pint_stream_t* streams)
{
    if (ngpu < 1 || !devices || !a || !b || !out || !counts) return PINT_E_ARG;
    int prev = 0;
    if (cudaError_t e = cudaGetDevice(&prev)) return (int)e;
    int rc = PINT_OK;
    for (int k = 0; k < ngpu && rc == PINT_OK; ++k) {
        if (cudaError_t e = cudaSetDevice(devices[k])) { rc = (int)e; break; }
        rc = binary_impl(op, true, layout, a[k], b[k], out[k], counts[k], streams ? (cudaStream_t)streams[k] : nullptr);
    }
    cudaSetDevice(prev);
    return rc;
}

int pint_cuda_sharded_unpack(const pint_layout* layout, int is_signed, int ngpu, const int* devices,
                             const void* const* packed, void* const* planes, int plane_bytes,
                             const size_t* counts, const pint_stream_t* streams)
{
    if (ngpu < 1 || !devices || !packed || !planes || !counts || !layout) return PINT_E_ARG;
    int prev = 0;
    if (cudaError_t e = cudaGetDevice(&prev)) return (int)e;
    int rc = PINT_OK;
    for (int k = 0; k < ngpu && rc == PINT_OK; ++k) {
        if (cudaError_t e = cudaSetDevice(devices[k])) { rc = (int)e; break; }
        rc = pint_cuda_unpack(layout, is_signed, packed[k], planes + (size_t)k * layout->nfields, plane_bytes, counts[k],
                              streams ? streams[k] : nullptr);
    }
    cudaSetDevice(prev);
    return rc;
}

int pint_cuda_sharded_repack(const pint_layout* layout, int ngpu, const int* devices, const void* const* planes,
                             int plane_bytes, void* const* packed, const size_t* counts, const pint_stream_t* streams)
{
    if (ngpu < 1 || !devices || !packed || !planes || !counts || !layout) return PINT_E_ARG;
    int prev = 0;
    if (cudaError_t e = cudaGetDevice(&prev)) return (int)e;
    int rc = PINT_OK;
    for (int k = 0; k < ngpu && rc == PINT_OK; ++k) {
        if (cudaError_t e = cudaSetDevice(devices[k])) { rc = (int)e; break; }
        rc = pint_cuda_repack(layout, planes + (size_t)k * layout->nfields, plane_bytes, packed[k], counts[k],
                              streams ? streams[k] : nullptr);
    }
    cudaSetDevice(prev);
    return rc;
}

int pint_cuda_host_alloc(void** ptr, size_t bytes)
{
    return pint_cuda_host_alloc_flags(ptr, bytes, 0);
}

int pint_cuda_host_alloc_flags(void** ptr, size_t bytes, unsigned flags)
{
    if (!ptr) return PINT_E_NULL;
    if (flags & ~(unsigned)(PINT_HOST_PORTABLE | PINT_HOST_WRITE_COMBINED)) return PINT_E_ARG;
    unsigned f = cudaHostAllocDefault;
    if (flags & PINT_HOST_PORTABLE) f |= cudaHostAllocPortable;
    if (flags & PINT_HOST_WRITE_COMBINED) f |= cudaHostAllocWriteCombined;
    return (int)cudaHostAlloc(ptr, bytes ? bytes : 1, f);
}

int pint_cuda_host_free(void* ptr) { return ptr ? (int)cudaFreeHost(ptr) : PINT_OK; }

int pint_cuda_enable_peer_access(int ngpu, const int* devices)
{
    if (ngpu < 1 || !devices) return PINT_E_ARG;
    int prev = 0;
    if (cudaError_t e = cudaGetDevice(&prev)) return (int)e;
    int rc = PINT_OK;
    for (int i = 0; i < ngpu && rc == PINT_OK; ++i) {
        if (cudaError_t e = cudaSetDevice(devices[i])) { rc = (int)e; break; }
        for (int j = 0; j < ngpu; ++j) {
            if (devices[j] == devices[i]) continue;
            int can = 0;
            if (cudaError_t e = cudaDeviceCanAccessPeer(&can, devices[i], devices[j])) { rc = (int)e; break; }
            if (!can) continue;  // copies between the two are staged by the driver instead
            const cudaError_t e = cudaDeviceEnablePeerAccess(devices[j], 0);
            if (e == cudaErrorPeerAccessAlreadyEnabled) cudaGetLastError();
            else if (e != cudaSuccess) { rc = (int)e; break; }
        }
    }
    cudaSetDevice(prev);
    return rc;
}

// One cudaMemcpyPeerAsync per shard; `to_shards` selects the direction.
static int move_shards(bool to_shards, void* whole, int whole_device, uint32_t elem_bytes, size_t n, int ngpu,
                       const int* devices, void* const* shards, const pint_stream_t* streams)
{
    if (ngpu < 1 || ngpu > 64 || !devices || !shards) return PINT_E_ARG;
    size_t begin[64], count[64];
    if (int rc = pint_cuda_plan_shards(n, elem_bytes, ngpu, begin, count)) return rc;
    if (n == 0) return PINT_OK;
    if (!whole) return PINT_E_NULL;
    for (int k = 0; k < ngpu; ++k) {
        if (!count[k]) continue;
        if (!shards[k]) return PINT_E_NULL;
        char* w = static_cast<char*>(whole) + begin[k] * elem_bytes;
        const size_t bytes = count[k] * elem_bytes;
        cudaStream_t s = streams ? (cudaStream_t)streams[k] : nullptr;
        const cudaError_t e = to_shards ? cudaMemcpyPeerAsync(shards[k], devices[k], w, whole_device, bytes, s)
                                        : cudaMemcpyPeerAsync(w, whole_device, shards[k], devices[k], bytes, s);
        if (e != cudaSuccess) return (int)e;
    }
    return PINT_OK;
}

int pint_cuda_scatter_shards(const void* src, int src_device, uint32_t elem_bytes, size_t n, int ngpu, const int* devices,
                             void* const* shards, const pint_stream_t* streams)
{
    return move_shards(true, const_cast<void*>(src), src_device, elem_bytes, n, ngpu, devices, shards, streams);
}

int pint_cuda_gather_shards(void* dst, int dst_device, uint32_t elem_bytes, size_t n, int ngpu, const int* devices,
                            const void* const* shards, const pint_stream_t* streams)
{
    return move_shards(false, dst, dst_device, elem_bytes, n, ngpu, devices, const_cast<void* const*>(shards), streams);
}

}  // extern "C"

// ---------------------------------------------------------------------------
// Host-buffer path: per device a ring of `host_slots` slots, each with its own stream and
// device staging for one chunk of a, b and every output.  On a slot's stream the order is
// H2D(a), H2D(b), kernel, D2H(out...); different slots overlap, so both copy engines and the
// SMs are busy at once.  Staging is allocated on first use per device and kept until
// pint_cuda_host_release (the device entry points above never allocate).  Several ops on the
// same inputs share one upload.  Each device's ring has its own lock: calls on different
// devices from different threads run concurrently; the sharded form drives all rings from
// the calling thread (every call it makes is asynchronous) and locks them in device order.
namespace pint_cuda_impl {
namespace {

constexpr int kMaxDevices = 64;

struct host_ring {
    std::mutex mutex;
    int device = -1;
    size_t chunk_bytes = 0;
    int slots = 0;
    cudaStream_t stream[8] = {};
    void* buf[8][2 + PINT_MAX_HOST_OPS] = {};  // a, b, one output per op
};

host_ring g_rings[kMaxDevices];

// caller holds r.mutex and has r.device current
void ring_release(host_ring& r)
{
    for (int i = 0; i < r.slots; ++i) {
        for (int j = 0; j < 2 + PINT_MAX_HOST_OPS; ++j) if (r.buf[i][j]) { cudaFree(r.buf[i][j]); r.buf[i][j] = nullptr; }
        if (r.stream[i]) { cudaStreamDestroy(r.stream[i]); r.stream[i] = nullptr; }
    }
    r.device = -1; r.chunk_bytes = 0; r.slots = 0;
}

// caller holds r.mutex and has `dev` current
cudaError_t ring_prepare(host_ring& r, int dev, int nops)
{
    const size_t want = (size_t)opts().host_chunk_bytes / 16 * 16;
    const int slots = (int)opts().host_slots;
    if (r.device != dev || r.chunk_bytes != want || r.slots != slots) {
        ring_release(r);
        r.device = dev; r.chunk_bytes = want; r.slots = slots;
        for (int i = 0; i < slots; ++i)
            if (cudaError_t e = cudaStreamCreateWithFlags(&r.stream[i], cudaStreamNonBlocking)) return e;
    }
    for (int i = 0; i < slots; ++i)
        for (int j = 0; j < 2 + nops; ++j)
            if (!r.buf[i][j])
                if (cudaError_t e = cudaMalloc(&r.buf[i][j], want)) return e;
    return cudaSuccess;
}

// One shard's pipeline: issue() queues the next chunk on the next slot (asynchronous).
struct host_job {
    host_ring* ring = nullptr;
    int device = 0;
    const char* a = nullptr;
    const char* b = nullptr;
    char* outs[PINT_MAX_HOST_OPS] = {};
    size_t n = 0, done = 0;
    int it = 0;

    bool finished() const { return done >= n; }

    // caller has `device` current
    int issue(int nops, const int* ops, const pint_layout* layout, uint32_t elem_bytes)
    {
        host_ring& r = *ring;
        const int slot = it++ % r.slots;
        cudaStream_t s = r.stream[slot];
        const size_t chunk_elems = r.chunk_bytes / elem_bytes;
        const size_t cnt = n - done < chunk_elems ? n - done : chunk_elems;
        const size_t bytes = cnt * elem_bytes, off = done * elem_bytes;
        // stream order makes the slot's previous D2H copies finish before its buffers are overwritten
        if (cudaError_t e = cudaMemcpyAsync(r.buf[slot][0], a + off, bytes, cudaMemcpyHostToDevice, s)) return (int)e;
        if (cudaError_t e = cudaMemcpyAsync(r.buf[slot][1], b + off, bytes, cudaMemcpyHostToDevice, s)) return (int)e;
        void* douts[PINT_MAX_HOST_OPS];
        for (int j = 0; j < nops; ++j) douts[j] = r.buf[slot][2 + j];
        if (int rc = binary_multi_impl(nops, ops, layout, r.buf[slot][0], r.buf[slot][1], douts, cnt, s)) return rc;  // one pass for all ops
        for (int j = 0; j < nops; ++j)
            if (cudaError_t e = cudaMemcpyAsync(outs[j] + off, r.buf[slot][2 + j], bytes, cudaMemcpyDeviceToHost, s)) return (int)e;
        done += cnt;
        return PINT_OK;
    }

    int drain()
    {
        int rc = PINT_OK;
        for (int i = 0; i < ring->slots; ++i) {
            const cudaError_t e = cudaStreamSynchronize(ring->stream[i]);
            if (e != cudaSuccess && rc == PINT_OK) rc = (int)e;
        }
        return rc;
    }
};

int check_host_args(int nops, const int* ops, const pint_layout* layout, layout_desc* d, const void* a_host,
                    const void* b_host, void* const* outs_host, size_t n)
{
    if (int rc = check_layout(layout, d)) return rc;
    if (nops < 1 || nops > PINT_MAX_HOST_OPS || !ops || !outs_host) return PINT_E_ARG;
    for (int j = 0; j < nops; ++j)
        if (ops[j] < 0 || ops[j] >= PINT_NUM_BINARY_OPS) return PINT_E_OP;
    if (n == 0) return PINT_OK;
    if (!a_host || !b_host) return PINT_E_NULL;
    for (int j = 0; j < nops; ++j)
        if (!outs_host[j]) return PINT_E_NULL;
    return PINT_OK;
}

}  // namespace
}  // namespace pint_cuda_impl

extern "C" int pint_cuda_binary_multi_host(int nops, const int* ops, const pint_layout* layout, const void* a_host,
                                           const void* b_host, void* const* outs_host, size_t n)
{
    layout_desc d;
    if (int rc = check_host_args(nops, ops, layout, &d, a_host, b_host, outs_host, n)) return rc;
    if (n == 0) return PINT_OK;
    int dev = 0;
    if (cudaError_t e = cudaGetDevice(&dev)) return (int)e;
    if (dev < 0 || dev >= kMaxDevices) return (int)cudaErrorInvalidDevice;
    host_ring& r = g_rings[dev];
    std::lock_guard<std::mutex> lock(r.mutex);
    if (cudaError_t e = ring_prepare(r, dev, nops)) return (int)e;
    host_job job;
    job.ring = &r; job.device = dev; job.a = (const char*)a_host; job.b = (const char*)b_host; job.n = n;
    for (int j = 0; j < nops; ++j) job.outs[j] = (char*)outs_host[j];
    int rc = PINT_OK;
    while (!job.finished() && rc == PINT_OK) rc = job.issue(nops, ops, layout, d.elem_bytes);
    const int rc2 = job.drain();
    return rc != PINT_OK ? rc : rc2;
}

extern "C" int pint_cuda_binary_host(int op, const pint_layout* layout, const void* a_host, const void* b_host,
                                     void* out_host, size_t n)
{
    void* outs[1] = {out_host};
    return pint_cuda_binary_multi_host(1, &op, layout, a_host, b_host, outs, n);
}

// The partitioner for HOST arrays (north_star (4); the reference's only array workload is a host
// vector, tests/pint_bench.cpp:22,75): [0, n) is split by pint_cuda_plan_shards, shard k goes through
// devices[k]'s ring, and the chunks of all shards are queued round-robin from this thread -- every
// link and copy engine of the node works at once.  No collective, no peer traffic.
extern "C" int pint_cuda_sharded_binary_multi_host(int nops, const int* ops, const pint_layout* layout, int ngpu,
                                                   const int* devices, const void* a_host, const void* b_host,
                                                   void* const* outs_host, size_t n)
{
    layout_desc d;
    if (int rc = check_host_args(nops, ops, layout, &d, a_host, b_host, outs_host, n)) return rc;
    if (ngpu < 1 || ngpu > kMaxDevices || !devices) return PINT_E_ARG;
    for (int k = 0; k < ngpu; ++k) {
        if (devices[k] < 0 || devices[k] >= kMaxDevices) return (int)cudaErrorInvalidDevice;
        for (int i = 0; i < k; ++i)
            if (devices[i] == devices[k]) return PINT_E_ARG;  // one ring per device: a device may own one shard
    }
    if (n == 0) return PINT_OK;
    size_t begin[kMaxDevices], count[kMaxDevices];
    if (int rc = pint_cuda_plan_shards(n, d.elem_bytes, ngpu, begin, count)) return rc;
    int prev = 0;
    if (cudaError_t e = cudaGetDevice(&prev)) return (int)e;
    // lock the rings in ascending device order (two sharded calls with overlapping device sets cannot deadlock)
    int order[kMaxDevices];
    for (int k = 0; k < ngpu; ++k) order[k] = k;
    for (int i = 1; i < ngpu; ++i)
        for (int j = i; j > 0 && devices[order[j]] < devices[order[j - 1]]; --j) { const int t = order[j]; order[j] = order[j - 1]; order[j - 1] = t; }
    for (int i = 0; i < ngpu; ++i) g_rings[devices[order[i]]].mutex.lock();
    host_job jobs[kMaxDevices];
    int rc = PINT_OK;
    for (int k = 0; k < ngpu && rc == PINT_OK; ++k) {
        host_ring& r = g_rings[devices[k]];
        if (cudaError_t e = cudaSetDevice(devices[k])) { rc = (int)e; break; }
        if (cudaError_t e = ring_prepare(r, devices[k], nops)) { rc = (int)e; break; }
        host_job& j = jobs[k];
        j.ring = &r; j.device = devices[k]; j.n = count[k];
        j.a = (const char*)a_host + begin[k] * d.elem_bytes;
        j.b = (const char*)b_host + begin[k] * d.elem_bytes;
        for (int o = 0; o < nops; ++o) j.outs[o] = (char*)outs_host[o] + begin[k] * d.elem_bytes;
    }
    for (bool any = true; any && rc == PINT_OK;) {
        any = false;
        for (int k = 0; k < ngpu && rc == PINT_OK; ++k) {
            if (jobs[k].finished()) continue;
            any = true;
            if (cudaError_t e = cudaSetDevice(devices[k])) { rc = (int)e; break; }
            rc = jobs[k].issue(nops, ops, layout, d.elem_bytes);
        }
    }
    for (int k = 0; k < ngpu; ++k) {
        if (!jobs[k].ring) continue;
        cudaSetDevice(devices[k]);
        const int rc2 = jobs[k].drain();
        if (rc == PINT_OK) rc = rc2;
    }
    for (int i = ngpu - 1; i >= 0; --i) g_rings[devices[order[i]]].mutex.unlock();
    cudaSetDevice(prev);
    return rc;
}

extern "C" int pint_cuda_host_release(void)
{
    int prev = 0, ndev = 0;
    if (cudaError_t e = cudaGetDevice(&prev)) return (int)e;
    if (cudaError_t e = cudaGetDeviceCount(&ndev)) return (int)e;
    for (int dev = 0; dev < ndev && dev < kMaxDevices; ++dev) {
        host_ring& r = g_rings[dev];
        std::lock_guard<std::mutex> lock(r.mutex);
        if (r.device < 0) continue;
        cudaSetDevice(dev);
        ring_release(r);
    }
    cudaSetDevice(prev);
    return PINT_OK;
}
