// api_static.cu -- pint_cuda_pairs_reduce_static: the benchmark-shaped pair reduction (api_aux.cu,
// tests/pint_bench.cpp:90-94) for callers that may not allocate: code under CUDA graph capture, tight
// serving loops, harnesses that forbid cudaMalloc; it has
// no counterpart in the reference.
//
// Two passes instead of one atomic per CTA: every CTA reduces a fixed chunk of pairs and writes its partial
// sum into SCRATCH, a second one-CTA kernel adds the partials into the caller's result.  The scratch is a
// file-scope static device array of a fixed, modest size (kStaticPartials * 8 bytes = 64 KiB): every GPU
// that loads the library holds it until the process exits, and nothing is ever cudaMalloc'ed or freed.
//
//   - In HOST code `g_static_partials` is only a host-side stand-in for the symbol, not a device address.
//     The address a kernel can use comes from cudaGetSymbolAddress, once per device with that device
//     current (static_partials() below); the kernels take it as an ordinary pointer argument.
//   - All calls on a device share the one array, and it keeps the last call's contents.  Nothing in it has
//     to start at zero: pass 1 writes partials [0, nctas) before pass 2 reads exactly those.  Calls that
//     OVERLAP on the GPU (two streams, two threads) race on it: serialise them.
//   - Arrays longer than PINT_STATIC_MAX_PAIRS are rejected (PINT_E_TOO_LARGE): use pint_cuda_pairs_reduce.
#include "api_common.hpp"

namespace pint_cuda_impl {
namespace {

using namespace pint::cuda;
using namespace pint::cuda::kernels;

constexpr int kStaticPartials = 8192;        // uint64_t slots: 64 KiB of static device memory per GPU
constexpr int kPairsPerCta = 16384;          // one CTA reduces this many pairs (128 KiB of uint32_t pairs)
static_assert(size_t(kStaticPartials) * kPairsPerCta == PINT_STATIC_MAX_PAIRS, "header and kernel geometry must agree");

static __device__ __align__(256) uint64_t g_static_partials[kStaticPartials];

// device address of the scratch on the CURRENT device (cached per device; the symbol lookup allocates nothing)
cudaError_t static_partials(uint64_t** out)
{
    static uint64_t* addr[64] = {};
    int dev = 0;
    if (cudaError_t e = cudaGetDevice(&dev)) return e;
    if (dev < 0 || dev >= 64) return cudaErrorInvalidDevice;
    uint64_t* p = __atomic_load_n(&addr[dev], __ATOMIC_ACQUIRE);
    if (!p) {
        void* sym = nullptr;
        if (cudaError_t e = cudaGetSymbolAddress(&sym, g_static_partials)) return e;
        p = static_cast<uint64_t*>(sym);
        __atomic_store_n(&addr[dev], p, __ATOMIC_RELEASE);
    }
    *out = p;
    return cudaSuccess;
}

__device__ __forceinline__ uint64_t block_sum(uint64_t s)
{
    __shared__ uint64_t ss[8];
    for (int o = 16; o; o >>= 1) s += __shfl_down_sync(0xffffffffu, s, o);
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    if (lane == 0) ss[warp] = s;
    __syncthreads();
    if (threadIdx.x == 0)
        for (int w = 1; w < int(blockDim.x >> 5); ++w) s += ss[w];
    return s;  // valid in thread 0
}

// pass 1: partials[blockIdx.x] = sum over this CTA's chunk of OP(pairs[2i], pairs[2i+1]) (pint_bench.cpp:90-94)
template <class Elem, int Op>
__global__ void __launch_bounds__(256)
pairs_partial_kernel(const Elem* pairs, size_t n, uint64_t* partials, const __grid_constant__ runtime_masks<word_of<Elem>> p)
{
    using W = word_of<Elem>;
    const size_t lo = size_t(blockIdx.x) * kPairsPerCta;
    const size_t hi = lo + kPairsPerCta < n ? lo + kPairsPerCta : n;
    uint64_t s = 0;
    constexpr int PER = 8 / int(sizeof(Elem));  // pairs per 16-byte vector
    const bool vec_ok = (reinterpret_cast<uintptr_t>(pairs) & 15u) == 0;  // lo is a multiple of PER
    size_t i = lo;
    if (vec_ok) {
        const uint4* vec = reinterpret_cast<const uint4*>(pairs);
        const size_t v0 = lo / PER, v1 = hi / PER;
        for (size_t v = v0 + threadIdx.x; v < v1; v += blockDim.x) {
            union { uint4 q; Elem e[2 * PER]; } u;
            u.q = ld16<CACHE_STREAMING>(vec + v);
#pragma unroll
            for (int j = 0; j < PER; ++j) s += Elem(apply_op<Op>(W(u.e[2 * j]), W(u.e[2 * j + 1]), p));
        }
        i = v1 * PER;
    }
    for (size_t k = i + threadIdx.x; k < hi; k += blockDim.x) s += Elem(apply_op<Op>(W(pairs[2 * k]), W(pairs[2 * k + 1]), p));
    s = block_sum(s);
    if (threadIdx.x == 0) partials[blockIdx.x] = s;
}

// pass 2: *sum = partials[0] + ... + partials[count - 1]
__global__ void __launch_bounds__(256) partials_sum_kernel(const uint64_t* partials, int count, uint64_t* sum)
{
    uint64_t s = 0;
    for (int i = threadIdx.x; i < count; i += blockDim.x) s += partials[i];
    s = block_sum(s);
    if (threadIdx.x == 0) *sum = s;
}

template <class Elem, int... Ops>
cudaError_t launch_partials(int op, const layout_desc& d, const void* pairs, size_t n, uint64_t* partials, unsigned nctas,
                            cudaStream_t s, std::integer_sequence<int, Ops...>)
{
    using W = word_of<Elem>;
    const runtime_masks<W> p(make_word_masks<W>(d));
    ((op == Ops ? (pairs_partial_kernel<Elem, Ops><<<nctas, 256, 0, s>>>((const Elem*)pairs, n, partials, p), 0) : 0), ...);
    return cudaGetLastError();
}

}  // namespace
}  // namespace pint_cuda_impl

using namespace pint_cuda_impl;

extern "C" int pint_cuda_pairs_reduce_static(int op, const pint_layout* layout, const void* pairs, size_t n, uint64_t* sum_dev,
                                             pint_stream_t stream)
{
    layout_desc d;
    if (int rc = check_layout(layout, &d)) return rc;
    if (op < 0 || op >= PINT_NUM_BINARY_OPS) return PINT_E_OP;
    if (!sum_dev) return PINT_E_NULL;
    if (n > (size_t)PINT_STATIC_MAX_PAIRS) return PINT_E_TOO_LARGE;
    cudaStream_t s = (cudaStream_t)stream;
    if (n == 0) return (int)cudaMemsetAsync(sum_dev, 0, sizeof(uint64_t), s);
    if (!pairs) return PINT_E_NULL;
    if ((uintptr_t)pairs % d.elem_bytes) return PINT_E_ARG;
    uint64_t* partials = nullptr;
    if (cudaError_t e = static_partials(&partials)) return (int)e;
    const unsigned nctas = unsigned((n + kPairsPerCta - 1) / kPairsPerCta);  // <= kStaticPartials by the check above
    const launch_geom g = geom();
    auto ops = std::make_integer_sequence<int, pint::cuda::OP_NUM_BINARY>();
    cudaError_t rc;
    switch (d.elem_bytes) {
    case 1: rc = launch_partials<uint8_t>(op, d, pairs, n, partials, nctas, s, ops); break;
    case 2: rc = launch_partials<uint16_t>(op, d, pairs, n, partials, nctas, s, ops); break;
    case 4: rc = launch_partials<uint32_t>(op, d, pairs, n, partials, nctas, s, ops); break;
    default: rc = launch_partials<uint64_t>(op, d, pairs, n, partials, nctas, s, ops); break;
    }
    if (rc != cudaSuccess) return (int)rc;
    count_launch(g);
    partials_sum_kernel<<<1, 256, 0, s>>>(partials, (int)nctas, sum_dev);
    count_launch(g);
    return (int)cudaGetLastError();
}
