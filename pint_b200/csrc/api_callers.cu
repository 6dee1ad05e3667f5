// api_callers.cu -- C ABI of the two extra caller shapes (kernels: pint/cuda_kernels_callers.cuh):
// pint_cuda_binary_scalar (one operand the same packed_int for every element) and
// pint_cuda_binary_pairs (interleaved std::pair operands, tests/pint_bench.cpp:22).  Layouts of
// registry_layouts.def run functors with compiled-in masks, every other layout the run-time-mask
// functors (masks in the constant bank); results are identical.
#include "api_common.hpp"

namespace pint_cuda_impl {
namespace {

using namespace pint::cuda;
using namespace pint::cuda::kernels;

template <class Elem, int Op, bool ScalarFirst>
cudaError_t scalar_one(const layout_desc& d, const void* a, uint64_t scalar, void* out, size_t n, cudaStream_t s,
                       const launch_geom& g)
{
    using W = word_of<Elem>;
    const auto m = make_word_masks<W>(d);
    auto run = [&](auto steps_tag) {
        using RM = runtime_masks<W, decltype(steps_tag)::value>;
        using Bin = swar_fn<Op, RM>;
        const scalar_fn<Bin, ScalarFirst> fn{Bin{RM(m)}, replicate_scalar<Elem>(scalar)};
        return launch_map1<Elem>(a, out, n, fn, s, g);
    };
    // ops that build a field-wide mask run only as many smear steps as the widest field needs (as generic.cu does):
    // at 2W bytes per element the ALU side weighs 1.5 x what it does for the array-array op
    constexpr bool fills = !(Op == OP_ADD_WRAP || Op == OP_SUB_WRAP || Op == OP_OR || Op == OP_AND || Op == OP_XOR);
    if constexpr (fills) {
        if (m.nsh > 1 && m.nsmear <= 2) return run(std::integral_constant<int, 2>{});
        if (m.nsh > 1 && m.nsmear <= 4) return run(std::integral_constant<int, 4>{});
    }
    return run(std::integral_constant<int, (sizeof(W) == 8 ? 6 : 5)>{});
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

template <class Elem, int Op>
cudaError_t pairs_one(const layout_desc& d, const void* pairs, void* out, size_t n, cudaStream_t s, const launch_geom& g)
{
    using W = word_of<Elem>;
    const auto m = make_word_masks<W>(d);
    auto run = [&](auto steps_tag) {
        using RM = runtime_masks<W, decltype(steps_tag)::value>;
        return launch_pairs_map<Elem>(pairs, out, n, swar_fn<Op, RM>{RM(m)}, s, g);
    };
    constexpr bool fills = !(Op == OP_ADD_WRAP || Op == OP_SUB_WRAP || Op == OP_OR || Op == OP_AND || Op == OP_XOR);
    if constexpr (fills) {
        if (m.nsh > 1 && m.nsmear <= 2) return run(std::integral_constant<int, 2>{});
        if (m.nsh > 1 && m.nsmear <= 4) return run(std::integral_constant<int, 4>{});
    }
    return run(std::integral_constant<int, (sizeof(W) == 8 ? 6 : 5)>{});
}

template <class Elem, int... Ops>
cudaError_t pairs_map_ops(int op, const layout_desc& d, const void* pairs, void* out, size_t n, cudaStream_t s,
                          const launch_geom& g, std::integer_sequence<int, Ops...>)
{
    cudaError_t rc = cudaErrorInvalidValue;
    ((op == Ops ? (rc = pairs_one<Elem, Ops>(d, pairs, out, n, s, g), 0) : 0), ...);
    return rc;
}

__global__ void zero_count_kernel(uint64_t* count)
{
    pdl_prologue();
    if (threadIdx.x == 0) *count = 0;
}

// packed_int::operator== / operator!= (pint.hpp:778-779) over arrays: the whole Integer is compared,
// unused high bits included.  equal[i] = (a[i] == b[i]) as one byte per element (optional) and the
// number of unequal elements (optional; integer atomics: exact and order-independent).
// One 16-byte vector of each operand per step, U steps in flight; a thread writes its PER flag bytes
// with one store, so a warp's flags are contiguous.
template <class Elem, bool Flags, bool Count>
__global__ void __launch_bounds__(256)
equal_kernel(const Elem* a, const Elem* b, size_t n, uint8_t* equal, unsigned long long* not_equal, bool vec_ok)
{
    pdl_prologue();
    constexpr int PER = 16 / sizeof(Elem), U = 4;
    using flags_t = typename std::conditional<PER == 16, uint4, typename std::conditional<PER == 8, uint2,
                    typename std::conditional<PER == 4, uint32_t, uint16_t>::type>::type>::type;
    unsigned long long bad = 0;
    const size_t nvec = vec_ok ? n / PER : 0;
    const uint4* va = reinterpret_cast<const uint4*>(a);
    const uint4* vb = reinterpret_cast<const uint4*>(b);
    auto one = [&](size_t v, const uint4& x, const uint4& y) {
        union { uint4 q; Elem e[PER]; } ux, uy;
        union { flags_t f; uint8_t c[PER]; } uf;
        ux.q = x; uy.q = y;
#pragma unroll
        for (int j = 0; j < PER; ++j) {
            const bool eq = ux.e[j] == uy.e[j];
            uf.c[j] = eq ? 1 : 0;
            if (Count) bad += eq ? 0 : 1;
        }
        if (Flags) reinterpret_cast<flags_t*>(equal)[v] = uf.f;
    };
    const size_t tile = size_t(U) * blockDim.x, nfull = nvec / tile;
    for (size_t t = blockIdx.x; t < nfull; t += gridDim.x) {
        uint4 x[U], y[U];
        const size_t base = t * tile + threadIdx.x;
#pragma unroll
        for (int k = 0; k < U; ++k) x[k] = ld16<CACHE_STREAMING>(va + base + size_t(k) * blockDim.x);
#pragma unroll
        for (int k = 0; k < U; ++k) y[k] = ld16<CACHE_STREAMING>(vb + base + size_t(k) * blockDim.x);
#pragma unroll
        for (int k = 0; k < U; ++k) one(base + size_t(k) * blockDim.x, x[k], y[k]);
    }
    for (size_t v = nfull * tile + size_t(blockIdx.x) * blockDim.x + threadIdx.x; v < nvec; v += size_t(gridDim.x) * blockDim.x)
        one(v, ld16<CACHE_STREAMING>(va + v), ld16<CACHE_STREAMING>(vb + v));
    for (size_t i = nvec * PER + size_t(blockIdx.x) * blockDim.x + threadIdx.x; i < n; i += size_t(gridDim.x) * blockDim.x) {
        const bool eq = a[i] == b[i];
        if (Flags) equal[i] = eq ? 1 : 0;
        if (Count) bad += eq ? 0 : 1;
    }
    if (Count) {
        __shared__ unsigned long long part[8];
        for (int o = 16; o; o >>= 1) bad += __shfl_down_sync(0xffffffffu, bad, o);
        if ((threadIdx.x & 31) == 0) part[threadIdx.x >> 5] = bad;
        __syncthreads();
        if (threadIdx.x == 0) {
            for (int w = 1; w < int(blockDim.x >> 5); ++w) bad += part[w];
            if (bad) atomicAdd(not_equal, bad);
        }
    }
}

template <class Elem>
cudaError_t equal_launch(const void* a, const void* b, size_t n, uint8_t* equal, uint64_t* count, cudaStream_t s, launch_geom g)
{
    constexpr size_t per = 16 / sizeof(Elem);
    // vectors need a and b 16-byte aligned and the flag bytes aligned to their per-vector store
    const bool vec_ok = (((uintptr_t)a | (uintptr_t)b) & 15u) == 0 && (!equal || ((uintptr_t)equal % per) == 0);
    if (count) g.ctas_per_sm = 8;  // one atomic per CTA at the end: persistent grid
    const unsigned grid = grid_for((n / per + 1023) / 1024 + 1, g);
    const Elem* ea = static_cast<const Elem*>(a);
    const Elem* eb = static_cast<const Elem*>(b);
    unsigned long long* c = reinterpret_cast<unsigned long long*>(count);
    const bool pdl = g.pdl != 0;
    cudaError_t rc;
    if (equal && count) rc = launch_stream_kernel(equal_kernel<Elem, true, true>, grid, 256, s, pdl, ea, eb, n, equal, c, vec_ok);
    else if (equal) rc = launch_stream_kernel(equal_kernel<Elem, true, false>, grid, 256, s, pdl, ea, eb, n, equal, c, vec_ok);
    else rc = launch_stream_kernel(equal_kernel<Elem, false, true>, grid, 256, s, pdl, ea, eb, n, equal, c, vec_ok);
    if (rc != cudaSuccess) return rc;
    count_launch(g);
    return cudaGetLastError();
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
    const bool first = scalar_first != 0;
    if (const layout_kernels* k = find_static(d)) return (int)k->scalar[first][op](a, scalar, out, n, s, g);
    auto ops = std::make_integer_sequence<int, pint::cuda::OP_NUM_BINARY>();
    auto runtime = [&](size_t start, size_t count) -> int {
        if (!count) return PINT_OK;
        const char* pa = static_cast<const char*>(a) + start * d.elem_bytes;
        char* po = static_cast<char*>(out) + start * d.elem_bytes;
        switch (d.elem_bytes) {
        case 1: return (int)scalar_ops<uint8_t>(op, first, d, pa, scalar, po, count, s, g, ops);
        case 2: return (int)scalar_ops<uint16_t>(op, first, d, pa, scalar, po, count, s, g, ops);
        case 4: return (int)scalar_ops<uint32_t>(op, first, d, pa, scalar, po, count, s, g, ops);
        default: return (int)scalar_ops<uint64_t>(op, first, d, pa, scalar, po, count, s, g, ops);
        }
    };
    // option "jit": the vector body with this layout's constants compiled in at run time (jit.cu), the rest as before
    if (const long jit = opts().jit) {
        size_t head = 0, body = 0;
        bool handled = false;
        const int rc = jit_scalar_body(op, first, d, a, scalar, out, n, s, g, &head, &body, &handled);
        if (handled) {
            if (int r = runtime(0, head)) return r;
            return runtime(head + body, n - head - body);
        }
        if (rc == PINT_E_JIT && jit >= 2) return rc;
        if (rc != PINT_OK && rc != PINT_E_JIT) return rc;
    }
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
    if (const layout_kernels* k = find_static(d)) return (int)k->pairs[op](pairs, out, n, s, g);
    auto ops = std::make_integer_sequence<int, pint::cuda::OP_NUM_BINARY>();
    auto runtime = [&](size_t start, size_t count) -> int {
        if (!count) return PINT_OK;
        const char* pp = static_cast<const char*>(pairs) + 2 * start * d.elem_bytes;
        char* po = static_cast<char*>(out) + start * d.elem_bytes;
        switch (d.elem_bytes) {
        case 1: return (int)pairs_map_ops<uint8_t>(op, d, pp, po, count, s, g, ops);
        case 2: return (int)pairs_map_ops<uint16_t>(op, d, pp, po, count, s, g, ops);
        case 4: return (int)pairs_map_ops<uint32_t>(op, d, pp, po, count, s, g, ops);
        default: return (int)pairs_map_ops<uint64_t>(op, d, pp, po, count, s, g, ops);
        }
    };
    if (const long jit = opts().jit) {
        size_t head = 0, body = 0;
        bool handled = false;
        const int rc = jit_pairs_body(op, d, pairs, out, n, s, g, &head, &body, &handled);
        if (handled) {
            if (int r = runtime(0, head)) return r;
            return runtime(head + body, n - head - body);
        }
        if (rc == PINT_E_JIT && jit >= 2) return rc;
        if (rc != PINT_OK && rc != PINT_E_JIT) return rc;
    }
    switch (d.elem_bytes) {
    case 1: return (int)pairs_map_ops<uint8_t>(op, d, pairs, out, n, s, g, ops);
    case 2: return (int)pairs_map_ops<uint16_t>(op, d, pairs, out, n, s, g, ops);
    case 4: return (int)pairs_map_ops<uint32_t>(op, d, pairs, out, n, s, g, ops);
    default: return (int)pairs_map_ops<uint64_t>(op, d, pairs, out, n, s, g, ops);
    }
}

int pint_cuda_equal(const pint_layout* layout, const void* a, const void* b, size_t n, uint8_t* equal_out,
                    uint64_t* not_equal_count_dev, pint_stream_t stream)
{
    layout_desc d;
    if (int rc = check_layout(layout, &d)) return rc;
    cudaStream_t s = (cudaStream_t)stream;
    if (n == 0) {  // (an empty flag array may well be a null pointer)
        if (not_equal_count_dev) return (int)cudaMemsetAsync(not_equal_count_dev, 0, sizeof(uint64_t), s);
        return PINT_OK;
    }
    if (!equal_out && !not_equal_count_dev) return PINT_E_NULL;
    if (!a || !b) return PINT_E_NULL;
    if (((uintptr_t)a | (uintptr_t)b) % d.elem_bytes) return PINT_E_ARG;
    if (equal_out) {
        const uintptr_t o = (uintptr_t)equal_out, bytes = n * d.elem_bytes;
        if ((o < (uintptr_t)a + bytes && (uintptr_t)a < o + n) || (o < (uintptr_t)b + bytes && (uintptr_t)b < o + n)) return PINT_E_ALIAS;
    }
    const launch_geom g = geom();
    // the count starts at 0: a one-thread kernel chained to the comparison by programmatic dependent launch, like
    // pint_cuda_pairs_reduce (a cudaMemsetAsync and its launch gap cost 10-20 us per pass on frame-sized arrays)
    if (not_equal_count_dev) {
        if (cudaError_t e = launch_stream_kernel(zero_count_kernel, 1, 32, s, g.pdl != 0, not_equal_count_dev)) return (int)e;
        count_launch(g);
    }
    switch (d.elem_bytes) {
    case 1: return (int)equal_launch<uint8_t>(a, b, n, equal_out, not_equal_count_dev, s, g);
    case 2: return (int)equal_launch<uint16_t>(a, b, n, equal_out, not_equal_count_dev, s, g);
    case 4: return (int)equal_launch<uint32_t>(a, b, n, equal_out, not_equal_count_dev, s, g);
    default: return (int)equal_launch<uint64_t>(a, b, n, equal_out, not_equal_count_dev, s, g);
    }
}

}  // extern "C"
