// pint/cuda_kernels_multi.cuh -- several binary ops on the SAME operands in one pass.
//
// out_j[i] = op_j(a[i], b[i]) for j < nops (nops <= 4): a and b are read once and
// every result is written, so a step such as BASELINE.json's C1
// (add_unsigned_saturate + add_signed_saturate over the same arrays) moves
// (2 + nops) * W bytes per element instead of nops * 3W -- 16 instead of 24 bytes
// for two ops on uint32 packs.  Same per-element functions as apply_op (the op
// code is a warp-uniform switch per 16-byte vector), same streaming structure as
// binary_stream_kernel.  Included by cuda_kernels.cuh.
#pragma once

namespace pint {
namespace cuda {
namespace kernels {

constexpr int kMaxMultiOps = 4;

struct multi_ops {
    int nops;
    int op[kMaxMultiOps];
    void* out[kMaxMultiOps];  // device pointers (vector-aligned part is passed separately)
};

// one 16-byte vector through op code `op` (uniform across the grid)
template <class P>
__device__ __forceinline__ uint4 apply_op_vec(int op, const uint4& a, const uint4& b, const P& p)
{
    using W = typename P::word;
    switch (op) {
#define PINT_MULTI_CASE(OP) \
    case OP: return map_words2<W>(a, b, [&](W x, W y) { return apply_op<OP>(x, y, p); });
        PINT_MULTI_CASE(OP_ADD_WRAP)
        PINT_MULTI_CASE(OP_ADD_UNSIGNED_SATURATE)
        PINT_MULTI_CASE(OP_ADD_SIGNED_SATURATE)
        PINT_MULTI_CASE(OP_SUB_WRAP)
        PINT_MULTI_CASE(OP_SUB_UNSIGNED_SATURATE)
        PINT_MULTI_CASE(OP_SUB_SIGNED_SATURATE)
        PINT_MULTI_CASE(OP_OR)
        PINT_MULTI_CASE(OP_AND)
        PINT_MULTI_CASE(OP_XOR)
        PINT_MULTI_CASE(OP_MIN_UNSIGNED)
        PINT_MULTI_CASE(OP_MAX_UNSIGNED)
        PINT_MULTI_CASE(OP_MIN_SIGNED)
    default:
        PINT_MULTI_CASE(OP_MAX_SIGNED)
#undef PINT_MULTI_CASE
    }
}

template <class P>
__device__ __forceinline__ typename P::word apply_op_word(int op, typename P::word a, typename P::word b, const P& p)
{
    switch (op) {
    case OP_ADD_WRAP: return apply_op<OP_ADD_WRAP>(a, b, p);
    case OP_ADD_UNSIGNED_SATURATE: return apply_op<OP_ADD_UNSIGNED_SATURATE>(a, b, p);
    case OP_ADD_SIGNED_SATURATE: return apply_op<OP_ADD_SIGNED_SATURATE>(a, b, p);
    case OP_SUB_WRAP: return apply_op<OP_SUB_WRAP>(a, b, p);
    case OP_SUB_UNSIGNED_SATURATE: return apply_op<OP_SUB_UNSIGNED_SATURATE>(a, b, p);
    case OP_SUB_SIGNED_SATURATE: return apply_op<OP_SUB_SIGNED_SATURATE>(a, b, p);
    case OP_OR: return a | b;
    case OP_AND: return a & b;
    case OP_XOR: return a ^ b;
    case OP_MIN_UNSIGNED: return apply_op<OP_MIN_UNSIGNED>(a, b, p);
    case OP_MAX_UNSIGNED: return apply_op<OP_MAX_UNSIGNED>(a, b, p);
    case OP_MIN_SIGNED: return apply_op<OP_MIN_SIGNED>(a, b, p);
    default: return apply_op<OP_MAX_SIGNED>(a, b, p);
    }
}

// One CTA per tile (or grid stride when the grid is capped), UNROLL vectors per
// operand per thread; the outputs of one vector are produced one op at a time so
// that only the inputs stay live in registers.
template <class P, class Cfg>
__global__ void __launch_bounds__(Cfg::BLOCK)
binary_multi_stream_kernel(const uint4* a, const uint4* b, size_t nvec, const __grid_constant__ multi_ops m,
                           const __grid_constant__ P p)
{
    constexpr int U = Cfg::UNROLL, B = Cfg::BLOCK;
    constexpr size_t tile = size_t(U) * B;
    const size_t nfull = nvec / tile;
    for (size_t t = blockIdx.x; t < nfull; t += gridDim.x) {
        const size_t base = t * tile + threadIdx.x;
        uint4 va[U], vb[U];
#pragma unroll
        for (int u = 0; u < U; ++u) va[u] = ld16<Cfg::LD>(a + base + size_t(u) * B);
#pragma unroll
        for (int u = 0; u < U; ++u) vb[u] = ld16<Cfg::LD>(b + base + size_t(u) * B);
        for (int j = 0; j < m.nops; ++j) {
            uint4* out = static_cast<uint4*>(m.out[j]);
            const int op = m.op[j];
#pragma unroll
            for (int u = 0; u < U; ++u) st16<Cfg::ST>(out + base + size_t(u) * B, apply_op_vec(op, va[u], vb[u], p));
        }
    }
    if (blockIdx.x == nfull % gridDim.x) {
        for (size_t i = nfull * tile + threadIdx.x; i < nvec; i += B) {
            const uint4 x = ld16<Cfg::LD>(a + i), y = ld16<Cfg::LD>(b + i);
            for (int j = 0; j < m.nops; ++j)
                st16<Cfg::ST>(static_cast<uint4*>(m.out[j]) + i, apply_op_vec(m.op[j], x, y, p));
        }
    }
}

template <class Elem, class P>
__global__ void __launch_bounds__(256)
binary_multi_elem_kernel(const Elem* a, const Elem* b, size_t n, const __grid_constant__ multi_ops m,
                         const __grid_constant__ P p)
{
    using W = typename P::word;
    for (size_t i = size_t(blockIdx.x) * blockDim.x + threadIdx.x; i < n; i += size_t(gridDim.x) * blockDim.x) {
        const W x = W(a[i]), y = W(b[i]);
        for (int j = 0; j < m.nops; ++j) static_cast<Elem*>(m.out[j])[i] = Elem(apply_op_word(m.op[j], x, y, p));
    }
}

// outs[j] may alias a or b only for the LAST op that reads them... to keep the
// contract simple the C ABI forbids any overlap between outputs and inputs here.
template <class Elem, class P, class Cfg = default_stream_cfg>
cudaError_t launch_binary_multi(const void* a, const void* b, const multi_ops& m, size_t n, const P& p,
                                cudaStream_t stream, const launch_geom& g)
{
    if (n == 0 || m.nops == 0) return cudaSuccess;
    const Elem* ea = static_cast<const Elem*>(a);
    const Elem* eb = static_cast<const Elem*>(b);
    const void* ptrs[2 + kMaxMultiOps] = {a, b};
    for (int j = 0; j < m.nops; ++j) ptrs[2 + j] = m.out[j];
    size_t head = 0, nvec = 0, tail = 0;
    const bool vec_ok = split_aligned(sizeof(Elem), n, ptrs, 2 + m.nops, &head, &nvec, &tail);
    auto shifted = [&](size_t first) {
        multi_ops s = m;
        for (int j = 0; j < m.nops; ++j) s.out[j] = static_cast<Elem*>(m.out[j]) + first;
        return s;
    };
    auto elems = [&](size_t first, size_t count) {
        if (!count) return;
        const unsigned grid = grid_for((count + 255) / 256, g);
        binary_multi_elem_kernel<Elem, P><<<grid, 256, 0, stream>>>(ea + first, eb + first, count, shifted(first), p);
        count_launch(g);
    };
    if (!vec_ok) {
        elems(0, n);
        return cudaGetLastError();
    }
    elems(0, head);
    if (nvec) {
        constexpr size_t tile = size_t(Cfg::UNROLL) * Cfg::BLOCK;
        const size_t ntiles = (nvec + tile - 1) / tile;
        binary_multi_stream_kernel<P, Cfg><<<grid_for(ntiles, g), Cfg::BLOCK, 0, stream>>>(
            reinterpret_cast<const uint4*>(ea + head), reinterpret_cast<const uint4*>(eb + head), nvec, shifted(head), p);
        count_launch(g);
    }
    elems(n - tail, tail);
    return cudaGetLastError();
}

}  // namespace kernels
}  // namespace cuda
}  // namespace pint
