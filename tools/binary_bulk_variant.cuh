// tools/binary_bulk_variant.cuh -- TMA-staged variant of the binary streaming kernel (sweep tool only).
//
// Same map as binary_stream_kernel, different data movement: tiles of a and b are brought into
// shared memory by the TMA engine, all threads compute from shared memory, write the result over
// the a-tile, and one thread hands the tile back to the TMA engine.  A ring of STAGES slots keeps
// (STAGES - 1) tile pairs in flight per CTA.  Measured and rejected for the library (6.1-6.8 TB/s
// against 7.0 for the vectorised kernel, profiles/r01_tune_sweep1.log); kept here so that
// tools/tune.cu can re-measure it.  Not part of libpint_cuda.so, hence no parity test of its own.
#pragma once

#include "pint/cuda_kernels_bulk.cuh"

namespace pint {
namespace cuda {
namespace kernels {
namespace bulk {

template <int Stages, int TileBytes, int Block>
struct bulk_cfg {
    static constexpr int STAGES = Stages;
    static constexpr int TILE_BYTES = TileBytes;  // per operand per stage
    static constexpr int BLOCK = Block;
    static constexpr int SMEM_BYTES = Stages * 2 * TileBytes;
};

// ntiles full tiles of TILE_BYTES each; the launcher sends the remainder to
// binary_stream_kernel.  `out` may alias a or b: a tile is stored only after
// both of its loads have completed.
template <class Fn, class Cfg>
__global__ void __launch_bounds__(Cfg::BLOCK)
binary_bulk_kernel(const uint4* a, const uint4* b, uint4* out, size_t ntiles, const __grid_constant__ Fn fn)
{
    using W = typename Fn::word;
    constexpr int S = Cfg::STAGES, TB = Cfg::TILE_BYTES, B = Cfg::BLOCK;
    constexpr int TV = TB / 16;  // vectors per tile
    extern __shared__ __align__(128) unsigned char smem[];
    __shared__ __align__(8) uint64_t full[S];
    auto tile_a = [&](int s) { return reinterpret_cast<uint4*>(smem + size_t(s) * 2 * TB); };
    auto tile_b = [&](int s) { return reinterpret_cast<uint4*>(smem + size_t(s) * 2 * TB + TB); };
    auto f = [&](W x, W y) { return fn(x, y); };

    // tiles of this CTA: blockIdx.x, blockIdx.x + gridDim.x, ...
    const size_t first = blockIdx.x, stride = gridDim.x;
    const size_t mine = first < ntiles ? (ntiles - first + stride - 1) / stride : 0;
    uint64_t pol = 0;
    if (threadIdx.x == 0) {
        pol = evict_first_policy();
        for (int s = 0; s < S; ++s) mbar_init(&full[s], 1);
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
        asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
    }
    __syncthreads();
    auto issue = [&](size_t k) {  // thread 0: load this CTA's k-th tile into slot k % S
        const int s = int(k % S);
        const size_t t = first + k * stride;
        mbar_expect_tx(&full[s], 2 * TB);
        bulk_load(tile_a(s), a + t * TV, TB, &full[s], pol);
        bulk_load(tile_b(s), b + t * TV, TB, &full[s], pol);
    };
    if (threadIdx.x == 0)
        for (size_t k = 0; k < mine && k < size_t(S); ++k) issue(k);

    for (size_t k = 0; k < mine; ++k) {
        const int s = int(k % S);
        mbar_wait(&full[s], uint32_t((k / S) & 1));
        uint4* ta = tile_a(s);
        const uint4* tb = tile_b(s);
#pragma unroll
        for (int v = threadIdx.x; v < TV; v += B) ta[v] = map_words2<W>(ta[v], tb[v], f);
        asm volatile("fence.proxy.async.shared::cta;" ::: "memory");  // generic-proxy writes -> async proxy
        __syncthreads();
        if (threadIdx.x == 0) {
            bulk_store(out + (first + k * stride) * TV, ta, TB, pol);
            // slot of the PREVIOUS tile: its store must have finished reading shared
            // memory (at most this tile's store may still be pending), then refill it
            if (k >= 1 && k - 1 + S < mine) {
                bulk_wait_read<1>();
                issue(k - 1 + S);
            }
        }
    }
    if (threadIdx.x == 0) bulk_wait_read<0>();
}

}  // namespace bulk
}  // namespace kernels
}  // namespace cuda
}  // namespace pint
