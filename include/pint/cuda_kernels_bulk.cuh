// pint/cuda_kernels_bulk.cuh -- the TMA building blocks the kernels share: 1-D bulk copies
// global -> shared (cp.async.bulk.shared::cluster.global + mbarrier complete_tx; SASS UBLKCP.S.G),
// shared -> global (cp.async.bulk.global.shared::cta; UBLKCP.G.S), mbarrier helpers and the L2
// evict-first policy.  Used by the dense bit-stream kernels (tiles staged in and out) and by the
// bulk-store unpack kernel (cuda_kernels_planes_bulk.cuh).  The TMA-staged variant of the plain
// binary streaming kernel, measured slower than the vectorised kernel and never launched by the
// library (6.1-6.8 against 7.0 TB/s, profiles/r01_tune_sweep1.log), lives with the sweep tool:
// tools/binary_bulk_variant.cuh.
#pragma once
#ifndef PINT_CUDA_KERNELS_BULK_CUH_INCLUDED
#define PINT_CUDA_KERNELS_BULK_CUH_INCLUDED

#include "cuda_kernels.cuh"

namespace pint {
namespace cuda {
namespace kernels {
namespace bulk {

__device__ __forceinline__ uint32_t smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }

__device__ __forceinline__ void mbar_init(uint64_t* bar, uint32_t count)
{
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(bar)), "r"(count));
}

__device__ __forceinline__ void mbar_expect_tx(uint64_t* bar, uint32_t bytes)
{
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(bar)), "r"(bytes) : "memory");
}

__device__ __forceinline__ void mbar_wait(uint64_t* bar, uint32_t parity)
{
    uint32_t done;
    do {
        asm volatile(
            "{\n"
            ".reg .pred p;\n"
            "mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n"
            "selp.u32 %0, 1, 0, p;\n"
            "}\n" : "=r"(done) : "r"(smem_u32(bar)), "r"(parity) : "memory");
    } while (!done);
}

__device__ __forceinline__ uint64_t evict_first_policy()
{
    uint64_t pol;
    asm volatile("createpolicy.fractional.L2::evict_first.b64 %0, 1.0;" : "=l"(pol));
    return pol;
}

__device__ __forceinline__ void bulk_load(void* smem_dst, const void* gmem_src, uint32_t bytes, uint64_t* bar, uint64_t pol)
{
    asm volatile(
        "cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes.L2::cache_hint [%0], [%1], %2, [%3], %4;"
        ::"r"(smem_u32(smem_dst)), "l"(gmem_src), "r"(bytes), "r"(smem_u32(bar)), "l"(pol) : "memory");
}

__device__ __forceinline__ void bulk_store(void* gmem_dst, const void* smem_src, uint32_t bytes, uint64_t pol)
{
    asm volatile("cp.async.bulk.global.shared::cta.bulk_group.L2::cache_hint [%0], [%1], %2, %3;"
                 ::"l"(gmem_dst), "r"(smem_u32(smem_src)), "r"(bytes), "l"(pol) : "memory");
    asm volatile("cp.async.bulk.commit_group;" ::: "memory");
}

template <int N>
__device__ __forceinline__ void bulk_wait_read()
{
    asm volatile("cp.async.bulk.wait_group.read %0;" ::"n"(N) : "memory");
}

}  // namespace bulk
}  // namespace kernels
}  // namespace cuda
}  // namespace pint

#endif  // PINT_CUDA_KERNELS_BULK_CUH_INCLUDED
