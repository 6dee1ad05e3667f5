// ab_reduce.cu -- variants of the benchmark-shaped pair reduction (pint_cuda_pairs_reduce, tests/pint_bench.cpp:90-94)
// timed next to each other on one array: which of (grid size, loads in flight, staging) holds the kernel back at
// the reference benchmark's own size (10^8 pairs = 0.8 GB, 6.3x the L2) and at 2^29 pairs.
//   build: nvcc -std=c++17 -O3 -gencode arch=compute_100a,code=sm_100a -lineinfo -Iinclude -o build/ab_reduce tools/ab_reduce.cu
//   run:   build/ab_reduce [pairs]
// Variants:  lib_old_grid  the round's kernel with its former grid (one CTA per 1024 pairs)
//            lib           the same kernel, one CTA per tile
//            fence         + a data dependency that keeps the arithmetic behind the last load of the tile
//            u8            8 loads of 32 bytes per thread and tile
//            tma<S>        persistent CTAs, tiles of 16 KiB staged through an S-deep ring of bulk (TMA) loads
// Every variant's sum is checked against `lib`.
#include <cstdint>
#include <cstdio>
#include <cstdlib>
#include <vector>

#include <cuda_runtime.h>

#include "pint/cuda_kernels.cuh"
#include "pint/cuda_kernels_bulk.cuh"

using namespace pint::cuda;
using namespace pint::cuda::kernels;

#define CK(x) do { cudaError_t e_ = (x); if (e_ != cudaSuccess) { std::printf("CUDA error %s at line %d\n", cudaGetErrorString(e_), __LINE__); std::exit(2); } } while (0)

__device__ __forceinline__ void block_sum_atomic(uint64_t s, uint64_t* result)
{
    __shared__ uint64_t ss[8];
    for (int o = 16; o; o >>= 1) s += __shfl_down_sync(0xffffffffu, s, o);
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    if (lane == 0) ss[warp] = s;
    __syncthreads();
    if (threadIdx.x == 0) {
        for (int w = 1; w < int(blockDim.x >> 5); ++w) s += ss[w];
        atomicAdd(reinterpret_cast<unsigned long long*>(result), (unsigned long long)s);
    }
}

template <int Op, class P>
__device__ __forceinline__ void add_vec(uint64_t& s, const uint4& q, const P& p)
{
    s += uint32_t(apply_op<Op>(q.x, q.y, p));
    s += uint32_t(apply_op<Op>(q.z, q.w, p));
}

// uint32_t pairs, n a multiple of the tile (the tool picks n so)
template <int Op, class P, int U, bool Fence>
__global__ void __launch_bounds__(256) reduce_ld(const uint32_t* pairs, size_t n, uint64_t* sum, const __grid_constant__ P p)
{
    uint64_t s = 0;
    const vec32* v32 = reinterpret_cast<const vec32*>(pairs);
    const size_t tile = size_t(U) * blockDim.x, ntiles = n / (4 * tile);  // 4 pairs per 32-byte vector
    for (size_t t = blockIdx.x; t < ntiles; t += gridDim.x) {
        vec32 q[U];
#pragma unroll
        for (int k = 0; k < U; ++k) q[k] = ld32(v32 + t * tile + size_t(k) * blockDim.x + threadIdx.x);
        if (Fence) {
            const uint32_t zero = uint32_t(n >> 63);  // 0, but only at run time
            uint32_t f = 0;
#pragma unroll
            for (int k = 0; k < U; ++k) f |= q[k].lo.x;
            f &= zero;
#pragma unroll
            for (int k = 0; k < U; ++k) { q[k].lo.x |= f; q[k].lo.z |= f; q[k].hi.x |= f; q[k].hi.z |= f; }
        }
#pragma unroll
        for (int k = 0; k < U; ++k) { add_vec<Op>(s, q[k].lo, p); add_vec<Op>(s, q[k].hi, p); }
    }
    block_sum_atomic(s, sum);
}

// persistent CTAs; tile = 256 threads x 64 bytes = 16 KiB, S stages
template <int Op, class P, int S>
__global__ void __launch_bounds__(256) reduce_tma(const uint32_t* pairs, size_t n, uint64_t* sum, const __grid_constant__ P p)
{
    constexpr uint32_t kTileBytes = 256 * 64;
    extern __shared__ __align__(128) unsigned char smem[];
    __shared__ uint64_t full[S];
    const size_t ntiles = n * 8 / kTileBytes;
    const size_t mine = blockIdx.x < ntiles ? (ntiles - blockIdx.x + gridDim.x - 1) / gridDim.x : 0;
    const unsigned char* src = reinterpret_cast<const unsigned char*>(pairs);
    uint64_t pol = 0;
    if (threadIdx.x == 0) {
        for (int i = 0; i < S; ++i) bulk::mbar_init(&full[i], 1);
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
        pol = bulk::evict_first_policy();
    }
    __syncthreads();
    if (threadIdx.x == 0)
        for (int i = 0; i < S && size_t(i) < mine; ++i) {
            bulk::mbar_expect_tx(&full[i], kTileBytes);
            bulk::bulk_load(smem + size_t(i) * kTileBytes, src + (blockIdx.x + size_t(i) * gridDim.x) * kTileBytes, kTileBytes, &full[i], pol);
        }
    uint64_t s = 0;
    for (size_t i = 0; i < mine; ++i) {
        const int st = int(i % S);
        bulk::mbar_wait(&full[st], uint32_t(i / S) & 1u);
        const uint4* v = reinterpret_cast<const uint4*>(smem + size_t(st) * kTileBytes);
        uint4 q[4];
#pragma unroll
        for (int k = 0; k < 4; ++k) q[k] = v[k * 256 + threadIdx.x];
#pragma unroll
        for (int k = 0; k < 4; ++k) add_vec<Op>(s, q[k], p);
        __syncthreads();  // every thread has its 64 bytes of this stage in registers
        if (threadIdx.x == 0 && i + S < mine) {
            bulk::mbar_expect_tx(&full[st], kTileBytes);
            bulk::bulk_load(smem + size_t(st) * kTileBytes, src + (blockIdx.x + (i + S) * gridDim.x) * kTileBytes, kTileBytes, &full[st], pol);
        }
    }
    block_sum_atomic(s, sum);
}

struct result { float ms; unsigned long long sum; };

template <class F>
static result time_it(F&& launch, uint64_t* dsum, int reps)
{
    cudaEvent_t e0, e1;
    CK(cudaEventCreate(&e0)); CK(cudaEventCreate(&e1));
    for (int i = 0; i < 3; ++i) launch();
    CK(cudaDeviceSynchronize());
    CK(cudaMemset(dsum, 0, 8));
    launch();
    unsigned long long sum = 0;
    CK(cudaMemcpy(&sum, dsum, 8, cudaMemcpyDeviceToHost));
    CK(cudaEventRecord(e0));
    for (int i = 0; i < reps; ++i) launch();
    CK(cudaEventRecord(e1));
    CK(cudaEventSynchronize(e1));
    CK(cudaGetLastError());
    float ms = 0;
    CK(cudaEventElapsedTime(&ms, e0, e1));
    return {ms / reps, sum};
}

__global__ void fill(uint32_t* x, size_t n)
{
    for (size_t i = size_t(blockIdx.x) * blockDim.x + threadIdx.x; i < n; i += size_t(gridDim.x) * blockDim.x) {
        uint64_t z = i + 0x9E3779B97F4A7C15ULL;
        z = (z ^ (z >> 30)) * 0xBF58476D1CE4E5B9ULL;
        z = (z ^ (z >> 27)) * 0x94D049BB133111EBULL;
        x[i] = uint32_t(z ^ (z >> 31));
    }
}

template <int Op, class P>
static void run_all(const char* what, const uint32_t* pairs, size_t n, uint64_t* dsum, int sms)
{
    const P p{};
    const int reps = 20;
    const double gb = 8.0 * double(n) / 1e9;
    unsigned long long want = 0;
    auto report = [&](const char* name, result r) {
        if (!want) want = r.sum;
        std::printf("{\"case\": \"%s\", \"pairs\": %zu, \"variant\": \"%s\", \"us\": %.1f, \"gbs\": %.0f, \"sum_ok\": %s}\n", what, n, name,
                    r.ms * 1e3, gb / (r.ms * 1e-3), r.sum == want ? "true" : "false");
        std::fflush(stdout);
    };
    const unsigned tiles4 = unsigned(n / (4 * 4 * 256)), tiles8 = unsigned(n / (4 * 8 * 256));
    report("lib", time_it([&] { reduce_ld<Op, P, 4, false><<<tiles4, 256>>>(pairs, n, dsum, p); }, dsum, reps));
    report("lib_old_grid", time_it([&] { reduce_ld<Op, P, 4, false><<<unsigned((n + 1023) / 1024), 256>>>(pairs, n, dsum, p); }, dsum, reps));
    report("fence", time_it([&] { reduce_ld<Op, P, 4, true><<<tiles4, 256>>>(pairs, n, dsum, p); }, dsum, reps));
    report("u8", time_it([&] { reduce_ld<Op, P, 8, false><<<tiles8, 256>>>(pairs, n, dsum, p); }, dsum, reps));
    report("u8_fence", time_it([&] { reduce_ld<Op, P, 8, true><<<tiles8, 256>>>(pairs, n, dsum, p); }, dsum, reps));
    for (int per_sm : {2, 4, 6, 8}) {
        char name[64];
        std::snprintf(name, sizeof name, "lib_persistent_%d_per_sm", per_sm);
        report(name, time_it([&] { reduce_ld<Op, P, 4, false><<<sms * per_sm, 256>>>(pairs, n, dsum, p); }, dsum, reps));
    }
    CK(cudaFuncSetAttribute(reduce_tma<Op, P, 3>, cudaFuncAttributeMaxDynamicSharedMemorySize, 3 * 16384));
    CK(cudaFuncSetAttribute(reduce_tma<Op, P, 4>, cudaFuncAttributeMaxDynamicSharedMemorySize, 4 * 16384));
    CK(cudaFuncSetAttribute(reduce_tma<Op, P, 6>, cudaFuncAttributeMaxDynamicSharedMemorySize, 6 * 16384));
    for (int per_sm : {2, 3, 4}) {
        char name[64];
        std::snprintf(name, sizeof name, "tma_3_stages_%d_per_sm", per_sm);
        report(name, time_it([&] { reduce_tma<Op, P, 3><<<sms * per_sm, 256, 3 * 16384>>>(pairs, n, dsum, p); }, dsum, reps));
        std::snprintf(name, sizeof name, "tma_4_stages_%d_per_sm", per_sm);
        report(name, time_it([&] { reduce_tma<Op, P, 4><<<sms * per_sm, 256, 4 * 16384>>>(pairs, n, dsum, p); }, dsum, reps));
    }
    report("tma_6_stages_2_per_sm", time_it([&] { reduce_tma<Op, P, 6><<<sms * 2, 256, 6 * 16384>>>(pairs, n, dsum, p); }, dsum, reps));
}

int main(int argc, char** argv)
{
    size_t want_pairs = argc > 1 ? std::strtoull(argv[1], nullptr, 10) : 100000000ull;
    const size_t unit = 4 * 8 * 256;  // pairs per tile of the widest variant; 16 KiB TMA tiles divide it
    const size_t n = want_pairs / unit * unit;
    int sms = 0;
    CK(cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, 0));
    uint32_t* pairs;
    uint64_t* dsum;
    CK(cudaMalloc((void**)&pairs, n * 8));
    CK(cudaMalloc((void**)&dsum, 8));
    fill<<<sms * 8, 256>>>(pairs, 2 * n);
    CK(cudaDeviceSynchronize());
    using rgba = static_masks<static_layout<uint32_t, 8, 8, 8, 8>>;
    using type2 = static_masks<static_layout<uint32_t, 1, 2, 3, 4, 5, 6, 11>>;
    run_all<OP_ADD_WRAP, rgba>("add_wrap <8,8,8,8>", pairs, n, dsum, sms);
    run_all<OP_ADD_UNSIGNED_SATURATE, rgba>("add_unsigned_saturate <8,8,8,8>", pairs, n, dsum, sms);
    run_all<OP_ADD_SIGNED_SATURATE, type2>("add_signed_saturate <1,2,3,4,5,6,11>", pairs, n, dsum, sms);
    return 0;
}
