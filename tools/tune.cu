// tools/tune.cu -- kernel-variant sweep for the streaming kernels (run on a B200
// through gpurun; results are copied to profiles/).  Not part of the library.
//   build/tune [log2_elems=30] [reps=10]
// Prints one line per variant: achieved algorithmic GB/s (12 B/elem for the
// binary op on uint32 RGBA8, 24 B/elem for uint64) measured with CUDA events,
// best and median over reps, buffers >> L2.
#include <algorithm>
#include <cstdio>
#include <cstdlib>
#include <vector>

#include "pint/cuda_kernels.cuh"
#include "pint/cuda_kernels_bulk.cuh"
#include "binary_bulk_variant.cuh"

using namespace pint::cuda;
using namespace pint::cuda::kernels;

#define CK(x) do { cudaError_t e_ = (x); if (e_ != cudaSuccess) { std::printf("CUDA error %s at %s:%d\n", cudaGetErrorString(e_), __FILE__, __LINE__); std::exit(1); } } while (0)

static int g_reps = 10;
static int g_sms = 148;

template <class F>
static void time_it(const char* name, double bytes, F launch)
{
    cudaEvent_t e0, e1;
    CK(cudaEventCreate(&e0)); CK(cudaEventCreate(&e1));
    for (int i = 0; i < 3; ++i) launch();
    CK(cudaDeviceSynchronize());
    std::vector<float> ms(g_reps);
    for (int i = 0; i < g_reps; ++i) {
        CK(cudaEventRecord(e0));
        launch();
        CK(cudaEventRecord(e1));
        CK(cudaEventSynchronize(e1));
        CK(cudaEventElapsedTime(&ms[i], e0, e1));
    }
    CK(cudaGetLastError());
    std::sort(ms.begin(), ms.end());
    std::printf("%-64s best %8.1f GB/s  median %8.1f GB/s  (%.3f ms)\n", name, bytes / ms[0] / 1e6, bytes / ms[g_reps / 2] / 1e6, ms[g_reps / 2]);
    std::fflush(stdout);
}

template <class Fn, int U, int B, int LD, int ST>
static void run_stream(const char* tag, const void* a, const void* b, void* out, size_t nvec, int ctas_per_sm)
{
    using Cfg = stream_cfg<U, B, LD, ST>;
    constexpr size_t tile = size_t(U) * B;
    const size_t ntiles = (nvec + tile - 1) / tile;
    const size_t grid = ctas_per_sm > 0 ? std::min(ntiles, size_t(g_sms) * ctas_per_sm) : ntiles;
    char name[128];
    std::snprintf(name, sizeof name, "%s stream U%d B%d ld%d st%d ctas/sm %d", tag, U, B, LD, ST, ctas_per_sm);
    time_it(name, double(nvec) * 48, [&] {
        binary_stream_kernel<Fn, Cfg><<<unsigned(grid), B>>>((const uint4*)a, (const uint4*)b, (uint4*)out, nvec, Fn{});
    });
}

template <class Fn, int S, int TB, int B>
static void run_bulk(const char* tag, const void* a, const void* b, void* out, size_t nvec, int ctas_per_sm)
{
    using Cfg = bulk::bulk_cfg<S, TB, B>;
    const size_t ntiles = nvec / (TB / 16);
    CK(cudaFuncSetAttribute(bulk::binary_bulk_kernel<Fn, Cfg>, cudaFuncAttributeMaxDynamicSharedMemorySize, Cfg::SMEM_BYTES));
    const size_t grid = std::min(ntiles, size_t(g_sms) * ctas_per_sm);
    char name[128];
    std::snprintf(name, sizeof name, "%s bulk  S%d tile%dK B%d ctas/sm %d (smem %d KB)", tag, S, TB / 1024, B, ctas_per_sm, Cfg::SMEM_BYTES / 1024);
    time_it(name, double(ntiles) * TB * 3, [&] {
        bulk::binary_bulk_kernel<Fn, Cfg><<<unsigned(grid), B, Cfg::SMEM_BYTES>>>((const uint4*)a, (const uint4*)b, (uint4*)out, ntiles, Fn{});
    });
}

template <class Fn, int U, int B, int LD, int ST>
static void run_pair(const char* tag, const void* a, const void* b, void* o0, void* o1, size_t nvec, int ctas_per_sm)
{
    using Cfg = stream_cfg<U, B, LD, ST>;
    constexpr size_t tile = size_t(U) * B;
    const size_t ntiles = (nvec + tile - 1) / tile;
    const size_t grid = ctas_per_sm > 0 ? std::min(ntiles, size_t(g_sms) * ctas_per_sm) : ntiles;
    char name[128];
    std::snprintf(name, sizeof name, "%s pair U%d B%d ld%d st%d ctas/sm %d", tag, U, B, LD, ST, ctas_per_sm);
    time_it(name, double(nvec) * 64, [&] {
        binary_pair_stream_kernel<Fn, Cfg><<<unsigned(grid), B>>>((const uint4*)a, (const uint4*)b, (uint4*)o0, (uint4*)o1, nvec, Fn{});
    });
}

// 256-bit global accesses (LDG.E.256 / STG.E.256, sm_100a): half the memory instructions per byte
struct u8v { uint32_t v[8]; };
__device__ __forceinline__ u8v ld32cs(const void* p)
{
    u8v r;
    asm volatile("ld.global.cs.v8.u32 {%0,%1,%2,%3,%4,%5,%6,%7}, [%8];"
                 : "=r"(r.v[0]), "=r"(r.v[1]), "=r"(r.v[2]), "=r"(r.v[3]), "=r"(r.v[4]), "=r"(r.v[5]), "=r"(r.v[6]), "=r"(r.v[7]) : "l"(p));
    return r;
}
__device__ __forceinline__ void st32cs(void* p, const u8v& r)
{
    asm volatile("st.global.cs.v8.u32 [%0], {%1,%2,%3,%4,%5,%6,%7,%8};"
                 :: "l"(p), "r"(r.v[0]), "r"(r.v[1]), "r"(r.v[2]), "r"(r.v[3]), "r"(r.v[4]), "r"(r.v[5]), "r"(r.v[6]), "r"(r.v[7]) : "memory");
}
template <class Fn, int U, int B, bool Pair>
__global__ void __launch_bounds__(B)
stream256_kernel(const char* a, const char* b, char* o0, char* o1, size_t nv32, const __grid_constant__ Fn fn)
{
    const size_t base = (size_t(blockIdx.x) * U * B + threadIdx.x);
    u8v va[U], vb[U];
#pragma unroll
    for (int u = 0; u < U; ++u) if (base + size_t(u) * B < nv32) va[u] = ld32cs(a + (base + size_t(u) * B) * 32);
#pragma unroll
    for (int u = 0; u < U; ++u) if (base + size_t(u) * B < nv32) vb[u] = ld32cs(b + (base + size_t(u) * B) * 32);
#pragma unroll
    for (int u = 0; u < U; ++u) {
        if (base + size_t(u) * B >= nv32) continue;
        u8v r0, r1;
#pragma unroll
        for (int j = 0; j < 8; ++j) {
            if constexpr (Pair) { r0.v[j] = fn.first(va[u].v[j], vb[u].v[j]); r1.v[j] = fn.second(va[u].v[j], vb[u].v[j]); }
            else r0.v[j] = fn(va[u].v[j], vb[u].v[j]);
        }
        st32cs(o0 + (base + size_t(u) * B) * 32, r0);
        if constexpr (Pair) st32cs(o1 + (base + size_t(u) * B) * 32, r1);
    }
}
template <class Fn, int U, int B, bool Pair>
static void run256(const char* tag, const void* a, const void* b, void* o0, void* o1, size_t nvec)
{
    const size_t nv32 = nvec / 2;
    const size_t grid = (nv32 + size_t(U) * B - 1) / (size_t(U) * B);
    char name[128];
    std::snprintf(name, sizeof name, "%s 256-bit accesses U%d B%d", tag, U, B);
    time_it(name, double(nvec) * (Pair ? 64 : 48), [&] {
        stream256_kernel<Fn, U, B, Pair><<<unsigned(grid), B>>>((const char*)a, (const char*)b, (char*)o0, (char*)o1, nv32, Fn{});
    });
}

// persistent CTAs that take tiles in address order from a global counter (instead of a grid stride)
template <class Fn, class Cfg>
__global__ void __launch_bounds__(Cfg::BLOCK)
binary_stream_dyn_kernel(const uint4* a, const uint4* b, uint4* out, size_t nvec, unsigned long long* counter, const __grid_constant__ Fn fn)
{
    using W = typename Fn::word;
    constexpr int U = Cfg::UNROLL, B = Cfg::BLOCK;
    constexpr size_t tile = size_t(U) * B;
    const size_t nfull = nvec / tile;
    auto f = [&](W x, W y) { return fn(x, y); };
    __shared__ unsigned long long next;
    for (;;) {
        if (threadIdx.x == 0) next = atomicAdd(counter, 1ULL);
        __syncthreads();
        const size_t t = next;
        __syncthreads();
        if (t >= nfull) break;
        const size_t base = t * tile + threadIdx.x;
        uint4 va[U], vb[U];
#pragma unroll
        for (int u = 0; u < U; ++u) va[u] = ld16<Cfg::LD>(a + base + size_t(u) * B);
#pragma unroll
        for (int u = 0; u < U; ++u) vb[u] = ld16<Cfg::LD>(b + base + size_t(u) * B);
#pragma unroll
        for (int u = 0; u < U; ++u) st16<Cfg::ST>(out + base + size_t(u) * B, map_words2<W>(va[u], vb[u], f));
    }
}

static unsigned long long* g_counter = nullptr;

template <class Fn, int U, int B>
static void run_dyn(const char* tag, const void* a, const void* b, void* out, size_t nvec, int ctas_per_sm)
{
    using Cfg = stream_cfg<U, B, 2, 2>;
    if (!g_counter) CK(cudaMalloc(&g_counter, 8));
    char name[128];
    std::snprintf(name, sizeof name, "%s dyn    U%d B%d ctas/sm %d (atomic tile counter)", tag, U, B, ctas_per_sm);
    time_it(name, double(nvec) * 48, [&] {
        CK(cudaMemsetAsync(g_counter, 0, 8));
        binary_stream_dyn_kernel<Fn, Cfg><<<unsigned(g_sms * ctas_per_sm), B>>>((const uint4*)a, (const uint4*)b, (uint4*)out, nvec, g_counter, Fn{});
    });
}

__global__ void fill(uint4* p, size_t nvec, uint32_t seed)
{
    for (size_t i = size_t(blockIdx.x) * blockDim.x + threadIdx.x; i < nvec; i += size_t(gridDim.x) * blockDim.x) {
        uint32_t x = uint32_t(i) * 2654435761u ^ seed;
        p[i] = make_uint4(x, x * 3 + 1, x ^ 0x9E3779B9u, x * 7 + 5);
    }
}

__global__ void compare(const uint4* x, const uint4* y, size_t nvec, unsigned long long* bad)
{
    for (size_t i = size_t(blockIdx.x) * blockDim.x + threadIdx.x; i < nvec; i += size_t(gridDim.x) * blockDim.x) {
        uint4 p = x[i], q = y[i];
        if (p.x != q.x || p.y != q.y || p.z != q.z || p.w != q.w) atomicAdd(bad, 1ULL);
    }
}

// ---- mode 10: L2 prefetch-size / eviction hints on the 256-bit loads and stores of the op-pair kernel ----
template <int H>
__device__ __forceinline__ vec32 ld32_hint(const vec32* p)
{
    vec32 r;
#define PINT_LD32(QUAL)                                                                                       \
    asm volatile("ld.global" QUAL ".v8.u32 {%0,%1,%2,%3,%4,%5,%6,%7}, [%8];"                                  \
                 : "=r"(r.lo.x), "=r"(r.lo.y), "=r"(r.lo.z), "=r"(r.lo.w), "=r"(r.hi.x), "=r"(r.hi.y), "=r"(r.hi.z), "=r"(r.hi.w) \
                 : "l"(p))
    if constexpr (H == 0) PINT_LD32(".cs");
    else if constexpr (H == 1) PINT_LD32(".cs.L2::256B");
    else if constexpr (H == 2) PINT_LD32(".L1::no_allocate.L2::evict_first.L2::256B");
    else if constexpr (H == 3) PINT_LD32(".nc.L1::no_allocate.L2::256B");
    else if constexpr (H == 4) PINT_LD32(".cs.L2::128B");
    else PINT_LD32("");
#undef PINT_LD32
    return r;
}

template <int H>
__device__ __forceinline__ void st32_hint(vec32* p, const vec32& v)
{
#define PINT_ST32(QUAL)                                                                                       \
    asm volatile("st.global" QUAL ".v8.u32 [%0], {%1,%2,%3,%4,%5,%6,%7,%8};"                                  \
                 :: "l"(p), "r"(v.lo.x), "r"(v.lo.y), "r"(v.lo.z), "r"(v.lo.w), "r"(v.hi.x), "r"(v.hi.y), "r"(v.hi.z), "r"(v.hi.w) \
                 : "memory")
    if constexpr (H == 0) PINT_ST32(".cs");
    else if constexpr (H == 1) PINT_ST32(".L1::no_allocate.L2::evict_first");
    else if constexpr (H == 2) PINT_ST32(".wt");
    else PINT_ST32("");
#undef PINT_ST32
}

template <class Fn, int LH, int SH>
__global__ void __launch_bounds__(256)
pair256_hint_kernel(const vec32* a, const vec32* b, vec32* o0, vec32* o1, size_t nvec, const __grid_constant__ Fn fn)
{
    using W = typename Fn::word;
    const size_t i = size_t(blockIdx.x) * 256 + threadIdx.x;
    if (i >= nvec) return;
    const vec32 x = ld32_hint<LH>(a + i), y = ld32_hint<LH>(b + i);
    vec32 r0, r1;
    auto f0 = [&](W p, W q) { return fn.first(p, q); };
    auto f1 = [&](W p, W q) { return fn.second(p, q); };
    r0.lo = map_words2<W>(x.lo, y.lo, f0); r0.hi = map_words2<W>(x.hi, y.hi, f0);
    r1.lo = map_words2<W>(x.lo, y.lo, f1); r1.hi = map_words2<W>(x.hi, y.hi, f1);
    st32_hint<SH>(o0 + i, r0);
    st32_hint<SH>(o1 + i, r1);
}

template <class Fn, int LH, int SH>
static void run_hint(const void* a, const void* b, void* o0, void* o1, size_t nvec16)
{
    const size_t nvec = nvec16 / 2;
    char name[128];
    std::snprintf(name, sizeof name, "usat+ssat rgba8 pair 256-bit, load hint %d, store hint %d", LH, SH);
    time_it(name, double(nvec) * 128, [&] {
        pair256_hint_kernel<Fn, LH, SH><<<unsigned((nvec + 255) / 256), 256>>>((const vec32*)a, (const vec32*)b, (vec32*)o0, (vec32*)o1, nvec, Fn{});
    });
}

// ---- read / write mix ceilings (mode 12): pure data movement with the stream counts of the library's kernels ----
// NIN input streams are XOR-combined into NOUT identical output streams, 32-byte accesses, one CTA per tile,
// U vectors per stream per thread: the same access pattern as the stream kernels, without their arithmetic.
template <int NIN, int NOUT, int U>
__global__ void __launch_bounds__(256) mix_kernel(const vec32* in, vec32* out, size_t nvec, size_t in_stride, size_t out_stride)
{
    constexpr int B = 256;
    const size_t base = size_t(blockIdx.x) * (size_t(U) * B) + threadIdx.x;
    vec32 v[NIN][U];
#pragma unroll
    for (int k = 0; k < NIN; ++k)
#pragma unroll
        for (int u = 0; u < U; ++u)
            if (base + size_t(u) * B < nvec) v[k][u] = ld32(in + size_t(k) * in_stride + base + size_t(u) * B);
#pragma unroll
    for (int u = 0; u < U; ++u) {
        if (base + size_t(u) * B >= nvec) continue;
        vec32 r = v[0][u];
#pragma unroll
        for (int k = 1; k < NIN; ++k) {
            r.lo.x ^= v[k][u].lo.x; r.lo.y ^= v[k][u].lo.y; r.lo.z ^= v[k][u].lo.z; r.lo.w ^= v[k][u].lo.w;
            r.hi.x ^= v[k][u].hi.x; r.hi.y ^= v[k][u].hi.y; r.hi.z ^= v[k][u].hi.z; r.hi.w ^= v[k][u].hi.w;
        }
#pragma unroll
        for (int k = 0; k < NOUT; ++k) {
            r.lo.x += k;  // the outputs differ, so no store can be merged away
            st32(out + size_t(k) * out_stride + base + size_t(u) * B, r);
        }
    }
}

template <int NIN, int NOUT, int U>
static void run_mix(const void* in, void* out, size_t stream_bytes)
{
    const size_t nvec = stream_bytes / 32, tile = size_t(U) * 256;
    char name[160];
    std::snprintf(name, sizeof name, "mix %d read : %d write streams (%2.0f %% writes), U%d", NIN, NOUT, 100.0 * NOUT / (NIN + NOUT), U);
    time_it(name, double(stream_bytes) * (NIN + NOUT), [&] {
        mix_kernel<NIN, NOUT, U><<<unsigned((nvec + tile - 1) / tile), 256>>>((const vec32*)in, (vec32*)out, nvec, nvec, nvec);
    });
}

int main(int argc, char** argv)
{
    const int log2n = argc > 1 ? std::atoi(argv[1]) : 30;
    g_reps = argc > 2 ? std::atoi(argv[2]) : 10;
    const size_t bytes = (size_t(1) << log2n) * 4;  // per array
    const size_t nvec = bytes / 16;
    cudaDeviceProp prop;
    CK(cudaGetDeviceProperties(&prop, 0));
    g_sms = prop.multiProcessorCount;
    std::printf("device %s, %d SMs, array %.2f GiB x3\n", prop.name, g_sms, bytes / 1073741824.0);
    void *a, *b, *out, *out2;
    CK(cudaMalloc(&a, bytes)); CK(cudaMalloc(&b, bytes)); CK(cudaMalloc(&out, bytes)); CK(cudaMalloc(&out2, bytes));
    fill<<<g_sms * 8, 256>>>((uint4*)a, nvec, 1u);
    fill<<<g_sms * 8, 256>>>((uint4*)b, nvec, 2u);
    CK(cudaDeviceSynchronize());

    time_it("cudaMemcpyAsync D2D (1R+1W; bytes = 2 x array)", double(bytes) * 2, [&] { CK(cudaMemcpyAsync(out, a, bytes, cudaMemcpyDeviceToDevice)); });

    using RGBA = static_layout<uint32_t, 8, 8, 8, 8>;
    using FnU = swar_fn<OP_ADD_UNSIGNED_SATURATE, static_masks<RGBA>>;
    using FnS = swar_fn<OP_ADD_SIGNED_SATURATE, static_masks<RGBA>>;
    using IRR = static_layout<uint64_t, 3, 7, 11, 13, 17, 13>;
    using FnI = swar_fn<OP_ADD_SIGNED_SATURATE, static_masks<IRR>>;
    using FnIF = fused_fn<OP_ADD_SIGNED_SATURATE, static_fields<IRR>>;
    using FnX = swar_fn<OP_XOR, static_masks<RGBA>>;

    if (argc > 3 && std::atoi(argv[3]) == 10) {
        using Pu = pair_fn<OP_ADD_UNSIGNED_SATURATE, OP_ADD_SIGNED_SATURATE, static_masks<RGBA>>;
        for (int rep = 0; rep < 3; ++rep) {
            run256<Pu, 1, 256, true>("usat+ssat rgba8 pair (library kernel)", a, b, out, out2, nvec);
            run_hint<Pu, 0, 0>(a, b, out, out2, nvec);
            run_hint<Pu, 1, 0>(a, b, out, out2, nvec);
            run_hint<Pu, 2, 0>(a, b, out, out2, nvec);
            run_hint<Pu, 3, 0>(a, b, out, out2, nvec);
            run_hint<Pu, 4, 0>(a, b, out, out2, nvec);
            run_hint<Pu, 5, 0>(a, b, out, out2, nvec);
            run_hint<Pu, 0, 1>(a, b, out, out2, nvec);
            run_hint<Pu, 0, 2>(a, b, out, out2, nvec);
            run_hint<Pu, 0, 3>(a, b, out, out2, nvec);
            run_hint<Pu, 1, 1>(a, b, out, out2, nvec);
        }
        std::printf("done\n");
        return 0;
    }
    if (argc > 3 && std::atoi(argv[3]) == 9) {
        using Pu = pair_fn<OP_ADD_UNSIGNED_SATURATE, OP_ADD_SIGNED_SATURATE, static_masks<RGBA>>;
        for (int rep = 0; rep < 2; ++rep) {
            run_stream<FnU, 2, 256, 2, 2>("usat rgba8 128-bit", a, b, out, nvec, 0);
            run256<FnU, 1, 256, false>("usat rgba8", a, b, out, out2, nvec);
            run256<FnU, 2, 256, false>("usat rgba8", a, b, out, out2, nvec);
            run256<FnU, 1, 128, false>("usat rgba8", a, b, out, out2, nvec);
            run256<FnU, 1, 512, false>("usat rgba8", a, b, out, out2, nvec);
            run_pair<Pu, 2, 256, 2, 2>("usat+ssat rgba8 128-bit", a, b, out, out2, nvec, 0);
            run256<Pu, 1, 256, true>("usat+ssat rgba8 pair", a, b, out, out2, nvec);
            run256<Pu, 2, 256, true>("usat+ssat rgba8 pair", a, b, out, out2, nvec);
            run256<Pu, 1, 128, true>("usat+ssat rgba8 pair", a, b, out, out2, nvec);
            run256<Pu, 1, 512, true>("usat+ssat rgba8 pair", a, b, out, out2, nvec);
        }
        std::printf("done\n");
        return 0;
    }
    if (argc > 3 && std::atoi(argv[3]) == 8) {
        // two ops in one pass (2 reads + 2 writes, 16 B per uint32 element): unroll / block / cache policy
        using Pu = pair_fn<OP_ADD_UNSIGNED_SATURATE, OP_ADD_SIGNED_SATURATE, static_masks<RGBA>>;
        using C16 = static_layout<uint64_t, 16, 16, 16, 16>;
        using Pc2 = pair_fn<OP_ADD_SIGNED_SATURATE, OP_SUB_SIGNED_SATURATE, static_masks<C16>>;
        for (int rep = 0; rep < 2; ++rep) {
            run_stream<FnU, 2, 256, 2, 2>("usat rgba8 single op (12 B/elem)", a, b, out, nvec, 0);
            run_pair<Pu, 1, 256, 2, 2>("usat+ssat rgba8", a, b, out, out2, nvec, 0);
            run_pair<Pu, 2, 256, 2, 2>("usat+ssat rgba8", a, b, out, out2, nvec, 0);
            run_pair<Pu, 4, 256, 2, 2>("usat+ssat rgba8", a, b, out, out2, nvec, 0);
            run_pair<Pu, 2, 128, 2, 2>("usat+ssat rgba8", a, b, out, out2, nvec, 0);
            run_pair<Pu, 2, 512, 2, 2>("usat+ssat rgba8", a, b, out, out2, nvec, 0);
            run_pair<Pu, 1, 512, 2, 2>("usat+ssat rgba8", a, b, out, out2, nvec, 0);
            run_pair<Pu, 2, 256, 2, 0>("usat+ssat rgba8", a, b, out, out2, nvec, 0);
            run_pair<Pu, 2, 256, 0, 0>("usat+ssat rgba8", a, b, out, out2, nvec, 0);
            run_pair<Pu, 2, 256, 1, 2>("usat+ssat rgba8", a, b, out, out2, nvec, 0);
            run_pair<Pu, 2, 256, 2, 2>("usat+ssat rgba8", a, b, out, out2, nvec, 4);
            run_pair<Pu, 2, 256, 2, 2>("usat+ssat rgba8", a, b, out, out2, nvec, 8);
            run_pair<Pc2, 2, 256, 2, 2>("ssat add+sub u64 16x4", a, b, out, out2, nvec, 0);
            run_pair<Pc2, 4, 256, 2, 2>("ssat add+sub u64 16x4", a, b, out, out2, nvec, 0);
        }
        std::printf("done\n");
        return 0;
    }
    if (argc > 3 && std::atoi(argv[3]) == 7) {
        // library ordered kernel (self re-arming counter ring) vs the tune-local dyn kernel (memset per launch)
        tile_counter* ring;
        CK(cudaMalloc(&ring, sizeof(tile_counter) * kTileCounterSlots));
        CK(cudaMemset(ring, 0, sizeof(tile_counter) * kTileCounterSlots));
        unsigned cursor = 0;
        auto lib_run = [&](auto fn_tag, const char* label, int cps, bool same_slot) {
            using Fn = decltype(fn_tag);
            using Cfg = stream_cfg<2, 256, 2, 2>;
            char name[160];
            std::snprintf(name, sizeof name, "%s library ordered kernel ctas/sm %d %s", label, cps, same_slot ? "(one slot)" : "(ring)");
            time_it(name, double(nvec) * 48, [&] {
                tile_counter* tc = ring + (same_slot ? 0 : (cursor++ % kTileCounterSlots));
                binary_stream_ordered_kernel<Fn, Cfg><<<unsigned(g_sms * cps), 256>>>((const uint4*)a, (const uint4*)b, (uint4*)out, nvec, tc, Fn{});
            });
        };
        for (int rep = 0; rep < 2; ++rep) {
            run_stream<FnU, 2, 256, 2, 2>("usat rgba8 (one CTA per tile)", a, b, out, nvec, 0);
            run_stream<FnS, 2, 256, 2, 2>("ssat rgba8 (one CTA per tile)", a, b, out, nvec, 0);
            run_dyn<FnU, 2, 256>("usat rgba8", a, b, out, nvec, 4);
            run_dyn<FnS, 2, 256>("ssat rgba8", a, b, out, nvec, 4);
            lib_run(FnU{}, "usat rgba8", 4, false);
            lib_run(FnS{}, "ssat rgba8", 4, false);
            lib_run(FnU{}, "usat rgba8", 4, true);
            lib_run(FnS{}, "ssat rgba8", 4, true);
            lib_run(FnU{}, "usat rgba8", 5, false);
            lib_run(FnS{}, "ssat rgba8", 5, false);
        }
        // different buffers: swap roles (out -> a) to change relative alignment of the three streams
        run_dyn<FnS, 2, 256>("ssat rgba8 (a=out2,b=a,out=b)", out2, a, b, nvec, 4);
        run_dyn<FnU, 2, 256>("usat rgba8 (a=out2,b=a,out=b)", out2, a, b, nvec, 4);
        run_dyn<FnS, 2, 256>("ssat rgba8 (a=b,b=out,out=a)", b, out, a, nvec, 4);
        run_stream<FnS, 2, 256, 2, 2>("ssat rgba8 one CTA per tile (a=out2,b=a,out=b)", out2, a, b, nvec, 0);
        std::printf("done\n");
        return 0;
    }
    if (argc > 3 && std::atoi(argv[3]) == 6) {
        // persistent CTAs + atomic tile counter (tiles handed out in address order): how robust is the
        // best CTAs/SM across ops, element types and array sizes?
        using C16 = static_layout<uint64_t, 16, 16, 16, 16>;
        using FnC2 = swar_fn<OP_SUB_SIGNED_SATURATE, static_masks<C16>>;
        using C565 = static_layout<uint16_t, 5, 6, 5>;
        using FnC0 = swar_fn<OP_ADD_WRAP, static_masks<C565>>;
        using T2 = static_layout<uint32_t, 1, 2, 3, 4, 5, 6, 11>;
        using FnT2 = swar_fn<OP_ADD_SIGNED_SATURATE, static_masks<T2>>;
        for (size_t nv : {nvec, nvec / 8}) {
            std::printf("--- %zu vectors per array (%.2f GiB)\n", nv, nv * 16 / 1073741824.0);
            run_stream<FnU, 2, 256, 2, 2>("usat rgba8 (one CTA per tile)", a, b, out, nv, 0);
            for (int cps : {3, 4, 5, 6}) run_dyn<FnU, 2, 256>("usat rgba8", a, b, out, nv, cps);
            for (int cps : {3, 4, 5, 6}) run_dyn<FnS, 2, 256>("ssat rgba8", a, b, out, nv, cps);
            for (int cps : {3, 4, 5, 6}) run_dyn<FnC2, 2, 256>("sub_ssat u64 16x4", a, b, out, nv, cps);
            for (int cps : {3, 4, 5, 6}) run_dyn<FnC0, 2, 256>("add_wrap u16 565", a, b, out, nv, cps);
            for (int cps : {4, 6, 8}) run_dyn<FnT2, 2, 256>("add_ssat u32 type2", a, b, out, nv, cps);
            for (int cps : {4, 6, 8}) run_dyn<FnI, 2, 256>("add_ssat irregular u64 swar", a, b, out, nv, cps);
            for (int cps : {4, 6, 8}) run_dyn<FnIF, 2, 256>("add_ssat irregular u64 fused", a, b, out, nv, cps);
            for (int cps : {2, 3, 4}) run_dyn<FnI, 4, 256>("add_ssat irregular u64 swar", a, b, out, nv, cps);
            for (int cps : {2, 3, 4}) run_dyn<FnIF, 4, 256>("add_ssat irregular u64 fused", a, b, out, nv, cps);
            for (int cps : {2, 3}) run_dyn<FnU, 2, 512>("usat rgba8", a, b, out, nv, cps);
            for (int cps : {1, 2}) run_dyn<FnU, 4, 512>("usat rgba8", a, b, out, nv, cps);
        }
        std::printf("done\n");
        return 0;
    }
    if (argc > 3 && std::atoi(argv[3]) == 5) {
        // unpack (C3): 2^28 irregular u64 elements -> 6 int32 planes; quads per thread, block, store policy
        const size_t n = size_t(1) << 28;            // 2 GiB packed in `a`, planes in b/out (4 GiB each = 3+3 planes of 1 GiB)
        plane_ptrs pp{};
        for (int f = 0; f < 6; ++f) pp.p[f] = (char*)(f < 3 ? b : out) + size_t(f % 3) * (n * 4);
        auto unpack_run = [&](auto q_tag, auto st_tag, int block) {
            constexpr int Q = decltype(q_tag)::value, ST = decltype(st_tag)::value;
            const size_t nquads = n / 4, tile = size_t(Q) * block;
            char name[160];
            std::snprintf(name, sizeof name, "unpack irregular u64 -> 6 x int32: Q%d B%d st%d", Q, block, ST);
            time_it(name, double(n) * 32, [&] {
                unpack_quad_kernel<uint64_t, 4, true, static_fields<IRR>, Q, ST><<<unsigned((nquads + tile - 1) / tile), block>>>((const uint64_t*)a, pp, nquads, static_fields<IRR>{});
            });
        };
        using I1 = std::integral_constant<int, 1>; using I2 = std::integral_constant<int, 2>; using I4 = std::integral_constant<int, 4>; using I0 = std::integral_constant<int, 0>;
        for (int block : {128, 256}) {
            unpack_run(I1{}, I2{}, block); unpack_run(I2{}, I2{}, block); unpack_run(I4{}, I2{}, block);
            unpack_run(I1{}, I0{}, block); unpack_run(I2{}, I0{}, block);
            unpack_run(I1{}, I1{}, block);
        }
        std::printf("done\n");
        return 0;
    }
    if (argc > 3 && std::atoi(argv[3]) == 12) {
        // what the DRAM side delivers for each read / write mix, 1 GiB per stream: `a`+`b` hold up to 8 input streams,
        // `out`+`out2` up to 8 output streams (the arrays are 4 GiB each at log2 = 30)
        const size_t sb = size_t(1) << 30;
        run_mix<1, 1, 2>(a, out, sb); run_mix<1, 1, 4>(a, out, sb);   // copy:        unary maps (slice, shift, binary_scalar)
        run_mix<2, 1, 2>(a, out, sb); run_mix<2, 1, 4>(a, out, sb);   // 67 % reads:  binary ops
        run_mix<2, 2, 1>(a, out, sb); run_mix<2, 2, 2>(a, out, sb);   // 50 / 50:     two ops per pass
        run_mix<1, 3, 1>(a, out, sb); run_mix<1, 3, 2>(a, out, sb);   // 75 % writes: unpack u64 -> 6 int32 planes (8 B in, 24 B out)
        run_mix<3, 1, 1>(a, out, sb); run_mix<3, 1, 2>(a, out, sb);   // 75 % reads:  repack
        run_mix<1, 4, 1>(a, out, sb); run_mix<4, 1, 1>(a, out, sb);   // RGBA8 <-> four int32 planes (20 % / 80 %)
        time_it("cudaMemsetAsync 4 GiB (write-only, constant data)", double(bytes), [&] { CK(cudaMemsetAsync(out, 1, bytes)); });
        time_it("cudaMemcpyAsync D2D 4 GiB (50 / 50)", double(bytes) * 2, [&] { CK(cudaMemcpyAsync(out, a, bytes, cudaMemcpyDeviceToDevice)); });
        std::printf("done\n");
        return 0;
    }
    if (argc > 3 && std::atoi(argv[3]) == 11) {
        // unpack (C3) through shared memory + bulk (TMA) stores against the register-to-global kernels:
        // tile size, CTA size, one tile per CTA or a persistent grid
        const size_t n = size_t(1) << 28;
        plane_ptrs pp{};
        for (int f = 0; f < 6; ++f) pp.p[f] = (char*)(f < 3 ? b : out) + size_t(f % 3) * (n * 4);
        time_it("unpack_oct_kernel K1 (library default)", double(n) * 32, [&] {
            unpack_oct_kernel<uint64_t, true, static_fields<IRR>><<<unsigned((n / 8 + 255) / 256), 256>>>((const uint64_t*)a, pp, n / 8, static_fields<IRR>{});
        });
        time_it("unpack_oct_kernel K2", double(n) * 32, [&] {
            unpack_oct_kernel<uint64_t, true, static_fields<IRR>, 2><<<unsigned((n / 8 + 511) / 512), 256>>>((const uint64_t*)a, pp, n / 8, static_fields<IRR>{});
        });
        auto bulk_run = [&](auto tile_tag, auto block_tag, int ctas_per_sm) {
            constexpr int TILE = decltype(tile_tag)::value, BLOCK = decltype(block_tag)::value;
            auto kern = unpack_bulk_kernel<uint64_t, true, static_fields<IRR>, TILE, BLOCK>;
            const int smem = 6 * TILE * 4;
            CK(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, smem));
            const size_t ntiles = n / TILE;
            const size_t grid = ctas_per_sm > 0 ? std::min(ntiles, size_t(g_sms) * ctas_per_sm) : ntiles;
            char name[160];
            std::snprintf(name, sizeof name, "unpack_bulk_kernel TILE %d B%d smem %d KB ctas/sm %d", TILE, BLOCK, smem / 1024, ctas_per_sm);
            time_it(name, double(n) * 32, [&] { kern<<<unsigned(grid), BLOCK, smem>>>((const uint64_t*)a, pp, ntiles, 6, static_fields<IRR>{}); });
        };
        using T512 = std::integral_constant<int, 512>; using T1K = std::integral_constant<int, 1024>; using T2K = std::integral_constant<int, 2048>;
        using T4K = std::integral_constant<int, 4096>; using T8K = std::integral_constant<int, 8192>;
        using B128 = std::integral_constant<int, 128>; using B256 = std::integral_constant<int, 256>; using B512 = std::integral_constant<int, 512>;
        for (int cps : {0, 2, 4, 8}) {
            bulk_run(T512{}, B128{}, cps);
            bulk_run(T1K{}, B128{}, cps); bulk_run(T1K{}, B256{}, cps);
            bulk_run(T2K{}, B128{}, cps); bulk_run(T2K{}, B256{}, cps); bulk_run(T2K{}, B512{}, cps);
            bulk_run(T4K{}, B256{}, cps); bulk_run(T4K{}, B512{}, cps);
            bulk_run(T8K{}, B512{}, cps);
        }
        // directional ceilings of this box on the same buffers
        time_it("cudaMemsetAsync 6 GiB (write-only, constant data)", double(n) * 24, [&] { CK(cudaMemsetAsync(b, 1, n * 12)); CK(cudaMemsetAsync(out, 1, n * 12)); });
        std::printf("done\n");
        return 0;
    }
    if (argc > 3 && std::atoi(argv[3]) == 4) {
        // one CTA per tile, residency capped by an unused dynamic shared-memory request
        // (227 KB per SM / request = CTAs per SM): is full occupancy the best occupancy?
        using C16 = static_layout<uint64_t, 16, 16, 16, 16>;
        using FnC2 = swar_fn<OP_ADD_SIGNED_SATURATE, static_masks<C16>>;
        auto occ_run = [&](auto fn_tag, auto cfg_tag, const char* label, int ctas) {
            using Fn = decltype(fn_tag);
            using Cfg = decltype(cfg_tag);
            constexpr size_t tile = size_t(Cfg::UNROLL) * Cfg::BLOCK;
            const size_t ntiles = (nvec + tile - 1) / tile;
            const int smem = ctas > 0 ? (227 * 1024 / ctas - 1024) / 1024 * 1024 : 0;
            CK(cudaFuncSetAttribute(binary_stream_kernel<Fn, Cfg>, cudaFuncAttributeMaxDynamicSharedMemorySize, smem));
            char name[160];
            std::snprintf(name, sizeof name, "%s U%d B%d one CTA/tile, <= %d CTAs/SM (smem pad %d KB)", label, Cfg::UNROLL, Cfg::BLOCK, ctas, smem / 1024);
            time_it(name, double(nvec) * 48, [&] {
                binary_stream_kernel<Fn, Cfg><<<unsigned(ntiles), Cfg::BLOCK, smem>>>((const uint4*)a, (const uint4*)b, (uint4*)out, nvec, Fn{});
            });
        };
        for (int ctas : {0, 7, 6, 5, 4, 3, 2}) occ_run(FnU{}, stream_cfg<2, 256, 2, 2>{}, "usat rgba8", ctas);
        for (int ctas : {0, 3, 2, 1}) occ_run(FnU{}, stream_cfg<2, 512, 2, 2>{}, "usat rgba8", ctas);
        for (int ctas : {0, 6, 4, 3, 2}) occ_run(FnU{}, stream_cfg<4, 256, 2, 2>{}, "usat rgba8", ctas);
        for (int ctas : {0, 6, 5, 4, 3}) occ_run(FnC2{}, stream_cfg<2, 256, 2, 2>{}, "add_ssat u64 16x4", ctas);
        for (int ctas : {0, 6, 4}) occ_run(FnI{}, stream_cfg<4, 256, 2, 2>{}, "add_ssat irregular u64 swar", ctas);
        for (int ctas : {0, 6, 4}) occ_run(FnIF{}, stream_cfg<4, 256, 2, 2>{}, "add_ssat irregular u64 fused", ctas);
        std::printf("done\n");
        return 0;
    }
    if (argc > 3 && std::atoi(argv[3]) == 3) {
        // small arrays (C0: 2^26 uint16 = 128 MiB per array): rotate over 8 disjoint windows of the
        // big buffers so that no launch finds its data in the 126 MB L2
        using C565 = static_layout<uint16_t, 5, 6, 5>;
        using FnC0 = swar_fn<OP_ADD_WRAP, static_masks<C565>>;
        const size_t small = size_t(1) << 23;  // vectors = 128 MiB
        int it = 0;
        auto small_run = [&](auto cfg_tag, const char* label, int cps) {
            using Cfg = decltype(cfg_tag);
            constexpr size_t tile = size_t(Cfg::UNROLL) * Cfg::BLOCK;
            const size_t ntiles = (small + tile - 1) / tile;
            const size_t grid = cps > 0 ? std::min(ntiles, size_t(g_sms) * cps) : ntiles;
            char name[128];
            std::snprintf(name, sizeof name, "add_wrap u16 565, 128 MiB arrays: %s ctas/sm %d", label, cps);
            time_it(name, double(small) * 48, [&] {
                const size_t off = size_t(it++ % 8) * small;
                binary_stream_kernel<FnC0, Cfg><<<unsigned(grid), Cfg::BLOCK>>>((const uint4*)a + off, (const uint4*)b + off, (uint4*)out + off, small, FnC0{});
            });
        };
        g_reps = 40;
        for (int cps : {0, 4, 8}) {
            small_run(stream_cfg<1, 256, 2, 2>{}, "U1 B256", cps);
            small_run(stream_cfg<2, 256, 2, 2>{}, "U2 B256", cps);
            small_run(stream_cfg<4, 256, 2, 2>{}, "U4 B256", cps);
            small_run(stream_cfg<2, 512, 2, 2>{}, "U2 B512", cps);
            small_run(stream_cfg<2, 1024, 2, 2>{}, "U2 B1024", cps);
            small_run(stream_cfg<8, 256, 2, 2>{}, "U8 B256", cps);
        }
        time_it("cudaMemcpyAsync D2D 128 MiB (bytes = 2 x array)", double(small) * 32, [&] {
            const size_t off = size_t(it++ % 8) * small;
            CK(cudaMemcpyAsync((uint4*)out + off, (const uint4*)a + off, small * 16, cudaMemcpyDeviceToDevice));
        });
        std::printf("done\n");
        return 0;
    }
    if (argc > 3 && std::atoi(argv[3]) == 2) {
        // second sweep: one CTA per tile for every representative op, and dynamic persistent grids
        using C16 = static_layout<uint64_t, 16, 16, 16, 16>;
        using FnC2a = swar_fn<OP_ADD_SIGNED_SATURATE, static_masks<C16>>;
        using FnC2s = swar_fn<OP_SUB_SIGNED_SATURATE, static_masks<C16>>;
        using C565 = static_layout<uint16_t, 5, 6, 5>;
        using FnC0 = swar_fn<OP_ADD_WRAP, static_masks<C565>>;
        using FnC0s = swar_fn<OP_ADD_SIGNED_SATURATE, static_masks<C565>>;
        using T2 = static_layout<uint32_t, 1, 2, 3, 4, 5, 6, 11>;
        using FnT2 = swar_fn<OP_ADD_SIGNED_SATURATE, static_masks<T2>>;
        using FnRT = swar_fn<OP_ADD_SIGNED_SATURATE, runtime_masks<uint64_t>>;
#define SWEEP(FN, TAG)                                                   \
        run_stream<FN, 1, 256, 2, 2>(TAG, a, b, out, nvec, 0);           \
        run_stream<FN, 2, 256, 2, 2>(TAG, a, b, out, nvec, 0);           \
        run_stream<FN, 4, 256, 2, 2>(TAG, a, b, out, nvec, 0);           \
        run_stream<FN, 2, 512, 2, 2>(TAG, a, b, out, nvec, 0);           \
        run_stream<FN, 1, 1024, 2, 2>(TAG, a, b, out, nvec, 0);          \
        run_stream<FN, 2, 1024, 2, 2>(TAG, a, b, out, nvec, 0);
        SWEEP(FnU, "usat rgba8")
        SWEEP(FnS, "ssat rgba8")
        SWEEP(FnC2a, "add_ssat u64 16x4")
        SWEEP(FnC2s, "sub_ssat u64 16x4")
        SWEEP(FnI, "add_ssat irregular u64 swar")
        SWEEP(FnIF, "add_ssat irregular u64 fused")
        SWEEP(FnC0, "add_wrap u16 565")
        SWEEP(FnC0s, "add_ssat u16 565")
        SWEEP(FnT2, "add_ssat u32 type2")
        for (int cps : {2, 4, 8}) run_dyn<FnU, 2, 256>("usat rgba8", a, b, out, nvec, cps);
        for (int cps : {1, 2}) run_dyn<FnU, 2, 1024>("usat rgba8", a, b, out, nvec, cps);
        for (int cps : {2, 4}) run_dyn<FnU, 4, 512>("usat rgba8", a, b, out, nvec, cps);
        for (int cps : {2, 4, 8}) run_dyn<FnI, 2, 256>("add_ssat irregular u64 swar", a, b, out, nvec, cps);
        std::printf("done\n");
        return 0;
    }
    // --- grid / unroll / block sweep with the streaming policy ---
    for (int cps : {0, 2, 4, 8, 16, 32}) run_stream<FnU, 4, 256, 2, 2>("usat rgba8", a, b, out, nvec, cps);
    for (int cps : {0, 4, 8, 16}) run_stream<FnU, 1, 256, 2, 2>("usat rgba8", a, b, out, nvec, cps);
    for (int cps : {0, 4, 8, 16}) run_stream<FnU, 2, 256, 2, 2>("usat rgba8", a, b, out, nvec, cps);
    for (int cps : {0, 2, 4, 8}) run_stream<FnU, 8, 256, 2, 2>("usat rgba8", a, b, out, nvec, cps);
    for (int cps : {0, 2, 4, 8}) run_stream<FnU, 4, 512, 2, 2>("usat rgba8", a, b, out, nvec, cps);
    for (int cps : {0, 8, 16, 32}) run_stream<FnU, 4, 128, 2, 2>("usat rgba8", a, b, out, nvec, cps);
    for (int cps : {0, 2, 4}) run_stream<FnU, 2, 1024, 2, 2>("usat rgba8", a, b, out, nvec, cps);
    // --- cache policy sweep ---
    run_stream<FnU, 4, 256, 0, 0>("usat rgba8", a, b, out, nvec, 8);
    run_stream<FnU, 4, 256, 1, 1>("usat rgba8", a, b, out, nvec, 8);
    run_stream<FnU, 4, 256, 2, 0>("usat rgba8", a, b, out, nvec, 8);
    run_stream<FnU, 4, 256, 0, 2>("usat rgba8", a, b, out, nvec, 8);
    run_stream<FnU, 4, 256, 1, 2>("usat rgba8", a, b, out, nvec, 8);
    run_stream<FnU, 4, 256, 1, 0>("usat rgba8", a, b, out, nvec, 0);
    run_stream<FnU, 4, 256, 0, 0>("usat rgba8", a, b, out, nvec, 0);
    // --- other ops: is anything ALU-bound? ---
    run_stream<FnX, 4, 256, 2, 2>("xor  rgba8", a, b, out, nvec, 8);
    run_stream<FnS, 4, 256, 2, 2>("ssat rgba8", a, b, out, nvec, 8);
    run_stream<FnI, 4, 256, 2, 2>("ssat irregular u64 swar ", a, b, out, nvec, 8);
    run_stream<FnIF, 4, 256, 2, 2>("ssat irregular u64 fused", a, b, out, nvec, 8);
    run_stream<FnI, 2, 256, 2, 2>("ssat irregular u64 swar ", a, b, out, nvec, 8);
    run_stream<FnIF, 2, 256, 2, 2>("ssat irregular u64 fused", a, b, out, nvec, 8);

    // --- TMA bulk variants ---
    run_bulk<FnU, 4, 8192, 256>("usat rgba8", a, b, out2, nvec, 3);
    // correctness of the bulk path against the stream path
    {
        run_stream<FnU, 4, 256, 2, 2>("usat rgba8 (reference for compare)", a, b, out, nvec, 8);
        unsigned long long* bad;
        CK(cudaMallocManaged(&bad, 8));
        *bad = 0;
        const size_t covered = nvec / 512 * 512;
        compare<<<g_sms * 8, 256>>>((const uint4*)out, (const uint4*)out2, covered, bad);
        CK(cudaDeviceSynchronize());
        std::printf("bulk vs stream mismatching vectors: %llu of %zu\n", *bad, covered);
    }
    run_bulk<FnU, 3, 8192, 256>("usat rgba8", a, b, out2, nvec, 4);
    run_bulk<FnU, 6, 4096, 256>("usat rgba8", a, b, out2, nvec, 4);
    run_bulk<FnU, 4, 16384, 256>("usat rgba8", a, b, out2, nvec, 1);
    run_bulk<FnU, 3, 16384, 512>("usat rgba8", a, b, out2, nvec, 2);
    run_bulk<FnU, 6, 16384, 512>("usat rgba8", a, b, out2, nvec, 1);
    run_bulk<FnU, 8, 4096, 128>("usat rgba8", a, b, out2, nvec, 3);
    run_bulk<FnU, 4, 4096, 128>("usat rgba8", a, b, out2, nvec, 6);
    run_bulk<FnU, 2, 32768, 512>("usat rgba8", a, b, out2, nvec, 1);
    run_bulk<FnI, 4, 8192, 256>("ssat irregular u64 swar ", a, b, out2, nvec, 3);
    std::printf("done\n");
    return 0;
}
