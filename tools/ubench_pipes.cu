// ubench_pipes.cu -- which pipe runs a right shift when it is written as a multiply-high, and at what rate?
// The mixed-width saturating ops are bound by the ALU pipe (LOP3 / SHF / LEA: 64 lanes per clock and SM); the FMA
// pipe (IMAD) idles next to it.  x >> k == umulhi(x, 2^(32-k)): if IMAD.HI.U32 issues at the IMAD rate, the shift
// plans' SHF instructions can move over.  Each variant runs 8 independent chains per thread, 256 steps per loop pass.
//   build: nvcc -O3 -gencode arch=compute_100a,code=sm_100a -o build/ubench_pipes tools/ubench_pipes.cu
#include <cstdint>
#include <cstdio>
#include <cstdlib>

#include <cuda_runtime.h>

#define CK(x) do { cudaError_t e_ = (x); if (e_ != cudaSuccess) { std::printf("CUDA error %s at line %d\n", cudaGetErrorString(e_), __LINE__); std::exit(2); } } while (0)

__constant__ uint32_t kMul[4] = {1u << 29, 1u << 27, 1u << 30, 1u << 28};  // >> 3, >> 5, >> 2, >> 4 as multiply-high

__device__ __forceinline__ uint32_t mulhi_c(uint32_t x, uint32_t c)
{
    uint32_t r;
    asm("mul.hi.u32 %0, %1, %2;" : "=r"(r) : "r"(x), "r"(c));
    return r;
}

// MODE 0: SHF + LOP3; 1: IMAD.HI + LOP3; 2: every other shift as IMAD.HI; 3: LOP3 only (same count as 0 without the shifts);
// 4 / 5: a plain IMAD / FFMA in the shift's place (is the FMA pipe's issue slot independent of the ALU pipe's at all?)
template <int MODE>
__global__ void __launch_bounds__(256) chains(uint32_t* out, uint32_t seed, int iters)
{
    uint32_t x[8];
#pragma unroll
    for (int j = 0; j < 8; ++j) x[j] = seed * (threadIdx.x + 1) + j * 0x9E3779B9u + blockIdx.x;
    const uint32_t m0 = 0x300618c8u ^ seed, m1 = 0x0f018400u ^ seed;
    const uint32_t c0 = kMul[0], c1 = kMul[1];
    for (int i = 0; i < iters; ++i) {
#pragma unroll
        for (int r = 0; r < 16; ++r) {
#pragma unroll
            for (int j = 0; j < 8; ++j) {
                uint32_t s;
                if (MODE == 0) s = x[j] >> 3;
                else if (MODE == 1) s = mulhi_c(x[j], c0);
                else if (MODE == 2) s = (j & 1) ? mulhi_c(x[j], c0) : (x[j] >> 3);
                else if (MODE == 4) s = x[j] * 5u + m1;          // IMAD
                else if (MODE == 5) s = __float_as_uint(__uint_as_float(x[j]) * 1.0001f + 0.5f);  // FFMA
                else s = x[j] ^ m1;
                x[j] = (s & m0) ^ x[j] ^ (r * 0x01010101u);
                (void)c1;
            }
        }
    }
    uint32_t acc = 0;
#pragma unroll
    for (int j = 0; j < 8; ++j) acc ^= x[j];
    if (acc == 0x12345678u) out[0] = acc;  // keeps the chains alive
}

template <int MODE>
static void run(const char* name, uint32_t* out, int sms)
{
    const int iters = 2048, grid = sms * 8;
    cudaEvent_t e0, e1;
    CK(cudaEventCreate(&e0)); CK(cudaEventCreate(&e1));
    chains<MODE><<<grid, 256>>>(out, 7, iters);
    CK(cudaDeviceSynchronize());
    CK(cudaEventRecord(e0));
    chains<MODE><<<grid, 256>>>(out, 7, iters);
    CK(cudaEventRecord(e1));
    CK(cudaEventSynchronize(e1));
    float ms = 0;
    CK(cudaEventElapsedTime(&ms, e0, e1));
    const double steps = double(grid) * 256 * iters * 16 * 8;  // one (shift-or-mulhi, LOP3) pair each
    int clk = 0;
    CK(cudaDeviceGetAttribute(&clk, cudaDevAttrClockRate, 0));
    std::printf("{\"variant\": \"%s\", \"ms\": %.3f, \"steps_per_clk_per_sm\": %.1f}\n", name, ms, steps / (ms * 1e-3) / (double(clk) * 1e3) / sms);
}

int main()
{
    int sms = 0;
    CK(cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, 0));
    uint32_t* out;
    CK(cudaMalloc((void**)&out, 4));
    run<0>("shf+lop3", out, sms);
    run<1>("imad.hi+lop3", out, sms);
    run<2>("half shf, half imad.hi (+lop3)", out, sms);
    run<3>("lop3+lop3", out, sms);
    run<4>("imad+lop3", out, sms);
    run<5>("ffma+lop3", out, sms);
    return 0;
}
