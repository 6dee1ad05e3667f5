// cuda_api_nvcc_test.cu -- include/pint/cuda.hpp compiled by NVCC: the template layer instantiates the
// kernels for static_layout<Integer, Bits...> in THIS translation unit (compiled-in mask constants for
// layouts libpint_cuda.so has none for).  For three unregistered layouts every operation is compared,
// bit for bit, with the library's C ABI on the same device arrays (run-time masks; itself parity-tested
// against the compiled reference in tests/test_gpu_parity.py) and, for the saturating adds, with a naive
// per-field model written here.  Built by `make tools` into build/cuda_api_nvcc_test.
#include <cstdio>
#include <cstdlib>
#include <vector>

#include "pint/cuda.hpp"

#if !PINT_CUDA_INLINE_KERNELS
#error "this test must be compiled by nvcc with C++17: the inline-kernel branch of cuda.hpp is what it tests"
#endif

#define CK(x) do { cudaError_t e_ = (x); if (e_ != cudaSuccess) { std::printf("CUDA error %s at line %d\n", cudaGetErrorString(e_), __LINE__); std::exit(2); } } while (0)

static uint64_t splitmix64(uint64_t x)
{
    x += 0x9E3779B97F4A7C15ULL;
    x = (x ^ (x >> 30)) * 0xBF58476D1CE4E5B9ULL;
    x = (x ^ (x >> 27)) * 0x94D049BB133111EBULL;
    return x ^ (x >> 31);
}

template <class Integer>
static Integer naive_add_sat(Integer a, Integer b, const unsigned* bits, int nf, bool is_signed)
{
    unsigned long long r = 0;
    unsigned off = 0;
    for (int f = 0; f < nf; ++f) {
        const unsigned w = bits[f];
        const unsigned long long m = w >= 64 ? ~0ULL : ((1ULL << w) - 1);
        const unsigned long long x = ((unsigned long long)a >> off) & m, y = ((unsigned long long)b >> off) & m;
        unsigned long long v;
        if (is_signed) {
            const long long sx = (long long)(x << (64 - w)) >> (64 - w), sy = (long long)(y << (64 - w)) >> (64 - w);
            long long s = sx + sy;
            const long long hi = (long long)(m >> 1), lo = -hi - 1;
            s = s > hi ? hi : (s < lo ? lo : s);
            v = (unsigned long long)s & m;
        } else {
            v = x + y > m ? m : x + y;
        }
        r |= v << off;
        off += w;
    }
    return (Integer)r;
}

template <class Integer, size_t... Bits>
static int run_layout(const char* name)
{
    using P = pint::packed_int<Integer, Bits...>;
    const unsigned bits[] = {(unsigned)Bits...};
    const int nf = (int)sizeof...(Bits);
    const pint_layout L = pint::cuda::layout<Integer, Bits...>();
    if (pint_cuda_layout_is_specialized(&L) == 1) { std::printf("%s is registered in the library: pick another layout\n", name); return 1; }
    int fails = 0;
    const size_t sizes[] = {(1u << 20) + 13, 4099, 7};
    for (size_t n : sizes) {
        std::vector<Integer> ha(n + 8), hb(n + 8), h1(n), h2(n);
        for (size_t i = 0; i < n + 8; ++i) { ha[i] = (Integer)splitmix64(3 + i); hb[i] = (Integer)splitmix64(0x2000000 + i); }
        Integer *a, *b, *o1, *o2;
        CK(cudaMalloc((void**)&a, (n + 8) * sizeof(Integer))); CK(cudaMalloc((void**)&b, (n + 8) * sizeof(Integer)));
        CK(cudaMalloc((void**)&o1, (n + 8) * sizeof(Integer))); CK(cudaMalloc((void**)&o2, (n + 8) * sizeof(Integer)));
        CK(cudaMemcpy(a, ha.data(), (n + 8) * sizeof(Integer), cudaMemcpyHostToDevice));
        CK(cudaMemcpy(b, hb.data(), (n + 8) * sizeof(Integer), cudaMemcpyHostToDevice));
        for (int off : {0, 1}) {  // off = 1: every array starts one element past a 256-byte boundary
            const P* pa = reinterpret_cast<const P*>(a + off);
            const P* pb = reinterpret_cast<const P*>(b + off);
            P* po = reinterpret_cast<P*>(o1 + off);
            auto compare = [&](int code, const char* what) {
                CK(cudaMemcpy(h1.data(), o1 + off, n * sizeof(Integer), cudaMemcpyDeviceToHost));
                const int rc = pint_cuda_binary(code, &L, a + off, b + off, o2 + off, n, nullptr);
                if (rc) { std::printf("pint_cuda_binary rc=%d\n", rc); std::exit(2); }
                CK(cudaMemcpy(h2.data(), o2 + off, n * sizeof(Integer), cudaMemcpyDeviceToHost));
                size_t bad = 0;
                for (size_t i = 0; i < n; ++i) bad += h1[i] != h2[i];
                if (code == PINT_ADD_UNSIGNED_SATURATE || code == PINT_ADD_SIGNED_SATURATE)
                    for (size_t i = 0; i < n; i += 97)
                        bad += h1[i] != naive_add_sat<Integer>(ha[i + off], hb[i + off], bits, nf, code == PINT_ADD_SIGNED_SATURATE);
                if (bad) { std::printf("FAIL %s %s n=%zu off=%d: %zu mismatches\n", name, what, n, off, bad); ++fails; }
            };
#define RUN(NAME, CODE) pint::cuda::NAME(pa, pb, po, n); compare(CODE, #NAME);
            RUN(add_wrap, PINT_ADD_WRAP) RUN(add_unsigned_saturate, PINT_ADD_UNSIGNED_SATURATE) RUN(add_signed_saturate, PINT_ADD_SIGNED_SATURATE)
            RUN(sub_wrap, PINT_SUB_WRAP) RUN(sub_unsigned_saturate, PINT_SUB_UNSIGNED_SATURATE) RUN(sub_signed_saturate, PINT_SUB_SIGNED_SATURATE)
            RUN(bit_or, PINT_OR) RUN(bit_and, PINT_AND) RUN(bit_xor, PINT_XOR)
            RUN(min_unsigned, PINT_MIN_UNSIGNED) RUN(max_unsigned, PINT_MAX_UNSIGNED) RUN(min_signed, PINT_MIN_SIGNED) RUN(max_signed, PINT_MAX_SIGNED)
#undef RUN
            pint::cuda::fused_unpack_arith_repack<PINT_ADD_SIGNED_SATURATE>(pa, pb, po, n);
            compare(PINT_ADD_SIGNED_SATURATE, "fused add_signed_saturate");
            pint::cuda::fused_unpack_arith_repack<PINT_SUB_UNSIGNED_SATURATE>(pa, pb, po, n);
            compare(PINT_SUB_UNSIGNED_SATURATE, "fused sub_unsigned_saturate");
        }
        // unpack (inline kernels) against the library's unpack, then repack(unpack(x)) == x & A
        {
            int32_t* pl[sizeof...(Bits)];
            int32_t* pl2[sizeof...(Bits)];
            void* pv[sizeof...(Bits)];
            for (int f = 0; f < nf; ++f) { CK(cudaMalloc((void**)&pl[f], n * 4)); CK(cudaMalloc((void**)&pl2[f], n * 4)); pv[f] = pl2[f]; }
            int32_t* const (&planes)[sizeof...(Bits)] = pl;
            pint::cuda::unpack(reinterpret_cast<const P*>(a), planes, n);
            if (int rc = pint_cuda_unpack(&L, 1, a, pv, 4, n, nullptr)) { std::printf("pint_cuda_unpack rc=%d\n", rc); std::exit(2); }
            std::vector<int32_t> x(n), y(n);
            for (int f = 0; f < nf; ++f) {
                CK(cudaMemcpy(x.data(), pl[f], n * 4, cudaMemcpyDeviceToHost));
                CK(cudaMemcpy(y.data(), pl2[f], n * 4, cudaMemcpyDeviceToHost));
                size_t bad = 0;
                for (size_t i = 0; i < n; ++i) bad += x[i] != y[i];
                if (bad) { std::printf("FAIL %s unpack field %d n=%zu: %zu mismatches\n", name, f, n, bad); ++fails; }
            }
            const int32_t* cpl[sizeof...(Bits)];
            for (int f = 0; f < nf; ++f) cpl[f] = pl[f];
            const int32_t* const (&cplanes)[sizeof...(Bits)] = cpl;
            pint::cuda::repack(cplanes, reinterpret_cast<P*>(o1), n);
            CK(cudaMemcpy(h1.data(), o1, n * sizeof(Integer), cudaMemcpyDeviceToHost));
            unsigned total = 0;
            for (int f = 0; f < nf; ++f) total += bits[f];
            const unsigned long long A = total >= 64 ? ~0ULL : ((1ULL << total) - 1);
            size_t bad = 0;
            for (size_t i = 0; i < n; ++i) bad += h1[i] != (Integer)(ha[i] & (Integer)A);
            if (bad) { std::printf("FAIL %s repack(unpack) n=%zu: %zu mismatches\n", name, n, bad); ++fails; }
            for (int f = 0; f < nf; ++f) { cudaFree(pl[f]); cudaFree(pl2[f]); }
        }
        cudaFree(a); cudaFree(b); cudaFree(o1); cudaFree(o2);
    }
    std::printf("%s: %d failures\n", name, fails);
    return fails;
}

int main()
{
    int ndev = 0;
    if (cudaGetDeviceCount(&ndev) != cudaSuccess || ndev == 0) { std::printf("no CUDA device\n"); return 3; }
    int fails = 0;
    fails += run_layout<uint32_t, 7, 9, 3>("packed_int<uint32_t,7,9,3>");
    fails += run_layout<uint32_t, 5, 6, 5, 3>("packed_int<uint32_t,5,6,5,3>");
    fails += run_layout<uint64_t, 10, 10, 10, 10>("packed_int<uint64_t,10,10,10,10>");
    fails += run_layout<uint16_t, 4, 4, 4>("packed_int<uint16_t,4,4,4>");
    // error behaviour of the inline path equals the ABI's
    try {
        pint::cuda::add_wrap((const pint::packed_int<uint32_t, 7, 9, 3>*)nullptr, (const pint::packed_int<uint32_t, 7, 9, 3>*)nullptr,
                             (pint::packed_int<uint32_t, 7, 9, 3>*)nullptr, 5);
        std::printf("FAIL: null arrays accepted\n");
        ++fails;
    } catch (const pint::cuda::error& e) {
        if (e.code != PINT_E_NULL) { std::printf("FAIL: wrong code %d\n", e.code); ++fails; }
    }
    std::printf("total: %d failures\n", fails);
    return fails ? 1 : 0;
}
