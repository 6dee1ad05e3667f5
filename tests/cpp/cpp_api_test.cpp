// cpp_api_test.cpp -- exercises include/pint/cuda.hpp (the pint::cuda template
// layer) the way a pint user would: device arrays of packed_int<...>, same
// function names as the scalar API.  C++11, no nvcc, no pint.hpp needed: the
// expected values come from a naive per-field model written here.
// Built by `make tools` into build/cpp_api_test; run on a GPU by tests/test_gpu_cpp_api.py.
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <vector>

#include <cuda_runtime.h>

#include "pint/cuda.hpp"

#define CK(x) do { cudaError_t e_ = (x); if (e_ != cudaSuccess) { std::printf("CUDA error %s at line %d\n", cudaGetErrorString(e_), __LINE__); return 2; } } while (0)

typedef pint::packed_int<uint32_t, 8, 8, 8, 8> rgba;
typedef pint::packed_int<uint64_t, 3, 7, 11, 13, 17, 13> irregular;
typedef pint::packed_int<uint16_t, 1, 2, 3, 4, 5> p5;

static uint64_t splitmix64(uint64_t x)
{
    x += 0x9E3779B97F4A7C15ULL;
    x = (x ^ (x >> 30)) * 0xBF58476D1CE4E5B9ULL;
    x = (x ^ (x >> 27)) * 0x94D049BB133111EBULL;
    return x ^ (x >> 31);
}

static uint32_t rgba_add_ssat(uint32_t a, uint32_t b)
{
    uint32_t r = 0;
    for (int f = 0; f < 4; ++f) {
        int x = (int8_t)(a >> (8 * f)), y = (int8_t)(b >> (8 * f)), s = x + y;
        s = s > 127 ? 127 : (s < -128 ? -128 : s);
        r |= (uint32_t)(uint8_t)s << (8 * f);
    }
    return r;
}

static uint32_t rgba_add_usat(uint32_t a, uint32_t b)
{
    uint32_t r = 0;
    for (int f = 0; f < 4; ++f) {
        unsigned s = ((a >> (8 * f)) & 255u) + ((b >> (8 * f)) & 255u);
        r |= (s > 255u ? 255u : s) << (8 * f);
    }
    return r;
}

int main()
{
    int ndev = 0;
    if (cudaGetDeviceCount(&ndev) != cudaSuccess || ndev == 0) { std::printf("no CUDA device\n"); return 3; }
    const size_t n = (1u << 20) + 13;
    int fails = 0;

    // ---- binary ops on RGBA8 ----
    std::vector<uint32_t> ha(n), hb(n), ho(n);
    for (size_t i = 0; i < n; ++i) { ha[i] = (uint32_t)splitmix64(1 + i); hb[i] = (uint32_t)splitmix64(0x1000000 + i); }
    rgba *a, *b, *out;
    CK(cudaMalloc((void**)&a, n * 4)); CK(cudaMalloc((void**)&b, n * 4)); CK(cudaMalloc((void**)&out, n * 4));
    CK(cudaMemcpy(a, ha.data(), n * 4, cudaMemcpyHostToDevice));
    CK(cudaMemcpy(b, hb.data(), n * 4, cudaMemcpyHostToDevice));
    cudaStream_t s;
    CK(cudaStreamCreate(&s));
    try {
        pint::cuda::add_signed_saturate(a, b, out, n, s);
        CK(cudaStreamSynchronize(s));
        CK(cudaMemcpy(ho.data(), out, n * 4, cudaMemcpyDeviceToHost));
        for (size_t i = 0; i < n; ++i) if (ho[i] != rgba_add_ssat(ha[i], hb[i])) { ++fails; if (fails < 5) std::printf("ssat mismatch at %zu\n", i); }
        pint::cuda::add_unsigned_saturate(a, b, out, n, s);
        CK(cudaStreamSynchronize(s));
        CK(cudaMemcpy(ho.data(), out, n * 4, cudaMemcpyDeviceToHost));
        for (size_t i = 0; i < n; ++i) if (ho[i] != rgba_add_usat(ha[i], hb[i])) { ++fails; if (fails < 5) std::printf("usat mismatch at %zu\n", i); }
        pint::cuda::fused_unpack_arith_repack(PINT_ADD_SIGNED_SATURATE, a, b, out, n, s);
        CK(cudaStreamSynchronize(s));
        CK(cudaMemcpy(ho.data(), out, n * 4, cudaMemcpyDeviceToHost));
        for (size_t i = 0; i < n; ++i) if (ho[i] != rgba_add_ssat(ha[i], hb[i])) { ++fails; if (fails < 5) std::printf("fused mismatch at %zu\n", i); }
        pint::cuda::bit_xor(a, b, out, n, s);
        CK(cudaStreamSynchronize(s));
        CK(cudaMemcpy(ho.data(), out, n * 4, cudaMemcpyDeviceToHost));
        for (size_t i = 0; i < n; ++i) if (ho[i] != (ha[i] ^ hb[i])) ++fails;

        // ---- unpack (get_signed) to int32 planes and repack on the irregular u64 layout ----
        const int widths[6] = {3, 7, 11, 13, 17, 13};
        std::vector<uint64_t> hx(n), hy(n);
        for (size_t i = 0; i < n; ++i) hx[i] = splitmix64(77 + i);
        irregular *x, *y;
        CK(cudaMalloc((void**)&x, n * 8)); CK(cudaMalloc((void**)&y, n * 8));
        CK(cudaMemcpy(x, hx.data(), n * 8, cudaMemcpyHostToDevice));
        int32_t* planes[6];
        for (int f = 0; f < 6; ++f) CK(cudaMalloc((void**)&planes[f], n * 4));
        pint::cuda::unpack(x, planes, n, s);
        const int32_t* cplanes[6];
        for (int f = 0; f < 6; ++f) cplanes[f] = planes[f];
        pint::cuda::repack(cplanes, y, n, s);
        CK(cudaStreamSynchronize(s));
        CK(cudaMemcpy(hy.data(), y, n * 8, cudaMemcpyDeviceToHost));
        for (size_t i = 0; i < n; ++i) if (hy[i] != hx[i]) { ++fails; if (fails < 5) std::printf("repack(unpack) mismatch at %zu\n", i); }
        std::vector<int32_t> hp(n);
        int off = 0;
        for (int f = 0; f < 6; ++f) {
            CK(cudaMemcpy(hp.data(), planes[f], n * 4, cudaMemcpyDeviceToHost));
            for (size_t i = 0; i < n; i += 97) {
                int64_t v = (int64_t)(hx[i] << (64 - off - widths[f])) >> (64 - widths[f]);
                if (hp[i] != (int32_t)v) { ++fails; if (fails < 5) std::printf("get_signed<%d> mismatch at %zu\n", f, i); }
            }
            off += widths[f];
        }

        // ---- slice<2,4> on <uint16_t,1,2,3,4,5> (tests/pint_test.cpp:135-146 shape) ----
        std::vector<uint16_t> hs(n), hr(n);
        for (size_t i = 0; i < n; ++i) hs[i] = (uint16_t)splitmix64(5 + i);
        p5* sv;
        pint::cuda::sliced_int<2, 4, uint16_t, 1, 2, 3, 4, 5>::type* sr;  // packed_int<uint16_t, 3, 4>
        CK(cudaMalloc((void**)&sv, n * 2)); CK(cudaMalloc((void**)&sr, n * 2));
        CK(cudaMemcpy(sv, hs.data(), n * 2, cudaMemcpyHostToDevice));
        pint::cuda::slice<2, 4>(sv, sr, n, s);
        pint::cuda::shift_left(sv, 2, sv, n, s);
        CK(cudaStreamSynchronize(s));
        CK(cudaMemcpy(hr.data(), sr, n * 2, cudaMemcpyDeviceToHost));
        for (size_t i = 0; i < n; ++i) if (hr[i] != ((hs[i] >> 3) & 0x7F)) { ++fails; if (fails < 5) std::printf("slice mismatch at %zu\n", i); }

        // ---- two ops in one pass (pint_cuda_binary_multi) ----
        {
            rgba *o0, *o1;
            CK(cudaMalloc((void**)&o0, n * 4)); CK(cudaMalloc((void**)&o1, n * 4));
            const pint_op two[2] = {PINT_ADD_UNSIGNED_SATURATE, PINT_ADD_SIGNED_SATURATE};
            rgba* const outs[2] = {o0, o1};
            pint::cuda::binary_multi(two, a, b, outs, n, s);
            CK(cudaStreamSynchronize(s));
            CK(cudaMemcpy(ho.data(), o0, n * 4, cudaMemcpyDeviceToHost));
            for (size_t i = 0; i < n; ++i) if (ho[i] != rgba_add_usat(ha[i], hb[i])) { ++fails; if (fails < 5) std::printf("multi usat mismatch at %zu\n", i); }
            CK(cudaMemcpy(ho.data(), o1, n * 4, cudaMemcpyDeviceToHost));
            for (size_t i = 0; i < n; ++i) if (ho[i] != rgba_add_ssat(ha[i], hb[i])) { ++fails; if (fails < 5) std::printf("multi ssat mismatch at %zu\n", i); }
        }

        // ---- one constant operand; interleaved std::pair operands (tests/pint_bench.cpp:22) ----
        {
            const uint32_t k = 0x7F80FF01u;
            pint::cuda::binary_scalar(PINT_ADD_SIGNED_SATURATE, a, k, out, n, s);
            CK(cudaStreamSynchronize(s));
            CK(cudaMemcpy(ho.data(), out, n * 4, cudaMemcpyDeviceToHost));
            for (size_t i = 0; i < n; ++i) if (ho[i] != rgba_add_ssat(ha[i], k)) { ++fails; if (fails < 5) std::printf("binary_scalar mismatch at %zu\n", i); }
            pint::cuda::scalar_binary(PINT_ADD_UNSIGNED_SATURATE, k, a, out, n, s);
            CK(cudaStreamSynchronize(s));
            CK(cudaMemcpy(ho.data(), out, n * 4, cudaMemcpyDeviceToHost));
            for (size_t i = 0; i < n; ++i) if (ho[i] != rgba_add_usat(k, ha[i])) { ++fails; if (fails < 5) std::printf("scalar_binary mismatch at %zu\n", i); }
            std::vector<std::pair<uint32_t, uint32_t> > hpairs(n);
            for (size_t i = 0; i < n; ++i) hpairs[i] = std::make_pair(ha[i], hb[i]);
            std::pair<uint32_t, uint32_t>* dpairs;
            CK(cudaMalloc((void**)&dpairs, n * 8));
            CK(cudaMemcpy(dpairs, hpairs.data(), n * 8, cudaMemcpyHostToDevice));
            pint::cuda::binary_pairs(PINT_ADD_SIGNED_SATURATE, dpairs, out, n, s);
            CK(cudaStreamSynchronize(s));
            CK(cudaMemcpy(ho.data(), out, n * 4, cudaMemcpyDeviceToHost));
            for (size_t i = 0; i < n; ++i) if (ho[i] != rgba_add_ssat(ha[i], hb[i])) { ++fails; if (fails < 5) std::printf("binary_pairs mismatch at %zu\n", i); }
            // the benchmark's loop as one call, and its no-allocation twin: both equal the host sum
            uint64_t want_sum = 0, got_sum[2] = {1, 2}, *dsum;
            for (size_t i = 0; i < n; ++i) want_sum += rgba_add_ssat(ha[i], hb[i]);
            CK(cudaMalloc((void**)&dsum, 16));
            pint::cuda::pairs_reduce<rgba>(PINT_ADD_SIGNED_SATURATE, dpairs, n, dsum, s);
            pint::cuda::pairs_reduce_no_alloc<rgba>(PINT_ADD_SIGNED_SATURATE, dpairs, n, dsum + 1, s);
            CK(cudaStreamSynchronize(s));
            CK(cudaMemcpy(got_sum, dsum, 16, cudaMemcpyDeviceToHost));
            if (got_sum[0] != want_sum || got_sum[1] != want_sum) { ++fails; std::printf("pairs_reduce %llu / %llu, want %llu\n", (unsigned long long)got_sum[0], (unsigned long long)got_sum[1], (unsigned long long)want_sum); }
            bool too_large = false;
            try { pint::cuda::pairs_reduce_no_alloc<rgba>(PINT_ADD_WRAP, dpairs, (size_t)pint::cuda::max_pairs_no_alloc + 1, dsum, s); }
            catch (const pint::cuda::error& e) { too_large = e.code == PINT_E_TOO_LARGE; }
            if (!too_large) { ++fails; std::printf("pairs_reduce_no_alloc beyond its maximum did not throw PINT_E_TOO_LARGE\n"); }
            CK(cudaFree(dsum));
            CK(cudaFree(dpairs));
        }

        // ---- host arrays: one device, then the same arrays split over every device of the box ----
        {
            namespace ph = pint::cuda::host;
            rgba* pa = ph::alloc<rgba>(n, PINT_HOST_PORTABLE);
            rgba* pb = ph::alloc<rgba>(n, PINT_HOST_PORTABLE);
            rgba* po[2] = {ph::alloc<rgba>(n, PINT_HOST_PORTABLE), ph::alloc<rgba>(n, PINT_HOST_PORTABLE)};
            std::memcpy(pa, ha.data(), n * 4);
            std::memcpy(pb, hb.data(), n * 4);
            const pint_op two[2] = {PINT_ADD_UNSIGNED_SATURATE, PINT_ADD_SIGNED_SATURATE};
            std::vector<int> devs(ndev);
            for (int d = 0; d < ndev; ++d) devs[d] = d;
            for (int pass = 0; pass < 2; ++pass) {
                std::memset(po[0], 0xEE, n * 4);
                std::memset(po[1], 0xEE, n * 4);
                if (pass == 0) ph::binary_multi(two, pa, pb, po, n);
                else ph::binary_multi(two, pa, pb, po, n, ndev, devs.data());
                const uint32_t *u = (const uint32_t*)po[0], *v = (const uint32_t*)po[1];
                for (size_t i = 0; i < n; ++i)
                    if (u[i] != rgba_add_usat(ha[i], hb[i]) || v[i] != rgba_add_ssat(ha[i], hb[i])) { ++fails; if (fails < 5) std::printf("host::binary_multi pass %d mismatch at %zu\n", pass, i); }
            }
            ph::binary(PINT_SUB_WRAP, pa, pb, po[0], n);
            for (size_t i = 0; i < n; ++i) {
                uint32_t want = 0;
                for (int f = 0; f < 4; ++f) want |= (((ha[i] >> (8 * f)) - (hb[i] >> (8 * f))) & 255u) << (8 * f);
                if (((const uint32_t*)po[0])[i] != want) { ++fails; if (fails < 5) std::printf("host::binary mismatch at %zu\n", i); }
            }
            ph::dealloc(pa); ph::dealloc(pb); ph::dealloc(po[0]); ph::dealloc(po[1]);
            ph::release_staging();
        }

        // ---- an array resident on device 0 split over every device, processed there and collected again ----
        {
            namespace sh = pint::cuda::sharded;
            std::vector<int> devs(ndev);
            for (int d = 0; d < ndev; ++d) devs[d] = d;
            sh::enable_peer_access(ndev, devs.data());
            std::vector<size_t> begin(ndev), count(ndev);
            sh::plan(n, 4, ndev, begin.data(), count.data());
            std::vector<rgba*> sa(ndev), sb(ndev), so(ndev);
            for (int d = 0; d < ndev; ++d) {
                CK(cudaSetDevice(d));
                CK(cudaMalloc((void**)&sa[d], count[d] * 4 + 4)); CK(cudaMalloc((void**)&sb[d], count[d] * 4 + 4)); CK(cudaMalloc((void**)&so[d], count[d] * 4 + 4));
            }
            CK(cudaSetDevice(0));
            CK(cudaStreamSynchronize(s));
            sh::scatter(a, 0, n, ndev, devs.data(), sa.data());
            sh::scatter(b, 0, n, ndev, devs.data(), sb.data());
            sh::binary(PINT_ADD_SIGNED_SATURATE, ndev, devs.data(), sa.data(), sb.data(), so.data(), count.data());
            CK(cudaMemset(out, 0, n * 4));
            CK(cudaDeviceSynchronize());
            sh::gather(out, 0, n, ndev, devs.data(), so.data());
            for (int d = 0; d < ndev; ++d) { CK(cudaSetDevice(d)); CK(cudaDeviceSynchronize()); }
            CK(cudaSetDevice(0));
            CK(cudaMemcpy(ho.data(), out, n * 4, cudaMemcpyDeviceToHost));
            for (size_t i = 0; i < n; ++i) if (ho[i] != rgba_add_ssat(ha[i], hb[i])) { ++fails; if (fails < 5) std::printf("scatter / sharded / gather mismatch at %zu\n", i); }
            for (int d = 0; d < ndev; ++d) { CK(cudaSetDevice(d)); CK(cudaFree(sa[d])); CK(cudaFree(sb[d])); CK(cudaFree(so[d])); }
            CK(cudaSetDevice(0));
        }

        // ---- constants for an unregistered layout at run time (NVRTC), then the same call again ----
        {
            typedef pint::packed_int<uint32_t, 7, 9, 3> odd;
            std::vector<uint32_t> h1(n);
            pint::cuda::add_signed_saturate((const odd*)a, (const odd*)b, (odd*)out, n, s);
            CK(cudaStreamSynchronize(s));
            CK(cudaMemcpy(ho.data(), out, n * 4, cudaMemcpyDeviceToHost));
            const int before = pint::cuda::specialization<odd>();
            bool have_nvrtc = true;
            try { pint::cuda::specialize<odd>(); } catch (const pint::cuda::error& e) { have_nvrtc = false; std::printf("specialize skipped: %s\n", e.what()); }
            if (have_nvrtc) {
                if (pint::cuda::specialization<odd>() != 2 && before != 1) { ++fails; std::printf("specialize<odd> did not load (%d)\n", pint::cuda::specialization<odd>()); }
                pint::cuda::add_signed_saturate((const odd*)a, (const odd*)b, (odd*)out, n, s);
                CK(cudaStreamSynchronize(s));
                CK(cudaMemcpy(h1.data(), out, n * 4, cudaMemcpyDeviceToHost));
                for (size_t i = 0; i < n; ++i) if (h1[i] != ho[i]) { ++fails; if (fails < 5) std::printf("specialised kernel differs at %zu\n", i); }
            }
            pint_cuda_set_option("jit", 0);
        }

        // ---- dense bit streams: RGB24 = packed_int<uint32_t,8,8,8> at 3 bytes per element ----
        {
            typedef pint::packed_int<uint32_t, 8, 8, 8> rgb;
            namespace pd = pint::cuda::dense;
            const size_t nb = pd::bytes<rgb>(n);
            if (nb != ((n * 24 + 127) / 128) * 16) { ++fails; std::printf("dense::bytes\n"); }
            unsigned char *da, *db, *dout;
            CK(cudaMalloc((void**)&da, nb)); CK(cudaMalloc((void**)&db, nb)); CK(cudaMalloc((void**)&dout, nb));
            pd::pack((const rgb*)a, da, n, s);   // the RGBA8 arrays reinterpreted: the top byte is dropped
            pd::pack((const rgb*)b, db, n, s);
            pd::add_unsigned_saturate<rgb>(da, db, dout, n, s);
            std::vector<unsigned char> hd(nb);
            CK(cudaStreamSynchronize(s));
            CK(cudaMemcpy(hd.data(), dout, nb, cudaMemcpyDeviceToHost));
            for (size_t i = 0; i < n; ++i) {
                const uint32_t want = rgba_add_usat(ha[i], hb[i]) & 0xFFFFFFu;
                const uint32_t got = hd[3 * i] | (hd[3 * i + 1] << 8) | ((uint32_t)hd[3 * i + 2] << 16);
                if (got != want) { ++fails; if (fails < 5) std::printf("dense usat mismatch at %zu\n", i); }
            }
            for (size_t i = 3 * n; i < nb; ++i) if (hd[i]) { ++fails; std::printf("dense tail not zero\n"); break; }
            int32_t* dpl[3];
            for (int f = 0; f < 3; ++f) CK(cudaMalloc((void**)&dpl[f], n * 4));
            pd::to_planes<rgb>(dout, dpl, n, s);
            const int32_t* cdpl[3] = {dpl[0], dpl[1], dpl[2]};
            pd::from_planes<rgb>(cdpl, da, n, s);
            pd::unpack(da, (rgb*)out, n, s);
            CK(cudaStreamSynchronize(s));
            CK(cudaMemcpy(ho.data(), out, n * 4, cudaMemcpyDeviceToHost));
            for (size_t i = 0; i < n; ++i) if (ho[i] != (rgba_add_usat(ha[i], hb[i]) & 0xFFFFFFu)) { ++fails; if (fails < 5) std::printf("dense planes round trip mismatch at %zu\n", i); }
            // uniform shift of every 8-bit field on the stream: (x >> 3) per byte
            pd::pack((const rgb*)a, da, n, s);
            pd::shift_right_unsigned<rgb>(da, dout, n, 3, s);
            CK(cudaStreamSynchronize(s));
            CK(cudaMemcpy(hd.data(), dout, nb, cudaMemcpyDeviceToHost));
            for (size_t i = 0; i < 3 * n; ++i) {
                const unsigned char want = (unsigned char)(((ha[i / 3] >> (8 * (i % 3))) & 0xFF) >> 3);
                if (hd[i] != want) { ++fails; if (fails < 5) std::printf("dense shift mismatch at byte %zu\n", i); }
            }
        }

        // ---- runtime errors throw ----
        bool threw = false;
        try { pint::cuda::add_wrap(a, b, (rgba*)((char*)a + 4), n - 1, s); } catch (const pint::cuda::error& e) { threw = e.code == PINT_E_ALIAS; }
        if (!threw) { ++fails; std::printf("partial overlap did not throw PINT_E_ALIAS\n"); }
    } catch (const pint::cuda::error& e) {
        std::printf("pint::cuda::error: %s\n", e.what());
        return 1;
    }
    std::printf("cpp_api_test: %d failures, %llu kernel launches\n", fails, pint_cuda_launch_count());
    return fails ? 1 : 0;
}
