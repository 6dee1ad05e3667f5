// tools/e2e_ceiling.cu -- what the host side of a box can feed N GPUs, and how close the host-buffer
// entry points get to it (VERDICT r1 "next" 3b/3c, 6).
//
//   build/e2e_ceiling [log2_bytes_per_device_per_direction=30] [reps=4]
//
// Part A (raw ceiling): plain cudaMemcpyAsync from/to pinned host memory on G = 1, 2, 4, 8 devices at
//   once -- H2D only, D2H only, both directions concurrently (the shape of the e2e path) -- for
//   pinned memory allocated with cudaHostAllocDefault / Portable / WriteCombined (inputs only), issued
//   from ONE host thread or from one thread per device (each thread bound to its own slice of the
//   cores, allocating and first-touching its buffers itself).
// Part B (the library): pint_cuda_sharded_binary_multi_host over G devices from one thread, and
//   pint_cuda_binary_multi_host from one thread per device, same bytes per device as part A's
//   "both" rows (RGBA8, add_unsigned_saturate + add_signed_saturate: 8 B/element up, 8 B/element down).
// One JSON line per measurement; GB/s are per direction, summed over the devices (GB = 1e9).
#include <sched.h>

#include <atomic>
#include <chrono>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <thread>
#include <vector>

#include <cuda_runtime.h>

#include "pint_cuda.h"

#define CK(x)                                                                                          \
    do {                                                                                               \
        cudaError_t e_ = (x);                                                                          \
        if (e_ != cudaSuccess) {                                                                       \
            std::fprintf(stderr, "%s:%d %s: %s\n", __FILE__, __LINE__, #x, cudaGetErrorString(e_));    \
            std::exit(2);                                                                              \
        }                                                                                              \
    } while (0)

static double now()
{
    return std::chrono::duration<double>(std::chrono::steady_clock::now().time_since_epoch()).count();
}

struct dev_buffers {
    void* h_in = nullptr;
    void* h_out = nullptr;
    void* d_in = nullptr;
    void* d_out = nullptr;
    cudaStream_t s_in = nullptr, s_out = nullptr;
};

static const char* kAllocName[] = {"default", "portable", "write_combined_inputs"};

static void alloc_dev(dev_buffers& b, int dev, size_t bytes, int alloc_kind)
{
    CK(cudaSetDevice(dev));
    const unsigned fin = alloc_kind == 1 ? cudaHostAllocPortable : alloc_kind == 2 ? cudaHostAllocWriteCombined : cudaHostAllocDefault;
    const unsigned fout = alloc_kind == 1 ? cudaHostAllocPortable : cudaHostAllocDefault;
    CK(cudaHostAlloc(&b.h_in, bytes, fin));
    CK(cudaHostAlloc(&b.h_out, bytes, fout));
    std::memset(b.h_in, 0x5A, bytes);   // first touch by the allocating thread
    std::memset(b.h_out, 0, bytes);
    CK(cudaMalloc(&b.d_in, bytes));
    CK(cudaMalloc(&b.d_out, bytes));
    CK(cudaMemset(b.d_out, 1, bytes));
    CK(cudaStreamCreateWithFlags(&b.s_in, cudaStreamNonBlocking));
    CK(cudaStreamCreateWithFlags(&b.s_out, cudaStreamNonBlocking));
}

static void free_dev(dev_buffers& b, int dev)
{
    CK(cudaSetDevice(dev));
    CK(cudaStreamDestroy(b.s_in));
    CK(cudaStreamDestroy(b.s_out));
    CK(cudaFree(b.d_in));
    CK(cudaFree(b.d_out));
    CK(cudaFreeHost(b.h_in));
    CK(cudaFreeHost(b.h_out));
    b = dev_buffers{};
}

// mode: 0 = H2D, 1 = D2H, 2 = both
static void issue(dev_buffers& b, int dev, size_t bytes, int mode, size_t chunk)
{
    CK(cudaSetDevice(dev));
    for (size_t off = 0; off < bytes; off += chunk) {
        const size_t c = bytes - off < chunk ? bytes - off : chunk;
        if (mode != 1) CK(cudaMemcpyAsync((char*)b.d_in + off, (char*)b.h_in + off, c, cudaMemcpyHostToDevice, b.s_in));
        if (mode != 0) CK(cudaMemcpyAsync((char*)b.h_out + off, (char*)b.d_out + off, c, cudaMemcpyDeviceToHost, b.s_out));
    }
}

static void drain(dev_buffers& b, int dev)
{
    CK(cudaSetDevice(dev));
    CK(cudaStreamSynchronize(b.s_in));
    CK(cudaStreamSynchronize(b.s_out));
}

static void bind_to_slice(int k, int G)
{
    const int ncpu = (int)std::thread::hardware_concurrency();
    if (ncpu <= 0) return;
    cpu_set_t set;
    CPU_ZERO(&set);
    const int lo = k * ncpu / G, hi = (k + 1) * ncpu / G;
    for (int c = lo; c < (hi > lo ? hi : lo + 1); ++c) CPU_SET(c % ncpu, &set);
    sched_setaffinity(0, sizeof(set), &set);
}

static void report_raw(int G, int alloc_kind, const char* threads, int mode, size_t bytes, int reps, double best)
{
    static const char* mode_name[] = {"h2d", "d2h", "h2d+d2h"};
    const double gbs = double(G) * double(bytes) / best / 1e9;
    std::printf("{\"part\": \"raw\", \"devices\": %d, \"pinned\": \"%s\", \"issue\": \"%s\", \"mode\": \"%s\", "
                "\"bytes_per_device_per_direction\": %zu, \"reps\": %d, \"best_s\": %.5f, \"gbs_per_direction_total\": %.1f, "
                "\"gbs_per_direction_per_device\": %.1f}\n",
                G, kAllocName[alloc_kind], threads, mode_name[mode], bytes, reps, best, gbs, gbs / G);
    std::fflush(stdout);
}

int main(int argc, char** argv)
{
    const int log2b = argc > 1 ? std::atoi(argv[1]) : 30;
    const int reps = argc > 2 ? std::atoi(argv[2]) : 4;
    const size_t bytes = size_t(1) << log2b;
    const size_t chunk = size_t(32) << 20;
    int ndev = 0;
    CK(cudaGetDeviceCount(&ndev));
    std::printf("{\"part\": \"box\", \"devices\": %d, \"host_threads\": %u}\n", ndev, std::thread::hardware_concurrency());

    // ---- part A ----
    for (int G = 1; G <= ndev; G *= 2) {
        for (int alloc_kind = 0; alloc_kind < 3; ++alloc_kind) {
            // (1) one issuing thread
            {
                std::vector<dev_buffers> bufs(G);
                for (int k = 0; k < G; ++k) alloc_dev(bufs[k], k, bytes, alloc_kind);
                for (int mode = 0; mode < 3; ++mode) {
                    double best = 1e30;
                    for (int r = 0; r <= reps; ++r) {
                        const double t0 = now();
                        for (int k = 0; k < G; ++k) issue(bufs[k], k, bytes, mode, chunk);
                        for (int k = 0; k < G; ++k) drain(bufs[k], k);
                        const double dt = now() - t0;
                        if (r > 0 && dt < best) best = dt;  // rep 0 is the warm-up
                    }
                    report_raw(G, alloc_kind, "one thread", mode, bytes, reps, best);
                }
                for (int k = 0; k < G; ++k) free_dev(bufs[k], k);
            }
            // (2) one thread per device, bound to its slice of the cores, allocating its own buffers
            if (G > 1) {
                std::vector<dev_buffers> bufs(G);
                std::vector<std::thread> th;
                for (int k = 0; k < G; ++k) th.emplace_back([&, k] { bind_to_slice(k, G); alloc_dev(bufs[k], k, bytes, alloc_kind); });
                for (auto& t : th) t.join();
                for (int mode = 0; mode < 3; ++mode) {
                    double best = 1e30;
                    for (int r = 0; r <= reps; ++r) {
                        std::atomic<int> ready{0};
                        std::atomic<bool> go{false};
                        std::vector<std::thread> w;
                        for (int k = 0; k < G; ++k)
                            w.emplace_back([&, k] {
                                bind_to_slice(k, G);
                                CK(cudaSetDevice(k));
                                ready.fetch_add(1);
                                while (!go.load(std::memory_order_acquire)) {}
                                issue(bufs[k], k, bytes, mode, chunk);
                                drain(bufs[k], k);
                            });
                        while (ready.load() < G) {}
                        const double t0 = now();
                        go.store(true, std::memory_order_release);
                        for (auto& t : w) t.join();
                        const double dt = now() - t0;
                        if (r > 0 && dt < best) best = dt;
                    }
                    report_raw(G, alloc_kind, "thread per device, bound, own first touch", mode, bytes, reps, best);
                }
                for (int k = 0; k < G; ++k) free_dev(bufs[k], k);
            }
        }
    }

    // ---- part B: the library's host-buffer entry points, same bytes per device ----
    pint_layout L{};
    L.elem_bytes = 4; L.nfields = 4; L.bits[0] = L.bits[1] = L.bits[2] = L.bits[3] = 8;
    const int ops[2] = {PINT_ADD_UNSIGNED_SATURATE, PINT_ADD_SIGNED_SATURATE};
    for (int G = 1; G <= ndev; G *= 2) {
        // per device: a and b of bytes/2 each up (= `bytes` up), two results of bytes/2 each down
        const size_t n_dev = bytes / 2 / 4, n = n_dev * G;
        void *a, *b, *o0, *o1;
        CK(cudaSetDevice(0));
        CK((cudaError_t)pint_cuda_host_alloc(&a, n * 4));
        CK((cudaError_t)pint_cuda_host_alloc(&b, n * 4));
        CK((cudaError_t)pint_cuda_host_alloc(&o0, n * 4));
        CK((cudaError_t)pint_cuda_host_alloc(&o1, n * 4));
        for (size_t i = 0; i < n; ++i) { ((uint32_t*)a)[i] = uint32_t(i * 2654435761u); ((uint32_t*)b)[i] = uint32_t(i * 40503u + 77u); }
        std::memset(o0, 0, n * 4);
        std::memset(o1, 0, n * 4);
        void* outs[2] = {o0, o1};
        int devices[64];
        for (int k = 0; k < G; ++k) devices[k] = k;
        double best = 1e30;
        for (int r = 0; r <= reps; ++r) {
            const double t0 = now();
            const int rc = pint_cuda_sharded_binary_multi_host(2, ops, &L, G, devices, a, b, outs, n);
            const double dt = now() - t0;
            if (rc) { std::fprintf(stderr, "sharded_binary_multi_host rc=%d\n", rc); return 3; }
            if (r > 0 && dt < best) best = dt;
        }
        // spot check: RGBA8 unsigned saturating add of element 12345 of the last shard
        {
            const size_t i = n - 1 - 12345 % n;
            const uint32_t x = ((uint32_t*)a)[i], y = ((uint32_t*)b)[i];
            uint32_t want = 0;
            for (int f = 0; f < 4; ++f) {
                const uint32_t s = ((x >> (8 * f)) & 255u) + ((y >> (8 * f)) & 255u);
                want |= (s > 255u ? 255u : s) << (8 * f);
            }
            if (((uint32_t*)o0)[i] != want) { std::fprintf(stderr, "sharded host result mismatch at %zu\n", i); return 4; }
        }
        std::printf("{\"part\": \"library\", \"api\": \"pint_cuda_sharded_binary_multi_host (one thread)\", \"devices\": %d, "
                    "\"elements\": %zu, \"best_s\": %.5f, \"gelem_results_s\": %.2f, \"gbs_per_direction_total\": %.1f, "
                    "\"gbs_per_direction_per_device\": %.1f}\n",
                    G, n, best, 2.0 * n / best / 1e9, 8.0 * n / best / 1e9, 8.0 * n / best / 1e9 / G);
        std::fflush(stdout);
        // one thread per device, each calling pint_cuda_binary_multi_host on its shard (per-device ring lock)
        if (G > 1) {
            size_t begin[64], count[64];
            pint_cuda_plan_shards(n, 4, G, begin, count);
            best = 1e30;
            for (int r = 0; r <= reps; ++r) {
                std::atomic<int> ready{0};
                std::atomic<bool> go{false};
                std::vector<std::thread> w;
                for (int k = 0; k < G; ++k)
                    w.emplace_back([&, k] {
                        bind_to_slice(k, G);
                        CK(cudaSetDevice(k));
                        void* so[2] = {(uint32_t*)o0 + begin[k], (uint32_t*)o1 + begin[k]};
                        ready.fetch_add(1);
                        while (!go.load(std::memory_order_acquire)) {}
                        const int rc = pint_cuda_binary_multi_host(2, ops, &L, (uint32_t*)a + begin[k], (uint32_t*)b + begin[k], so, count[k]);
                        if (rc) { std::fprintf(stderr, "binary_multi_host rc=%d on device %d\n", rc, k); std::exit(3); }
                    });
                while (ready.load() < G) {}
                const double t0 = now();
                go.store(true, std::memory_order_release);
                for (auto& t : w) t.join();
                const double dt = now() - t0;
                if (r > 0 && dt < best) best = dt;
            }
            std::printf("{\"part\": \"library\", \"api\": \"pint_cuda_binary_multi_host, one thread per device\", \"devices\": %d, "
                        "\"elements\": %zu, \"best_s\": %.5f, \"gelem_results_s\": %.2f, \"gbs_per_direction_total\": %.1f, "
                        "\"gbs_per_direction_per_device\": %.1f}\n",
                        G, n, best, 2.0 * n / best / 1e9, 8.0 * n / best / 1e9, 8.0 * n / best / 1e9 / G);
            std::fflush(stdout);
        }
        pint_cuda_host_release();
        pint_cuda_host_free(a); pint_cuda_host_free(b); pint_cuda_host_free(o0); pint_cuda_host_free(o1);
    }
    return 0;
}
