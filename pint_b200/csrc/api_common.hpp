// api_common.hpp -- shared by api.cu and api_aux.cu.
#pragma once

#include <cstdint>
#include <cstring>

#include "pint_cuda.h"
#include "registry.hpp"

namespace pint_cuda_impl {

struct options {
    long force_generic = 0;
    long ctas_per_sm = 0;  // 0 = one CTA per tile (fastest, see launch_geom)
    long host_chunk_bytes = 32L << 20;  // per operand per pipeline slot
    long host_slots = 3;
    long vec256 = 1;            // 32-byte vector kernels: 0 never, 1 auto (light functors), 2 wherever aligned
    long dense_windows = 0;     // 1 = run-time dense layouts always use the element-by-element window kernels (testing)
    long dense_threads = 0;     // dense bit-stream kernels: threads per CTA (0 = from the chunk size)
    long dense_tile_bytes = 0;  // ... and tile bytes per operand (0 = 16 KB target)
    long ordered_ctas_per_sm = 0;  // ordered persistent grid for bandwidth-bound functors; 0 = off (default, see cuda_kernels_stream.cuh)
};
options& opts();

extern unsigned long long g_launches;

int check_layout(const pint_layout* l, layout_desc* d);
int current_sms();
launch_geom geom();
const layout_kernels* find_static(const layout_desc& d);
int binary_multi_impl(int nops, const int* ops, const pint_layout* layout, const void* a, const void* b,
                      void* const* outs, size_t n, cudaStream_t s);
int binary_impl(int op, bool fused, const pint_layout* layout, const void* a, const void* b, void* out, size_t n,
                cudaStream_t s);

}  // namespace pint_cuda_impl
