// api_common.hpp -- shared by api.cu and api_aux.cu.
#pragma once

#include <cstdint>
#include <cstring>
#include <string>

#include "pint_cuda.h"
#include "registry.hpp"

namespace pint { namespace cuda { namespace kernels { namespace dense { struct dense_tuning; } } } }

namespace pint_cuda_impl {

// process-wide option: relaxed atomic reads and writes (pint_cuda_set_option may race with launches)
struct option_value {
    long v;
    constexpr option_value(long x = 0) : v(x) {}
    operator long() const { return __atomic_load_n(&v, __ATOMIC_RELAXED); }
    option_value& operator=(long x) { __atomic_store_n(&v, x, __ATOMIC_RELAXED); return *this; }
};

struct options {
    option_value force_generic = 0;
    option_value ctas_per_sm = 0;  // 0 = one CTA per tile (fastest, see launch_geom)
    option_value host_chunk_bytes = 32L << 20;  // per operand per pipeline slot
    option_value host_slots = 3;
    option_value vec256 = 1;            // 32-byte vector kernels: 0 never, 1 auto (light functors), 2 wherever aligned
    option_value dense_windows = 0;     // 1 = run-time dense layouts always use the element-by-element window kernels (testing)
    option_value dense_threads = 0;     // dense bit-stream kernels: threads per CTA (0 = from the chunk size)
    option_value dense_tile_bytes = 0;  // ... and tile bytes per operand (0 = 16 KB target)
    option_value dense_shift_stream = 1;  // dense shifts by less than 32 bits: the word-stream kernel (cuda_kernels_dense_shift.cuh); 0 = whole-chunk kernels
    option_value unpack_bulk = 1;   // unpack to int32 planes: 1 = plane tiles staged in shared memory, bulk (TMA) stores; 0 = register kernels
    option_value pdl = 1;           // packed -> packed stream kernels launched with programmatic stream serialization (default; +4...14 % on arrays up to 2^26 elements, profiles/r19_ab_pdl.log)
    option_value jit = 0;           // 0 off; 1 = specialise unregistered layouts with NVRTC, fall back on failure; 2 = fail instead
    option_value ordered_ctas_per_sm = 0;  // ordered persistent grid for bandwidth-bound functors; 0 = off (default, see cuda_kernels_stream.cuh)
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


// jit.cu: the kernel templates specialised at run time for a layout without pre-built constants.
// *handled = false means the call was not (or could not be) served and nothing was launched.
int jit_binary(int op, bool fused, const pint_layout* layout, const layout_desc& d, const void* a, const void* b, void* out,
               size_t n, cudaStream_t s, const launch_geom& g, bool* handled);
int jit_binary_pair(int pair, const layout_desc& d, const void* a, const void* b, void* out0, void* out1, size_t n,
                    cudaStream_t s, const launch_geom& g, bool* handled);
// dense bit streams (whole-chunk layouts): op >= 0 one op, pair >= 0 one of kOpPairs with both results
int jit_dense(int op, int pair, const layout_desc& d, const void* a, const void* b, void* out, void* out1, size_t n,
              cudaStream_t s, const launch_geom& g, const pint::cuda::kernels::dense::dense_tuning& tune, bool* handled);
// caller shapes: only the vector body is launched; *head elements before it and n - *head - *body after it are the caller's
int jit_scalar_body(int op, bool first, const layout_desc& d, const void* a, uint64_t scalar, void* out, size_t n, cudaStream_t s,
                    const launch_geom& g, size_t* head, size_t* body, bool* handled);
int jit_pairs_body(int op, const layout_desc& d, const void* pairs, void* out, size_t n, cudaStream_t s, const launch_geom& g,
                   size_t* head, size_t* body, bool* handled);
int jit_compile_all(const layout_desc& d);
int jit_prebuild(const layout_desc& d, int major, int minor);
bool jit_is_loaded(const layout_desc& d);

}  // namespace pint_cuda_impl
