// api.cu -- the C ABI of libpint_cuda.so (include/pint_cuda.h): argument
// checking (the reference's compile-time static_asserts become return codes),
// kernel selection (compiled-in masks when the layout is in the registry,
// runtime masks otherwise) and launches.  No CPU fallback anywhere.
#include "api_common.hpp"

namespace pint_cuda_impl {

options& opts()
{
    static options o;
    return o;
}

unsigned long long g_launches = 0;

static_assert(sizeof(pint_layout) == sizeof(layout_desc), "pint_layout and layout_desc must be layout-compatible");

int check_layout(const pint_layout* l, layout_desc* d)
{
    if (!l) return PINT_E_NULL;
    if (l->nfields > PINT_MAX_FIELDS) return PINT_E_NFIELDS;
    std::memcpy(d, l, sizeof(*d));
    return pint::cuda::validate(*d);  // same codes as PINT_E_*
}

// SM count of the current device, cached per device.
int current_sms()
{
    static int sms[64] = {};
    int dev = 0;
    if (cudaGetDevice(&dev) != cudaSuccess || dev < 0 || dev >= 64) return 148;
    int v = __atomic_load_n(&sms[dev], __ATOMIC_RELAXED);
    if (!v) {
        if (cudaDeviceGetAttribute(&v, cudaDevAttrMultiProcessorCount, dev) != cudaSuccess || v <= 0) v = 148;
        __atomic_store_n(&sms[dev], v, __ATOMIC_RELAXED);
    }
    return v;
}

// Counter ring of the ordered persistent kernels: static device memory (zero at
// module load, re-armed by the kernels themselves), so no call ever allocates.
__device__ pint::cuda::kernels::tile_counter g_tile_counters[pint::cuda::kernels::kTileCounterSlots];
unsigned g_counter_cursor = 0;

pint::cuda::kernels::tile_counter* current_counters()
{
    static pint::cuda::kernels::tile_counter* addr[64] = {};
    int dev = 0;
    if (cudaGetDevice(&dev) != cudaSuccess || dev < 0 || dev >= 64) return nullptr;
    pint::cuda::kernels::tile_counter* p = __atomic_load_n(&addr[dev], __ATOMIC_ACQUIRE);
    if (!p) {
        void* sym = nullptr;
        if (cudaGetSymbolAddress(&sym, g_tile_counters) != cudaSuccess) { cudaGetLastError(); return nullptr; }
        p = static_cast<pint::cuda::kernels::tile_counter*>(sym);
        __atomic_store_n(&addr[dev], p, __ATOMIC_RELEASE);
    }
    return p;
}

launch_geom geom()
{
    launch_geom g;
    g.num_sms = current_sms();
    g.ctas_per_sm = (int)opts().ctas_per_sm;
    g.launch_counter = &g_launches;
    g.ordered_ctas_per_sm = (int)opts().ordered_ctas_per_sm;
    g.vec256 = (int)opts().vec256;
    g.unpack_bulk = (int)opts().unpack_bulk;
    g.pdl = (int)opts().pdl;
    if (g.ordered_ctas_per_sm > 0) {
        g.counters = current_counters();
        g.counter_cursor = &g_counter_cursor;
    }
    return g;
}

const layout_kernels* find_static(const layout_desc& d)
{
    if (opts().force_generic) return nullptr;
    using part_fn = const layout_kernels* (*)(int*);
    static const part_fn parts[] = {registry_part0, registry_part1, registry_part2, registry_part3};
    for (part_fn p : parts) {
        int n = 0;
        const layout_kernels* tab = p(&n);
        for (int i = 0; i < n; ++i)
            if (same_layout(tab[i].desc, d)) return &tab[i];
    }
    return nullptr;
}

// exact aliasing is allowed, partial overlap is not
bool bad_overlap(const void* in, const void* out, size_t bytes)
{
    if (in == out || !bytes) return false;
    const uintptr_t i = (uintptr_t)in, o = (uintptr_t)out;
    return i < o + bytes && o < i + bytes;
}

int binary_impl(int op, bool fused, const pint_layout* layout, const void* a, const void* b, void* out, size_t n,
                cudaStream_t s)
{
    layout_desc d;
    if (int rc = check_layout(layout, &d)) return rc;
    if (op < 0 || op >= PINT_NUM_BINARY_OPS) return PINT_E_OP;
    if (n == 0) return PINT_OK;
    if (!a || !b || !out) return PINT_E_NULL;
    const size_t bytes = n * d.elem_bytes;
    if (bad_overlap(a, out, bytes) || bad_overlap(b, out, bytes)) return PINT_E_ALIAS;
    if (((uintptr_t)a | (uintptr_t)b | (uintptr_t)out) % d.elem_bytes) return PINT_E_ARG;
    const launch_geom g = geom();
    if (const layout_kernels* k = find_static(d))
        return (int)(fused ? k->fused[op] : k->binary[op])(a, b, out, n, s, g);
    if (const long jit = opts().jit) {
        bool handled = false;
        const int rc = jit_binary(op, fused, layout, d, a, b, out, n, s, g, &handled);
        if (handled || (rc == PINT_E_JIT && jit >= 2)) return rc;
        if (rc != PINT_OK && rc != PINT_E_JIT) return rc;
    }
    return (int)generic_binary(op, fused, d, a, b, out, n, s, g);
}

int binary_multi_impl(int nops, const int* ops, const pint_layout* layout, const void* a, const void* b,
                      void* const* outs, size_t n, cudaStream_t s)
{
    layout_desc d;
    if (int rc = check_layout(layout, &d)) return rc;
    if (nops < 1 || nops > pint::cuda::kernels::kMaxMultiOps || !ops || !outs) return PINT_E_ARG;
    struct { int op[pint::cuda::kernels::kMaxMultiOps]; void* out[pint::cuda::kernels::kMaxMultiOps]; } m{};
    for (int j = 0; j < nops; ++j) {
        if (ops[j] < 0 || ops[j] >= PINT_NUM_BINARY_OPS) return PINT_E_OP;
        m.op[j] = ops[j];
        m.out[j] = outs[j];
    }
    if (n == 0) return PINT_OK;
    if (!a || !b) return PINT_E_NULL;
    const size_t bytes = n * d.elem_bytes;
    for (int j = 0; j < nops; ++j) {
        if (!outs[j]) return PINT_E_NULL;
        if ((uintptr_t)outs[j] % d.elem_bytes) return PINT_E_ARG;
        // every input vector is read once and then each output is written: outputs must not
        // overlap the inputs or each other at all
        auto overlaps = [&](const void* x) {
            const uintptr_t i = (uintptr_t)x, o = (uintptr_t)outs[j];
            return i < o + bytes && o < i + bytes;
        };
        if (overlaps(a) || overlaps(b)) return PINT_E_ALIAS;
        for (int k = 0; k < j; ++k)
            if (overlaps(outs[k])) return PINT_E_ALIAS;
    }
    if (((uintptr_t)a | (uintptr_t)b) % d.elem_bytes) return PINT_E_ARG;
    // Decompose the op list into pairs that have a one-pass kernel (kOpPairs) and single-op passes.
    const launch_geom g = geom();
    const layout_kernels* k = find_static(d);
    bool done[pint::cuda::kernels::kMaxMultiOps] = {};
    for (int j = 0; j < nops; ++j) {
        if (done[j]) continue;
        done[j] = true;
        int rc = PINT_OK;
        bool paired = false;
        for (int i = j + 1; i < nops && !paired; ++i) {
            bool swapped = false;
            const int pi = done[i] ? -1 : pint::cuda::kernels::find_op_pair(m.op[j], m.op[i], &swapped);
            if (pi < 0) continue;
            done[i] = true;
            paired = true;
            void* o0 = swapped ? m.out[i] : m.out[j];
            void* o1 = swapped ? m.out[j] : m.out[i];
            if (k) {
                rc = (int)k->pair[pi](a, b, o0, o1, n, s, g);
            } else {
                bool handled = false;
                const long jit = opts().jit;
                rc = jit ? jit_binary_pair(pi, d, a, b, o0, o1, n, s, g, &handled) : PINT_OK;
                if (!handled && !(rc == PINT_E_JIT && jit >= 2) && (rc == PINT_OK || rc == PINT_E_JIT))
                    rc = (int)generic_binary_pair(pi, d, a, b, o0, o1, n, s, g);
            }
        }
        if (!paired)
            rc = k ? (int)k->binary[m.op[j]](a, b, m.out[j], n, s, g) : binary_impl(m.op[j], false, layout, a, b, m.out[j], n, s);
        if (rc) return rc;
    }
    return PINT_OK;
}

template <class Elem>
cudaError_t unary_launch(const void* in, void* out, size_t n, uint64_t pre, uint64_t post, uint32_t amount, bool left,
                         cudaStream_t s)
{
    using W = pint::cuda::word_of<Elem>;
    pint::cuda::kernels::unary_params<W> u{W(pre), W(post), amount, left ? 1u : 0u};
    return pint::cuda::kernels::launch_unary<Elem, W>(in, out, n, u, s, geom());
}

int unary_impl(const layout_desc& d, const void* in, void* out, size_t n, uint64_t pre, uint64_t post,
               uint32_t amount, bool left, cudaStream_t s)
{
    if (n == 0) return PINT_OK;
    if (!in || !out) return PINT_E_NULL;
    if (bad_overlap(in, out, n * d.elem_bytes)) return PINT_E_ALIAS;
    if (((uintptr_t)in | (uintptr_t)out) % d.elem_bytes) return PINT_E_ARG;
    switch (d.elem_bytes) {
    case 1: return (int)unary_launch<uint8_t>(in, out, n, pre, post, amount, left, s);
    case 2: return (int)unary_launch<uint16_t>(in, out, n, pre, post, amount, left, s);
    case 4: return (int)unary_launch<uint32_t>(in, out, n, pre, post, amount, left, s);
    default: return (int)unary_launch<uint64_t>(in, out, n, pre, post, amount, left, s);
    }
}

// masks of one Word, replicated, as uint64_t (Word = uint32_t for Integers up to 4 bytes)
template <class F>
auto with_masks(const layout_desc& d, F f)
{
    if (d.elem_bytes == 8) return f(pint::cuda::make_word_masks<uint64_t>(d));
    return f(pint::cuda::make_word_masks<uint32_t>(d));
}

}  // namespace pint_cuda_impl

using namespace pint_cuda_impl;

extern "C" {

int pint_cuda_abi_version(void) { return PINT_CUDA_ABI_VERSION; }

const char* pint_cuda_error_string(int rc)
{
    switch (rc) {
    case PINT_OK: return "ok";
    case PINT_E_NULL: return "null layout or array";
    case PINT_E_ELEM_BYTES: return "Integer must be uint8_t, uint16_t, uint32_t or uint64_t";
    case PINT_E_NFIELDS: return "layout needs 1..64 fields";
    case PINT_E_ZERO_WIDTH: return "field of 0 bits";
    case PINT_E_TOO_WIDE: return "Integer won't fit given number of bits";
    case PINT_E_OP: return "unknown operation";
    case PINT_E_INDEX: return "incorrect index or slice bounds";
    case PINT_E_PLANE_BYTES: return "plane_bytes must be 4 or 8";
    case PINT_E_ALIAS: return "partially overlapping input and output";
    case PINT_E_ARG: return "invalid argument";
    case PINT_E_JIT: return "layout could not be specialised at run time (pint_cuda_jit_last_error)";
    case PINT_E_TOO_LARGE: return "more elements than this fixed-scratch entry point takes (PINT_STATIC_MAX_PAIRS): use the general entry point";
    default: return rc > 0 ? cudaGetErrorString((cudaError_t)rc) : "unknown error";
    }
}

unsigned long long pint_cuda_launch_count(void) { return __atomic_load_n(&g_launches, __ATOMIC_RELAXED); }

int pint_cuda_set_option(const char* key, long value)
{
    if (!key) return PINT_E_ARG;
    if (!std::strcmp(key, "force_generic")) { opts().force_generic = value != 0; return PINT_OK; }
    if (!std::strcmp(key, "ctas_per_sm")) { if (value < 0 || value > 64) return PINT_E_ARG; opts().ctas_per_sm = value; return PINT_OK; }
    if (!std::strcmp(key, "ordered_ctas_per_sm")) { if (value < 0 || value > 32) return PINT_E_ARG; opts().ordered_ctas_per_sm = value; return PINT_OK; }
    if (!std::strcmp(key, "host_chunk_bytes")) { if (value < 4096 || value > (1L << 30)) return PINT_E_ARG; opts().host_chunk_bytes = value; return PINT_OK; }
    if (!std::strcmp(key, "host_slots")) { if (value < 1 || value > 8) return PINT_E_ARG; opts().host_slots = value; return PINT_OK; }
    if (!std::strcmp(key, "vec256")) { if (value < 0 || value > 2) return PINT_E_ARG; opts().vec256 = value; return PINT_OK; }
    if (!std::strcmp(key, "dense_windows")) { opts().dense_windows = value != 0; return PINT_OK; }
    if (!std::strcmp(key, "dense_shift_stream")) { opts().dense_shift_stream = value != 0; return PINT_OK; }
    if (!std::strcmp(key, "pdl")) { opts().pdl = value != 0; return PINT_OK; }
    if (!std::strcmp(key, "jit")) { if (value < 0 || value > 2) return PINT_E_ARG; opts().jit = value; return PINT_OK; }
    if (!std::strcmp(key, "unpack_bulk")) { if (value < 0 || value > 1) return PINT_E_ARG; opts().unpack_bulk = value; return PINT_OK; }
    if (!std::strcmp(key, "dense_threads")) { if (value < 0 || value > 1024 || value % 32) return PINT_E_ARG; opts().dense_threads = value; return PINT_OK; }
    if (!std::strcmp(key, "dense_tile_bytes")) { if (value < 0 || value > 100 * 1024) return PINT_E_ARG; opts().dense_tile_bytes = value; return PINT_OK; }
    return PINT_E_ARG;
}

long pint_cuda_get_option(const char* key)
{
    if (!key) return -1;
    if (!std::strcmp(key, "force_generic")) return opts().force_generic;
    if (!std::strcmp(key, "ctas_per_sm")) return opts().ctas_per_sm;
    if (!std::strcmp(key, "ordered_ctas_per_sm")) return opts().ordered_ctas_per_sm;
    if (!std::strcmp(key, "host_chunk_bytes")) return opts().host_chunk_bytes;
    if (!std::strcmp(key, "host_slots")) return opts().host_slots;
    if (!std::strcmp(key, "vec256")) return opts().vec256;
    if (!std::strcmp(key, "dense_windows")) return opts().dense_windows;
    if (!std::strcmp(key, "dense_shift_stream")) return opts().dense_shift_stream;
    if (!std::strcmp(key, "unpack_bulk")) return opts().unpack_bulk;
    if (!std::strcmp(key, "jit")) return opts().jit;
    if (!std::strcmp(key, "pdl")) return opts().pdl;
    if (!std::strcmp(key, "dense_threads")) return opts().dense_threads;
    if (!std::strcmp(key, "dense_tile_bytes")) return opts().dense_tile_bytes;
    return -1;
}

int pint_cuda_layout_validate(const pint_layout* layout)
{
    layout_desc d;
    return check_layout(layout, &d);
}

uint32_t pint_cuda_storage_bytes(uint32_t total_bits) { return pint::cuda::storage_bytes(total_bits); }

int pint_cuda_layout_masks(const pint_layout* layout, uint64_t* H, uint64_t* L, uint64_t* A, uint32_t* sat_type)
{
    layout_desc d;
    if (int rc = check_layout(layout, &d)) return rc;
    // element-level view: build with a 64-bit word over a 64-bit "element" so nothing is replicated
    layout_desc e = d;
    e.elem_bytes = 8;
    const auto m = pint::cuda::make_word_masks<uint64_t>(e);
    if (H) *H = m.H;
    if (L) *L = m.L;
    if (A) *A = m.A;
    if (sat_type) *sat_type = m.ref_sat_type;
    return PINT_OK;
}

int pint_cuda_layout_is_specialized(const pint_layout* layout)
{
    layout_desc d;
    if (int rc = check_layout(layout, &d)) return rc;
    if (find_static(d)) return 1;
    return jit_is_loaded(d) ? 2 : 0;
}

int pint_cuda_slice_layout(const pint_layout* layout, uint32_t start, uint32_t end, pint_layout* result)
{
    layout_desc d;
    if (int rc = check_layout(layout, &d)) return rc;
    if (!(start < end && end <= d.nfields)) return PINT_E_INDEX;  // pint.hpp:748
    if (!result) return PINT_E_NULL;
    std::memset(result, 0, sizeof(*result));
    result->elem_bytes = d.elem_bytes;
    result->nfields = end - start;
    for (uint32_t i = start; i < end; ++i) result->bits[i - start] = d.bits[i];
    return PINT_OK;
}

int pint_cuda_binary(int op, const pint_layout* layout, const void* a, const void* b, void* out, size_t n,
                     pint_stream_t stream)
{
    return binary_impl(op, false, layout, a, b, out, n, (cudaStream_t)stream);
}

int pint_cuda_binary_multi(int nops, const int* ops, const pint_layout* layout, const void* a, const void* b,
                           void* const* outs, size_t n, pint_stream_t stream)
{
    return binary_multi_impl(nops, ops, layout, a, b, outs, n, (cudaStream_t)stream);
}

int pint_cuda_fused_unpack_arith_repack(int op, const pint_layout* layout, const void* a, const void* b, void* out,
                                        size_t n, pint_stream_t stream)
{
    return binary_impl(op, true, layout, a, b, out, n, (cudaStream_t)stream);
}

int pint_cuda_shift(int left, const pint_layout* layout, uint64_t amount, const void* in, void* out, size_t n,
                    pint_stream_t stream)
{
    layout_desc d;
    if (int rc = check_layout(layout, &d)) return rc;
    // The masks depend only on the layout and the uniform amount: build them here.
    // shift_left: keep the low (w - amount) bits of every field, then shift
    // (pint.hpp:630-668); shift_right_unsigned: drop the low min(w, amount) bits,
    // then shift (pint.hpp:670-705); amount >= max(Bits) gives 0 (the outer guard
    // at pint.hpp:1013, 1025).
    uint64_t pre = 0;
    uint32_t amt = 0;
    with_masks(d, [&](const auto& m) {
        if (amount >= m.max_bits) return 0;
        amt = (uint32_t)amount;
        uint64_t mask = 0;
        for (uint32_t k = 0; k < m.nuniq; ++k) {
            const uint64_t lo = m.ulo[k];
            const uint32_t w = m.uw[k];
            const uint32_t s = left ? (w > amt ? w - amt : 0) : (w < amt ? w : amt);
            mask |= (s >= 64 ? 0 : (lo << s)) - lo;
        }
        pre = left ? mask : ~mask;
        return 0;
    });
    return unary_impl(d, in, out, n, pre, ~0ULL, amt, left != 0, (cudaStream_t)stream);
}

int pint_cuda_slice(const pint_layout* layout, uint32_t start, uint32_t end, const void* in, void* out, size_t n,
                    pint_stream_t stream)
{
    layout_desc d;
    if (int rc = check_layout(layout, &d)) return rc;
    if (!(start < end && end <= d.nfields)) return PINT_E_INDEX;  // pint.hpp:748
    uint32_t lo = 0, mid = 0;
    for (uint32_t i = 0; i < start; ++i) lo += d.bits[i];
    for (uint32_t i = start; i < end; ++i) mid += d.bits[i];
    // (v >> lo) & ones(mid), pint.hpp:930-931, replicated over the elements of a Word
    const uint64_t e_ones = pint::cuda::detail::ones<uint64_t>(mid);
    uint64_t post = 0;
    const uint32_t ebits = 8 * d.elem_bytes, wbits = d.elem_bytes == 8 ? 64 : 32;
    for (uint32_t r = 0; r < wbits / ebits; ++r) post |= e_ones << (r * ebits);
    return unary_impl(d, in, out, n, ~0ULL, post, lo, false, (cudaStream_t)stream);
}

int pint_cuda_unpack(const pint_layout* layout, int is_signed, const void* packed, void* const* planes,
                     int plane_bytes, size_t n, pint_stream_t stream)
{
    layout_desc d;
    if (int rc = check_layout(layout, &d)) return rc;
    if (plane_bytes != 1 && plane_bytes != 2 && plane_bytes != 4 && plane_bytes != 8) return PINT_E_PLANE_BYTES;
    if (n == 0) return PINT_OK;
    if (!packed || !planes) return PINT_E_NULL;
    if ((uintptr_t)packed % d.elem_bytes) return PINT_E_ARG;
    plane_ptrs pp{};
    for (uint32_t f = 0; f < d.nfields; ++f) {
        pp.p[f] = planes[f];
        if ((uintptr_t)planes[f] % plane_bytes) return PINT_E_ARG;
    }
    const launch_geom g = geom();
    if (const layout_kernels* k = find_static(d))
        return (int)k->unpack[is_signed ? 1 : 0][plane_index(plane_bytes)](packed, pp, n, (cudaStream_t)stream, g);
    return (int)generic_unpack(d, is_signed != 0, plane_bytes, packed, pp, n, (cudaStream_t)stream, g);
}

int pint_cuda_repack(const pint_layout* layout, const void* const* planes, int plane_bytes, void* packed, size_t n,
                     pint_stream_t stream)
{
    layout_desc d;
    if (int rc = check_layout(layout, &d)) return rc;
    if (plane_bytes != 1 && plane_bytes != 2 && plane_bytes != 4 && plane_bytes != 8) return PINT_E_PLANE_BYTES;
    if (n == 0) return PINT_OK;
    if (!packed || !planes) return PINT_E_NULL;
    if ((uintptr_t)packed % d.elem_bytes) return PINT_E_ARG;
    plane_ptrs pp{};
    for (uint32_t f = 0; f < d.nfields; ++f) {
        if (!planes[f]) return PINT_E_NULL;
        if ((uintptr_t)planes[f] % plane_bytes) return PINT_E_ARG;
        pp.p[f] = const_cast<void*>(planes[f]);
    }
    const launch_geom g = geom();
    if (const layout_kernels* k = find_static(d))
        return (int)k->repack[plane_index(plane_bytes)](pp, packed, n, (cudaStream_t)stream, g);
    return (int)generic_repack(d, plane_bytes, pp, packed, n, (cudaStream_t)stream, g);
}

}  // extern "C"
