// api_dense.cu -- C ABI of the dense bit-stream storage format (include/pint_cuda.h,
// "dense" section): argument checking, kernel selection, launches.
#include "api_common.hpp"
#include "dense_registry.hpp"
#include "pint/cuda_kernels_dense_convert.cuh"
#include "pint/cuda_kernels_dense_bytes.cuh"
#include "pint/cuda_kernels_dense_shift.cuh"

namespace pint_cuda_impl {
namespace {

using namespace pint::cuda;
using namespace pint::cuda::kernels;

uint32_t total_bits(const layout_desc& d)
{
    uint32_t s = 0;
    for (uint32_t i = 0; i < d.nfields; ++i) s += d.bits[i];
    return s;
}

size_t dense_bytes(uint32_t S, size_t n)
{
    const unsigned __int128 bits = (unsigned __int128)n * S;
    return size_t((bits + 127) / 128) * 16;
}

// The dense stream's elements as packed_int<make_packed_int's Integer, Bits...>: narrow
// elements are processed several to a word, one per Integer-sized slot (replicated masks).
layout_desc slot_desc(const layout_desc& d, uint32_t S)
{
    layout_desc e = d;
    e.elem_bytes = storage_bytes(S);
    return e;
}

dense_tuning tuning()
{
    dense_tuning t;
    t.threads = (int)opts().dense_threads;
    t.tile_bytes = (int)opts().dense_tile_bytes;
    return t;
}

template <class W, int... Ops>
cudaError_t dense_generic(int op, const layout_desc& d, uint32_t S, const void* a, const void* b, void* out, size_t n,
                          cudaStream_t s, const launch_geom& g, std::integer_sequence<int, Ops...>)
{
    const layout_desc e = slot_desc(d, S);
    const runtime_masks<W> p(make_word_masks<W>(e));
    cudaError_t rc = cudaErrorInvalidValue;
    if constexpr (sizeof(W) == 8)  // 33..63-bit elements: 64-bit functor over a tile of 32-bit words
        ((op == Ops ? (rc = dense::launch_dense_binary(a, b, out, n, swar_fn<Ops, runtime_masks<W>>{p}, dense::wide_geom(S), s, g,
                                                       tuning()), 0) : 0), ...);
    else
        ((op == Ops ? (rc = dense::launch_dense_binary(a, b, out, n, swar_fn<Ops, runtime_masks<W>>{p},
                                                       dense::runtime_geom<W>(S, 8 * e.elem_bytes), s, g, tuning()), 0) : 0), ...);
    return rc;
}

dense::dense_fields make_fields(const layout_desc& d)
{
    dense::dense_fields f{};
    layout_desc e = d;
    e.elem_bytes = 8;  // offsets within one element, nothing replicated
    f.t = make_field_table<uint64_t>(e);
    return f;
}

// Conversions INTO a dense stream: the kernels with S compiled in (dense_fixed.cu) take the whole
// tiles when the inputs are 16-byte aligned and a kernel exists for (S, Integer); *done = elements
// converted, the rest is left to the run-time-geometry kernel.  Option dense_windows = 1 turns them off.
cudaError_t insert_fixed(uint32_t S, uint32_t elem_bytes, int plane_bytes, const void* other, void* dn, size_t n,
                         const dense::dense_fields& f, cudaStream_t s, size_t* done)
{
    *done = 0;
    if (opts().dense_windows) return cudaSuccess;
    uintptr_t align = reinterpret_cast<uintptr_t>(other);
    if (!elem_bytes)
        for (uint32_t i = 0; i < f.t.nfields; ++i) align |= reinterpret_cast<uintptr_t>(f.planes.p[i]);
    if (align % 16) return cudaSuccess;
    const launch_geom g = geom();
    cudaError_t rc = cudaSuccess;
    if (dense_insert_fixed_part0(S, elem_bytes, plane_bytes, other, dn, n, f, s, g, done, &rc) ||
        dense_insert_fixed_part1(S, elem_bytes, plane_bytes, other, dn, n, f, s, g, done, &rc) ||
        dense_insert_fixed_part2(S, elem_bytes, plane_bytes, other, dn, n, f, s, g, done, &rc) ||
        dense_insert_fixed_part3(S, elem_bytes, plane_bytes, other, dn, n, f, s, g, done, &rc))
        return rc;
    return cudaSuccess;
}

// dense -> padded with S compiled in: uint8_t / uint16_t / uint64_t elements (dense_fixed.cu says why)
cudaError_t extract_fixed(uint32_t S, uint32_t elem_bytes, const void* dn, void* other, size_t n, cudaStream_t s, size_t* done,
                          int plane_bytes = 0, bool is_signed = false, const dense::dense_fields* planes = nullptr)
{
    *done = 0;
    if (opts().dense_windows) return cudaSuccess;
    uintptr_t align = reinterpret_cast<uintptr_t>(other);
    const dense::dense_fields none{};
    const dense::dense_fields& f = planes ? *planes : none;
    if (planes)
        for (uint32_t i = 0; i < f.t.nfields; ++i) align |= reinterpret_cast<uintptr_t>(f.planes.p[i]);
    if (align % 16) return cudaSuccess;
    const launch_geom g = geom();
    cudaError_t rc = cudaSuccess;
    if (dense_extract_fixed_part0(S, elem_bytes, plane_bytes, is_signed, dn, other, n, f, s, g, done, &rc) ||
        dense_extract_fixed_part1(S, elem_bytes, plane_bytes, is_signed, dn, other, n, f, s, g, done, &rc) ||
        dense_extract_fixed_part2(S, elem_bytes, plane_bytes, is_signed, dn, other, n, f, s, g, done, &rc) ||
        dense_extract_fixed_part3(S, elem_bytes, plane_bytes, is_signed, dn, other, n, f, s, g, done, &rc))
        return rc;
    return cudaSuccess;
}

template <class W, class Elem>
cudaError_t pack_unpack(bool to_dense, const void* padded, void* dense, size_t n, uint32_t S, cudaStream_t s)
{
    const dense::dense_fields f{};
    if (to_dense) return dense::launch_dense_insert<W, Elem, dense::SIDE_PADDED, 4>(padded, dense, n, S, f, s, geom());
    return dense::launch_dense_extract<W, Elem, dense::SIDE_PADDED, 4, false>(dense, const_cast<void*>(padded), n, S, f, s, geom());
}

template <class W>
cudaError_t pack_unpack_w(bool to_dense, uint32_t elem_bytes, const void* padded, void* dense, size_t n, uint32_t S,
                          cudaStream_t s)
{
    switch (elem_bytes) {
    case 1: return pack_unpack<W, uint8_t>(to_dense, padded, dense, n, S, s);
    case 2: return pack_unpack<W, uint16_t>(to_dense, padded, dense, n, S, s);
    case 4: return pack_unpack<W, uint32_t>(to_dense, padded, dense, n, S, s);
    default: return pack_unpack<W, uint64_t>(to_dense, padded, dense, n, S, s);
    }
}

int convert_impl(bool to_dense, const pint_layout* layout, const void* padded, void* dense, size_t n, cudaStream_t s)
{
    layout_desc d;
    if (int rc = check_layout(layout, &d)) return rc;
    if (n == 0) return PINT_OK;
    if (!padded || !dense) return PINT_E_NULL;
    if ((uintptr_t)dense % 16 || (uintptr_t)padded % d.elem_bytes) return PINT_E_ARG;
    const uint32_t S = total_bits(d);
    const uintptr_t p0 = (uintptr_t)padded, d0 = (uintptr_t)dense;
    if (p0 < d0 + dense_bytes(S, n) && d0 < p0 + n * d.elem_bytes) return PINT_E_ALIAS;  // no in-place conversion
    if (to_dense) {
        size_t done = 0;
        if (cudaError_t rc = insert_fixed(S, d.elem_bytes, 0, padded, dense, n, dense::dense_fields{}, s, &done)) return (int)rc;
        if (done == n) return PINT_OK;
        padded = static_cast<const char*>(padded) + done * d.elem_bytes;  // done is a multiple of 128 elements
        dense = static_cast<char*>(dense) + done / 8 * S;
        n -= done;
    } else {
        size_t done = 0;
        if (cudaError_t rc = extract_fixed(S, d.elem_bytes, dense, const_cast<void*>(padded), n, s, &done)) return (int)rc;
        if (done == n) return PINT_OK;
        padded = static_cast<const char*>(padded) + done * d.elem_bytes;
        dense = static_cast<char*>(dense) + done / 8 * S;
        n -= done;
    }
    return (int)(S <= 32 ? pack_unpack_w<uint32_t>(to_dense, d.elem_bytes, padded, dense, n, S, s)
                         : pack_unpack_w<uint64_t>(to_dense, d.elem_bytes, padded, dense, n, S, s));
}

template <class W>
cudaError_t planes_launch(bool to_dense, bool is_signed, int pb, const void* dense, const dense::dense_fields& f, size_t n,
                          uint32_t S, cudaStream_t s)
{
    const launch_geom g = geom();
    void* dn = const_cast<void*>(dense);
    if (to_dense) {
        switch (pb) {
        case 1: return dense::launch_dense_insert<W, W, dense::SIDE_PLANES, 1>(nullptr, dn, n, S, f, s, g);
        case 2: return dense::launch_dense_insert<W, W, dense::SIDE_PLANES, 2>(nullptr, dn, n, S, f, s, g);
        case 4: return dense::launch_dense_insert<W, W, dense::SIDE_PLANES, 4>(nullptr, dn, n, S, f, s, g);
        default: return dense::launch_dense_insert<W, W, dense::SIDE_PLANES, 8>(nullptr, dn, n, S, f, s, g);
        }
    }
    if (is_signed) {
        switch (pb) {
        case 1: return dense::launch_dense_extract<W, W, dense::SIDE_PLANES, 1, true>(dense, nullptr, n, S, f, s, g);
        case 2: return dense::launch_dense_extract<W, W, dense::SIDE_PLANES, 2, true>(dense, nullptr, n, S, f, s, g);
        case 4: return dense::launch_dense_extract<W, W, dense::SIDE_PLANES, 4, true>(dense, nullptr, n, S, f, s, g);
        default: return dense::launch_dense_extract<W, W, dense::SIDE_PLANES, 8, true>(dense, nullptr, n, S, f, s, g);
        }
    }
    switch (pb) {
    case 1: return dense::launch_dense_extract<W, W, dense::SIDE_PLANES, 1, false>(dense, nullptr, n, S, f, s, g);
    case 2: return dense::launch_dense_extract<W, W, dense::SIDE_PLANES, 2, false>(dense, nullptr, n, S, f, s, g);
    case 4: return dense::launch_dense_extract<W, W, dense::SIDE_PLANES, 4, false>(dense, nullptr, n, S, f, s, g);
    default: return dense::launch_dense_extract<W, W, dense::SIDE_PLANES, 8, false>(dense, nullptr, n, S, f, s, g);
    }
}

int planes_impl(bool to_dense, const pint_layout* layout, int is_signed, const void* dense, const void* const* planes,
                int plane_bytes, size_t n, cudaStream_t s)
{
    layout_desc d;
    if (int rc = check_layout(layout, &d)) return rc;
    if (plane_bytes != 1 && plane_bytes != 2 && plane_bytes != 4 && plane_bytes != 8) return PINT_E_PLANE_BYTES;
    if (n == 0) return PINT_OK;
    if (!dense || !planes) return PINT_E_NULL;
    if ((uintptr_t)dense % 16) return PINT_E_ARG;
    dense::dense_fields f = make_fields(d);
    for (uint32_t i = 0; i < d.nfields; ++i) {
        if (!planes[i] && to_dense) return PINT_E_NULL;
        if ((uintptr_t)planes[i] % plane_bytes) return PINT_E_ARG;
        f.planes.p[i] = const_cast<void*>(planes[i]);
    }
    const uint32_t S = total_bits(d);
    if (S == 8 || S == 16 || S == 32 || S == 64) {
        // no padding to remove: the stream IS the padded array of make_packed_int's Integer, and the padded
        // unpack / repack kernels (group form for narrow planes) serve it; the stream's tail past n stays zero
        pint_layout p = *layout;
        p.elem_bytes = S / 8;
        if (!to_dense) return pint_cuda_unpack(&p, is_signed, dense, const_cast<void* const*>(planes), plane_bytes, n, (pint_stream_t)s);
        if (int rc = pint_cuda_repack(&p, planes, plane_bytes, const_cast<void*>(dense), n, (pint_stream_t)s)) return rc;
        const size_t used = n * p.elem_bytes, total = dense_bytes(S, n);
        if (total > used) return (int)cudaMemsetAsync(static_cast<char*>(const_cast<void*>(dense)) + used, 0, total - used, s);
        return PINT_OK;
    }
    // every field exactly one plane value wide (RGB24 <-> three uint8_t planes, RGB48 <-> three uint16_t
    // planes ...): interleaved <-> planar is a byte permutation (cuda_kernels_dense_bytes.cuh)
    {
        bool bytes_only = plane_bytes <= 2 && !opts().dense_windows;
        uintptr_t align = 0;
        for (uint32_t i = 0; i < d.nfields; ++i) {
            bytes_only = bytes_only && d.bits[i] == 8 * plane_bytes;
            align |= reinterpret_cast<uintptr_t>(planes[i]);
        }
        if (bytes_only && align % 16 == 0) {
            size_t done = 0;
            cudaError_t rc = cudaSuccess;
            if (dense::dense_bytes_dispatch(d.nfields, plane_bytes, to_dense, f.planes, const_cast<void*>(dense), n, s, geom(), &done, &rc)) {
                if (rc != cudaSuccess) return (int)rc;
                if (done == n) return PINT_OK;
                for (uint32_t i = 0; i < d.nfields; ++i)
                    if (f.planes.p[i]) f.planes.p[i] = static_cast<char*>(f.planes.p[i]) + done * plane_bytes;
                dense = static_cast<const char*>(dense) + done / 8 * S;
                n -= done;
                planes = nullptr;  // (f.planes carries the advanced pointers from here on)
            }
        }
    }
    // kernels with S compiled in take the whole tiles: every plane width into a stream, 1- and 2-byte planes
    // out of one (for 4- and 8-byte planes the run-time-geometry kernel is as fast, profiles/r03m_ab_dense_convert.log)
    if (to_dense || plane_bytes <= 2) {
        size_t done = 0;
        bool all = true;
        for (uint32_t i = 0; i < d.nfields; ++i) all = all && f.planes.p[i];
        if (to_dense) {
            if (cudaError_t rc = insert_fixed(S, 0, plane_bytes, nullptr, const_cast<void*>(dense), n, f, s, &done)) return (int)rc;
        } else if (all) {
            if (cudaError_t rc = extract_fixed(S, 0, dense, nullptr, n, s, &done, plane_bytes, is_signed != 0, &f)) return (int)rc;
        }
        if (done == n) return PINT_OK;
        for (uint32_t i = 0; i < d.nfields; ++i)
            if (f.planes.p[i]) f.planes.p[i] = static_cast<char*>(f.planes.p[i]) + done * plane_bytes;
        dense = static_cast<const char*>(dense) + done / 8 * S;
        n -= done;
    }
    return (int)(S <= 32 ? planes_launch<uint32_t>(to_dense, is_signed != 0, plane_bytes, dense, f, n, S, s)
                         : planes_launch<uint64_t>(to_dense, is_signed != 0, plane_bytes, dense, f, n, S, s));
}

}  // namespace
}  // namespace pint_cuda_impl

using namespace pint_cuda_impl;

extern "C" {

size_t pint_cuda_dense_bytes(const pint_layout* layout, size_t n)
{
    layout_desc d;
    if (check_layout(layout, &d)) return 0;
    return dense_bytes(total_bits(d), n);
}

int pint_cuda_dense_binary_multi(int nops, const int* ops, const pint_layout* layout, const void* a, const void* b,
                                 void* const* outs, size_t n, pint_stream_t stream)
{
    layout_desc d;
    if (int rc = check_layout(layout, &d)) return rc;
    if (nops < 1 || nops > PINT_MAX_MULTI_OPS || !ops || !outs) return PINT_E_ARG;
    for (int j = 0; j < nops; ++j)
        if (ops[j] < 0 || ops[j] >= PINT_NUM_BINARY_OPS) return PINT_E_OP;
    if (n == 0) return PINT_OK;
    if (!a || !b) return PINT_E_NULL;
    const uint32_t S = total_bits(d);
    const size_t bytes = dense_bytes(S, n);
    if (((uintptr_t)a | (uintptr_t)b) % 16) return PINT_E_ARG;
    for (int j = 0; j < nops; ++j) {
        if (!outs[j]) return PINT_E_NULL;
        if ((uintptr_t)outs[j] % 16) return PINT_E_ARG;
        auto overlaps = [&](const void* x) {
            const uintptr_t i = (uintptr_t)x, o = (uintptr_t)outs[j];
            return i < o + bytes && o < i + bytes;
        };
        if (overlaps(a) || overlaps(b)) return PINT_E_ALIAS;
        for (int k = 0; k < j; ++k)
            if (overlaps(outs[k])) return PINT_E_ALIAS;
    }
    cudaStream_t s = (cudaStream_t)stream;
    const bool padded_format = S == 8 || S == 16 || S == 32 || S == 64;
    if (padded_format) {
        // no padding to remove: the streams ARE padded arrays of make_packed_int's Integer, and the padded multi-op
        // entry (one-pass pair kernels) serves them; every output's tail past n stays zero
        pint_layout p = *layout;
        p.elem_bytes = S / 8;
        if (int rc = binary_multi_impl(nops, ops, &p, a, b, outs, n, s)) return rc;
        const size_t used = n * p.elem_bytes;
        for (int j = 0; j < nops && bytes > used; ++j)
            if (cudaError_t e = cudaMemsetAsync(static_cast<char*>(outs[j]) + used, 0, bytes - used, s)) return (int)e;
        return PINT_OK;
    }
    const launch_geom g = geom();
    const dense_layout_kernels* k = opts().force_generic ? nullptr : find_dense_static(d);
    bool done[PINT_MAX_MULTI_OPS] = {};
    for (int j = 0; j < nops; ++j) {
        if (done[j]) continue;
        done[j] = true;
        int rc = -1;
        for (int i = j + 1; i < nops && rc < 0 && !padded_format; ++i) {
            bool swapped = false;
            const int pi = done[i] ? -1 : find_op_pair(ops[j], ops[i], &swapped);
            if (pi < 0) continue;
            done[i] = true;
            void* o0 = swapped ? outs[i] : outs[j];
            void* o1 = swapped ? outs[j] : outs[i];
            const int N = int(S / dense::gcd_u32(S, 32));
            cudaError_t mrc = cudaSuccess;
            // one pass for whole-chunk kernels only; window kernels (chunks above 19 words, or the
            // dense_windows option) run the two ops as two launches, which is faster there
            bool jit_done = false;
            if (!k && opts().jit && !opts().dense_windows) {
                const int jrc = jit_dense(-1, pi, d, a, b, o0, o1, n, s, g, tuning(), &jit_done);
                if (jit_done) rc = jrc;
                else if (jrc == PINT_E_JIT && opts().jit >= 2) return jrc;
                else if (jrc != PINT_OK && jrc != PINT_E_JIT) return jrc;
            }
            if (jit_done)
                ;
            else if (k && k->pair[pi])
                rc = (int)k->pair[pi](a, b, o0, o1, n, s, g, tuning());
            else if (!k && !opts().dense_windows &&
                     (dense_mw_pair_part0(N, pi, d, S, a, b, o0, o1, n, s, g, tuning(), &mrc) ||
                      dense_mw_pair_part1(N, pi, d, S, a, b, o0, o1, n, s, g, tuning(), &mrc) ||
                      dense_mw_pair_part2(N, pi, d, S, a, b, o0, o1, n, s, g, tuning(), &mrc)))
                rc = (int)mrc;
            else
                done[i] = false;  // not paired after all
        }
        if (rc < 0) rc = pint_cuda_dense_binary(ops[j], layout, a, b, outs[j], n, stream);
        if (rc) return rc;
    }
    return PINT_OK;
}

int pint_cuda_dense_binary(int op, const pint_layout* layout, const void* a, const void* b, void* out, size_t n,
                           pint_stream_t stream)
{
    layout_desc d;
    if (int rc = check_layout(layout, &d)) return rc;
    if (op < 0 || op >= PINT_NUM_BINARY_OPS) return PINT_E_OP;
    if (n == 0) return PINT_OK;
    if (!a || !b || !out) return PINT_E_NULL;
    if (((uintptr_t)a | (uintptr_t)b | (uintptr_t)out) % 16) return PINT_E_ARG;
    const uint32_t S = total_bits(d);
    const size_t bytes = dense_bytes(S, n);
    auto partial = [&](const void* in) {
        const uintptr_t i = (uintptr_t)in, o = (uintptr_t)out;
        return in != out && i < o + bytes && o < i + bytes;
    };
    if (partial(a) || partial(b)) return PINT_E_ALIAS;
    cudaStream_t s = (cudaStream_t)stream;
    if (S == 8 || S == 16 || S == 32 || S == 64) {
        // no padding to remove: the dense stream IS the padded array of make_packed_int's Integer;
        // run the vector kernels over the whole buffer (the zero tail stays zero: op(0, 0) == 0)
        pint_layout p = *layout;
        p.elem_bytes = S / 8;
        const size_t total = bytes / p.elem_bytes;
        // inputs may carry garbage above n*S in the last vector: the padded kernels would keep it
        // for | & ^; only whole elements exist here (S divides 128), so process n and zero the rest
        if (int rc = binary_impl(op, false, &p, a, b, out, n, s)) return rc;
        if (total > n) return (int)cudaMemsetAsync((char*)out + n * p.elem_bytes, 0, (total - n) * p.elem_bytes, s);
        return PINT_OK;
    }
    const launch_geom g = geom();
    if (!opts().force_generic)
        if (const dense_layout_kernels* k = find_dense_static(d)) return (int)k->binary[op](a, b, out, n, s, g, tuning());
    const int N = int(S / dense::gcd_u32(S, 32));
    // option "jit": the whole-chunk kernel with this layout's chunk masks compiled in, specialised at run time (jit.cu)
    if (const long jit = opts().jit) {
        if (!opts().dense_windows) {
            bool handled = false;
            const int rc = jit_dense(op, -1, d, a, b, out, nullptr, n, s, g, tuning(), &handled);
            if (handled || (rc == PINT_E_JIT && jit >= 2)) return rc;
            if (rc != PINT_OK && rc != PINT_E_JIT) return rc;
        }
    }
    // any layout whose chunk is short enough computes whole chunks with run-time masks ...
    if (!opts().dense_windows) {
        cudaError_t rc = cudaSuccess;
        if (dense_mw_binary_part0(N, op, d, S, a, b, out, n, s, g, tuning(), &rc) ||
            dense_mw_binary_part1(N, op, d, S, a, b, out, n, s, g, tuning(), &rc) ||
            dense_mw_binary_part2(N, op, d, S, a, b, out, n, s, g, tuning(), &rc))
            return (int)rc;
    }
    // ... the rest slides windows over the chunk element by element
    auto ops = std::make_integer_sequence<int, pint::cuda::OP_NUM_BINARY>();
    return (int)(S <= 32 ? dense_generic<uint32_t>(op, d, S, a, b, out, n, s, g, ops)
                         : dense_generic<uint64_t>(op, d, S, a, b, out, n, s, g, ops));
}

int pint_cuda_dense_shift(int left, const pint_layout* layout, uint64_t amount, const void* in, void* out, size_t n,
                          pint_stream_t stream)
{
    layout_desc d;
    if (int rc = check_layout(layout, &d)) return rc;
    if (n == 0) return PINT_OK;
    if (!in || !out) return PINT_E_NULL;
    if (((uintptr_t)in | (uintptr_t)out) % 16) return PINT_E_ARG;
    const uint32_t S = total_bits(d);
    const size_t bytes = dense_bytes(S, n);
    {
        const uintptr_t i = (uintptr_t)in, o = (uintptr_t)out;
        if (in != out && i < o + bytes && o < i + bytes) return PINT_E_ALIAS;
    }
    cudaStream_t s = (cudaStream_t)stream;
    if (S == 8 || S == 16 || S == 32 || S == 64) {
        pint_layout p = *layout;
        p.elem_bytes = S / 8;
        const size_t total = bytes / p.elem_bytes;
        if (int rc = pint_cuda_shift(left, &p, amount, in, out, n, stream)) return rc;
        if (total > n) return (int)cudaMemsetAsync((char*)out + n * p.elem_bytes, 0, (total - n) * p.elem_bytes, s);
        return PINT_OK;
    }
    // per-field keep mask of ONE element, as pint_cuda_shift builds it (pint.hpp:630-705, 1006-1029):
    // shift_left keeps the low (w - amount) bits of every field, shift_right_unsigned drops the low
    // min(w, amount) bits; amount >= max(Bits) gives 0
    layout_desc e = d;
    e.elem_bytes = 8;
    const auto m = make_word_masks<uint64_t>(e);
    uint64_t pre = 0;
    uint32_t amt = 0;
    if (amount < m.max_bits) {
        amt = (uint32_t)amount;
        uint64_t mask = 0;
        for (uint32_t k = 0; k < m.nuniq; ++k) {
            const uint64_t lo = m.ulo[k];
            const uint32_t w = m.uw[k];
            const uint32_t sh = left ? (w > amt ? w - amt : 0) : (w < amt ? w : amt);
            mask |= (sh >= 64 ? 0 : (lo << sh)) - lo;
        }
        pre = (left ? mask : ~mask) & m.A;
    }
    const launch_geom g = geom();
    // the stream as one long integer: mask, then funnel-shift neighbouring words (cuda_kernels_dense_shift.cuh), in place
    // or not; only below 32 bits, the chunk kernels keep wider shifts
    if (opts().dense_shift_stream && !opts().dense_windows && amt < 32)
        return (int)dense::launch_dense_shift_stream(in, out, bytes, (unsigned long long)n * S, S,
                                                     dense::make_dense_shift_table(pre, S, amt, left != 0), s, g);
    const int N = int(S / dense::gcd_u32(S, 32));
    if (!opts().dense_windows) {
        cudaError_t rc = cudaSuccess;
        if (dense_mw_unary_part0(N, pre, amt, left != 0, S, in, out, n, s, g, tuning(), &rc) ||
            dense_mw_unary_part1(N, pre, amt, left != 0, S, in, out, n, s, g, tuning(), &rc) ||
            dense_mw_unary_part2(N, pre, amt, left != 0, S, in, out, n, s, g, tuning(), &rc))
            return (int)rc;
    }
    if (S <= 32) {
        dense::window_unary_fn<uint32_t> fn{{uint32_t(pre), ~0u, amt, left ? 1u : 0u}};
        return (int)dense::launch_dense_binary<dense::window_unary_fn<uint32_t>, dense::runtime_geom<uint32_t>, 2>(
            in, nullptr, out, n, fn, dense::runtime_geom<uint32_t>(S, 32), s, g, tuning());
    }
    dense::window_unary_fn<uint64_t> fn{{pre, ~0ULL, amt, left ? 1u : 0u}};
    return (int)dense::launch_dense_binary<dense::window_unary_fn<uint64_t>, dense::runtime_geom<uint64_t>, 2>(
        in, nullptr, out, n, fn, dense::runtime_geom<uint64_t>(S, 64), s, g, tuning());
}

int pint_cuda_dense_pack(const pint_layout* layout, const void* padded, void* dense, size_t n, pint_stream_t stream)
{
    return convert_impl(true, layout, padded, dense, n, (cudaStream_t)stream);
}

int pint_cuda_dense_unpack(const pint_layout* layout, const void* dense, void* padded, size_t n, pint_stream_t stream)
{
    return convert_impl(false, layout, padded, const_cast<void*>(dense), n, (cudaStream_t)stream);
}

int pint_cuda_dense_to_planes(const pint_layout* layout, int is_signed, const void* dense, void* const* planes,
                              int plane_bytes, size_t n, pint_stream_t stream)
{
    return planes_impl(false, layout, is_signed, dense, planes, plane_bytes, n, (cudaStream_t)stream);
}

int pint_cuda_dense_from_planes(const pint_layout* layout, const void* const* planes, int plane_bytes, void* dense,
                                size_t n, pint_stream_t stream)
{
    return planes_impl(true, layout, 0, dense, planes, plane_bytes, n, (cudaStream_t)stream);
}

int pint_cuda_dense_plan_shards(size_t n, int nshards, size_t* begin, size_t* count)
{
    if (nshards < 1 || !begin || !count) return PINT_E_ARG;
    const size_t base = (n / (size_t)nshards) / 128 * 128;  // shard length rounded down to 128 elements
    size_t pos = 0;
    for (int k = 0; k < nshards; ++k) {
        begin[k] = pos;
        count[k] = (k == nshards - 1) ? n - pos : base;
        pos += count[k];
    }
    return PINT_OK;
}

int pint_cuda_sharded_dense_binary(int op, const pint_layout* layout, int ngpu, const int* devices,
                                   const void* const* a, const void* const* b, void* const* out,
                                   const size_t* counts, const pint_stream_t* streams)
{
    if (ngpu < 1 || !devices || !a || !b || !out || !counts) return PINT_E_ARG;
    int prev = 0;
    if (cudaError_t e = cudaGetDevice(&prev)) return (int)e;
    int rc = PINT_OK;
    for (int k = 0; k < ngpu && rc == PINT_OK; ++k) {
        if (cudaError_t e = cudaSetDevice(devices[k])) { rc = (int)e; break; }
        rc = pint_cuda_dense_binary(op, layout, a[k], b[k], out[k], counts[k], streams ? streams[k] : nullptr);
    }
    cudaSetDevice(prev);
    return rc;
}

}  // extern "C"
