// ref_driver.cpp -- drives the UNMODIFIED reference header over arrays.
//
// TEST INFRASTRUCTURE ONLY.  This translation unit is ours; the only reference
// code it touches is `#include "pint/pint.hpp"` resolved with
// -I/root/reference/include at build time (oracle/Makefile).  The output,
// oracle/_ref/libpint_ref.so, is the reference's own per-element functions
// (pint.hpp:799-1029) instantiated for the layouts in ref_layouts.def and
// wrapped in the loop shape of tests/pint_bench.cpp:90-94 written over SoA
// arrays (out[i] = op(P(a[i]), P(b[i])).value()).  It validates the C
// restatement in pint_oracle.c and serves as the CPU baseline ("kind":
// "reference") in bench.py.  The product never loads it.
#include <cstddef>
#include <cstdint>
#include <cstring>
#include <thread>
#include <type_traits>
#include <utility>
#include <vector>

#include "pint/pint.hpp"

namespace {

using std::size_t;

enum {
    OP_ADD_WRAP = 0, OP_ADD_USAT, OP_ADD_SSAT, OP_SUB_WRAP, OP_SUB_USAT, OP_SUB_SSAT,
    OP_OR, OP_AND, OP_XOR, OP_MIN_U, OP_MAX_U, OP_MIN_S, OP_MAX_S, OP_NUM
};

template <class F>
void parallel_for(size_t n, int nthreads, F f)
{
    if (nthreads <= 1 || n < 4096) { f(size_t(0), n); return; }
    std::vector<std::thread> th;
    size_t chunk = (n + nthreads - 1) / nthreads;
    for (int t = 0; t < nthreads; ++t) {
        size_t lo = size_t(t) * chunk, hi = lo + chunk < n ? lo + chunk : n;
        if (lo >= hi) break;
        th.emplace_back([=] { f(lo, hi); });
    }
    for (auto& t : th) t.join();
}

// the synthetic-input generator of SURVEY.md 8(d) (same function as pint_oracle_splitmix64 and the device fill)
inline uint64_t splitmix64(uint64_t x)
{
    x += 0x9E3779B97F4A7C15ULL;
    x = (x ^ (x >> 30)) * 0xBF58476D1CE4E5B9ULL;
    x = (x ^ (x >> 27)) * 0x94D049BB133111EBULL;
    return x ^ (x >> 31);
}

template <int Op, class P>
inline P apply(P a, P b)
{
    if constexpr (Op == OP_ADD_WRAP) return pint::add_wrap(a, b);
    else if constexpr (Op == OP_ADD_USAT) return pint::add_unsigned_saturate(a, b);
    else if constexpr (Op == OP_ADD_SSAT) return pint::add_signed_saturate(a, b);
    else if constexpr (Op == OP_SUB_WRAP) return pint::sub_wrap(a, b);
    else if constexpr (Op == OP_SUB_USAT) return pint::sub_unsigned_saturate(a, b);
    else if constexpr (Op == OP_SUB_SSAT) return pint::sub_signed_saturate(a, b);
    else if constexpr (Op == OP_OR) return a | b;
    else if constexpr (Op == OP_AND) return a & b;
    else if constexpr (Op == OP_XOR) return a ^ b;
    else if constexpr (Op == OP_MIN_U) return pint::min_unsigned(a, b);
    else if constexpr (Op == OP_MAX_U) return pint::max_unsigned(a, b);
    else if constexpr (Op == OP_MIN_S) return pint::min_signed(a, b);
    else return pint::max_signed(a, b);
}

template <class P> struct Entry;

template <class I, size_t... B>
struct Entry<pint::packed_int<I, B...>> {
    using P = pint::packed_int<I, B...>;
    static constexpr size_t N = sizeof...(B);
    using S = typename std::make_signed<I>::type;

    static bool match(uint32_t eb, uint32_t nf, const uint8_t* bits)
    {
        const uint8_t mine[] = {uint8_t(B)...};
        if (eb != sizeof(I) || nf != N) return false;
        return std::memcmp(mine, bits, N) == 0;
    }

    template <int Op>
    static void binary_loop(const I* a, const I* b, I* out, size_t n, int nt)
    {
        parallel_for(n, nt, [=](size_t lo, size_t hi) {
            for (size_t i = lo; i < hi; ++i) out[i] = apply<Op>(P(a[i]), P(b[i])).value();
        });
    }

    template <int... Ops>
    static int binary_dispatch(int op, const I* a, const I* b, I* out, size_t n, int nt,
                               std::integer_sequence<int, Ops...>)
    {
        int hit = 0;
        ((op == Ops ? (binary_loop<Ops>(a, b, out, n, nt), hit = 1) : 0), ...);
        return hit ? 0 : -10;
    }

    static int binary(int op, const void* a, const void* b, void* out, size_t n, int nt)
    {
        return binary_dispatch(op, (const I*)a, (const I*)b, (I*)out, n, nt,
                               std::make_integer_sequence<int, OP_NUM>());
    }

    // tests/pint_bench.cpp:90-94 verbatim shape: AoS pairs, reduce into a sum.
    template <int Op>
    static uint64_t pairs_loop(const I* pairs, size_t n)
    {
        uint64_t sum = 0;
        for (size_t i = 0; i < n; ++i) sum += apply<Op>(P(pairs[2 * i]), P(pairs[2 * i + 1])).value();
        return sum;
    }

    template <int... Ops>
    static int pairs_dispatch(int op, const I* pairs, size_t n, uint64_t* sum,
                              std::integer_sequence<int, Ops...>)
    {
        int hit = 0;
        ((op == Ops ? (*sum = pairs_loop<Ops>(pairs, n), hit = 1) : 0), ...);
        return hit ? 0 : -10;
    }

    // Whole-array parity without the arrays: the reference op over the synthetic inputs
    // a[i] = (I) splitmix64(seed_a + start + i), b likewise, reduced to (sum, xor) of the results
    // zero-extended to 64 bits -- what pint_cuda_checksum yields for the GPU's output range
    // [start, start + n).  Used by bench.py's parity flag and tests/test_gpu_fullsize.py.
    template <int Op>
    static void stream_checksum_loop(uint64_t seed_a, uint64_t seed_b, uint64_t start, size_t n, int nt, uint64_t* out2)
    {
        const int slots = nt > 1 ? nt : 1;
        std::vector<uint64_t> part(2 * size_t(slots), 0);
        const size_t chunk = (n + slots - 1) / slots;
        auto body = [&](int t) {
            const size_t lo = size_t(t) * chunk, hi = lo + chunk < n ? lo + chunk : n;
            uint64_t s = 0, x = 0;
            for (size_t i = lo; i < hi; ++i) {
                const I a = (I)splitmix64(seed_a + start + i), b = (I)splitmix64(seed_b + start + i);
                const uint64_t v = apply<Op>(P(a), P(b)).value();
                s += v;
                x ^= v;
            }
            part[2 * t] = s;
            part[2 * t + 1] = x;
        };
        if (slots == 1) body(0);
        else {
            std::vector<std::thread> th;
            for (int t = 0; t < slots; ++t) th.emplace_back(body, t);
            for (auto& t : th) t.join();
        }
        out2[0] = out2[1] = 0;
        for (int t = 0; t < slots; ++t) { out2[0] += part[2 * t]; out2[1] ^= part[2 * t + 1]; }
    }

    template <int... Ops>
    static int stream_checksum_dispatch(int op, uint64_t sa, uint64_t sb, uint64_t start, size_t n, int nt, uint64_t* out2,
                                        std::integer_sequence<int, Ops...>)
    {
        int hit = 0;
        ((op == Ops ? (stream_checksum_loop<Ops>(sa, sb, start, n, nt, out2), hit = 1) : 0), ...);
        return hit ? 0 : -10;
    }

    static int stream_checksum(int op, uint64_t sa, uint64_t sb, uint64_t start, size_t n, int nt, uint64_t* out2)
    {
        return stream_checksum_dispatch(op, sa, sb, start, n, nt, out2, std::make_integer_sequence<int, OP_NUM>());
    }

    static int pairs_reduce(int op, const void* pairs, size_t n, uint64_t* sum)
    {
        return pairs_dispatch(op, (const I*)pairs, n, sum, std::make_integer_sequence<int, OP_NUM>());
    }

    static int shift(int left, uint64_t amount, const void* in_, void* out_, size_t n, int nt)
    {
        const I* in = (const I*)in_;
        I* out = (I*)out_;
        // volatile defeats constant folding, as tests/pint_test.cpp:667-670 does
        volatile size_t amt = (size_t)amount;
        parallel_for(n, nt, [=, &amt](size_t lo, size_t hi) {
            const size_t s = amt;
            for (size_t i = lo; i < hi; ++i)
                out[i] = left ? pint::shift_left(P(in[i]), s).value()
                              : pint::shift_right_unsigned(P(in[i]), s).value();
        });
        return 0;
    }

    template <size_t Idx>
    static void get_loop(int is_signed, const I* in, void* plane, int pb, size_t n)
    {
        for (size_t i = 0; i < n; ++i) {
            if (is_signed) {
                S v = pint::get_signed<Idx>(P(in[i]));
                if (pb == 4) ((int32_t*)plane)[i] = (int32_t)v; else ((int64_t*)plane)[i] = (int64_t)v;
            } else {
                I v = pint::get<Idx>(P(in[i]));
                if (pb == 4) ((uint32_t*)plane)[i] = (uint32_t)v; else ((uint64_t*)plane)[i] = (uint64_t)v;
            }
        }
    }

    template <size_t... Is>
    static void unpack_all(int is_signed, const I* in, void* const* planes, int pb, size_t n,
                           std::index_sequence<Is...>)
    {
        ((planes[Is] ? get_loop<Is>(is_signed, in, planes[Is], pb, n) : void()), ...);
    }

    static int unpack(int is_signed, const void* in, void* const* planes, int pb, size_t n)
    {
        unpack_all(is_signed, (const I*)in, planes, pb, n, std::make_index_sequence<N>());
        return 0;
    }

    template <size_t... Is>
    static void repack_all(const void* const* planes, int pb, I* out, size_t n, std::index_sequence<Is...>)
    {
        for (size_t i = 0; i < n; ++i) {
            // the field constructor's body (pint.hpp:773): make_truncate over value_type arguments
            out[i] = pint::detail::make_truncate<I, B...>(
                (pb == 4 ? (I)((const int32_t*)planes[Is])[i] : (I)((const uint64_t*)planes[Is])[i])...);
        }
    }

    static int repack(const void* const* planes, int pb, void* out, size_t n)
    {
        repack_all(planes, pb, (I*)out, n, std::make_index_sequence<N>());
        return 0;
    }

    template <size_t S0, size_t E0>
    static void slice_loop(const I* in, I* out, size_t n)
    {
        if constexpr (S0 < E0 && E0 <= N) {
            for (size_t i = 0; i < n; ++i) out[i] = pint::slice<S0, E0>(P(in[i])).value();
        }
    }

    template <size_t... Ks>
    static int slice_dispatch(uint32_t s, uint32_t e, const I* in, I* out, size_t n, std::index_sequence<Ks...>)
    {
        int hit = 0;
        const size_t key = size_t(s) * (N + 1) + e;
        ((key == Ks ? (slice_loop<Ks / (N + 1), Ks % (N + 1)>(in, out, n), hit = 1) : 0), ...);
        return hit ? 0 : -12;
    }

    static int slice(uint32_t s, uint32_t e, const void* in, void* out, size_t n)
    {
        if (!(s < e && e <= N)) return -12;
        return slice_dispatch(s, e, (const I*)in, (I*)out, n, std::make_index_sequence<N*(N + 1) + 1>());
    }

    static void constants(uint64_t* out)
    {
        out[0] = pint::detail::mask_hiorder<I, B...>::value;
        out[1] = pint::detail::mask_loorder<I, B...>::value;
        out[2] = pint::detail::all_ones<I, pint::detail::sum<B...>::value>::value;
        out[3] = pint::detail::detect_saturation_mask_type<I, B...>::value;
    }
};

struct VTable {
    bool (*match)(uint32_t, uint32_t, const uint8_t*);
    int (*binary)(int, const void*, const void*, void*, size_t, int);
    int (*pairs_reduce)(int, const void*, size_t, uint64_t*);
    int (*stream_checksum)(int, uint64_t, uint64_t, uint64_t, size_t, int, uint64_t*);
    int (*shift)(int, uint64_t, const void*, void*, size_t, int);
    int (*unpack)(int, const void*, void* const*, int, size_t);
    int (*repack)(const void* const*, int, void*, size_t);
    int (*slice)(uint32_t, uint32_t, const void*, void*, size_t);
    void (*constants)(uint64_t*);
    uint32_t elem_bytes, nfields;
    uint8_t bits[64];
};

template <class I, size_t... B>
VTable make_vtable()
{
    using E = Entry<pint::packed_int<I, B...>>;
    VTable v{&E::match, &E::binary, &E::pairs_reduce, &E::stream_checksum, &E::shift, &E::unpack, &E::repack, &E::slice,
             &E::constants, (uint32_t)sizeof(I), (uint32_t)sizeof...(B), {uint8_t(B)...}};
    return v;
}

const std::vector<VTable>& registry()
{
    static const std::vector<VTable> r = {
#define X(I, ...) make_vtable<I, __VA_ARGS__>(),
#include "ref_layouts.def"
#undef X
    };
    return r;
}

const VTable* find(uint32_t eb, uint32_t nf, const uint8_t* bits)
{
    for (const auto& v : registry())
        if (v.match(eb, nf, bits)) return &v;
    return nullptr;
}

// storage type make_packed_int would pick for a single field of `total` bits
// (pint.hpp:789-795); mirrors tests/pint_test.cpp:58-87.
template <size_t... T>
uint32_t storage_table(uint32_t total, std::index_sequence<T...>)
{
    uint32_t r = 0;
    ((total == T + 1 ? (r = (uint32_t)sizeof(typename pint::make_packed_int<T + 1>::value_type)) : 0), ...);
    return r;
}


} // namespace

extern "C" {

int pint_ref_num_layouts() { return (int)registry().size(); }

int pint_ref_layout_at(int i, uint32_t* elem_bytes, uint32_t* nfields, uint8_t* bits)
{
    if (i < 0 || i >= (int)registry().size()) return -1;
    const VTable& v = registry()[i];
    *elem_bytes = v.elem_bytes; *nfields = v.nfields;
    std::memcpy(bits, v.bits, v.nfields);
    return 0;
}

int pint_ref_has_layout(uint32_t eb, uint32_t nf, const uint8_t* bits) { return find(eb, nf, bits) != nullptr; }

// H, L, A, saturation-mask type as the header's metafunctions compute them
int pint_ref_constants(uint32_t eb, uint32_t nf, const uint8_t* bits, uint64_t* out4)
{
    const VTable* v = find(eb, nf, bits);
    if (!v) return -1;
    v->constants(out4);
    return 0;
}

int pint_ref_binary(int op, uint32_t eb, uint32_t nf, const uint8_t* bits,
                    const void* a, const void* b, void* out, size_t n, int nthreads)
{
    const VTable* v = find(eb, nf, bits);
    return v ? v->binary(op, a, b, out, n, nthreads) : -1;
}

int pint_ref_pairs_reduce(int op, uint32_t eb, uint32_t nf, const uint8_t* bits,
                          const void* pairs, size_t n, uint64_t* sum)
{
    const VTable* v = find(eb, nf, bits);
    return v ? v->pairs_reduce(op, pairs, n, sum) : -1;
}

// out2[0] = sum, out2[1] = xor over i in [start, start + n) of OP(a[i], b[i]) with the synthetic inputs
int pint_ref_stream_checksum(int op, uint32_t eb, uint32_t nf, const uint8_t* bits, uint64_t seed_a, uint64_t seed_b,
                             uint64_t start, size_t n, int nthreads, uint64_t* out2)
{
    const VTable* v = find(eb, nf, bits);
    return v ? v->stream_checksum(op, seed_a, seed_b, start, n, nthreads, out2) : -1;
}

int pint_ref_shift(int left, uint32_t eb, uint32_t nf, const uint8_t* bits, uint64_t amount,
                   const void* in, void* out, size_t n, int nthreads)
{
    const VTable* v = find(eb, nf, bits);
    return v ? v->shift(left, amount, in, out, n, nthreads) : -1;
}

int pint_ref_unpack(uint32_t eb, uint32_t nf, const uint8_t* bits, int is_signed,
                    const void* packed, void* const* planes, int plane_bytes, size_t n)
{
    const VTable* v = find(eb, nf, bits);
    return v ? v->unpack(is_signed, packed, planes, plane_bytes, n) : -1;
}

int pint_ref_repack(uint32_t eb, uint32_t nf, const uint8_t* bits,
                    const void* const* planes, int plane_bytes, void* packed, size_t n)
{
    const VTable* v = find(eb, nf, bits);
    return v ? v->repack(planes, plane_bytes, packed, n) : -1;
}

int pint_ref_slice(uint32_t eb, uint32_t nf, const uint8_t* bits, uint32_t start, uint32_t end,
                   const void* in, void* out, size_t n)
{
    const VTable* v = find(eb, nf, bits);
    return v ? v->slice(start, end, in, out, n) : -1;
}

uint32_t pint_ref_storage_bytes(uint32_t total_bits)
{
    return storage_table(total_bits, std::make_index_sequence<64>());
}

int pint_ref_hardware_threads() { return (int)std::thread::hardware_concurrency(); }

} // extern "C"
