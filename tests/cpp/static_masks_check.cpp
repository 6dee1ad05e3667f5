// static_masks_check.cpp -- compile-time proof that the pint::cuda mask generator
// (include/pint/cuda_layout.hpp) yields exactly the constants of the reference's
// metafunctions, for every layout in oracle/ref_layouts.def, and that
// include/pint/cuda.hpp coexists with the unmodified pint.hpp.
// Compiled (not run) in the build container:
//   g++ -std=c++17 -fsyntax-only -I/root/reference/include -Iinclude -Ioracle tests/cpp/static_masks_check.cpp
#include "pint/pint.hpp"       // the reference, unmodified
#include "pint/cuda.hpp"       // host template layer (C++11)
#include "pint/cuda_layout.hpp"  // mask generator (C++17)

namespace {

template <class I, std::size_t... B>
constexpr bool check_layout()
{
    using SL = pint::cuda::static_layout<I, B...>;
    // element-level view (64-bit word over a 64-bit element: nothing replicated)
    constexpr pint::cuda::layout_desc e{8, (uint32_t)sizeof...(B), {uint8_t(B)...}};
    constexpr auto m = pint::cuda::make_word_masks<uint64_t>(e);
    static_assert(m.H == (uint64_t)pint::detail::mask_hiorder<I, B...>::value, "H != mask_hiorder (pint.hpp:324)");
    static_assert(m.L == (uint64_t)pint::detail::mask_loorder<I, B...>::value, "L != mask_loorder (pint.hpp:332)");
    static_assert(m.A == (uint64_t)pint::detail::all_ones<uint64_t, pint::detail::sum<B...>::value>::value,
                  "A != all_ones<sum> (pint.hpp:340)");
    static_assert(m.ref_sat_type == pint::detail::detect_saturation_mask_type<I, B...>::value,
                  "saturation-mask type != detect_saturation_mask_type (pint.hpp:455)");
    static_assert(m.max_bits == pint::detail::find_max<B...>::value, "max != find_max (pint.hpp:277)");
    // the word the kernels use replicates the element masks
    using W = typename SL::word;
    constexpr unsigned reps = sizeof(W) / sizeof(I);
    constexpr W wH = SL::masks.H;
    constexpr uint64_t eH = (uint64_t)pint::detail::mask_hiorder<I, B...>::value;
    static_assert((wH & (W)(I)(~(I)0)) == (W)eH, "low element of the word mask");
    static_assert(reps == 1 || ((uint64_t)wH >> (8 * sizeof(I))) == ((uint64_t)wH & (~0ULL >> (64 - 8 * (sizeof(W) - sizeof(I))))),
                  "replication");
    // descriptor of cuda.hpp == descriptor of cuda_layout.hpp
    return true;
}

#define X(I, ...) static_assert(check_layout<I, __VA_ARGS__>(), "layout " #I " <" #__VA_ARGS__ ">");
#include "ref_layouts.def"
#undef X

// make_packed_int's storage choice (pint.hpp:789-795; tests/pint_test.cpp:61-87)
template <std::size_t N>
constexpr bool check_storage()
{
    static_assert(pint::cuda::storage_bytes(N) == sizeof(typename pint::make_packed_int<N>::value_type), "storage");
    if constexpr (N > 1) return check_storage<N - 1>();
    return true;
}
static_assert(check_storage<64>(), "storage_bytes == make_packed_int for 1..64 bits");

// cuda.hpp: slice result type == the reference's sliced_int (pint.hpp:746-754)
using P5 = pint::packed_int<uint16_t, 1, 2, 3, 4, 5>;
static_assert(std::is_same<pint::cuda::sliced_int<2, 4, uint16_t, 1, 2, 3, 4, 5>::type,
                           pint::detail::sliced_int<2, 4, uint16_t, 1, 2, 3, 4, 5>>::value, "slice<2,4> type");
static_assert(std::is_same<pint::cuda::sliced_int<0, 5, uint16_t, 1, 2, 3, 4, 5>::type, P5>::value, "slice<0,5> type");
static_assert(std::is_same<pint::cuda::sliced_int<4, 5, uint16_t, 1, 2, 3, 4, 5>::type,
                           pint::packed_int<uint16_t, 5>>::value, "slice<4,5> type");
// packed_int arrays are Integer arrays (pint.hpp:758-787)
static_assert(sizeof(pint::packed_int<uint32_t, 8, 8, 8, 8>) == 4 && std::is_trivially_copyable<P5>::value, "layout");

}  // namespace

int main() { return 0; }
