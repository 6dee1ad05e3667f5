// readme_examples_check.cpp -- the worked examples of /root/reference/README.md (lines 100-410),
// static_asserted against the UNMODIFIED reference header.  tests/golden/readme_examples.json holds the
// same statements as data; tests/test_oracle_golden.py and tests/test_gpu_parity.py run them through
// the oracle and through the CUDA kernels.  Compile-only (the build container, where the reference is):
//   g++ -std=c++17 -fsyntax-only -I/root/reference/include tests/cpp/readme_examples_check.cpp
#include <type_traits>

#include "pint/pint.hpp"

using namespace pint;

namespace get_example {  // README.md:100-103
using MyPack = make_packed_int<3, 5>;
static_assert(get<0>(MyPack(4, 23)) == 4, "");
static_assert(get<1>(MyPack(4, 23)) == 23, "");
static_assert(sizeof(MyPack::value_type) == 1, "");
}
namespace get_signed_example {  // README.md:118-123
using MyPack = make_packed_int<4, 6>;
static_assert(get_signed<0>(MyPack(-4, 23)) == -4, "");
static_assert(get_signed<1>(MyPack(-4, 23)) == 23, "");
static_assert(get<0>(MyPack(-4, 23)) == 12, "");
static_assert(sizeof(MyPack::value_type) == 2, "");
}
namespace slice_example {  // README.md:139-144
using MyPack = make_packed_int<2, 3, 4, 5, 6>;
constexpr auto a = slice<2, 4>(MyPack(3, 7, 15, 31, 63));
static_assert(std::is_same<std::remove_const<decltype(a)>::type, packed_int<MyPack::value_type, 4, 5>>::value, "");
static_assert(get<0>(a) == 15 && get<1>(a) == 31, "");
static_assert(sizeof(MyPack::value_type) == 4, "");
}
namespace add_examples {  // README.md:163-168, 185-190, 211-216
using MyPack = make_packed_int<5, 6, 5>;
static_assert(add_wrap(MyPack(2, 10, 20), MyPack(1, 58, 12)) == MyPack(3, 4, 0), "");
static_assert(add_unsigned_saturate(MyPack(2, 10, 20), MyPack(1, 58, 12)) == MyPack(3, 63, 31), "");
static_assert(add_signed_saturate(MyPack(2, 10, -10), MyPack(1, 28, -12)) == MyPack(3, 31, -16), "");
}
namespace sub_examples {  // README.md:233-234, 251-252, 273-278
using P444 = make_packed_int<4, 4, 4>;
static_assert(sub_wrap(P444(1, 4, 2), P444(7, 2, 6)) == P444(10, 2, 12), "");
static_assert(sub_unsigned_saturate(P444(1, 4, 2), P444(7, 2, 6)) == P444(0, 2, 0), "");
using P464 = make_packed_int<4, 6, 4>;
static_assert(sub_signed_saturate(P464(4, -3, 7), P464(-6, 30, 1)) == P464(7, -32, 6), "");
}
namespace min_max_examples {  // README.md:297-367
using MyPack = make_packed_int<4, 6, 4>;
static_assert(min_unsigned(MyPack(4, 5, 3), MyPack(1, 15, 3)) == MyPack(1, 5, 3), "");
static_assert(max_unsigned(MyPack(4, 5, 3), MyPack(1, 15, 3)) == MyPack(4, 15, 3), "");
static_assert(min_signed(MyPack(-1, 5, 0), MyPack(4, -2, 7)) == MyPack(-1, -2, 0), "");
static_assert(max_signed(MyPack(-1, 5, 0), MyPack(4, -2, 7)) == MyPack(4, 5, 7), "");
}
namespace shift_examples {  // README.md:387-389, 407-409
using MyPack = make_packed_int<3, 7, 6>;
static_assert(shift_left(MyPack(1, 2, 3), 3) == MyPack(0, 16, 24), "");
static_assert(shift_right_unsigned(MyPack(5, 106, 42), 4) == MyPack(0, 6, 2), "");
}

int main() { return 0; }
