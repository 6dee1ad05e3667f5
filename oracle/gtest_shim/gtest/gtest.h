// Minimal stand-in for <gtest/gtest.h> (GoogleTest is not installed in this
// image).  TEST INFRASTRUCTURE ONLY: lets the reference's own unit-test TU,
// /root/reference/tests/pint_test.cpp, compile UNMODIFIED (oracle/Makefile
// target _ref/pint_test) so its 61 known-answer tests run against the
// reference header here.  Only the macros that file uses are provided.
#pragma once
#include <cstdio>
#include <ostream>
#include <vector>

namespace gtest_shim {
struct Case { const char* suite; const char* name; void (*fn)(int&); };
inline std::vector<Case>& cases() { static std::vector<Case> c; return c; }
struct Reg { Reg(const char* s, const char* n, void (*f)(int&)) { cases().push_back({s, n, f}); } };
template <class A, class B> bool eq(const A& a, const B& b) { return a == b; }
}

#define TEST(suite, name)                                                            \
    static void suite##_##name##_body(int& gtest_shim_failures);                     \
    static gtest_shim::Reg suite##_##name##_reg(#suite, #name, suite##_##name##_body); \
    static void suite##_##name##_body(int& gtest_shim_failures)

#define ASSERT_EQ(a, b)                                                              \
    do { if (!gtest_shim::eq((a), (b))) { ++gtest_shim_failures;                     \
        std::printf("  ASSERT_EQ failed %s:%d: %s == %s\n", __FILE__, __LINE__, #a, #b); return; } } while (0)
#define ASSERT_NE(a, b)                                                              \
    do { if (gtest_shim::eq((a), (b))) { ++gtest_shim_failures;                      \
        std::printf("  ASSERT_NE failed %s:%d: %s != %s\n", __FILE__, __LINE__, #a, #b); return; } } while (0)

#ifndef GTEST_SHIM_NO_MAIN
int main()
{
    int failed = 0;
    for (const auto& c : gtest_shim::cases()) {
        int f = 0;
        c.fn(f);
        std::printf("[%s] %s.%s\n", f ? "FAILED" : "  OK  ", c.suite, c.name);
        failed += f ? 1 : 0;
    }
    std::printf("ran %zu tests, %d failures\n", gtest_shim::cases().size(), failed);
    return failed ? 1 : 0;
}
#endif
