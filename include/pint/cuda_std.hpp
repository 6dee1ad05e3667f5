// pint/cuda_std.hpp -- the few standard-library names the pint::cuda headers use.
// A normal compile (nvcc, g++) takes them from the host's standard library.  Under NVRTC -- the
// run-time specialisation of the kernels for a layout that has no pre-built constants
// (pint_b200/csrc/jit.cu) -- there are no host headers, so the same names come from libcu++
// (<cuda/std/...>, shipped with the CUDA toolkit) and are made visible as std::.
#pragma once
#ifndef PINT_CUDA_STD_HPP_INCLUDED
#define PINT_CUDA_STD_HPP_INCLUDED

#if defined(__CUDACC_RTC__)
#include <cuda/std/cstddef>
#include <cuda/std/cstdint>
#include <cuda/std/type_traits>
#include <cuda/std/utility>
namespace std {
using namespace ::cuda::std;
}
using ::cuda::std::int16_t;
using ::cuda::std::int32_t;
using ::cuda::std::int64_t;
using ::cuda::std::int8_t;
using ::cuda::std::size_t;
using ::cuda::std::uint16_t;
using ::cuda::std::uint32_t;
using ::cuda::std::uint64_t;
using ::cuda::std::uint8_t;
using ::cuda::std::uintptr_t;
#else
#include <cstddef>
#include <cstdint>
#include <type_traits>
#include <utility>
#endif

#endif  // PINT_CUDA_STD_HPP_INCLUDED
