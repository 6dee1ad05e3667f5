#!/bin/bash
# Builds build/r1_regress/libpint_cuda_r1.so: today's library with round 1's pairs_reduce reduction
# (commit 8e7f7e4: atomicXor(sum_dev + 1, 0), 8 bytes past the caller's buffer).
# tests/test_gpu_mapping_edge.py::test_round1_library_is_caught runs the VMM probe against it.
# Build here (nvcc cross-compiles); build/ travels to the GPU box with gpurun.
set -e
cd "$(dirname "$0")/.."
make -j8 all >/dev/null
mkdir -p build/r1_regress
# today's api_aux.cu with round 1's reduction put back: pairs_reduce runs the sum+xor form on ONE uint64_t
sed 's/block_reduce_sum_xor<false>(s, 0, sum);/block_reduce_sum_xor<true>(s, 0, sum);  \/\/ round 1: atomicXor(sum + 1, 0)/' \
    pint_b200/csrc/api_aux.cu > build/r1_regress/api_aux_r1.cu
grep -q 'block_reduce_sum_xor<true>(s, 0, sum)' build/r1_regress/api_aux_r1.cu
nvcc -std=c++17 -O3 -gencode arch=compute_100a,code=sm_100a -lineinfo -Xcompiler -fPIC -Iinclude -Ipint_b200/csrc \
     -c build/r1_regress/api_aux_r1.cu -o build/r1_regress/api_aux_r1.o
objs=$(ls build/obj/*.o | grep -v '/api_aux.o')
nvcc -gencode arch=compute_100a,code=sm_100a -shared -cudart static -o build/r1_regress/libpint_cuda_r1.so $objs build/r1_regress/api_aux_r1.o
ls -la build/r1_regress/libpint_cuda_r1.so
