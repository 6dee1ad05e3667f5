#!/bin/bash
# Builds build/r1_regress/libpint_cuda_r1.so: today's library with round 1's api_aux.cu
# (commit 8e7f7e4: pairs_reduce does atomicXor(sum_dev + 1, 0), 8 bytes past the caller's buffer).
# tests/test_gpu_mapping_edge.py::test_round1_library_is_caught runs the VMM probe against it.
# Needs git history (not present on the GPU box: build here, build/ travels with gpurun).
set -e
cd "$(dirname "$0")/.."
make -j8 all >/dev/null
mkdir -p build/r1_regress
git show 8e7f7e4:pint_b200/csrc/api_aux.cu > build/r1_regress/api_aux_r1.cu
nvcc -std=c++17 -O3 -gencode arch=compute_100a,code=sm_100a -lineinfo -Xcompiler -fPIC -Iinclude -Ipint_b200/csrc \
     -c build/r1_regress/api_aux_r1.cu -o build/r1_regress/api_aux_r1.o
objs=$(ls build/obj/*.o | grep -v '/api_aux.o')
nvcc -gencode arch=compute_100a,code=sm_100a -shared -cudart static -o build/r1_regress/libpint_cuda_r1.so $objs build/r1_regress/api_aux_r1.o
ls -la build/r1_regress/libpint_cuda_r1.so
