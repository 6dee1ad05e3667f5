#!/bin/bash
# C3 unpack: parity of the bulk-store variant, tile / CTA sweep, same-box A/B, ncu --set full of both
# unpack kernels, repack, the narrow-plane group kernel and the 64-bit stream kernels (C2, C4).
tag=${1:-r12}
out=gpurun_out
mkdir -p $out
timeout 600 python -m pytest tests/test_gpu_narrow_planes.py tests/test_gpu_parity.py -x -q -m gpu > $out/${tag}_pytest_planes.log 2>&1; echo "pytest rc=$?"; tail -3 $out/${tag}_pytest_planes.log
timeout 600 build/tune 30 10 11 > $out/${tag}_tune_unpack_bulk.log 2>&1; echo "tune rc=$?"; cat $out/${tag}_tune_unpack_bulk.log
timeout 300 python tools/ab_unpack_bulk.py > $out/${tag}_ab_unpack_bulk.log 2>&1; cat $out/${tag}_ab_unpack_bulk.log
PK="python tools/profile_planes.py 27"
timeout 300 $PK > $out/${tag}_profile_planes.log 2>&1 && \
timeout 1500 ncu --set full --clock-control none --import-source on -k regex:"unpack_oct|unpack_bulk|repack_quad|unpack_group|binary_stream|fused" -c 6 -o $out/${tag}_ncu_full_planes $PK > $out/${tag}_ncu_full_planes.log 2>&1
echo "ncu rc=$?"; tail -3 $out/${tag}_ncu_full_planes.log
