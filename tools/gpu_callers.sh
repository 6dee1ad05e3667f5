#!/bin/bash
# One gpurun call for the extra caller shapes: their parity tests first, then the whole GPU suite,
# the C++ API test, the per-config table (with the new rows and the directional ceilings) and an
# ncu --set full capture of the two new kernels.
tag=${1:-r05}
out=gpurun_out
mkdir -p $out
timeout 600 python -m pytest tests/test_gpu_callers.py -x -q > $out/${tag}_pytest_callers.log 2>&1; echo "callers rc=$?" >> $out/${tag}_pytest_callers.log
tail -15 $out/${tag}_pytest_callers.log
timeout 1500 python -m pytest tests -m gpu -x -q > $out/${tag}_pytest_gpu.log 2>&1; echo "pytest rc=$?" >> $out/${tag}_pytest_gpu.log
tail -3 $out/${tag}_pytest_gpu.log
timeout 600 python tools/bench_configs.py > $out/${tag}_configs.jsonl 2> $out/${tag}_configs.err; echo "configs rc=$?"
grep -E "binary_pairs|binary_scalar|ceiling|equal|byte planes" $out/${tag}_configs.jsonl | cut -c1-330
tail -3 $out/${tag}_configs.err
PK="python tools/profile_callers.py 28"
timeout 300 $PK > $out/${tag}_profile_callers.log 2>&1 && \
timeout 900 ncu --set full --clock-control none --import-source on -k regex:"map1_stream|pairs_map_kernel" -c 6 -o $out/${tag}_ncu_full_callers $PK > $out/${tag}_ncu_full_callers.log 2>&1
echo "ncu callers rc=$?"
