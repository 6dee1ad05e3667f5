#!/bin/bash
# One gpurun call: GPU tests, smoke, bench N=1, reference arm, launch list, ncu --set full captures.
# Usage (on the GPU box, from the repo root): bash tools/gpu_round.sh <tag>
tag=${1:-r01x}
out=gpurun_out
mkdir -p $out
nvidia-smi --query-gpu=name,memory.total,clocks.max.sm,clocks.max.mem,power.limit --format=csv > $out/${tag}_box.txt 2>&1
nproc >> $out/${tag}_box.txt
timeout 1500 python -m pytest tests -m gpu -x -q > $out/${tag}_pytest_gpu.log 2>&1; echo "pytest rc=$?" >> $out/${tag}_pytest_gpu.log
tail -3 $out/${tag}_pytest_gpu.log
timeout 300 python -c "import __graft_entry__ as g; g.smoke()" > $out/${tag}_smoke.log 2>&1; echo "smoke rc=$?" >> $out/${tag}_smoke.log
tail -2 $out/${tag}_smoke.log
timeout 600 python bench.py --impl reference --steps 5 --warmup 1 > $out/${tag}_bench_ref.json 2> $out/${tag}_bench_ref.err
timeout 900 python bench.py > $out/${tag}_bench.json 2> $out/${tag}_bench.err; echo "bench rc=$?"
cat $out/${tag}_bench.json
timeout 600 python tools/bench_configs.py > $out/${tag}_configs.jsonl 2> $out/${tag}_configs.err; echo "configs rc=$?"
timeout 600 python tools/bench_dense.py 29 10 > $out/${tag}_dense.jsonl 2> $out/${tag}_dense.err; echo "dense rc=$?"
# profiler passes: only after the same command exited 0 without ncu
PB="python bench.py --steps 2 --warmup 3 --no-cpu --no-e2e --log2 28"
timeout 300 $PB > $out/${tag}_bench_small.json 2>&1 && \
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file $out/${tag}_ncu_launches.csv $PB > $out/${tag}_ncu_launches.log 2>&1
timeout 900 ncu --set full --clock-control none --import-source on -k regex:binary_pair -c 2 -o $out/${tag}_ncu_full_pair $PB > $out/${tag}_ncu_full.log 2>&1
echo "ncu pair rc=$?"
PK="python tools/profile_kernels.py 26"
timeout 300 $PK > $out/${tag}_profile_kernels.log 2>&1 && \
timeout 1200 ncu --set full --clock-control none --import-source on -k regex:dense_binary -c 14 -o $out/${tag}_ncu_full_dense $PK > $out/${tag}_ncu_full_dense.log 2>&1
echo "ncu dense rc=$?"
PC="python tools/profile_callers.py 28"
timeout 300 $PC > $out/${tag}_profile_callers.log 2>&1 && \
timeout 900 ncu --set full --clock-control none --import-source on -k regex:"map1_stream|pairs_map_kernel" -c 6 -o $out/${tag}_ncu_full_callers $PC > $out/${tag}_ncu_full_callers.log 2>&1
echo "ncu callers rc=$?"
ls -la $out | tail -24
