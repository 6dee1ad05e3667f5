#!/bin/bash
# Round 2, full single-GPU round: parity suite, smoke, reference arm + bench, the reference benchmark's fixtures
# replayed, per-config tables, launch list and ncu --set full captures (only after the same command exited 0).
# The .ncu-rep files are summarised on the box (tools/ncu_summary.py) and removed: gpurun brings back <= 64 MiB.
tag=${1:-r15}
out=gpurun_out
mkdir -p $out
nvidia-smi --query-gpu=name,memory.total,clocks.max.sm,clocks.max.mem,power.limit --format=csv > $out/${tag}_box.txt 2>&1
nproc >> $out/${tag}_box.txt
timeout 1800 python -m pytest tests -m gpu -x -q > $out/${tag}_pytest_gpu.log 2>&1; echo "pytest rc=$?" >> $out/${tag}_pytest_gpu.log
tail -4 $out/${tag}_pytest_gpu.log
timeout 300 python -c "import __graft_entry__ as g; g.smoke()" > $out/${tag}_smoke.log 2>&1; echo "smoke rc=$?" >> $out/${tag}_smoke.log
tail -2 $out/${tag}_smoke.log
timeout 600 python bench.py --impl reference --steps 5 --warmup 1 > $out/${tag}_bench_ref.json 2> $out/${tag}_bench_ref.err
timeout 900 python bench.py > $out/${tag}_bench.json 2> $out/${tag}_bench.err; echo "bench rc=$?"
python - <<PY
import json
d = json.loads(open("$out/${tag}_bench.json").read().strip().splitlines()[-1])
print("value", round(d["value"] / 1e9, 1), "Gelem/s", round(d["roofline"]["achieved"], 1), "GB/s parity_ok", d["parity_ok"], "e2e", round(d["e2e"]["value"] / 1e9, 2), d["e2e"]["roofline"]["frac"], d["clocks"])
for c in d["configs"]:
    print("   ", c.get("config"), c.get("op", ""), c.get("gbs_per_gpu"), c.get("frac_spec_8000"), c.get("parity_ok"), c.get("error", ""))
PY
timeout 600 python tools/replay_pint_bench.py > $out/${tag}_replay_pint_bench.jsonl 2> $out/${tag}_replay_pint_bench.err; echo "replay rc=$?"
tail -1 $out/${tag}_replay_pint_bench.jsonl
timeout 900 python tools/bench_configs.py C0 C1 C2 C3 C4 > $out/${tag}_configs.jsonl 2> $out/${tag}_configs.err; echo "configs rc=$?"
# profiler passes
PB="python bench.py --steps 2 --warmup 3 --no-cpu --no-e2e --no-configs --no-parity --log2 28"
timeout 300 $PB > $out/${tag}_bench_small.json 2>&1 && \
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file $out/${tag}_ncu_launches.csv $PB > $out/${tag}_ncu_launches.log 2>&1
timeout 900 ncu --set full --clock-control none --import-source on -k regex:binary_pair -c 2 -o $out/${tag}_ncu_full_pair $PB > $out/${tag}_ncu_full.log 2>&1
echo "ncu pair rc=$?"
python tools/ncu_summary.py $out/${tag}_ncu_full_pair.ncu-rep $out/${tag}_ncu_full_pair_summary.txt
PK="python tools/profile_planes.py 27"
timeout 300 $PK > $out/${tag}_profile_planes.log 2>&1 && \
timeout 1500 ncu --set full --clock-control none --import-source on -k regex:"unpack_oct|unpack_bulk|repack_quad|unpack_group|binary_stream|fused|pairs_reduce" -c 8 -o $out/${tag}_ncu_full_planes $PK > $out/${tag}_ncu_full_planes.log 2>&1
echo "ncu planes rc=$?"
python tools/ncu_summary.py $out/${tag}_ncu_full_planes.ncu-rep $out/${tag}_ncu_full_planes_summary.txt
rm -f $out/*.ncu-rep
ls $out | grep ${tag} | wc -l
