#!/bin/bash
# Round 2, first GPU call: parity suite (incl. the mapping-edge probe and its regression check), smoke, bench.
tag=${1:-r10}
out=gpurun_out
mkdir -p $out
nvidia-smi --query-gpu=name,memory.total,clocks.max.sm,clocks.max.mem,power.limit --format=csv > $out/${tag}_box.txt 2>&1
nproc >> $out/${tag}_box.txt; free -g >> $out/${tag}_box.txt
timeout 600 python tests/vmm_probe.py > $out/${tag}_vmm_probe.log 2>&1; echo "probe rc=$?" >> $out/${tag}_vmm_probe.log
timeout 300 python tests/vmm_probe.py --lib build/r1_regress/libpint_cuda_r1.so > $out/${tag}_vmm_probe_r1lib.log 2>&1; echo "probe(r1 lib) rc=$?" >> $out/${tag}_vmm_probe_r1lib.log
timeout 300 python tests/vmm_probe.py --inject > $out/${tag}_vmm_probe_inject.log 2>&1; echo "probe(inject) rc=$?" >> $out/${tag}_vmm_probe_inject.log
tail -2 $out/${tag}_vmm_probe.log $out/${tag}_vmm_probe_r1lib.log $out/${tag}_vmm_probe_inject.log
timeout 1500 python -m pytest tests -m gpu -x -q > $out/${tag}_pytest_gpu.log 2>&1; echo "pytest rc=$?" >> $out/${tag}_pytest_gpu.log
tail -15 $out/${tag}_pytest_gpu.log
timeout 300 python -c "import __graft_entry__ as g; g.smoke()" > $out/${tag}_smoke.log 2>&1; echo "smoke rc=$?" >> $out/${tag}_smoke.log
tail -2 $out/${tag}_smoke.log
timeout 600 python bench.py --impl reference --steps 5 --warmup 1 > $out/${tag}_bench_ref.json 2> $out/${tag}_bench_ref.err
timeout 900 python bench.py > $out/${tag}_bench.json 2> $out/${tag}_bench.err; echo "bench rc=$?"
cat $out/${tag}_bench.json; tail -5 $out/${tag}_bench.err
