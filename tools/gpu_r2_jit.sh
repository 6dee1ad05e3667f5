#!/bin/bash
# Round 2: whole GPU suite (JIT, nvcc template layer, bulk unpack default ...), then the jit / runtime-mask
# throughput table and the C3 / extra rows of tools/bench_configs.py.
tag=${1:-r13}
out=gpurun_out
mkdir -p $out
timeout 1500 python -m pytest tests -m gpu -x -q > $out/${tag}_pytest_gpu.log 2>&1; echo "pytest rc=$?" >> $out/${tag}_pytest_gpu.log
tail -6 $out/${tag}_pytest_gpu.log
timeout 120 build/cuda_api_nvcc_test > $out/${tag}_cuda_api_nvcc_test.log 2>&1; tail -6 $out/${tag}_cuda_api_nvcc_test.log
timeout 900 python tools/bench_configs.py jit > $out/${tag}_configs_jit.jsonl 2> $out/${tag}_configs_jit.err; echo "jit configs rc=$?"
python - <<PY
import json
for l in open("$out/${tag}_configs_jit.jsonl"):
    d = json.loads(l); print(d["layout"], d["op"], d["kernel"][:22], d["gbs_median"])
PY
timeout 600 python tools/bench_configs.py C0 C3 > $out/${tag}_configs_c0c3.jsonl 2> $out/${tag}_configs_c0c3.err
python - <<PY
import json
for l in open("$out/${tag}_configs_c0c3.jsonl"):
    d = json.loads(l); print(d["config"], d.get("op"), d["gbs_median"])
PY
