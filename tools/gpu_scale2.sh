#!/bin/bash
# N-GPU box (default 2): bench.py and the reference arm under torchrun at N, plus the multi-device tests.
# Usage: bash tools/gpu_scale2.sh <tag> [N]
tag=${1:-r09s}
N=${2:-2}
out=gpurun_out
mkdir -p $out
nvidia-smi -L > $out/${tag}_gpus.txt 2>&1
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29512"
timeout 600 $TR bench.py --impl reference --gpus $N --steps 5 --warmup 1 > $out/${tag}_bench_ref_n$N.json 2> $out/${tag}_bench_ref_n$N.err; echo "ref rc=$?"
timeout 600 $TR bench.py --gpus $N --steps 20 --warmup 3 > $out/${tag}_bench_n$N.json 2> $out/${tag}_bench_n$N.err; echo "bench rc=$?"
timeout 600 python -m pytest tests -q -m gpu -k "sharded or multi_device or devices" > $out/${tag}_pytest_multi.log 2>&1
tail -3 $out/${tag}_pytest_multi.log
python - <<PY
import json
for f in ("$out/${tag}_bench_ref_n$N.json", "$out/${tag}_bench_n$N.json"):
    for l in open(f):
        if l.startswith("{"):
            d = json.loads(l); print(d.get("impl", "ours"), d["n_gpus"], round(d["value"] / 1e9, 1), "Gelem/s", d.get("roofline", {}).get("achieved"), "e2e", round(d["e2e"]["value"] / 1e9, 1), d.get("clocks"))
PY
