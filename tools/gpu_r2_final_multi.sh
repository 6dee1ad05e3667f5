#!/bin/bash
# Final-code multi-GPU evidence: bench.py + the reference arm under torchrun at the given Ns (the launch line the
# driver uses), then the multi-device tests.  Usage: bash tools/gpu_r2_final_multi.sh <tag> "<N list>"
tag=${1:-r24m}
NS=${2:-2}
out=gpurun_out
mkdir -p $out
nvidia-smi -L > $out/${tag}_gpus.txt 2>&1; nproc >> $out/${tag}_gpus.txt
for N in $NS; do
  TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29512"
  timeout 400 $TR bench.py --impl reference --gpus $N --steps 5 --warmup 1 > $out/${tag}_bench_ref_n$N.json 2> $out/${tag}_bench_ref_n$N.err; echo "ref N=$N rc=$?"
  timeout 600 $TR bench.py --gpus $N --steps 20 --warmup 3 > $out/${tag}_bench_n$N.json 2> $out/${tag}_bench_n$N.err; echo "bench N=$N rc=$?"
  python - <<PY
import json
for f in ("$out/${tag}_bench_ref_n$N.json", "$out/${tag}_bench_n$N.json"):
    for l in open(f):
        if l.startswith("{"):
            d = json.loads(l)
            print(d.get("impl", "ours"), d["n_gpus"], round(d["value"] / 1e9, 1), "Gelem/s", d.get("roofline", {}).get("achieved"),
                  "e2e", round(d["e2e"]["value"] / 1e9, 2), d["e2e"].get("roofline"), "parity_ok", d.get("parity_ok"), d.get("clocks"))
            for c in d.get("configs", []):
                print("   ", c.get("config"), c.get("op", ""), c.get("gbs_per_gpu"), c.get("frac_spec_8000"), c.get("parity_ok"))
PY
done
timeout 400 python -m pytest tests -q -m gpu -k "sharded or multi_device or devices or scatter or threads" > $out/${tag}_pytest_multi.log 2>&1
tail -2 $out/${tag}_pytest_multi.log
