#!/bin/bash
# 8-GPU box: bench.py at N=1,2,4,8 (torchrun for N>1), plus the multi-device tests.
tag=${1:-r01s}
out=gpurun_out
mkdir -p $out
nvidia-smi topo -m > $out/${tag}_topo.txt 2>&1
timeout 300 python bench.py --no-cpu > $out/${tag}_bench_n1.json 2> $out/${tag}_bench_n1.err
for n in 2 4 8; do
  timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port $((29500+n)) \
     bench.py --gpus $n --steps 20 --warmup 3 > $out/${tag}_bench_n$n.json 2> $out/${tag}_bench_n$n.err
done
timeout 600 python -m pytest tests/test_gpu_dense.py tests/test_gpu_host_and_cpp.py -q -m gpu -k "sharded or multi_device or devices" > $out/${tag}_pytest_multi.log 2>&1
tail -3 $out/${tag}_pytest_multi.log
for n in 1 2 4 8; do python - <<PY
import json
for l in open("$out/${tag}_bench_n$n.json"):
    if l.startswith("{"):
        d=json.loads(l); print(d["n_gpus"], round(d["value"]/1e9,1), "Gelem/s", round(d["roofline"]["achieved"]), "GB/s/GPU", "e2e", round(d["e2e"]["value"]/1e9,1), d["clocks"])
PY
done
