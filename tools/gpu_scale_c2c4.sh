#!/bin/bash
tag=${1:-r02m}
out=gpurun_out
mkdir -p $out
: > $out/${tag}_c2c4.jsonl
timeout 300 python tools/bench_multi.py >> $out/${tag}_c2c4.jsonl 2> $out/${tag}_c2c4_n1.err
for n in 2 4 8; do
  timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port $((29600+n)) \
     tools/bench_multi.py >> $out/${tag}_c2c4.jsonl 2> $out/${tag}_c2c4_n$n.err
done
cat $out/${tag}_c2c4.jsonl | cut -c1-260
