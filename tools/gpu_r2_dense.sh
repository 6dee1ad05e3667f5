#!/bin/bash
# dense side after round 2's changes: parity (all dense tests + narrow planes), bench_dense table, the narrow-plane rows
tag=${1:-r14}
out=gpurun_out
mkdir -p $out
timeout 1500 python -m pytest tests/test_gpu_dense.py tests/test_gpu_narrow_planes.py tests/test_gpu_dense_fullsize.py -x -q -m gpu > $out/${tag}_pytest_dense.log 2>&1; echo "pytest rc=$?" >> $out/${tag}_pytest_dense.log
tail -5 $out/${tag}_pytest_dense.log
timeout 900 python tools/bench_dense.py 29 10 > $out/${tag}_dense.jsonl 2> $out/${tag}_dense.err; echo "dense rc=$?"
python - <<PY
import json
for l in open("$out/${tag}_dense.jsonl"):
    try: d = json.loads(l)
    except Exception: continue
    print({k: d[k] for k in d if k in ("config","op","layout","kernel","gbs_median","gelem_s_median")})
PY
timeout 900 python tools/bench_configs.py extra > $out/${tag}_configs_extra.jsonl 2> $out/${tag}_configs_extra.err; echo "extra rc=$?"
python - <<PY
import json
for l in open("$out/${tag}_configs_extra.jsonl"):
    d = json.loads(l); print(d["config"][:60], d.get("op",""), d.get("layout",""), d["gbs_median"])
PY
