#!/bin/bash
set -u
O=gpurun_out; mkdir -p $O
timeout 900 python -m pytest tests -m gpu -x -q > $O/r2_tests16.log 2>&1; echo "pytest rc=$?"
tail -8 $O/r2_tests16.log
for v in 0 1; do
GSC_GNN_BWD=$v timeout 300 python bench.py > $O/r2_bench_c3_x3_f$v.json 2> $O/r2_bench_c3_x3_f$v.err; echo "bench variant $v rc=$?"
done
python - <<'PY'
import json
for f in ("r2_bench_c3_x3_f0","r2_bench_c3_x3_f1"):
    try:
        d = json.loads(open(f"gpurun_out/{f}.json").read().strip().splitlines()[-1])
        print(f, d["ms_per_step"], d["value"], d["e2e"]["value"], d["gpu_launches"], d["roofline"]["per_entry_point_family_ms"])
    except Exception as e:
        print(f, "failed", e)
PY
