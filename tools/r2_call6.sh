#!/bin/bash
set -u
O=gpurun_out; mkdir -p $O
timeout 900 python -m pytest tests -m gpu -x -q > $O/r2_tests6.log 2>&1; echo "pytest rc=$?"
tail -8 $O/r2_tests6.log
timeout 300 python bench.py > $O/r2_bench_c3_x3_e.json 2> $O/r2_bench_c3_x3_e.err; echo "bench x3 rc=$?"
python - <<'PY'
import json
for f in ("r2_bench_c3_x3_e",):
    try:
        d = json.loads(open(f"gpurun_out/{f}.json").read().strip().splitlines()[-1])
        print(f, d["ms_per_step"], d["value"], d["e2e"]["value"], d["gpu_launches"], d["roofline"]["per_entry_point_family_ms"]); print(d["per_op_ms"])
    except Exception as e:
        print(f, "failed", e)
PY
tail -5 $O/r2_bench_c3_x3_e.err
