#!/bin/bash
set -u
O=gpurun_out; mkdir -p $O
timeout 900 python -m pytest tests -m gpu -x -q > $O/r2_tests5.log 2>&1; echo "pytest rc=$?"
tail -8 $O/r2_tests5.log
timeout 120 python tools/tf32_trunc_check.py; timeout 300 python tools/gemm_probe.py > $O/r2_gemm_probe5.log 2>&1; echo "probe rc=$?"
cat $O/r2_gemm_probe5.log
timeout 300 python bench.py --precision tf32 > $O/r2_bench_c3_tf32_d.json 2> $O/r2_bench_c3_tf32_d.err; echo "bench tf32 rc=$?"
timeout 300 python bench.py --precision tf32x3 > $O/r2_bench_c3_x3_d.json 2> $O/r2_bench_c3_x3_d.err; echo "bench x3 rc=$?"
python - <<'PY'
import json
for f in ("r2_bench_c3_tf32_d", "r2_bench_c3_x3_d"):
    try:
        d = json.loads(open(f"gpurun_out/{f}.json").read().strip().splitlines()[-1])
        print(f, d["ms_per_step"], d["value"], d["e2e"]["value"], d["roofline"]["per_entry_point_family_ms"])
    except Exception as e:
        print(f, "failed", e)
PY
tail -5 $O/r2_bench_c3_x3_d.err
