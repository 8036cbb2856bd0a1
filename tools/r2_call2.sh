#!/bin/bash
set -u
O=gpurun_out; mkdir -p $O
timeout 900 python -m pytest tests -m gpu -x -q > $O/r2_tests2.log 2>&1; echo "pytest rc=$?"
tail -8 $O/r2_tests2.log
timeout 300 python tools/gemm_probe.py > $O/r2_gemm_probe2.log 2>&1; echo "probe rc=$?"
cat $O/r2_gemm_probe2.log
timeout 300 python bench.py --precision tf32 > $O/r2_bench_c3_tf32_b.json 2> $O/r2_bench_c3_tf32_b.err; echo "bench tf32 rc=$?"
timeout 300 python bench.py --precision tf32x3 > $O/r2_bench_c3_x3_b.json 2> $O/r2_bench_c3_x3_b.err; echo "bench x3 rc=$?"
python - <<'PY'
import json
for f in ("r2_bench_c3_tf32_b", "r2_bench_c3_x3_b"):
    try:
        d = json.loads(open(f"gpurun_out/{f}.json").read().strip().splitlines()[-1])
        print(f, d["ms_per_step"], d["value"], d["e2e"]["value"], d["roofline"]["per_kernel_ms"])
    except Exception as e:
        print(f, "failed", e)
PY
tail -5 $O/r2_bench_c3_x3_b.err
