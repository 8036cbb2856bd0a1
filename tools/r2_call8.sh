#!/bin/bash
set -u
O=gpurun_out; mkdir -p $O
for wl in c2 c5 c4; do
  timeout 600 python bench.py --workload $wl --steps 10 --warmup 3 > $O/r2_bench_${wl}_a.json 2> $O/r2_bench_${wl}_a.err; echo "bench $wl rc=$?"
  tail -3 $O/r2_bench_${wl}_a.err
done
python - <<'PY'
import json
for f in ("r2_bench_c2_a", "r2_bench_c5_a", "r2_bench_c4_a"):
    try:
        d = json.loads(open(f"gpurun_out/{f}.json").read().strip().splitlines()[-1])
        print(f, d["ms_per_step"], d["value"], d["e2e"]["value"], d["gpu_launches"], d["roofline"]["kernel"], d["roofline"]["frac"], d["roofline"]["step_frac_of_hbm_roofline"]); print(d["per_op_ms"]); print(d["cpu_baseline"])
    except Exception as e:
        print(f, "failed", e)
PY
nvidia-smi --query-gpu=memory.used,memory.total --format=csv
