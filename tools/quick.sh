#!/bin/bash
set -u
O=gpurun_out; mkdir -p $O
python tools/gnn_bwd_only.py
GSC_GNN_FWD=1 python tools/gnn_bwd_only.py
timeout 900 python -m pytest tests -m gpu -x -q 2>&1 | tail -3
timeout 300 python bench.py > $O/r2_bench_quick.json 2> $O/r2_bench_quick.err; echo "bench rc=$?"
timeout 300 python bench.py --workload c5 --steps 10 > $O/r2_bench_quick_c5.json 2> $O/r2_bench_quick_c5.err; echo "bench c5 rc=$?"
timeout 300 python bench.py --workload c2 > $O/r2_bench_quick_c2.json 2> $O/r2_bench_quick_c2.err; echo "bench c2 rc=$?"
python - <<'PY'
import json
for f in ("r2_bench_quick", "r2_bench_quick_c5", "r2_bench_quick_c2"):
    d = json.loads(open(f"gpurun_out/{f}.json").read().strip().splitlines()[-1])
    print(f, d["ms_per_step"], d["value"], d["e2e"]["value"], d["gpu_launches"], d["roofline"]["per_entry_point_family_ms"])
PY
