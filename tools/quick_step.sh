#!/bin/bash
# update-step loop: all GPU tests, then the c3 bench line with / without the given extra flags
set -u
O=gpurun_out; mkdir -p $O
timeout 900 python -m pytest tests -m gpu -x -q 2>&1 | tail -3
timeout 300 python bench.py > $O/q_a.json 2> $O/q_a.err; echo "bench rc=$?"
timeout 300 python bench.py "$@" > $O/q_b.json 2> $O/q_b.err; echo "bench $* rc=$?"
python - <<'PY'
import json
for f in ("q_a", "q_b"):
    d = json.loads(open(f"gpurun_out/{f}.json").read().strip().splitlines()[-1])
    print(f, d["ms_per_step"], d["value"], d["e2e"]["value"], d["gpu_launches"], d["roofline"]["per_entry_point_family_ms"])
PY
