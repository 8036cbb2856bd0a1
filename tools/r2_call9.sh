#!/bin/bash
set -u
O=gpurun_out; mkdir -p $O
timeout 600 python -m pytest tests/test_gpu_dp.py -x -q > $O/r2_tests_dp.log 2>&1; echo "pytest dp rc=$?"; tail -5 $O/r2_tests_dp.log
for wl in c3 c4; do
  timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus 2 --workload $wl --steps 10 --warmup 3 > $O/r2_bench_${wl}_dp2.json 2> $O/r2_bench_${wl}_dp2.err; echo "bench $wl dp2 rc=$?"
  tail -3 $O/r2_bench_${wl}_dp2.err
done
python - <<'PY'
import json
for f in ("r2_bench_c3_dp2", "r2_bench_c4_dp2"):
    try:
        d = json.loads(open(f"gpurun_out/{f}.json").read().strip().splitlines()[-1])
        print(f, d["ms_per_step"], d["value"], d["e2e"]["value"], d["gpu_launches"], d.get("dp_parity"), d["config"]["parallelism"])
    except Exception as e:
        print(f, "failed", e)
PY
