#!/bin/bash
# 2-GPU checks: NCCL parity tests, then c3 / c4 bench lines in the data-parallel variants
set -u
O=gpurun_out; mkdir -p $O
timeout 240 python -m pytest tests/test_gpu_dp.py -x -q 2>&1 | tail -3
run() { tag=$1; shift; timeout 400 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus 2 "$@" > $O/r2_bench_${tag}_dp2.json 2> $O/r2_bench_${tag}_dp2.err; echo "$tag rc=$?"; }
run c3
if [ "${DP_ALL:-0}" = "1" ]; then
  run c3_mlp --bucket-allreduce mlp
  run c3_shard --shard-optimizer
fi
run c4 --workload c4 --steps 10
python - <<'PY'
import json
for f in ("c3", "c3_mlp", "c3_shard", "c4"):
    try:
        d = json.loads(open(f"gpurun_out/r2_bench_{f}_dp2.json").read().strip().splitlines()[-1])
        print(f, round(d["ms_per_step"], 4), round(d["value"]), d.get("dp_parity", {}).get("ok"), d["config"]["parallelism"])
    except Exception as e:
        print(f, "failed", e)
PY
