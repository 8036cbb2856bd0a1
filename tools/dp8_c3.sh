#!/bin/bash
# 8-GPU c3 bench line (default data-parallel mode), result in gpurun_out/
set -u
N=${1:-8}
O=gpurun_out; mkdir -p $O
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29533 bench.py --gpus $N ${DP_FLAGS:-} > $O/r2_bench_c3_dp$N.json 2> $O/r2_bench_c3_dp$N.err; echo "rc=$?"
python - <<PY
import json
d = json.loads(open("gpurun_out/r2_bench_c3_dp$N.json").read().strip().splitlines()[-1])
print(d["n_gpus"], round(d["ms_per_step"], 4), round(d["value"]), round(d["e2e"]["value"]), d.get("dp_parity", {}).get("ok"))
PY
