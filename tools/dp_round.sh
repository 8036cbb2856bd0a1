#!/bin/bash
# N-GPU bench lines of the c3 workload (default: one all-reduce per network; --bucket-allreduce: per-layer
# all-reduces behind the embedder backward), results in gpurun_out/
set -u
N=${1:-8}
O=gpurun_out; mkdir -p $O
python __graft_entry__.py --smoke 2>&1 | tail -2
run() { tag=$1; shift; timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29533 bench.py --gpus $N "$@" > $O/r2_bench_${tag}_dp$N.json 2> $O/r2_bench_${tag}_dp$N.err; echo "$tag rc=$?"; }
run c3
run c3_bucket --bucket-allreduce
run c3_bucket16 --bucket-allreduce --sm-reserve 16
python - <<PY
import json
for f in ("c3", "c3_bucket", "c3_bucket16"):
    try:
        d = json.loads(open(f"gpurun_out/r2_bench_{f}_dp$N.json").read().strip().splitlines()[-1])
        print(f, d["n_gpus"], round(d["ms_per_step"], 4), round(d["value"]), round(d["e2e"]["value"]), d.get("dp_parity", {}).get("ok"))
    except Exception as e:
        print(f, "failed", e)
PY
