#!/bin/bash
set -u
O=gpurun_out; mkdir -p $O
timeout 300 python bench.py --timeline > $O/r2_bench_c3_tl.json 2> $O/r2_bench_c3_tl.err; echo "rc=$?"
grep timeline $O/r2_bench_c3_tl.err
