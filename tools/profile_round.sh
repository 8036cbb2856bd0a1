#!/bin/bash
# Round profile: bench lines (no profiler), then ncu launch lists and one --set full capture of the
# hot kernels of the same commands.  Everything lands in gpurun_out/ (scratch); summaries are copied
# into profiles/ by hand.
set -u
O=gpurun_out
mkdir -p $O
python bench.py --batch 1024 > $O/bench_c3_b1024.json 2> $O/bench_c3_b1024.err || exit 1
python bench.py --batch 100 > $O/bench_c3_b100.json 2>/dev/null
python bench.py --batch 4096 > $O/bench_c3_b4096.json 2>/dev/null
python bench.py --workload c2 > $O/bench_c2.json 2>/dev/null
python bench.py --batch 1024 --precision fp32 > $O/bench_c3_b1024_fp32.json 2>/dev/null
python bench.py --impl reference --steps 3 --warmup 1 > $O/bench_reference.json 2>/dev/null
# launch lists (cold-cache, serialised: compare shares, not absolutes)
python bench.py --no-graph --steps 2 --warmup 3 > $O/plain_c3.log 2>&1 && \
  timeout 500 ncu --metrics gpu__time_duration.sum --clock-control none -c 1500 --csv --log-file $O/launches_c3.csv \
    python bench.py --no-graph --steps 2 --warmup 3 > $O/ncu_c3.log 2>&1
python bench.py --workload c2 --no-graph --steps 2 --warmup 3 > $O/plain_c2.log 2>&1 && \
  timeout 300 ncu --metrics gpu__time_duration.sum --clock-control none -c 600 --csv --log-file $O/launches_c2.csv \
    python bench.py --workload c2 --no-graph --steps 2 --warmup 3 > $O/ncu_c2.log 2>&1
# full capture of the hot kernels of one c3 step (after the three warm-up steps)
timeout 900 ncu --set full --clock-control none --import-source on \
  -k regex:"gemm_tc_kernel|gnn_bwd_tc_kernel|gnn_fwd_tc_kernel|adam_soft_dev|mask_table|mask_wgrad|colsum" -s 150 -c 50 \
  -o $O/full_c3 -f python bench.py --no-graph --steps 2 --warmup 3 > $O/ncu_full_c3.log 2>&1
# the report itself is too large to travel back (64 MiB limit): keep its CSV pages and the summary
ncu -i $O/full_c3.ncu-rep --page raw --csv > $O/full_c3_raw.csv 2>/dev/null
python tools/ncu_traffic.py $O/full_c3.ncu-rep > $O/full_c3_summary.md 2>&1
ncu -i $O/full_c3.ncu-rep --page source --csv -k regex:gnn_bwd_tc_kernel -c 1 > $O/full_c3_src_gnn_bwd.csv 2>/dev/null
ncu -i $O/full_c3.ncu-rep --page source --csv -k regex:gnn_fwd_tc_kernel -c 1 > $O/full_c3_src_gnn_fwd.csv 2>/dev/null
rm -f $O/full_c3.ncu-rep
ls -la $O | tail -24
