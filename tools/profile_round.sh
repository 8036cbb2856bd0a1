#!/bin/bash
# Round-2 profile: bench lines (no profiler), then the ncu launch list and one --set full capture of the hot
# kernels of the same command.  Everything lands in gpurun_out/ (scratch); the summaries are copied into
# profiles/ by hand (see profiles/r2_summary_c3.md).
set -u
O=gpurun_out
mkdir -p $O
python bench.py > $O/r2_bench_c3_b1024.json 2> $O/r2_bench_c3_b1024.err || exit 1
python bench.py --batch 100 > $O/r2_bench_c3_b100.json 2>/dev/null
python bench.py --batch 4096 > $O/r2_bench_c3_b4096.json 2>/dev/null
python bench.py --precision tf32 > $O/r2_bench_c3_b1024_tf32.json 2>/dev/null
python bench.py --precision fp32 --steps 10 > $O/r2_bench_c3_b1024_fp32.json 2>/dev/null
python bench.py --workload c2 > $O/r2_bench_c2.json 2>/dev/null
python bench.py --workload c5 --steps 10 > $O/r2_bench_c5.json 2>/dev/null
python bench.py --workload c4 --steps 10 > $O/r2_bench_c4.json 2>/dev/null
python bench.py --impl reference --steps 5 --warmup 1 > $O/r2_bench_reference.json 2>/dev/null
# launch list (cold-cache, serialised: compare shares, not absolutes)
python bench.py --no-graph --steps 2 --warmup 3 > $O/plain_c3.log 2>&1 && \
  timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 1500 --csv --log-file $O/r2_launches_c3.csv \
    python bench.py --no-graph --steps 2 --warmup 3 > $O/ncu_c3.log 2>&1
# full capture of the hot kernels of one c3 step (after the three warm-up steps)
timeout 1200 ncu --set full --clock-control none --import-source on \
  -k regex:"gemm_tc_kernel|gnn_bwd_gather_kernel|gnn_wgrad_kernel|gnn_fwd_tc_kernel|adam_soft_dev|mask_table|mask_wgrad|colsum" -s 108 -c 72 \
  -o $O/r2_full_c3 -f python bench.py --no-graph --steps 2 --warmup 3 > $O/ncu_full_c3.log 2>&1
# the report itself is too large to travel back (64 MiB limit): keep its CSV pages and the summary
ncu -i $O/r2_full_c3.ncu-rep --page raw --csv > $O/r2_ncu_full_c3_raw.csv 2>/dev/null
python tools/ncu_traffic.py $O/r2_ncu_full_c3_raw.csv c3_b1024_tf32x3 > $O/r2_full_c3_summary.md 2>&1
cp profiles/r2_traffic.json $O/r2_traffic.json
ncu -i $O/r2_full_c3.ncu-rep --page source --csv -k regex:gnn_bwd_gather_kernel -c 1 > $O/r2_src_gnn_bwd_gather.csv 2>/dev/null
rm -f $O/r2_full_c3.ncu-rep
python tools/gemm_probe.py none > $O/r2_gemm_roles.log 2>&1
python tools/gnn_bwd_only.py > $O/r2_gnn_entry.log 2>&1
ls -la $O | tail -24
