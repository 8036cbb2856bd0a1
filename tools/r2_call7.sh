#!/bin/bash
set -u
O=gpurun_out; mkdir -p $O
python tools/gnn_bwd_only.py > $O/gnn_only.log 2>&1 && \
ncu --metrics gpu__time_duration.sum,launch__registers_per_thread,smsp__inst_executed.sum,sm__inst_executed.avg.per_cycle_elapsed,l1tex__data_bank_conflicts_pipe_lsu_mem_shared.sum,smsp__average_warps_issue_stalled_barrier_per_issue_active.ratio,smsp__average_warps_issue_stalled_short_scoreboard_per_issue_active.ratio,smsp__average_warps_issue_stalled_long_scoreboard_per_issue_active.ratio,smsp__average_warps_issue_stalled_wait_per_issue_active.ratio,smsp__average_warps_issue_stalled_math_pipe_throttle_per_issue_active.ratio,l1tex__t_bytes_pipe_lsu_mem_local_op_ld.sum,l1tex__t_bytes_pipe_lsu_mem_local_op_st.sum,sm__warps_active.avg.pct_of_peak_sustained_active,smsp__average_warps_issue_stalled_mio_throttle_per_issue_active.ratio,smsp__average_warps_issue_stalled_lg_throttle_per_issue_active.ratio,smsp__average_warps_issue_stalled_no_instruction_per_issue_active.ratio,smsp__average_warps_issue_stalled_dispatch_stall_per_issue_active.ratio,smsp__average_warps_issue_stalled_not_selected_per_issue_active.ratio,launch__occupancy_limit_registers,launch__occupancy_limit_shared_mem,launch__waves_per_multiprocessor,l1tex__data_pipe_lsu_wavefronts_mem_shared.sum --clock-control none --csv --log-file $O/r2_gnn_bwd_launches.csv python tools/gnn_bwd_only.py > $O/ncu_gnn.log 2>&1
python - <<'PY'
import csv
rows = list(csv.DictReader(l for l in open("gpurun_out/r2_gnn_bwd_launches.csv") if l.startswith('"')))
agg = {}
for r in rows:
    k = (r["Kernel Name"][:50], r["Metric Name"])
    agg.setdefault(k, []).append(r["Metric Value"])
for k, v in agg.items():
    if "gnn_bwd_gather" in k[0] or "wgrad" in k[0]:
        print(k[0], k[1], v[-3:])
PY
