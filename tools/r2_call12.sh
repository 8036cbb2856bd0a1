#!/bin/bash
set -u
O=gpurun_out; mkdir -p $O
timeout 1500 python compat/run_reference.py --reference baseline/_ref/GSC --episodes 2 > $O/r2_reference_loop.log 2>&1; echo "run_reference rc=$?"
grep -E "compat\]|Error|error|Traceback" $O/r2_reference_loop.log | head -20; tail -15 $O/r2_reference_loop.log
timeout 900 python -m pytest tests/test_gpu_reference_loop.py -x -q 2>&1 | tail -5
