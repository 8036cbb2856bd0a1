#!/bin/bash
# GEMM work loop: dense-layer parity tests, then per-role timings (tools/gemm_probe.py)
set -u
timeout 600 python -m pytest tests -m gpu -x -q -k "gemm or linear or tf32 or dense or smask or configs" 2>&1 | tail -3
python tools/gemm_probe.py "$@" > gpurun_out/probe.log 2>&1
grep -v "dbg= [1-9]\|dbg=[1-9][0-9]" gpurun_out/probe.log | tail -40
