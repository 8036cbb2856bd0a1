#!/bin/bash
set -u
O=gpurun_out; mkdir -p $O
echo "== x3 tests, truncation split"; timeout 600 python -m pytest tests/test_gpu_tf32x3.py -q -s 2>&1 | grep -E "err|passed|failed"
echo "== x3 tests, rounded in-place split"; GSC_TC_DBG=16 timeout 600 python -m pytest tests/test_gpu_tf32x3.py -q -s 2>&1 | grep -E "err|passed|failed"
timeout 300 python tools/gemm_probe.py 2>&1 | grep -E "handshake|in-place|tf32x3  "
