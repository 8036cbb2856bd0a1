#!/bin/bash
set -u
echo "unroll 2 (default)"; python tools/gnn_bwd_only.py
for u in 1 4; do cp gsc_b200/libgsc_b200.so /tmp/keep.so; cp tools/_variants/libgsc_u$u.so gsc_b200/libgsc_b200.so; echo "unroll $u"; python tools/gnn_bwd_only.py; cp /tmp/keep.so gsc_b200/libgsc_b200.so; done
