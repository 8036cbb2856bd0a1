#!/bin/bash
# gnn backward only: timing, then the embedder parity tests
set -u
python tools/gnn_bwd_only.py
timeout 600 python -m pytest tests -m gpu -x -q -k "gnn or embed or backward or update" 2>&1 | tail -3
