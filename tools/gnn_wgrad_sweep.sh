#!/bin/bash
# gnn_wgrad_kernel grid sweep (CTAs per SM; 0 = one CTA per gather-kernel CTA) + embedder parity tests
for c in 0 1 2 3; do echo "GSC_GNN_WGRAD_CTAS=$c"; GSC_GNN_WGRAD_CTAS=$c python tools/gnn_bwd_only.py 2>&1 | grep gnn_bwd; done
timeout 600 python -m pytest tests -m gpu -x -q -k "gnn or embed or backward or update" 2>&1 | tail -2
