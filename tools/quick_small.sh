#!/bin/bash
# small-batch loop: DDPG parity tests, then c3 at B = 100 and B = 1024
set -u
timeout 600 python -m pytest tests -m gpu -x -q -k "ddpg or update or smask or table or graph" 2>&1 | tail -2
for b in 100 1024; do timeout 300 python bench.py --batch $b 2>/dev/null | python -c "import json,sys; d=json.loads(sys.stdin.read().strip().splitlines()[-1]); print('B=$b', d['ms_per_step'], d['gpu_launches'], d['roofline']['per_entry_point_family_ms'].get('mask_table'))"; done
