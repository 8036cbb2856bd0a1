#!/bin/bash
# one bench line per GEMM ring configuration (env knobs of csrc/gemm_tc.cu)
for ares in 1 0; do for epi in 4 5; do for sp in 5 6 7; do
  GSC_GEMM_NO_ARES=$ares GSC_GEMM_STAGES_EPI=$epi GSC_GEMM_STAGES_SPLIT=$sp timeout 100 python bench.py --batch 1024 --steps 10 2>&1 | tail -1 | python -c "import sys,json; d=json.loads(sys.stdin.read()); po=d['per_op_ms']; print('no_ares=$ares epi=$epi split=$sp', round(d['ms_per_step'],4), {k:v for k,v in po.items() if k.startswith('linear')})"
done; done; done
