#!/bin/bash
OUT=gpurun_out/${1:-exp}; mkdir -p $OUT
for n in 1 2 4 8 16 32; do
  echo "== cfg2 split=$n" | tee -a $OUT/exp.log
  KLM_TC_SPLIT=$n timeout 200 python scripts/tc_trace.py cfg2 3000 2>&1 | grep -E "grid|REC l2|IN l2" | tee -a $OUT/exp.log
done
