#!/bin/bash
OUT=gpurun_out/${1:-exp5}; mkdir -p $OUT
export KLM_TC_PUB=0 KLM_TC_CONSUMER_FENCE=0 KLM_TC_PMAX=10
for k in 1 2 3; do
  export KLM_TC_DEBUG_SKIP=$k
  echo "== debug_skip $k" | tee -a $OUT/exp.log
  timeout 200 python scripts/tc_trace.py cfg2 3000 2>&1 | grep -E "grid|REC|IN l|PROJ" | tee -a $OUT/exp.log
  KLM_TC_TRACE_MODE=1 timeout 200 python scripts/tc_trace.py cfg2 3000 2>&1 | grep -E "REC l1|IN l1" | tee -a $OUT/exp.log
  KLM_TC_TRACE_MODE=2 timeout 200 python scripts/tc_trace.py cfg2 3000 2>&1 | grep -E "REC l1|IN l1" | tee -a $OUT/exp.log
done
