#!/bin/bash
OUT=gpurun_out/${1:-exp2}; mkdir -p $OUT
timeout 100 scripts/ubench/mma_rate | tee $OUT/mma_rate.txt
export KLM_TC_PUB=2 KLM_TC_CONSUMER_FENCE=0
for mode in 1 2; do
  echo "== trace mode $mode" | tee -a $OUT/exp.log
  KLM_TC_TRACE_MODE=$mode timeout 200 python scripts/tc_trace.py cfg2 3000 2>&1 | grep -E "grid|REC|IN l|PROJ" | tee -a $OUT/exp.log
done
echo "== pmax 11" | tee -a $OUT/exp.log
KLM_TC_PMAX=11 KLM_TC_PLAN_DEBUG=1 timeout 200 python scripts/tc_trace.py cfg2 3000 2>&1 | grep -E "plan|grid|REC|IN l|PROJ" | tee -a $OUT/exp.log
