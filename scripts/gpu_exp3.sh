#!/bin/bash
OUT=gpurun_out/${1:-exp3}; mkdir -p $OUT
timeout 300 python -m pytest tests/test_gpu_tc.py -m gpu -x -q 2>&1 | tail -3 | tee -a $OUT/exp.log
for cfg in "1 1 12" "2 0 12" "2 0 11" "2 1 11" "1 1 11" "0 1 10" "0 0 10"; do
  set -- $cfg
  export KLM_TC_PUB=$1 KLM_TC_CONSUMER_FENCE=$2 KLM_TC_PMAX=$3
  echo "== pub=$1 consumer_fence=$2 pmax=$3" | tee -a $OUT/exp.log
  KLM_TC_PLAN_DEBUG=1 timeout 200 python scripts/tc_trace.py cfg2 3000 2>&1 | grep -E "plan: [RIPL]|grid|REC|IN l|PROJ" | tee -a $OUT/exp.log
done
export KLM_TC_PUB=2 KLM_TC_CONSUMER_FENCE=0 KLM_TC_PMAX=11
for mode in 1 2; do
  echo "== trace mode $mode" | tee -a $OUT/exp.log
  KLM_TC_TRACE_MODE=$mode timeout 200 python scripts/tc_trace.py cfg2 3000 2>&1 | grep -E "grid|REC|IN l|PROJ" | tee -a $OUT/exp.log
done
