#!/bin/bash
# publication modes of the hidden vector (KLM_TC_PUB) x consumer-side proxy fence: parity tests + timeline
OUT=gpurun_out/${1:-pub}; mkdir -p $OUT
for cfg in "1 1" "2 1" "2 0" "1 0" "0 1"; do
  set -- $cfg
  export KLM_TC_PUB=$1 KLM_TC_CONSUMER_FENCE=$2
  echo "== pub=$1 consumer_fence=$2" | tee -a $OUT/exp.log
  timeout 300 python -m pytest tests/test_gpu_tc.py -m gpu -x -q 2>&1 | tail -3 | tee -a $OUT/exp.log
  KLM_TC_PLAN_DEBUG=1 timeout 200 python scripts/tc_trace.py cfg2 3000 2>&1 | grep -E "plan|grid|REC|IN l|PROJ" | tee -a $OUT/exp.log
done
