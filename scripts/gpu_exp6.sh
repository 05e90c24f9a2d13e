#!/bin/bash
OUT=gpurun_out/${1:-exp6}; mkdir -p $OUT
export KLM_TC_CONSUMER_FENCE=0
timeout 300 python -m pytest tests/test_gpu_tc.py -m gpu -x -q 2>&1 | tail -3 | tee -a $OUT/exp.log
for cfg in "0 10 0" "2 11 0" "2 11 1" "2 12 0" "1 11 0"; do
  set -- $cfg
  export KLM_TC_PUB=$1 KLM_TC_PMAX=$2 KLM_TC_KC128=$3
  echo "== pub=$1 pmax=$2 kc128=$3" | tee -a $OUT/exp.log
  timeout 300 python -m pytest tests/test_gpu_tc.py -m gpu -x -q 2>&1 | tail -1 | tee -a $OUT/exp.log
  timeout 200 python scripts/tc_trace.py cfg2 3000 2>&1 | grep -E "grid|REC|IN l|PROJ" | tee -a $OUT/exp.log
  KLM_TC_TRACE_MODE=1 timeout 200 python scripts/tc_trace.py cfg2 3000 2>&1 | grep -E "REC l1|IN l1" | tee -a $OUT/exp.log
done
