#!/bin/bash
OUT=gpurun_out/${1:-exp9}; mkdir -p $OUT
export KLM_TC_CONSUMER_FENCE=0
timeout 300 python -m pytest tests/test_gpu_tc.py -m gpu -x -q 2>&1 | tail -3 | tee -a $OUT/exp.log
for cfg in "2 11 1 1" "2 11 1 2" "2 11 0 1" "2 12 1 2" "0 10 0 2"; do
  set -- $cfg
  export KLM_TC_PUB=$1 KLM_TC_PMAX=$2 KLM_TC_KC128=$3 KLM_TC_NLSE=$4
  echo "== pub=$1 pmax=$2 kc128=$3 nlse=$4" | tee -a $OUT/exp.log
  timeout 300 python -m pytest tests/test_gpu_tc.py -m gpu -x -q 2>&1 | tail -1 | tee -a $OUT/exp.log
  KLM_TC_PLAN_DEBUG=1 timeout 200 python scripts/tc_trace.py cfg2 3000 2>&1 | grep -E "plan: [RIPL]|grid|REC|IN l|PROJ|LSE" | tee -a $OUT/exp.log
done
