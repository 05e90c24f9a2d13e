#!/bin/bash
OUT=gpurun_out/${1:-exp10}; mkdir -p $OUT
for nl in 1 2; do
  export KLM_TC_NLSE=$nl
  echo "== defaults, nlse=$nl" | tee -a $OUT/exp.log
  KLM_TC_PLAN_DEBUG=1 timeout 200 python scripts/tc_trace.py cfg2 3000 2>&1 | grep -E "plan: [RIPL]|grid|REC|IN l|PROJ|LSE" | tee -a $OUT/exp.log
done
unset KLM_TC_NLSE
timeout 900 python -m pytest tests -m gpu -x -q 2>&1 | tail -3 | tee -a $OUT/exp.log
timeout 600 python bench.py --steps 2 --warmup 3 > $OUT/bench.json 2> $OUT/bench.err; echo "bench exit $?"; cat $OUT/bench.json
