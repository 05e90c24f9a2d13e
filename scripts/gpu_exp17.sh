#!/bin/bash
OUT=gpurun_out/exp17; mkdir -p $OUT
timeout 900 python -m pytest tests -m gpu -x -q > $OUT/pytest.log 2>&1; echo "pytest exit $?" >> $OUT/pytest.log
KLM_STEP_TIMING=1 timeout 600 python bench.py --workload cfg3 --steps 1 --warmup 3 --docs 2 > $OUT/cfg3.json 2> $OUT/cfg3.err
tail -3 $OUT/pytest.log; grep "klm timing" $OUT/cfg3.err; cut -c1-160 $OUT/cfg3.json
