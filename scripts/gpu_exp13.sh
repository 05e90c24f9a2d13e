#!/bin/bash
# skew 3 (early grid-barrier signal of the IN / PROJ CTAs): parity tests, cfg2 / cfg4 / cfg5 / cfg3 against skew 2
OUT=gpurun_out/exp13; mkdir -p $OUT
timeout 900 python -m pytest tests -m gpu -x -q > $OUT/pytest.log 2>&1; echo "pytest exit $?" >> $OUT/pytest.log
for k in 3 2; do
  KLM_TC_SKEW=$k timeout 600 python bench.py --steps 3 --warmup 3 --no-cpu > $OUT/cfg2_k$k.json 2> $OUT/cfg2_k$k.err
  KLM_TC_SKEW=$k timeout 600 python bench.py --workload cfg5 --steps 3 --warmup 3 --no-cpu > $OUT/cfg5_k$k.json 2> $OUT/cfg5_k$k.err
done
KLM_TC_ZPUB=0 timeout 600 python bench.py --steps 3 --warmup 3 --no-cpu > $OUT/cfg2_k3z0.json 2> $OUT/cfg2_k3z0.err
timeout 600 python bench.py --workload cfg4 --steps 3 --warmup 3 --no-cpu > $OUT/cfg4_k3.json 2> $OUT/cfg4_k3.err
timeout 300 python scripts/tc_trace.py cfg2 3000 > $OUT/timeline.txt 2>&1
tail -5 $OUT/pytest.log
for f in $OUT/cfg*_k*.json; do echo $f; python - "$f" <<'PY'
import json,sys
try:
    j=json.load(open(sys.argv[1])); print(j['value'], j['e2e']['value'], j['roofline']['kernel_ms'], j['checksum_logp'])
except Exception as e: print('bad', e)
PY
done
tail -8 $OUT/timeline.txt
