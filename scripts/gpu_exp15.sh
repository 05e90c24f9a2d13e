#!/bin/bash
OUT=gpurun_out/exp15; mkdir -p $OUT
timeout 900 python -m pytest tests -m gpu -x -q > $OUT/pytest.log 2>&1; echo "pytest exit $?" >> $OUT/pytest.log
for k in 1 0; do
  KLM_TC_GO_LATE=$k timeout 600 python bench.py --steps 3 --warmup 3 --no-cpu > $OUT/cfg2_g$k.json 2> $OUT/cfg2_g$k.err
  KLM_TC_GO_LATE=$k timeout 600 python bench.py --workload cfg5 --steps 3 --warmup 3 --no-cpu > $OUT/cfg5_g$k.json 2> $OUT/cfg5_g$k.err
done
timeout 300 python scripts/tc_trace.py cfg2 3000 > $OUT/timeline.txt 2>&1
tail -4 $OUT/pytest.log
for f in $OUT/cfg*_g*.json; do echo $f; python - "$f" <<'PY'
import json,sys
try:
    j=json.load(open(sys.argv[1])); print(j['value'], j['e2e']['value'], j['roofline']['kernel_ms'], j['checksum_logp'])
except Exception as e: print('bad', e)
PY
done
tail -8 $OUT/timeline.txt
