#!/bin/bash
# bulk-store publication of Z / logits (IN, PROJ): parity tests, cfg2 / cfg4 with and without, timeline
OUT=gpurun_out/exp12; mkdir -p $OUT
timeout 900 python -m pytest tests -m gpu -x -q > $OUT/pytest.log 2>&1; echo "pytest exit $?" >> $OUT/pytest.log
for z in 1 0; do
  KLM_TC_ZPUB=$z timeout 600 python bench.py --steps 3 --warmup 3 --no-cpu > $OUT/cfg2_z$z.json 2> $OUT/cfg2_z$z.err
done
timeout 600 python bench.py --workload cfg4 --steps 3 --warmup 3 --no-cpu > $OUT/cfg4_z1.json 2> $OUT/cfg4_z1.err
KLM_TC_PLAN_DEBUG=1 timeout 300 python scripts/tc_trace.py cfg2 3000 > $OUT/timeline.txt 2>&1
tail -3 $OUT/pytest.log
for f in $OUT/cfg*_z*.json; do echo $f; python - "$f" <<'PY'
import json,sys
try:
    j=json.load(open(sys.argv[1])); print(j['value'], j['e2e']['value'], j['roofline']['kernel_ms'], j['checksum_logp'])
except Exception as e: print('bad', e)
PY
done
tail -9 $OUT/timeline.txt
