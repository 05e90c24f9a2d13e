#!/bin/bash
# relaxed vs release signalling of the REC publication: parity tests, then cfg2 / cfg5 with both settings
OUT=gpurun_out/exp11; mkdir -p $OUT
timeout 900 python -m pytest tests -m gpu -x -q > $OUT/pytest.log 2>&1; echo "pytest exit $?" >> $OUT/pytest.log
for rel in 0 1; do
  KLM_TC_PUB_RELEASE=$rel timeout 600 python bench.py --steps 3 --warmup 3 --no-cpu > $OUT/cfg2_rel$rel.json 2> $OUT/cfg2_rel$rel.err
  KLM_TC_PUB_RELEASE=$rel timeout 600 python bench.py --workload cfg5 --steps 3 --warmup 3 --no-cpu > $OUT/cfg5_rel$rel.json 2> $OUT/cfg5_rel$rel.err
done
tail -3 $OUT/pytest.log
for f in $OUT/cfg*_rel*.json; do echo $f; python - "$f" <<'PY'
import json,sys
try:
    j=json.load(open(sys.argv[1])); print(j['value'], j['e2e']['value'], j['roofline']['kernel_ms'], j['checksum_logp'])
except Exception as e: print('bad', e)
PY
done
