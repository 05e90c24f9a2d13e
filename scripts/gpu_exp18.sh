#!/bin/bash
OUT=gpurun_out/exp18; mkdir -p $OUT
KLM_TC_KC128=0 timeout 600 python bench.py --steps 3 --warmup 3 --no-cpu > $OUT/cfg2_kc64.json 2> $OUT/cfg2_kc64.err
timeout 600 python bench.py --steps 3 --warmup 3 --no-cpu > $OUT/cfg2_kc128.json 2> $OUT/cfg2_kc128.err
KLM_TC_KC128=0 timeout 300 python scripts/tc_trace.py cfg2 3000 > $OUT/timeline_kc64.txt 2>&1
for f in $OUT/cfg2_kc*.json; do echo $f; python - "$f" <<'PY'
import json,sys
try:
    j=json.load(open(sys.argv[1])); print(j['value'], j['e2e']['value'], j['roofline']['kernel_ms'], j['checksum_logp'])
except Exception as e: print('bad', e)
PY
done
tail -8 $OUT/timeline_kc64.txt
