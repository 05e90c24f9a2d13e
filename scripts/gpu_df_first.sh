#!/bin/bash
# scratch: cfg2 under combinations of environment switches (one bench line each)
OUT=gpurun_out/$1; shift
mkdir -p $OUT
B="python bench.py --steps 2 --warmup 3 --no-cpu"
i=0
for V in "$@"; do
  i=$((i+1))
  env $V timeout 300 $B > $OUT/v$i.json 2> $OUT/v$i.err
  python - "$V" $OUT/v$i.json <<'PY'
import json,sys
try:
    j=json.loads(open(sys.argv[2]).read().strip().splitlines()[-1])
    print('%-60s %.3f M chars/s  e2e %.3f  ms %.1f  checksum %.3f' % (sys.argv[1], j['value']/1e6, j['e2e']['value']/1e6, j['ms_per_step'], j['checksum_logp']))
except Exception as e:
    print(sys.argv[1], 'FAILED', e)
PY
done
