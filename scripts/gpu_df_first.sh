#!/bin/bash
# scratch: a workload under combinations of environment switches (one bench line each)
#   usage: gpu_df_first.sh <tag> <workload> "<ENV=..>" ...
OUT=gpurun_out/$1; shift
WL=$1; shift
mkdir -p $OUT
B="python bench.py --workload $WL --steps 2 --warmup 3 --no-cpu"
i=0
for V in "$@"; do
  i=$((i+1))
  env $V timeout 300 $B > $OUT/${WL}_v$i.json 2> $OUT/${WL}_v$i.err
  python - "$V" $OUT/${WL}_v$i.json <<'PY'
import json,sys
try:
    j=json.loads(open(sys.argv[2]).read().strip().splitlines()[-1])
    print('%-60s %.3f M  e2e %.3f  ms %.1f  checksum %.3f' % (sys.argv[1], j['value']/1e6, j['e2e']['value']/1e6, j['ms_per_step'], j.get('checksum_logp', 0)))
except Exception as e:
    print(sys.argv[1], 'FAILED', e)
PY
done
