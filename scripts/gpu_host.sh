#!/bin/bash
# host-path change: full GPU suite + cfg2 / cfg5 / cfg1 / cfg4 bench
OUT=gpurun_out/${1:-host}; mkdir -p $OUT
timeout 900 python -m pytest tests -m gpu -x -q 2>&1 | tail -4
for wl in cfg2 cfg5 cfg1 cfg4; do
  timeout 600 python bench.py --workload $wl --steps 2 --warmup 3 --no-cpu > $OUT/bench_$wl.json 2> $OUT/bench_$wl.err; echo "$wl exit $?"
done
python - <<PY
import json
for wl in ('cfg2','cfg5','cfg1','cfg4'):
    try:
        d=json.load(open('$OUT/bench_%s.json'%wl)); print(wl, 'value %.4g'%d['value'], 'e2e %.4g'%d['e2e']['value'], d['roofline']['kernel'], '%.4f'%d['roofline']['frac'], d['checksum_logp'], d['e2e_checksum_logp'])
    except Exception as e: print(wl, 'ERR', e)
PY
