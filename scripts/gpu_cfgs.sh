#!/bin/bash
# the other BASELINE configs on one GPU (short runs)
OUT=gpurun_out/${1:-cfgs}; mkdir -p $OUT
for wl in cfg1 cfg4 cfg5; do
  timeout 600 python bench.py --workload $wl --steps 2 --warmup 3 > $OUT/bench_$wl.json 2> $OUT/bench_$wl.err; echo "$wl exit $?"; tail -2 $OUT/bench_$wl.err
done
timeout 600 python bench.py --workload cfg3 --steps 1 --warmup 3 --docs 2 > $OUT/bench_cfg3.json 2> $OUT/bench_cfg3.err; echo "cfg3 exit $?"; tail -3 $OUT/bench_cfg3.err
timeout 600 python bench.py --workload cfg3 --impl reference --steps 1 > $OUT/bench_cfg3_ref.json 2> $OUT/bench_cfg3_ref.err
python - <<'PY'
import json,glob,sys
for f in sorted(glob.glob(sys.argv[1] if len(sys.argv)>1 else 'gpurun_out/*/bench_cfg*.json')):
    try:
        d=json.load(open(f))
        print(f.split('/')[-1], d['metric'], '%.4g'%d['value'], d['unit'], 'e2e %.4g'%d['e2e']['value'], 'cpu', d.get('cpu_baseline',{}).get('value'), 'roof', d.get('roofline',{}).get('kernel'), d.get('roofline',{}).get('frac'), d.get('mean_hypotheses_per_step'))
    except Exception as e:
        print(f, 'ERR', e)
PY
