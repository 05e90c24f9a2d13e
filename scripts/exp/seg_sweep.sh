#!/bin/bash
# cfg5 end to end under different segment sizes of Rater.rate_many (host/device overlap vs launches)
for M in 250000 120000 60000 30000; do
python - $M <<'PY' 2>/dev/null | python -c "
import sys, json
j = json.loads(sys.stdin.read().strip().splitlines()[-1])
print('MIN_SEGMENT_POSITIONS $M: device %.3f M, e2e %.3f M' % (j['value'] / 1e6, j['e2e']['value'] / 1e6))"
import sys
sys.path.insert(0, '.')
import ocrd_keraslm_b200.lib.rating as r
r.Rater.MIN_SEGMENT_POSITIONS = int(sys.argv[1])
import bench
sys.argv = ['bench.py', '--workload', 'cfg5', '--steps', '3', '--warmup', '3', '--no-cpu']
bench.main()
PY
done
