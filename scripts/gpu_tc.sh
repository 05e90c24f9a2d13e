#!/bin/bash
# tensor-core kernel iteration: parity tests, timeline, bench -- each under its own timeout
TAG=${1:-tc}
OUT=gpurun_out/$TAG
mkdir -p $OUT
timeout 300 python -m pytest tests/test_gpu_tc.py -m gpu -x -q > $OUT/pytest_tc.log 2>&1; echo "pytest_tc exit $?" >> $OUT/pytest_tc.log
tail -12 $OUT/pytest_tc.log
timeout 900 python -m pytest tests -m gpu -x -q > $OUT/pytest.log 2>&1; echo "pytest exit $?" >> $OUT/pytest.log
tail -12 $OUT/pytest.log
timeout 300 python scripts/tc_trace.py cfg2 3000 > $OUT/trace.log 2>&1; cat $OUT/trace.log | tail -12
timeout 600 python bench.py --steps 2 --warmup 3 > $OUT/bench.json 2> $OUT/bench.err; echo "bench exit $?" >> $OUT/bench.err
python -c "
import json; d=json.load(open('$OUT/bench.json')); print('value', d['value'], 'e2e', d['e2e']['value'], 'kernel_ms', d['roofline']['kernel_ms'], 'frac', d['roofline']['frac'], 'checksum', d['checksum_logp'])"; tail -5 $OUT/bench.err
