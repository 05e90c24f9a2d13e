#!/bin/bash
# tensor-core kernel iteration: parity tests, timeline, bench -- each under its own timeout
TAG=${1:-tc}
OUT=gpurun_out/$TAG
mkdir -p $OUT
timeout 900 python -m pytest tests -m gpu -x -q > $OUT/pytest.log 2>&1; echo "pytest exit $?" >> $OUT/pytest.log
tail -12 $OUT/pytest.log
timeout 300 python scripts/tc_trace.py cfg2 3000 > $OUT/trace.log 2>&1; cat $OUT/trace.log | tail -12
timeout 600 python bench.py --steps 2 --warmup 3 > $OUT/bench.json 2> $OUT/bench.err; echo "bench exit $?" >> $OUT/bench.err
cat $OUT/bench.json; tail -5 $OUT/bench.err
