#!/bin/bash
# tensor-core kernel bring-up: self-test, parity tests, bench -- each under its own timeout
TAG=${1:-tc}
OUT=gpurun_out/$TAG
mkdir -p $OUT
timeout 300 python -m pytest tests/test_gpu_tc.py -m gpu -x -q -k umma > $OUT/selftest.log 2>&1; echo "selftest exit $?" >> $OUT/selftest.log
tail -15 $OUT/selftest.log
timeout 900 python -m pytest tests -m gpu -x -q > $OUT/pytest.log 2>&1; echo "pytest exit $?" >> $OUT/pytest.log
tail -30 $OUT/pytest.log
timeout 600 python bench.py --steps 2 --warmup 3 > $OUT/bench.json 2> $OUT/bench.err; echo "bench exit $?" >> $OUT/bench.err
cat $OUT/bench.json; tail -5 $OUT/bench.err
