#!/bin/bash
OUT=gpurun_out/exp14; mkdir -p $OUT
timeout 900 python -m pytest tests/test_gpu_tc.py -m gpu -x -q -k "skew3" > $OUT/pytest.log 2>&1; echo "pytest exit $?" >> $OUT/pytest.log
tail -15 $OUT/pytest.log
