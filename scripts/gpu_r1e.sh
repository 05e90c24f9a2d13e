#!/bin/bash
# round-1 re-entry check: micro-benchmark of the fetch paths, then the tensor-core iteration script
OUT=gpurun_out/${1:-r1e}; mkdir -p $OUT
timeout 120 scripts/ubench/fetch_paths > $OUT/fetch_paths.txt 2>&1; cat $OUT/fetch_paths.txt
bash scripts/gpu_tc.sh ${1:-r1e}
