#!/bin/bash
# streamed-weights kernel: tc tests, plan, cfg5 bench, cfg2 regression check
OUT=gpurun_out/${1:-tcs}; mkdir -p $OUT
timeout 600 python -m pytest tests/test_gpu_tc.py -x -q 2>&1 | tail -15
KLM_TC_PLAN_DEBUG=1 timeout 600 python bench.py --workload cfg5 --steps 2 --warmup 3 --no-cpu > $OUT/bench_cfg5.json 2> $OUT/bench_cfg5.err; echo "cfg5 exit $?"; grep "klm tc plan" $OUT/bench_cfg5.err | head -20; tail -2 $OUT/bench_cfg5.err; cat $OUT/bench_cfg5.json
timeout 600 python bench.py --steps 2 --warmup 3 --no-cpu > $OUT/bench_cfg2.json 2> $OUT/bench_cfg2.err; echo "cfg2 exit $?"; cat $OUT/bench_cfg2.json
