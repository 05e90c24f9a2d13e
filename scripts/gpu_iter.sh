#!/bin/bash
# one iteration on the tensor-core kernel: parity tests of the kernel, timeline, bench (cfg2)
OUT=gpurun_out/${1:-iter}; mkdir -p $OUT
timeout 300 python -m pytest tests/test_gpu_tc.py tests/test_gpu_parity.py -m gpu -x -q 2>&1 | tail -3 | tee -a $OUT/exp.log
KLM_TC_PLAN_DEBUG=1 timeout 200 python scripts/tc_trace.py cfg2 3000 2>&1 | grep -E "plan: [RIPL]|grid|REC|IN l|PROJ|LSE" | tee -a $OUT/exp.log
timeout 600 python bench.py --steps 2 --warmup 3 --no-cpu > $OUT/bench.json 2> $OUT/bench.err; echo "bench exit $?"
python -c "
import json; d=json.load(open('$OUT/bench.json')); print('value', d['value'], 'e2e', d['e2e']['value'], 'kernel_ms', d['roofline']['kernel_ms'], 'frac', d['roofline']['frac'], 'checksum', d['checksum_logp'])" | tee -a $OUT/exp.log; tail -3 $OUT/bench.err
