#!/bin/bash
# One GPU-box session: parity tests, smoke, bench (both arms), ncu launch list + full capture of the sequence kernel.
# Usage (from the repo root, under gpurun): bash scripts/gpu_check.sh [tag]
TAG=${1:-run}
OUT=gpurun_out/$TAG
mkdir -p $OUT
nproc > $OUT/host.txt; free -g >> $OUT/host.txt; nvidia-smi -L >> $OUT/host.txt
timeout 900 python -m pytest tests -m gpu -x -q > $OUT/pytest.log 2>&1; echo "pytest exit $?" >> $OUT/pytest.log
timeout 300 python __graft_entry__.py smoke > $OUT/smoke.log 2>&1; echo "smoke exit $?" >> $OUT/smoke.log
timeout 900 python bench.py --steps 3 --warmup 3 > $OUT/bench.json 2> $OUT/bench.err; echo "bench exit $?" >> $OUT/bench.err
timeout 600 python bench.py --impl reference --steps 2 --warmup 1 > $OUT/bench_ref.json 2> $OUT/bench_ref.err
SMALL="python bench.py --steps 1 --warmup 3 --chars 3000 --no-cpu"
timeout 300 $SMALL > $OUT/plain.log 2>&1 &&
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 300 --csv --log-file $OUT/launches.csv $SMALL > $OUT/ncu_launches.log 2>&1
timeout 300 $SMALL > $OUT/plain2.log 2>&1 &&
timeout 900 ncu --set full --clock-control none --import-source on -k regex:klm_seq -s 3 -c 1 -o $OUT/prof $SMALL > $OUT/ncu_full.log 2>&1
tail -3 $OUT/pytest.log; tail -2 $OUT/smoke.log; cat $OUT/bench.json; tail -2 $OUT/bench.err; cat $OUT/bench_ref.json; tail -3 $OUT/ncu_full.log
