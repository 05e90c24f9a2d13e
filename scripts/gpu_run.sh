#!/bin/bash
# One GPU-box session, parameterised: bash scripts/gpu_run.sh <tag> <step> [<step> ...]   (from the repo root, under gpurun)
#   tests      the whole -m gpu suite            long       tests/test_gpu_long.py only (parity at BASELINE's shapes)
#   smoke      __graft_entry__.smoke()           bench      bench.py (cfg2, default steps) + the reference arm
#   cfgs       bench lines of cfg1/3/4/5         launches   ncu launch list of a short bench
#   ncu        ncu --set full of the sequence kernel (one launch)        trace   in-kernel timeline (scripts/tc_trace.py)
#   drift      16 x 99 999 positions against the fp64 windows (scripts/drift_1e5.py)
#   relaxed    cfg2 with the round-1 relaxed hand-over signal (KLM_TC_PUB_RELEASE=0; measures what the release costs)
#   memcheck | synccheck | racecheck   compute-sanitizer over the reduced selection of scripts/sanitize_select.txt (ONE tool per call)
TAG=$1; shift
OUT=gpurun_out/$TAG
mkdir -p $OUT
{ nproc; free -g; nvidia-smi -L; } > $OUT/host.txt 2>&1
SMALL="python bench.py --steps 1 --warmup 3 --chars 3000 --no-cpu"
SEL=$(grep -v '^#' scripts/sanitize_select.txt 2>/dev/null | tr '\n' ' ')
for STEP in "$@"; do
  case $STEP in
    tests) timeout 1500 python -m pytest tests -m gpu -x -q > $OUT/pytest.log 2>&1; echo "pytest exit $?" >> $OUT/pytest.log; tail -5 $OUT/pytest.log ;;
    long) timeout 1200 python -m pytest tests/test_gpu_long.py -m gpu -q > $OUT/pytest_long.log 2>&1; echo "pytest exit $?" >> $OUT/pytest_long.log; tail -15 $OUT/pytest_long.log; mv gpurun_out/parity_*.json $OUT/ 2>/dev/null ;;
    smoke) timeout 300 python __graft_entry__.py smoke > $OUT/smoke.log 2>&1; echo "smoke exit $?" >> $OUT/smoke.log; tail -2 $OUT/smoke.log ;;
    bench) timeout 900 python bench.py --steps 3 --warmup 3 > $OUT/bench.json 2> $OUT/bench.err; echo "bench exit $?" >> $OUT/bench.err; cat $OUT/bench.json; tail -2 $OUT/bench.err
           timeout 600 python bench.py --impl reference --steps 2 --warmup 1 > $OUT/bench_ref.json 2> $OUT/bench_ref.err; cat $OUT/bench_ref.json ;;
    benchq) timeout 600 python bench.py --steps 3 --warmup 3 --no-cpu > $OUT/bench_quick.json 2> $OUT/bench_quick.err; echo "bench exit $?" >> $OUT/bench_quick.err; cat $OUT/bench_quick.json; tail -2 $OUT/bench_quick.err ;;
    cfgs) for W in cfg1 cfg3 cfg4 cfg5; do timeout 900 python bench.py --workload $W --steps 3 --warmup 3 > $OUT/bench_$W.json 2> $OUT/bench_$W.err; echo "exit $?" >> $OUT/bench_$W.err; cat $OUT/bench_$W.json; tail -1 $OUT/bench_$W.err; done
          timeout 600 python bench.py --workload cfg3 --impl reference --steps 1 --warmup 1 > $OUT/bench_cfg3_ref.json 2> $OUT/bench_cfg3_ref.err ;;
    launches) timeout 300 $SMALL > $OUT/plain.log 2>&1 &&
              timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 300 --csv --log-file $OUT/launches.csv $SMALL > $OUT/ncu_launches.log 2>&1; tail -3 $OUT/ncu_launches.log ;;
    ncu) timeout 300 $SMALL > $OUT/plain2.log 2>&1 &&
         timeout 900 ncu --set full --clock-control none --import-source on -k regex:klm_seq -s 3 -c 1 -o $OUT/prof $SMALL > $OUT/ncu_full.log 2>&1; tail -3 $OUT/ncu_full.log ;;
    drift) timeout 900 python scripts/drift_1e5.py > $OUT/drift_1e5.json 2> $OUT/drift_1e5.err; echo "drift exit $?" >> $OUT/drift_1e5.err; cat $OUT/drift_1e5.json | tr -d '\n' | cut -c1-900; echo ;;
    relaxed) KLM_TC_PUB_RELEASE=0 timeout 600 python bench.py --steps 3 --warmup 3 --no-cpu > $OUT/bench_relaxed.json 2> $OUT/bench_relaxed.err; cat $OUT/bench_relaxed.json ;;
    trace) timeout 600 python scripts/tc_trace.py > $OUT/timeline.txt 2>&1; tail -30 $OUT/timeline.txt ;;
    memcheck|synccheck|racecheck)
         timeout 2400 compute-sanitizer --tool $STEP --error-exitcode 9 --log-file $OUT/sanitizer_$STEP.log python -m pytest $SEL -m gpu -x -q > $OUT/sanitizer_${STEP}_pytest.log 2>&1
         echo "sanitizer exit $?" >> $OUT/sanitizer_${STEP}_pytest.log; tail -4 $OUT/sanitizer_${STEP}_pytest.log; tail -5 $OUT/sanitizer_$STEP.log ;;
    *) echo "unknown step $STEP" ;;
  esac
done
