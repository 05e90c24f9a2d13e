OUT=gpurun_out/r2m; mkdir -p $OUT
timeout 300 python scripts/exp/df_step_trace.py 117 > $OUT/df_step_trace.txt 2>&1; head -4 $OUT/df_step_trace.txt | cut -c1-250; tail -5 $OUT/df_step_trace.txt
KLM_STEP_TIMING=1 timeout 600 python bench.py --workload cfg3 --steps 2 --warmup 2 > $OUT/bench_cfg3.json 2> $OUT/bench_cfg3.err; cut -c1-200 $OUT/bench_cfg3.json; tail -3 $OUT/bench_cfg3.err
timeout 900 python -m pytest tests -m gpu -q -x > $OUT/pytest.log 2>&1; echo "exit $?" >> $OUT/pytest.log; tail -4 $OUT/pytest.log
