OUT=gpurun_out/r2h; mkdir -p $OUT
timeout 900 python -m pytest tests/test_gpu_tc.py tests/test_gpu_long.py tests/test_gpu_parity.py tests/test_page_glue.py -m gpu -q > $OUT/pytest.log 2>&1; echo "exit $?" >> $OUT/pytest.log; tail -8 $OUT/pytest.log
timeout 600 python bench.py --steps 3 --warmup 3 --no-cpu > $OUT/bench_quick.json 2> $OUT/bench_quick.err; echo "bench exit $?" >> $OUT/bench_quick.err; cut -c1-140 $OUT/bench_quick.json; tail -2 $OUT/bench_quick.err
KLM_TC_PUB_RELEASE=0 timeout 600 python bench.py --steps 3 --warmup 3 --no-cpu > $OUT/bench_quick_relaxed.json 2> $OUT/bench_quick_relaxed.err; cut -c1-140 $OUT/bench_quick_relaxed.json
KLM_STEP_TIMING=1 timeout 600 python bench.py --workload cfg3 --steps 2 --warmup 2 > $OUT/bench_cfg3.json 2> $OUT/bench_cfg3.err; cut -c1-700 $OUT/bench_cfg3.json; tail -3 $OUT/bench_cfg3.err
KLM_TC_DF=0 timeout 600 python bench.py --workload cfg3 --steps 2 --warmup 2 > $OUT/bench_cfg3_nodf.json 2> $OUT/bench_cfg3_nodf.err; cut -c1-200 $OUT/bench_cfg3_nodf.json
