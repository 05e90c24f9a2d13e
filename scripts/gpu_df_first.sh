OUT=gpurun_out/r2d; mkdir -p $OUT
timeout 900 python -m pytest tests/test_gpu_tc.py tests/test_gpu_long.py tests/test_gpu_parity.py -m gpu -q > $OUT/pytest_df.log 2>&1; echo "exit $?" >> $OUT/pytest_df.log; tail -40 $OUT/pytest_df.log
timeout 600 python bench.py --steps 3 --warmup 3 --no-cpu > $OUT/bench_quick.json 2> $OUT/bench_quick.err; echo "bench exit $?" >> $OUT/bench_quick.err; cat $OUT/bench_quick.json; tail -2 $OUT/bench_quick.err
