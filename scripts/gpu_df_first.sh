OUT=gpurun_out/r2s; mkdir -p $OUT
timeout 300 python -m pytest tests/test_gpu_parity.py -m gpu -q -x -k "register_resident or golden or degenerate" > $OUT/pytest_rg.log 2>&1; echo "exit $?" >> $OUT/pytest_rg.log; tail -5 $OUT/pytest_rg.log
KLM_LL_DEBUG=1 timeout 600 python bench.py --workload cfg1 --steps 1 --warmup 3 --no-cpu > $OUT/bench_cfg1_dbg.json 2> $OUT/bench_cfg1_dbg.err; cut -c1-120 $OUT/bench_cfg1_dbg.json; grep "klm_seq_rg" $OUT/bench_cfg1_dbg.err | tail -7
