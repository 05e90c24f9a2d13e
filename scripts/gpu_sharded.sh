#!/bin/bash
# Multi-GPU lines on one box (under gpurun --gpus N):
#   * cfg5 the way BASELINE words it: 10 000 lines sharded over N GPUs, strong scaling, scores gathered (bench.py --sharded)
#   * the default bench (cfg2, weak scaling: 64 documents per GPU) at the same N
# usage: bash scripts/gpu_sharded.sh <tag> [test] <N> [<N> ...]      ("test": tests/test_gpu_multi.py first)
TAG=$1; shift
OUT=gpurun_out/$TAG; mkdir -p $OUT
nvidia-smi -L > $OUT/gpus.txt
if [ "$1" = test ]; then
  shift
  timeout 600 python -m pytest tests/test_gpu_multi.py -m gpu -q > $OUT/pytest_multi.log 2>&1; echo "exit $?" >> $OUT/pytest_multi.log; tail -3 $OUT/pytest_multi.log
fi
for N in "$@"; do
  if [ "$N" = 1 ]; then
    timeout 600 python bench.py --workload cfg5 --sharded --steps 3 --warmup 3 > $OUT/sharded_1.json 2> $OUT/sharded_1.err
    timeout 600 python bench.py --steps 3 --warmup 3 --no-cpu > $OUT/weak_1.json 2> $OUT/weak_1.err
  else
    RUN="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1"
    timeout 600 $RUN --master-port 29555 bench.py --workload cfg5 --sharded --gpus $N --steps 3 --warmup 3 > $OUT/sharded_$N.json 2> $OUT/sharded_$N.err
    timeout 600 $RUN --master-port 29556 bench.py --gpus $N --steps 3 --warmup 3 --no-cpu > $OUT/weak_$N.json 2> $OUT/weak_$N.err
  fi
  echo "N=$N exit $?"; cut -c1-200 $OUT/sharded_$N.json; tail -1 $OUT/sharded_$N.err | cut -c1-200; cut -c1-200 $OUT/weak_$N.json; tail -1 $OUT/weak_$N.err | cut -c1-200
done
