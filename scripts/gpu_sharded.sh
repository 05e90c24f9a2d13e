#!/bin/bash
# cfg5 the way BASELINE words it: 10 000 lines sharded over N GPUs of one box, strong scaling, scores gathered.
# usage (under gpurun --gpus N): bash scripts/gpu_sharded.sh <tag> <N> [<N> ...]
TAG=$1; shift
OUT=gpurun_out/$TAG; mkdir -p $OUT
nvidia-smi -L > $OUT/gpus.txt
timeout 600 python -m pytest tests/test_gpu_multi.py -m gpu -q > $OUT/pytest_multi.log 2>&1; echo "exit $?" >> $OUT/pytest_multi.log; tail -3 $OUT/pytest_multi.log
for N in "$@"; do
  if [ "$N" = 1 ]; then
    timeout 600 python bench.py --workload cfg5 --sharded --steps 3 --warmup 3 > $OUT/sharded_1.json 2> $OUT/sharded_1.err
  else
    timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29555 bench.py --workload cfg5 --sharded --gpus $N --steps 3 --warmup 3 > $OUT/sharded_$N.json 2> $OUT/sharded_$N.err
  fi
  echo "N=$N exit $?"; cut -c1-330 $OUT/sharded_$N.json; tail -2 $OUT/sharded_$N.err | cut -c1-200
done
