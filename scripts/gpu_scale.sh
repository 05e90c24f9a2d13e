#!/bin/bash
# weak scaling of cfg2 on N GPUs of one box (one rank per GPU, no data-path collective). Usage: gpu_scale.sh N [tag]
N=${1:-2}; OUT=gpurun_out/${2:-scale}; mkdir -p $OUT
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29511 \
  bench.py --gpus $N --steps 2 --warmup 3 > $OUT/bench_cfg2_${N}gpu.json 2> $OUT/bench_cfg2_${N}gpu.err
echo "exit $?"; tail -3 $OUT/bench_cfg2_${N}gpu.err; cut -c1-700 $OUT/bench_cfg2_${N}gpu.json
