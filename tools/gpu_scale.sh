#!/bin/bash
# Scaling runs on one 8-GPU box: inference bench at N=4 and N=8, training step at N=8 (cfg3 / cfg5).
#   gpurun --gpus 8 -- bash tools/gpu_scale.sh <tag>
tag=${1:-scale}
mkdir -p gpurun_out
run() {  # n, extra args, out name
  timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $1 --master-addr 127.0.0.1 \
    --master-port $((29500 + $1 + RANDOM % 100)) bench.py --gpus $1 --steps 10 --warmup 3 $2 \
    > gpurun_out/${tag}_$3.json 2> gpurun_out/${tag}_$3.err
  echo "$3 rc=$?"; head -c 400 gpurun_out/${tag}_$3.json; echo
}
run 8 "" n8
run 4 "" n4
run 8 "--train" train_n8
nvidia-smi topo -m > gpurun_out/${tag}_topo.txt 2>&1
