#!/bin/bash
# Measurement hygiene run on a 2-GPU box: N=2 line, the reference's modules on the GPU (torch-gpu arm), cfg4 at 64 crops.
set -u
TAG=${1:-hyg}
mkdir -p gpurun_out
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29533 \
  bench.py --gpus 2 --steps 10 --warmup 3 > gpurun_out/${TAG}_n2.json 2> gpurun_out/${TAG}_n2.err; echo "n2 rc=$?"; head -c 300 gpurun_out/${TAG}_n2.json; echo
timeout 900 python bench.py --impl torch-gpu --steps 3 --warmup 3 > gpurun_out/${TAG}_torch_gpu.json 2> gpurun_out/${TAG}_torch_gpu.err; echo "torch-gpu rc=$?"; head -c 1500 gpurun_out/${TAG}_torch_gpu.json; echo
timeout 900 python bench.py --image 512 --batch 64 --steps 5 > gpurun_out/${TAG}_cfg4_n64.json 2> gpurun_out/${TAG}_cfg4_n64.err; echo "cfg4 n64 rc=$?"; head -c 300 gpurun_out/${TAG}_cfg4_n64.json; echo
