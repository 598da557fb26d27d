#!/bin/bash
# Scaling runs on one multi-GPU box: weak scaling (64 crops per GPU) and configs[2] as written (512 crops split over
# the ranks) at N = 2/4/8 as available, training step at the largest N.   gpurun --gpus 8 -- bash tools/gpu_scale.sh <tag>
tag=${1:-scale}
NG=$(nvidia-smi -L | wc -l)
mkdir -p gpurun_out
run() {  # n, extra args, out name
  timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $1 --master-addr 127.0.0.1 \
    --master-port $((29500 + $1 + RANDOM % 100)) bench.py --gpus $1 --steps 10 --warmup 3 --no-cpu-baseline $2 \
    > gpurun_out/${tag}_$3.json 2> gpurun_out/${tag}_$3.err
  echo "$3 rc=$?"; head -c 300 gpurun_out/${tag}_$3.json; echo
}
timeout 600 python bench.py --steps 10 --warmup 3 --no-cpu-baseline > gpurun_out/${tag}_n1.json 2> gpurun_out/${tag}_n1.err; echo "n1 rc=$?"
timeout 600 python bench.py --steps 10 --warmup 3 --no-cpu-baseline --strong 512 > gpurun_out/${tag}_strong_n1.json 2> gpurun_out/${tag}_strong_n1.err; echo "strong n1 rc=$?"
for N in 2 4 8; do
  if [ $N -le $NG ]; then run $N "" n$N; run $N "--strong 512" strong_n$N; fi
done
run $NG "--train" train_n$NG
nvidia-smi topo -m > gpurun_out/${tag}_topo.txt 2>&1
