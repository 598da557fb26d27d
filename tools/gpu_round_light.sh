#!/bin/bash
# gpu_round.sh without the ncu --set full captures: tests, smoke, the four bench lines and the ncu launch list.
set -u
TAG=${1:-x}
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -q > gpurun_out/${TAG}_pytest.log 2>&1
echo "pytest rc=$?"; tail -2 gpurun_out/${TAG}_pytest.log
timeout 300 python __graft_entry__.py --smoke > gpurun_out/${TAG}_smoke.log 2>&1; echo "smoke rc=$?"; tail -1 gpurun_out/${TAG}_smoke.log
timeout 600 python bench.py --impl reference --steps 2 --warmup 1 > gpurun_out/${TAG}_bench_ref.json 2> gpurun_out/${TAG}_bench_ref.err
echo "bench reference rc=$?"
timeout 600 python bench.py --steps 10 --warmup 3 > gpurun_out/${TAG}_bench.json 2> gpurun_out/${TAG}_bench.err
echo "bench rc=$?"; cut -c1-300 gpurun_out/${TAG}_bench.json
timeout 600 python bench.py --image 512 --batch 16 --steps 5 > gpurun_out/${TAG}_bench_cfg4.json 2> gpurun_out/${TAG}_bench_cfg4.err
echo "bench cfg4 rc=$?"
timeout 600 python bench.py --train --steps 5 > gpurun_out/${TAG}_bench_train.json 2> gpurun_out/${TAG}_bench_train.err
echo "bench train rc=$?"
CMD="python bench.py --steps 1 --warmup 3 --no-cpu-baseline --no-graph"
SCFLOW_PROFILE=1 timeout 600 $CMD > gpurun_out/${TAG}_plain.log 2>&1 &&
SCFLOW_PROFILE=1 timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none --profile-from-start off \
  --csv --log-file gpurun_out/${TAG}_launches.csv $CMD > gpurun_out/${TAG}_ncu_launches.log 2>&1
echo "ncu launches rc=$?"
