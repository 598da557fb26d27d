#!/bin/bash
# One gpurun call: GPU parity tests, bench, then ncu captures of the hand-written kernels.
# usage: tools/gpu_round.sh <tag> [skip_tests]
set -u
TAG=${1:-x}
mkdir -p gpurun_out
if [ "${2:-}" != "skip_tests" ]; then
  timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/${TAG}_pytest.log 2>&1
  echo "pytest rc=$?" ; tail -3 gpurun_out/${TAG}_pytest.log
fi
timeout 600 python bench.py --steps 10 --warmup 3 > gpurun_out/${TAG}_bench.json 2> gpurun_out/${TAG}_bench.err
echo "bench rc=$?"; cat gpurun_out/${TAG}_bench.json
SCFLOW_PROFILE=1 timeout 600 python bench.py --steps 1 --warmup 3 --no-cpu-baseline > gpurun_out/${TAG}_plain.log 2>&1 &&
SCFLOW_PROFILE=1 timeout 900 ncu --set full --clock-control none --import-source on --profile-from-start off \
  -k regex:'corr_lookup_kernel|corr_tf32_kernel|convgru_pass_kernel|pose_flow_kernel' -c 33 \
  -o gpurun_out/${TAG}_prof -f python bench.py --steps 1 --warmup 3 --no-cpu-baseline > gpurun_out/${TAG}_ncu.log 2>&1
echo "ncu full rc=$?"
SCFLOW_PROFILE=1 timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none --profile-from-start off \
  --csv --log-file gpurun_out/${TAG}_launches.csv python bench.py --steps 1 --warmup 3 --no-cpu-baseline > gpurun_out/${TAG}_ncu2.log 2>&1
echo "ncu launches rc=$?"
