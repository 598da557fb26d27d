#!/bin/bash
# One gpurun call: GPU parity tests, smoke, bench (both arms), then the ncu launch list of the same
# bench command and ncu --set full captures of the hand-written kernels (exported to CSV here so the
# bulky .ncu-rep files stay on the box).  usage: tools/gpu_round.sh <tag>
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
timeout 600 python bench.py --impl torch-gpu --steps 5 --warmup 3 > gpurun_out/${TAG}_bench_torch_gpu.json 2> gpurun_out/${TAG}_bench_torch_gpu.err
echo "bench torch-gpu rc=$?"
timeout 300 python tools/convbench.py --n 64 > gpurun_out/${TAG}_convbench.txt 2>&1; echo "convbench rc=$?"
timeout 120 python tools/corrbwd_bench.py > gpurun_out/${TAG}_corrbwd.json 2>/dev/null; echo "corrbwd rc=$?"
timeout 120 python tools/lookupbwd_bench.py > gpurun_out/${TAG}_lookupbwd.json 2>/dev/null
timeout 120 python tools/lookupbwd_bench.py 16 64 64 >> gpurun_out/${TAG}_lookupbwd.json 2>/dev/null; echo "lookupbwd rc=$?"
CMD="python bench.py --steps 1 --warmup 3 --no-cpu-baseline --no-graph"
SCFLOW_PROFILE=1 timeout 600 $CMD > gpurun_out/${TAG}_plain.log 2>&1 &&
SCFLOW_PROFILE=1 timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none --profile-from-start off \
  --csv --log-file gpurun_out/${TAG}_launches.csv $CMD > gpurun_out/${TAG}_ncu_launches.log 2>&1
echo "ncu launches rc=$?"
for K in corr_lookup_v2_kernel:2:2 corr_tf32_kernel:1:0 conv_patch_kernel:8:40 conv_tc_kernel:4:10 norm_apply_split_kernel:2:4 pose_flow_kernel:1:2 resize4_up_kernel:2:2 group_norm4_act_kernel:1:3 expand_taps_kernel:1:3 tap_sum_kernel:2:2 pose_tail_kernel:1:2 stem_s2d_split_kernel:1:0 split_planar_kernel:1:2; do
  RE=${K%%:*}; R=${K#*:}; CNT=${R%%:*}; SKIP=${R#*:}
  SCFLOW_PROFILE=1 timeout 900 ncu --set full --clock-control none --import-source on --profile-from-start off \
    -k regex:"$RE" -s $SKIP -c $CNT -o /tmp/${TAG}_${RE} -f $CMD > gpurun_out/${TAG}_ncu_${RE}.log 2>&1
  echo "ncu $RE rc=$?"
  ncu -i /tmp/${TAG}_${RE}.ncu-rep --page raw --csv > gpurun_out/${TAG}_${RE}_raw.csv 2>/dev/null
  ncu -i /tmp/${TAG}_${RE}.ncu-rep --page source --csv > gpurun_out/${TAG}_${RE}_source.csv 2>/dev/null
done
ls -la gpurun_out | tail -30
