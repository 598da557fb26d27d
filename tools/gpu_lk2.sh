#!/bin/bash
# lookup v2 bring-up: parity tests for both generations, then A/B timing
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_gpu_parity.py -m gpu -q -k "lookup" > gpurun_out/lk2_pytest_v2.log 2>&1; echo "pytest v2 rc=$?"; tail -5 gpurun_out/lk2_pytest_v2.log
SCFLOW_LK=v1 timeout 600 python -m pytest tests/test_gpu_parity.py -m gpu -q -k "lookup" > gpurun_out/lk2_pytest_v1.log 2>&1; echo "pytest v1 rc=$?"; tail -2 gpurun_out/lk2_pytest_v1.log
timeout 300 python tools/kbench.py lookup > gpurun_out/lk2_kbench.log 2>&1; echo "kbench rc=$?"; cat gpurun_out/lk2_kbench.log
timeout 300 python tools/kbench.py lookup --s 512 --n 16 > gpurun_out/lk2_kbench512.log 2>&1; echo "kbench512 rc=$?"; cat gpurun_out/lk2_kbench512.log
