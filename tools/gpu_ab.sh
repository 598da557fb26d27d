#!/bin/bash
# A/B of one environment switch through the bench: tools/gpu_ab.sh VAR val1 val2 ...
mkdir -p gpurun_out
V=$1; shift
for C in "$@"; do
  env $V=$C timeout 600 python bench.py --steps 10 --warmup 3 --no-cpu-baseline > gpurun_out/ab_${V}_$C.json 2> gpurun_out/ab_${V}_$C.err
  echo "$V=$C rc=$?"; python - <<P
import json
try:
    d=json.load(open("gpurun_out/ab_${V}_$C.json"))
    print("$V=$C", "value %.1f  ms %.3f  e2e %.1f  eager_ms %.3f" % (d["value"], d["ms_per_step"], d["e2e"]["value"], d.get("eager_ms_per_step") or 0))
except Exception as e:
    print("failed", e); print(open("gpurun_out/ab_${V}_$C.err").read()[-1500:])
P
done
