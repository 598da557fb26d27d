"""Print a readable summary of a bench.py JSON line: python tools/bench_summary.py gpurun_out/x.json"""
import json, sys
d = json.load(open(sys.argv[1]))
for k in ("value", "ms_per_step", "e2e", "gpu_launches", "eager_ms_per_step", "cpu_baseline", "clocks", "conv_family"):
    print(k, d.get(k))
print("deviation", d["precision"]["deviation_from_oracle"])
print("roof", {k: v for k, v in d["roofline"].items() if k not in ("note", "peak_source")})
print("lookup", d["roofline_lookup"])
ks = sorted(d["kernels"].items(), key=lambda kv: -kv[1]["avg_ms"] * kv[1]["calls_per_step"])
tot = 0
for k, v in ks:
    t = v["avg_ms"] * v["calls_per_step"]; tot += t
    print(f'{k:58s} x{v["calls_per_step"]:5.1f} {v["avg_ms"]*1e3:7.1f}us {t:6.3f}ms  TF {v["tflops"] and round(v["tflops"],1)}  GB/s {v["gbs"] and round(v["gbs"])}')
print("sum of kernels ms", tot)
