"""gpurun_out/<tag>_ct_<i>_raw.csv + .label -> gpurun_out/<tag>_conv_traffic.json:
{bench shape label: {"traffic_bytes", "dram_read_bytes", "dram_write_bytes", "duration_us", "kernel"}}"""
import csv, glob, json, os, sys
tag = sys.argv[1]
go = os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "gpurun_out")
out = {}
U = {"byte": 1.0, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9}
for lab in sorted(glob.glob(os.path.join(go, f"{tag}_ct_*.label"))):
    raw = lab[:-6] + "_raw.csv"
    rows = list(csv.reader(open(raw))) if os.path.exists(raw) else []
    if len(rows) < 3:
        continue
    hdr, units, r = rows[0], rows[1], rows[2]
    ix = {h: i for i, h in enumerate(hdr)}
    val = lambda k: float(r[ix[k]].replace(",", ""))
    rd = val("dram__bytes_read.sum") * U[units[ix["dram__bytes_read.sum"]]]
    wr = val("dram__bytes_write.sum") * U[units[ix["dram__bytes_write.sum"]]]
    dur = val("gpu__time_duration.sum") * {"ns": 1e-3, "us": 1.0, "ms": 1e3}[units[ix["gpu__time_duration.sum"]]]
    out[open(lab).read().strip()] = {"traffic_bytes": rd + wr, "dram_read_bytes": rd, "dram_write_bytes": wr,
                                     "duration_us": dur, "kernel": r[ix["Kernel Name"]][:60],
                                     "tensor_pipe_pct": val("sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_active")}
json.dump(out, open(os.path.join(go, f"{tag}_conv_traffic.json"), "w"), indent=1)
print(json.dumps(out, indent=1))
