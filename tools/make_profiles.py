"""Collect the judged evidence from gpurun_out/<tag>_* into profiles/ (tracked):
   python tools/make_profiles.py <tag> <round-name>
- <round>_bench.json / <round>_bench_reference.json: the bench lines of that GPU call
- <round>_launches.csv + _summary.txt: ncu launch list of `bench.py --steps 1 --no-graph` (timed region)
- <round>_ncu_full_summary.txt: per hand-written kernel, the `ncu --set full` metrics the roofline uses
  (duration, DRAM bytes = traffic, DRAM/L2/tensor/issue utilisation, registers, occupancy)"""
import csv, glob, os, shutil, subprocess, sys
tag, rnd = sys.argv[1], sys.argv[2]
root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
go, pr = os.path.join(root, "gpurun_out"), os.path.join(root, "profiles")
os.makedirs(pr, exist_ok=True)
for src, dst in ((f"{tag}_bench.json", f"{rnd}_bench.json"), (f"{tag}_bench_ref.json", f"{rnd}_bench_reference.json"),
                 (f"{tag}_launches.csv", f"{rnd}_launches.csv"), (f"{tag}_bench_cfg4.json", f"{rnd}_bench_cfg4_512.json"),
                 (f"{tag}_bench_train.json", f"{rnd}_bench_train.json"),
                 (f"{tag}_bench_torch_gpu.json", f"{rnd}_bench_torch_gpu.json"),
                 (f"{tag}_convbench.txt", f"{rnd}_convbench.txt"), (f"{tag}_corrbwd.json", f"{rnd}_corr_bwd_gemm.json"),
                 (f"{tag}_lookupbwd.json", f"{rnd}_lookup_bwd.json")):
    if os.path.exists(os.path.join(go, src)):
        shutil.copy(os.path.join(go, src), os.path.join(pr, dst))
out = subprocess.run([sys.executable, os.path.join(root, "tools", "launch_summary.py"),
                      os.path.join(go, f"{tag}_launches.csv"), "40"], capture_output=True, text=True).stdout
open(os.path.join(pr, f"{rnd}_launches_summary.txt"), "w").write(
    "ncu --metrics gpu__time_duration.sum --clock-control none, timed region of\n"
    "`python bench.py --steps 1 --warmup 3 --no-cpu-baseline --no-graph` (cold-cache, serialised launches:\n"
    "compare SHARES with bench.py's `kernels[*].share_of_eager_step`, not absolutes)\n\n" + out)
KEYS = ["gpu__time_duration.sum", "dram__bytes_read.sum", "dram__bytes_write.sum",
        "gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed", "lts__throughput.avg.pct_of_peak_sustained_elapsed",
        "l1tex__throughput.avg.pct_of_peak_sustained_active", "sm__throughput.avg.pct_of_peak_sustained_elapsed",
        "sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_active",
        "smsp__issue_active.avg.pct_of_peak_sustained_active", "sm__warps_active.avg.pct_of_peak_sustained_active",
        "sm__inst_executed.sum", "launch__registers_per_thread", "launch__grid_size", "launch__block_size",
        "launch__occupancy_limit_registers", "launch__occupancy_limit_shared_mem"]
lines = ["ncu --set full --clock-control none --import-source on, same bench command, kernels of the timed region.",
         "traffic = dram__bytes_read.sum + dram__bytes_write.sum per launch (what bench.py's roofline.traffic quotes).", ""]
for f in sorted(glob.glob(os.path.join(go, f"{tag}_*_raw.csv"))):
    rows = list(csv.reader(open(f)))
    if len(rows) < 3:
        continue
    hdr, units = rows[0], rows[1]
    ix = {h: i for i, h in enumerate(hdr)}
    for r in rows[2:]:
        lines.append(f"== {r[ix['Kernel Name']][:90]}  (launch id {r[ix['ID']]}, grid {r[ix['launch__grid_size']]} x {r[ix['launch__block_size']]})")
        for k in KEYS:
            if k in ix:
                lines.append(f"   {k:66s} {r[ix[k]]:>16s} {units[ix[k]]}")
        def mb(k):
            v = float(r[ix[k]].replace(",", "")); u = units[ix[k]]
            return v * {"byte": 1e-6, "Kbyte": 1e-3, "Mbyte": 1.0, "Gbyte": 1e3}[u]
        lines.append(f"   traffic (DRAM read + write)                                        {mb('dram__bytes_read.sum') + mb('dram__bytes_write.sum'):16.3f} MB")
        lines.append("")
open(os.path.join(pr, f"{rnd}_ncu_full_summary.txt"), "w").write("\n".join(lines))
print("wrote", sorted(os.listdir(pr)))
