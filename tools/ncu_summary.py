"""Summarise an `ncu --page raw --csv` export: one line per kernel with the metrics the
roofline needs.  usage: python tools/ncu_summary.py gpurun_out/x_raw.csv"""
import csv, sys
KEYS = ["gpu__time_duration.sum", "dram__bytes_read.sum", "dram__bytes_write.sum",
        "gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed",
        "lts__t_bytes.sum", "lts__throughput.avg.pct_of_peak_sustained_elapsed",
        "l1tex__data_pipe_lsu_wavefronts_mem_shared.sum", "l1tex__throughput.avg.pct_of_peak_sustained_active",
        "sm__throughput.avg.pct_of_peak_sustained_elapsed", "sm__inst_executed.sum",
        "sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_active",
        "sm__inst_executed_pipe_tensor.sum",
        "sm__warps_active.avg.pct_of_peak_sustained_active", "launch__registers_per_thread",
        "launch__occupancy_limit_shared_mem", "launch__occupancy_limit_registers",
        "smsp__issue_active.avg.pct_of_peak_sustained_active", "smsp__inst_executed.avg.per_cycle_active",
        "sm__cycles_elapsed.max", "launch__grid_size", "launch__block_size"]
rows = list(csv.reader(open(sys.argv[1])))
hdr, units = rows[0], rows[1]
idx = {h: i for i, h in enumerate(hdr)}
for r in rows[2:]:
    print("==", r[idx["Kernel Name"]][:70], "id", r[idx["ID"]])
    for k in KEYS:
        if k in idx:
            print(f"   {k:70s} {r[idx[k]]:>16s} {units[idx[k]]}")
    if len(sys.argv) > 2:
        for h in hdr:
            if sys.argv[2] in h:
                print(f"   {h:70s} {r[idx[h]]:>16s} {units[idx[h]]}")
