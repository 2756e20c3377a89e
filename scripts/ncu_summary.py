"""Summarise an .ncu-rep (raw page) into the handful of metrics the roofline argument uses.
usage: python scripts/ncu_summary.py report.ncu-rep > profiles/<name>.txt"""
import csv, io, subprocess, sys
KEYS = ["gpu__time_duration.sum", "dram__bytes_read.sum", "dram__bytes_write.sum", "gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed",
        "dram__cycles_active.avg.pct_of_peak_sustained_elapsed", "lts__t_sector_hit_rate.pct", "lts__throughput.avg.pct_of_peak_sustained_elapsed",
        "l1tex__m_xbar2l1tex_read_bytes.sum", "sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_active",
        "sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_elapsed", "sm__throughput.avg.pct_of_peak_sustained_elapsed",
        "sm__warps_active.avg.pct_of_peak_sustained_active", "sm__cycles_elapsed.avg.per_second", "launch__registers_per_thread",
        "launch__grid_size", "launch__block_size", "launch__cluster_dim_x", "launch__shared_mem_per_block_dynamic", "smsp__inst_executed.sum"]
out = subprocess.run(["ncu", "-i", sys.argv[1], "--page", "raw", "--csv"], capture_output=True, text=True).stdout
rows = list(csv.reader(io.StringIO(out)))
hdr, units = rows[0], rows[1]
ci = {h: i for i, h in enumerate(hdr)}
for r in rows[2:]:
    print("kernel:", r[ci["Kernel Name"]][:100])
    for k in KEYS:
        if k in ci:
            print(f"  {k:75s} {r[ci[k]]:>18s} {units[ci[k]]}")
