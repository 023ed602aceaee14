#!/usr/bin/env python
"""Summarise an .ncu-rep (read on the CPU box): per launch duration, DRAM bytes, DRAM/L2/tensor utilisation.
    python tools/ncu_summary.py gpurun_out/prof.ncu-rep > profiles/xxx_summary.csv"""
import csv, io, re, subprocess, sys
rep = sys.argv[1]
out = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
rows = list(csv.reader(io.StringIO(out)))
hdr = rows[0]
want = ["Kernel Name", "Grid Size", "Block Size", "gpu__time_duration.sum", "dram__bytes_read.sum", "dram__bytes_write.sum",
        "gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed", "lts__t_bytes.sum", "sm__warps_active.avg.pct_of_peak_sustained_active",
        "sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_active", "sm__inst_executed_pipe_tensor.sum",
        "launch__registers_per_thread", "launch__shared_mem_per_block_dynamic", "sm__throughput.avg.pct_of_peak_sustained_elapsed",
        "launch__cluster_dim_x"]
idx = [(w, hdr.index(w)) for w in want if w in hdr]
units = rows[1]
w = csv.writer(sys.stdout)
w.writerow([n + (f" [{units[i]}]" if units[i] else "") for n, i in idx])
for r in rows[2:]:
    if len(r) < len(hdr):
        continue
    vals = []
    for n, i in idx:
        v = r[i]
        if n == "Kernel Name":
            v = re.sub(r"\(.*", "", v).replace("capdec::", "").replace("__nv_bfloat16", "bf16")[:70]
        vals.append(v)
    w.writerow(vals)
