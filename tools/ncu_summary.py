#!/usr/bin/env python
"""Condenses an .ncu-rep (ncu --set full) into one text line per launch: usage: ncu_summary.py <rep> [title]"""
import csv, subprocess, sys, io
rep = sys.argv[1]
raw = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
rows = list(csv.reader(io.StringIO(raw)))
hdr, units = rows[0], rows[1]
cols = [("Kernel Name", "kernel"), ("Grid Size", "grid"), ("Block Size", "block"), ("gpu__time_duration.sum", "us"),
        ("sm__cycles_elapsed.max", "cycles"), ("dram__bytes_read.sum", "dram_rd"), ("dram__bytes_write.sum", "dram_wr"),
        ("dram__throughput.avg.pct_of_peak_sustained_elapsed", "dram%"), ("lts__throughput.avg.pct_of_peak_sustained_elapsed", "l2%"),
        ("lts__t_sectors_srcunit_tex_op_read.sum", "l2_rd_sectors"),
        ("sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_elapsed", "tensor%"),
        ("sm__inst_executed_pipe_uniform.sum", "uniform_inst"),
        ("sm__warps_active.avg.pct_of_peak_sustained_active", "warps%"), ("launch__registers_per_thread", "regs"),
        ("sm__inst_issued.avg.pct_of_peak_sustained_active", "issue%"), ("launch__cluster_size", "cluster")]
if len(sys.argv) > 2: print("#", sys.argv[2])
for r in rows[2:]:
    out = []
    for k, name in cols:
        if k in hdr:
            i = hdr.index(k); v = r[i]
            if name == "kernel": v = v[:48]
            out.append("%s=%s%s" % (name, v, (" " + units[i]) if units[i] and name in ("dram_rd", "dram_wr", "us") else ""))
    print("  ".join(out))
