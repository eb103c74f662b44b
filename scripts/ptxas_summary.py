#!/usr/bin/env python
"""Compiles one instantiation unit with -Xptxas -v and prints one line per kernel: template arguments, registers, spills.
    python scripts/ptxas_summary.py KIND METRIC [ROUTE] [filter-substring]"""
import re, subprocess, sys, os
root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
kind, metric = sys.argv[1], sys.argv[2]
route = sys.argv[3] if len(sys.argv) > 3 else "0"
flt = sys.argv[4] if len(sys.argv) > 4 else ""
cmd = ["nvcc", "-gencode", "arch=compute_100a,code=sm_100a", "-lineinfo", "-O3", "-std=c++17", "-DLMD_INST_KIND=" + kind,
       "-DLMD_INST_METRIC=" + metric, "-DLMD_INST_ROUTE=" + route, "-Xptxas", "-v", "-c",
       os.path.join(root, "duckdb-lm-diskann_b200/csrc/lmd_kernel_inst.cu"), "-o", "/tmp/ptxas_summary.o"]
err = subprocess.run(cmd, capture_output=True, text=True).stderr
cur = None
for line in err.splitlines():
    m = re.search(r"Compiling entry function '(\S+)'", line)
    if m:
        cur = subprocess.run(["c++filt", m.group(1)], capture_output=True, text=True).stdout.strip()
        cur = re.sub(r"\(lmd::KParams\)|\(lmd::PruneParams\)|void lmd::", "", cur)
        spill = ""
    m = re.search(r"(\d+) bytes stack frame, (\d+) bytes spill stores, (\d+) bytes spill loads", line)
    if m:
        spill = "stack %s spill st %s ld %s" % m.groups()
    m = re.search(r"Used (\d+) registers", line)
    if m and cur and flt in cur:
        print("%-48s regs %3s  %s" % (cur, m.group(1), spill))
