#!/usr/bin/env python
"""Prints the handful of `ncu --set full` metrics the kernel work is steered by, from a .ncu-rep (details page)."""
import csv, subprocess, sys
WANT = ["Duration", "DRAM Throughput", "Memory Throughput", "Registers Per Thread", "Achieved Occupancy",
        "Theoretical Occupancy", "Issue Slots Busy", "No Eligible", "Eligible Warps Per Scheduler", "L2 Hit Rate",
        "L1/TEX Hit Rate", "Executed Ipc Active", "Dynamic Shared Memory Per Block", "Block Limit Registers",
        "Block Limit Shared Mem", "Compute (SM) Throughput", "Local Load", "Local Store", "Warp Cycles Per Issued Instruction",
        "Avg. Active Threads Per Warp", "Avg. Not Predicated Off Threads Per Warp", "Executed Instructions", "Grid Size", "Block Size"]
for rep in sys.argv[1:]:
    out = subprocess.run(["ncu", "-i", rep, "--page", "details", "--csv"], capture_output=True, text=True).stdout
    rows = list(csv.reader(out.splitlines()))
    h = rows[0]
    print("==", rep)
    seen = set()
    for r in rows[1:]:
        d = dict(zip(h, r))
        n = d.get("Metric Name")
        if n in WANT and (n, d["Metric Unit"]) not in seen:
            seen.add((n, d["Metric Unit"]))
            print("  %-45s %s %s" % (n, d["Metric Value"], d["Metric Unit"]))
    for r in rows[1:]:
        d = dict(zip(h, r))
        if d.get("Rule Name") and d.get("Rule Description") and ("Est. Speedup" in d.get("Rule Description", "") or d.get("Estimated Speedup")):
            pass
