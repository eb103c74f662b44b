#!/usr/bin/env python
"""Small-batch latency of the search call (pinned host buffers in, pinned results out, one synchronous
lmd_search_batch per measurement): throughput kernels vs latency kernels, 1 ... 4096 queries per call.
    python scripts/latency_bench.py [--n 1000000 --dims 128 --L 56]"""
import argparse
import importlib
import json
import os
import sys
import time

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
ap = argparse.ArgumentParser()
ap.add_argument("--rows", dest="n", type=int, default=1_000_000)
ap.add_argument("--dims", type=int, default=128)
ap.add_argument("--r", type=int, default=32)
ap.add_argument("--L", type=int, default=56)
ap.add_argument("--metric", default="L2")
ap.add_argument("--reps", type=int, default=200)
a = ap.parse_args()
lmd = importlib.import_module("duckdb-lm-diskann_b200")
sys.path.insert(0, os.path.join(ROOT, "scripts"))
import importlib as _il  # noqa: E402
_wl = _il.import_module("duckdb-lm-diskann_b200.workloads")
gen = _wl.gen

dev = torch.device("cuda", 0)
x = gen(a.n, a.dims, 42, dev, a.metric == "COSINE")
store = lmd.Store(a.dims, a.r, metric=lmd.METRIC_BY_NAME[a.metric], capacity=a.n)
store.upload_vectors(x)
store.build_graph(l_insert=64)
del x
qall = gen(8192, a.dims, 43, dev, a.metric == "COSINE").cpu().pin_memory()
ids = torch.empty((8192, 10), dtype=torch.int64).pin_memory()
rows = []
for nq in (1, 4, 16, 64, 256, 592, 1024, 1480, 2048, 4096):
    row = {"nq": nq}
    for mode, name in ((0, "throughput_kernels"), (1, "latency_kernels"), (-1, "auto")):
        store.set_tuning("latency_mode", mode)
        for _ in range(10):
            store.search_raw(qall, nq, 10, a.L, out_ids=ids, stats=False)
        ts = []
        for i in range(a.reps):
            off = (i * 131) % (8192 - nq + 1)
            t0 = time.perf_counter()
            store.search_raw(qall[off:off + nq], nq, 10, a.L, out_ids=ids, stats=False)
            ts.append(time.perf_counter() - t0)
        ts.sort()
        row[name + "_us_median"] = round(ts[len(ts) // 2] * 1e6, 1)
        row[name + "_us_p95"] = round(ts[int(len(ts) * 0.95)] * 1e6, 1)
    row["speedup_median"] = round(row["throughput_kernels_us_median"] / row["latency_kernels_us_median"], 2)
    row["auto_qps"] = round(nq / (row["auto_us_median"] * 1e-6))
    rows.append(row)
    print(row, file=sys.stderr, flush=True)
print(json.dumps({"driver": "scripts/latency_bench.py", "n": a.n, "dims": a.dims, "r": a.r, "L": a.L, "metric": a.metric,
                  "what": "wall time of one synchronous lmd_search_batch call, pinned host queries and results",
                  "rows": rows}))
