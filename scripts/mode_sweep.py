#!/usr/bin/env python
"""Kernel time of the traversal, throughput vs latency variants, over batch sizes (device-resident buffers).
    python scripts/mode_sweep.py [--dims 128 --metric L2 --L 56]"""
import argparse, importlib, json, os, sys
import torch
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "scripts"))
import importlib as _il  # noqa: E402
_wl = _il.import_module("duckdb-lm-diskann_b200.workloads")
gen = _wl.gen
ap = argparse.ArgumentParser()
ap.add_argument("--rows", dest="n", type=int, default=1_000_000)
ap.add_argument("--dims", type=int, default=128)
ap.add_argument("--L", type=int, default=56)
ap.add_argument("--metric", default="L2")
ap.add_argument("--nqs", type=int, nargs="*", default=[1024, 1480, 1776, 2048, 2368, 3000, 4096, 6000, 10000])
ap.add_argument("--filter-log2", type=int, nargs="*", default=[0], help="visited-filter sizes (0 = the library's choice)")
a = ap.parse_args()
lmd = importlib.import_module("duckdb-lm-diskann_b200")
dev = torch.device("cuda", 0)
x = gen(a.n, a.dims, 42, dev, a.metric == "COSINE")
store = lmd.Store(a.dims, 32, metric=lmd.METRIC_BY_NAME[a.metric], capacity=a.n)
store.upload_vectors(x); store.build_graph(l_insert=64); del x
q = gen(16384, a.dims, 43, dev, a.metric == "COSINE")
ids = torch.empty((16384, 10), dtype=torch.int64, device=dev)
rows = []
for nq, flg in [(n, f) for n in a.nqs for f in a.filter_log2]:
    row = {"nq": nq, "filter_log2": flg}
    store.set_tuning("filter_log2", flg)
    for mode, name in ((0, "throughput"), (1, "latency"), (2, "mid"), (-1, "auto")):
        store.set_tuning("latency_mode", mode)
        best = 1e9
        for i in range(8):
            st = store.search_raw(q, nq, 10, a.L, out_ids=ids)
            if i >= 2:
                best = min(best, st.traverse_ms)
        row[name + "_ms"] = round(best, 4)
    row["best"] = min(("throughput", "latency", "mid"), key=lambda k: row[k + "_ms"])
    rows.append(row); print(row, file=sys.stderr, flush=True)
print(json.dumps({"driver": "scripts/mode_sweep.py", "dims": a.dims, "metric": a.metric, "L": a.L, "rows": rows}))
