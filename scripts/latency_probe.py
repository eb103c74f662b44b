#!/usr/bin/env python
"""One query at a time through lmd_search_batch: pinned against pageable buffers, with / without the code sidecar, row
prefetch on / off. Median of 300 calls after 50 warm-ups."""
import importlib, json, os, sys, time
import numpy as np, torch
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
lmd = importlib.import_module("duckdb-lm-diskann_b200")
wl = importlib.import_module("duckdb-lm-diskann_b200.workloads")
dev = torch.device("cuda", 0)
x = wl.gen(1_000_000, 128, 42, dev)
store, _ = wl.build_store(lmd, x, 32, "L2", 0)
q = wl.gen(512, 128, 43, dev).cpu()
qp = q.pin_memory()
qn = q.numpy().copy()
ids_p = torch.empty((1, 10), dtype=torch.int64).pin_memory(); d_p = torch.empty((1, 10), dtype=torch.float32).pin_memory()
ids_n = np.empty((1, 10), dtype=np.int64); d_n = np.empty((1, 10), dtype=np.float32)
rows = []
def med(fn):
    for i in range(50): fn(i)
    t = []
    for i in range(300):
        t0 = time.perf_counter(); fn(i); t.append(time.perf_counter() - t0)
    return round(sorted(t)[150] * 1e6, 1)
for codec in (0, 1):
    if codec:
        store.enable_codec()
    for pre in ((1, 2) if codec else (1,)):
        store.set_tuning("prefilter", pre)
        for rp, nbp in ((2, 0), (2, 1), (0, 0)):
            store.set_tuning("row_prefetch", rp)
            store.set_tuning("nbr_prefetch", nbp)
            a = med(lambda i: store.search_raw(qp[i % 512:i % 512 + 1], 1, 10, 56, out_ids=ids_p, out_dist=d_p, stats=False))
            b = med(lambda i: store.search_raw(qn[i % 512:i % 512 + 1], 1, 10, 56, out_ids=ids_n, out_dist=d_n, stats=False))
            st = store.search_raw(qp[0:1], 1, 10, 56, out_ids=ids_p, out_dist=d_p)
            rows.append({"codec": codec, "prefilter": pre, "row_prefetch": rp, "nbr_prefetch": nbp, "pinned_us": a, "pageable_us": b,
                         "kernel_us": round(st.traverse_ms * 1e3, 1)})
            print(rows[-1], file=sys.stderr, flush=True)
print(json.dumps({"probe": "single-query latency, 1M x 128, L=56", "rows": rows}))
