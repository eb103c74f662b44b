#!/usr/bin/env python
"""Cost of keeping the mirror live (row f2) on the headline shape: lmd_store_apply for dirty / new blocks (pageable host
memory, as a host's buffer-managed blocks are), lmd_store_remove, and the first search after each (which rebuilds the
neighbour-slot sidecar when row ids were added or removed).  python scripts/live_bench.py [--rows 1000000]"""
import argparse, importlib, json, os, sys, time
import numpy as np, torch
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "scripts"))
import importlib as _il  # noqa: E402
_wl = _il.import_module("duckdb-lm-diskann_b200.workloads")
gen = _wl.gen
ap = argparse.ArgumentParser()
ap.add_argument("--rows", dest="n", type=int, default=1_000_000)
ap.add_argument("--dims", type=int, default=128)
a = ap.parse_args()
lmd = importlib.import_module("duckdb-lm-diskann_b200")
dev = torch.device("cuda", 0)
extra = 20_000
x = gen(a.n + extra, a.dims, 42, dev, False)
store = lmd.Store(a.dims, 32, capacity=a.n + extra)
store.upload_vectors(x[:a.n]); store.build_graph(l_insert=64)
q = gen(1000, a.dims, 43, dev, False)
ids = torch.empty((1000, 10), dtype=torch.int64, device=dev)
bb = store.block_bytes
host = store.download_blocks(0, 20_000)          # 20 k real blocks of the index as the host would hold them
rows = []


def timed(f):
    torch.cuda.synchronize(); t0 = time.perf_counter(); r = f(); torch.cuda.synchronize()
    return (time.perf_counter() - t0) * 1e3, r


def search_ms():
    return timed(lambda: store.search_raw(q, 1000, 10, 56, out_ids=ids, stats=False))[0]


search_ms()
base = search_ms()
for m in (1, 32, 1000, 20_000):                   # dirty blocks of EXISTING rows (MarkBlockDirty)
    rid = np.arange(m, dtype=np.int64)
    ms, _ = timed(lambda: store.apply(rid, host[:m]))
    rows.append({"op": "apply (dirty, existing rows)", "blocks": m, "ms": round(ms, 3),
                 "blocks_per_s": round(m / (ms * 1e-3)), "GB_per_s": round(m * bb / ms / 1e6, 2),
                 "next_search_ms": round(search_ms(), 3)})
newblk = host.copy()                               # NEW rows (AllocateNodeBlock): same blocks under new row ids
for m in (1, 1000, 18_000):
    start = a.n + sum(r["blocks"] for r in rows if r["op"].startswith("apply (new"))
    rid = np.arange(start, start + m, dtype=np.int64)
    ms, app = timed(lambda: store.apply(rid, newblk[:m]))
    assert app == m
    rows.append({"op": "apply (new rows)", "blocks": m, "ms": round(ms, 3), "blocks_per_s": round(m / (ms * 1e-3)),
                 "next_search_ms": round(search_ms(), 3), "note": "next search rebuilds the whole neighbour-slot sidecar"})
for m in (1, 1000):
    rid = np.arange(5000 + 10 * m, 5000 + 11 * m, dtype=np.int64)
    ms, rem = timed(lambda: store.remove(rid))
    assert rem == m
    rows.append({"op": "remove", "row_ids": m, "ms": round(ms, 3), "next_search_ms": round(search_ms(), 3)})
rows.append({"op": "search, nothing pending (1000 queries, L = 56)", "ms": round(base, 3), "again_ms": round(search_ms(), 3)})
for r in rows:
    print(r, file=sys.stderr, flush=True)
print(json.dumps({"driver": "scripts/live_bench.py", "n": a.n, "dims": a.dims, "block_bytes": bb, "rows": rows}))
