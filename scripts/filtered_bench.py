#!/usr/bin/env python
"""Filtered search (row f4) on the headline shape: QPS and recall against brute force over the ALLOWED rows, for a few
selectivities. Device-resident buffers; kernel time from lmd_stats.
    python scripts/filtered_bench.py [--rows 1000000 --dims 128]"""
import argparse, importlib, json, os, sys
import torch
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "scripts"))
import importlib as _il  # noqa: E402
_wl = _il.import_module("duckdb-lm-diskann_b200.workloads")
gen = _wl.gen
topk_scores = _wl.topk_scores
recall = _wl.recall
ap = argparse.ArgumentParser()
ap.add_argument("--rows", dest="n", type=int, default=1_000_000)
ap.add_argument("--dims", type=int, default=128)
ap.add_argument("--nq", type=int, default=10_000)
ap.add_argument("--filter-log2", type=int, nargs="*", default=[0], help="visited-filter sizes to compare (0 = the library's choice)")
ap.add_argument("--fracs", type=float, nargs="*", default=[1.0, 0.5, 0.1, 0.01])
ap.add_argument("--Ls", type=int, nargs="*", default=[64, 128, 256])
ap.add_argument("--tune", nargs="*", default=[], help="lmd_set_tuning pairs, e.g. filter_mode=3 filter_entries=3072")
a = ap.parse_args()
lmd = importlib.import_module("duckdb-lm-diskann_b200")
dev = torch.device("cuda", 0)
x = gen(a.n, a.dims, 42, dev, False)
q = gen(a.nq, a.dims, 43, dev, False)
store = lmd.Store(a.dims, 32, capacity=a.n)
store.upload_vectors(x); store.build_graph(l_insert=64)
for kv in a.tune:
    store.set_tuning(kv.split("=")[0], int(kv.split("=")[1]))
ids = torch.empty((a.nq, 10), dtype=torch.int64, device=dev)
cnt = torch.empty((a.nq,), dtype=torch.int32, device=dev)
rows = []
g = torch.Generator(device=dev); g.manual_seed(5)
for frac in a.fracs:
    mask = torch.rand(a.n, generator=g, device=dev) < frac
    allowed = torch.nonzero(mask).squeeze(1)
    xs = x[allowed]
    _, ti = topk_scores(xs, q, 10, "L2", 0)
    truth = allowed[ti]
    del xs
    with store.make_filter(allowed) as f:
        for L, flg in [(L, v) for L in a.Ls for v in a.filter_log2]:
            store.set_tuning("filter_log2", flg)
            best = None
            for i in range(5):
                st = store.search_raw(q, a.nq, 10, L, out_ids=ids, out_count=cnt, filter=f)
                if i >= 2 and (best is None or st.traverse_ms < best.traverse_ms):
                    best = st
            rows.append({"allowed_fraction": frac, "allowed_rows": int(allowed.numel()), "L": L, "filter_log2": flg,
                         "recall_at_10_vs_allowed_truth": round(recall(ids, truth), 4),
                         "queries_with_fewer_than_k": int((cnt < 10).sum()),
                         "traverse_ms": round(best.traverse_ms, 4), "kernel_qps": round(a.nq / (best.traverse_ms * 1e-3)),
                         "E_per_q": round(best.expansions / a.nq, 1), "V_per_q": round(best.distance_evals / a.nq, 1)})
            print(rows[-1], file=sys.stderr, flush=True)
store.set_tuning("filter_log2", 0)
for L in (64, 128):
    st = None
    for i in range(4):
        st = store.search_raw(q, a.nq, 10, L, out_ids=ids)
    rows.append({"allowed_fraction": "unfiltered kernel", "L": L, "traverse_ms": round(st.traverse_ms, 4)})
    print(rows[-1], file=sys.stderr, flush=True)
print(json.dumps({"driver": "scripts/filtered_bench.py", "n": a.n, "dims": a.dims, "nq": a.nq, "rows": rows}))
