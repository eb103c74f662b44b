#!/usr/bin/env python
"""GPU experiment driver (not part of the product): lmd_build_graph on a synthetic data set, build statistics and the
recall / kernel time of searches on the graph it built.
    python scripts/build_bench.py [--n 1000000 --dims 128 --r 32 --l-insert 128 --alpha 1.2 --max-batch 0]"""
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
ap.add_argument("--n", type=int, default=1_000_000)
ap.add_argument("--dims", type=int, default=128)
ap.add_argument("--r", type=int, default=32)
ap.add_argument("--l-insert", type=int, default=128)
ap.add_argument("--alpha", type=float, default=1.2)
ap.add_argument("--max-batch", type=int, default=0)
ap.add_argument("--nq", type=int, default=10_000)
ap.add_argument("--kind", default="clustered")
ap.add_argument("--metric", default="L2")
ap.add_argument("--Ls", type=int, nargs="+", default=[32, 48, 56, 64, 96, 128])
a = ap.parse_args()
fx = importlib.import_module("duckdb-lm-diskann_b200.fixtures")
lmd = importlib.import_module("duckdb-lm-diskann_b200")
x = fx.make_dataset(a.n, a.dims, kind=a.kind, seed=42, normalize=(a.metric == "COSINE"))
q = fx.make_dataset(a.nq, a.dims, kind=a.kind, seed=43, normalize=(a.metric == "COSINE"))
t0 = time.time()
truth = fx.brute_force_topk(x, q, 10, a.metric, device="cuda")
print("truth %.1f s" % (time.time() - t0), flush=True)
store = lmd.Store(a.dims, a.r, metric=lmd.METRIC_BY_NAME[a.metric], capacity=a.n)
t0 = time.time()
store.upload_vectors(x)
t1 = time.time()
st = store.build_graph(l_insert=a.l_insert, alpha=a.alpha, max_batch=a.max_batch)
t2 = time.time()
out = {"n": a.n, "dims": a.dims, "r": a.r, "l_insert": a.l_insert, "alpha": a.alpha, "upload_s": t1 - t0,
       "build_wall_s": t2 - t1, "build": st.as_dict(), "search": []}
blocks0 = store.download_blocks(0, 1000)
cnt = blocks0[:, 0:2].copy().view(np.uint16)[:, 0]
out["mean_degree_first_1000"] = float(cnt.mean())
dq = torch.from_numpy(q).cuda()
d_ids = torch.empty((a.nq, 10), dtype=torch.int64, device="cuda")
for L in a.Ls:
    store.search_raw(dq, a.nq, 10, L, out_ids=d_ids)
    s2 = store.search_raw(dq, a.nq, 10, L, out_ids=d_ids)
    rec = fx.recall_at_k(d_ids.cpu().numpy(), truth)
    out["search"].append({"L": L, "recall": round(rec, 4), "traverse_ms": round(s2.traverse_ms, 4),
                          "E_per_q": round(s2.expansions / a.nq, 1), "V_per_q": round(s2.distance_evals / a.nq, 1)})
    print(out["search"][-1], flush=True)
print(json.dumps(out))
