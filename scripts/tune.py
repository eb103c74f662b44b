#!/usr/bin/env python
"""GPU experiment driver (not part of the product): one 1M x 128 index, a grid of tuning knobs, kernel time per setting.
    python scripts/tune.py [--L 64] [--n 1000000]"""
import argparse
import importlib
import itertools
import os
import sys

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "oracle"))
import bench  # noqa: E402

ap = argparse.ArgumentParser()
ap.add_argument("--n", type=int, default=1_000_000)
ap.add_argument("--dims", type=int, default=128)
ap.add_argument("--r", type=int, default=32)
ap.add_argument("--nq", type=int, default=10_000)
ap.add_argument("--k", type=int, default=10)
ap.add_argument("--L", type=int, nargs="+", default=[64])
ap.add_argument("--kind", default="clustered")
ap.add_argument("--grid", default="filter_log2=0,10,11,12,13;ctas_per_sm=0")
ap.add_argument("--oracle-sample", type=int, default=256)
ap.add_argument("--metric", default="L2")
ap.add_argument("--route", default="exact")
a = ap.parse_args()
fx = importlib.import_module("duckdb-lm-diskann_b200.fixtures")
lmd = importlib.import_module("duckdb-lm-diskann_b200")
x, nbrs, counts = bench.load_or_build_index(a, fx, "cuda")
blocks = fx.pack_blocks(x, nbrs, counts, a.r, device="cuda")
store = lmd.Store(a.dims, a.r, metric=lmd.METRIC_BY_NAME[a.metric], capacity=a.n)
route = 1 if a.route == "ternary" else 0
if route:
    a.oracle_sample = 0
store.upload(blocks)
store.set_entry(0)
q = [fx.make_dataset(a.nq, a.dims, kind=a.kind, seed=43 + 17 * b, normalize=(a.metric == "COSINE")) for b in range(3)]
dq = [torch.from_numpy(b).cuda() for b in q]
d_ids = torch.empty((a.nq, a.k), dtype=torch.int64, device="cuda")
d_dist = torch.empty((a.nq, a.k), dtype=torch.float32, device="cuda")
if a.oracle_sample:
    import pyoracle
    bnp = blocks.cpu().numpy()
    for L in a.L:
        o = pyoracle.search(bnp, bnp.shape[1], a.dims, a.r, q[0][:a.oracle_sample], a.k, L, entry_row_id=0,
                            impl="fixed" if pyoracle.have_ref("fixed") else "port")
        g = store.search(q[0][:a.oracle_sample], a.k, L, qstats=True)
        v_or = o[3][:, 1].astype(np.int64) - np.minimum(L, o[3][:, 1]) - o[2]  # minus rerank + out_dist fetches
        print("L=%d oracle V/q %.1f  gpu V/q %.1f  E oracle %.1f gpu %.1f ids equal %.4f" % (
            L, v_or.mean(), g[4][:, 1].mean(), o[3][:, 0].mean(), g[4][:, 0].mean(),
            (o[0] == g[0]).all(axis=1).mean()), flush=True)
    del bnp
del blocks
keys, vals = [], []
for part in a.grid.split(";"):
    k, v = part.split("=")
    keys.append(k)
    vals.append([int(t) for t in v.split(",")])
for L in a.L:
    for combo in itertools.product(*vals):
        for k, v in zip(keys, combo):
            store.set_tuning(k, v)
        ms = []
        try:
            for it in range(6):
                st = store.search_raw(dq[it % 3], a.nq, a.k, L, route=route, out_ids=d_ids, out_dist=d_dist)
                if it >= 2:
                    ms.append(st.traverse_ms)
        except Exception as e:
            print("L=%d %s  FAILED: %s" % (L, dict(zip(keys, combo)), e), flush=True)
            continue
        print("L=%d %s  rerank %.4f ms" % (L, dict(zip(keys, combo)), st.rerank_ms), flush=True) if route else None
        print("L=%d %s  traverse %.4f ms (min %.4f)  V/q %.1f evict/q %.1f fallbacks %d  alg GB/s %.0f" % (
            L, dict(zip(keys, combo)), np.mean(ms), min(ms), st.distance_evals / a.nq, st.filter_evictions / a.nq,
            st.merge_fallbacks, st.algorithmic_bytes / (np.mean(ms) * 1e-3) / 1e9), flush=True)
