"""Debug: raw bits of the special-value case (NaN / Inf rows) through the C ABI."""
import importlib, os, sys
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "oracle"))
lmd = importlib.import_module("duckdb-lm-diskann_b200")
fx = importlib.import_module("duckdb-lm-diskann_b200.fixtures")
import pyoracle
r, n, dims = 32, 33, 128
rng = np.random.default_rng(11)
x = rng.standard_normal((n, dims)).astype(np.float32)
x[8, 5] = np.inf; x[10, 0] = np.nan
q = rng.standard_normal((2, dims)).astype(np.float32); q[0] = 0.0
nbrs = np.full((n, r), -1, dtype=np.int64); nbrs[0, :n - 1] = np.arange(1, n)
counts = np.zeros(n, dtype=np.int64); counts[0] = n - 1
blocks = fx.pack_blocks(x, nbrs, counts, r).numpy()
lay = fx.layout(dims, r)
TUNE = [t.split("=") for t in sys.argv[1:]]
for metric in (1, 3):
    for flags in (0, lmd.SEARCH_SEPARATE_RERANK, lmd.SEARCH_STRICT_FP):
        with lmd.Store(dims, r, metric=metric, capacity=n) as s:
            s.upload(blocks); s.set_entry(0)
            for key, val in TUNE:
                s.set_tuning(key, int(val))
            ids, dist, cnt, st = s.search(q, 40, 40, flags=flags)
        o = pyoracle.search(blocks, lay["block_bytes"], dims, r, q, 40, 40, metric=metric, entry_row_id=0)
        print("metric", metric, "flags", flags, "cnt", cnt, o[2])
        for qi in range(2):
            print(" ours  ", [(int(i), hex(int(b))) for i, b in zip(ids[qi, 28:34], dist[qi, 28:34].view(np.uint32))])
            print(" oracle", [(int(i), hex(int(b))) for i, b in zip(o[0][qi, 28:34], o[1][qi, 28:34].view(np.uint32))])
            w = {int(i): d for i, d in zip(ids[qi, :n], dist[qi, :n])}
            print("  row 8, 10 ours:", w.get(8), w.get(10), "positions", list(ids[qi, :n]).index(8), list(ids[qi, :n]).index(10))
