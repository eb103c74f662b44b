#!/usr/bin/env python
"""One 10k-query batch ALONE (no next batch to fill the persistent kernel's last, partly empty round): does it pay to give
the queries beyond the last full round to the small-batch kernels on a second stream? Device-resident buffers, CUDA
events, every step synchronised (nothing overlaps across steps)."""
import argparse, importlib, json, os, sys
import torch
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
lmd = importlib.import_module("duckdb-lm-diskann_b200")
wl = importlib.import_module("duckdb-lm-diskann_b200.workloads")
ap = argparse.ArgumentParser()
ap.add_argument("--rows", type=int, default=1_000_000)
ap.add_argument("--dims", type=int, default=128)
ap.add_argument("--nq", type=int, default=10_000)
ap.add_argument("--L", type=int, default=56)
ap.add_argument("--steps", type=int, default=40)
ap.add_argument("--splits", type=int, nargs="*", default=[0, 8288, 8400, 8800, 9200, 7600])
a = ap.parse_args()
dev = torch.device("cuda", 0)
x = wl.gen(a.rows, a.dims, 42, dev)
store, _ = wl.build_store(lmd, x, 32, "L2", 0)
store.enable_codec()
q = wl.gen(a.nq, a.dims, 43, dev)
ctxs = [store.context() for _ in range(2)]
streams = [torch.cuda.Stream(device=dev) for _ in range(2)]
for c, s in zip(ctxs, streams):
    c.set_stream(s.cuda_stream)
ids = torch.empty((a.nq, 10), dtype=torch.int64, device=dev)
dd = torch.empty((a.nq, 10), dtype=torch.float32, device=dev)
ref = None
out = []
for split in a.splits:
    parts = [(0, a.nq)] if split == 0 else [(0, split), (split, a.nq)]

    def step():
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(streams[0])
        streams[1].wait_event(e0)
        for j, (lo, hi) in enumerate(parts):
            store.search_raw(q[lo:hi], hi - lo, 10, a.L, flags=lmd.SEARCH_ASYNC, out_ids=ids[lo:hi], out_dist=dd[lo:hi],
                             ctx=ctxs[j])
        ev = torch.cuda.Event()
        ev.record(streams[1])
        streams[0].wait_event(ev)
        e1.record(streams[0])
        torch.cuda.synchronize()
        return e0.elapsed_time(e1)
    for _ in range(5):
        step()
    ts = sorted(step() for _ in range(a.steps))
    if ref is None:
        ref = (ids.clone(), dd.clone())
    same = bool(torch.equal(ids, ref[0]) and torch.equal(dd, ref[1]))
    out.append({"split": split, "ms_median": round(ts[len(ts) // 2], 4), "ms_min": round(ts[0], 4), "same_results": same})
    print(out[-1], file=sys.stderr, flush=True)
print(json.dumps({"probe": "tail split", "nq": a.nq, "L": a.L, "results": out}))
