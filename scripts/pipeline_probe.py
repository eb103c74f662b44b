#!/usr/bin/env python
"""Do consecutive 10k-query batches overlap when they are issued on different lmd_ctx streams? (the tail of a
persistent-kernel launch leaves SMs idle; the next batch's kernel can fill them). Device-resident buffers, CUDA events."""
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
ap.add_argument("--steps", type=int, default=60)
ap.add_argument("--codec", action="store_true")
ap.add_argument("--route", default="exact")
ap.add_argument("--tune", nargs="*", default=[])
a = ap.parse_args()
dev = torch.device("cuda", 0)
x = wl.gen(a.rows, a.dims, 42, dev)
store, _ = wl.build_store(lmd, x, 32, "L2", 0)
if a.codec:
    store.enable_codec()
for kv in a.tune:
    store.set_tuning(kv.split("=")[0], int(kv.split("=")[1]))
route = {"exact": lmd.ROUTE_EXACT, "int8": lmd.ROUTE_INT8}[a.route]
qs = [wl.gen(a.nq, a.dims, 43 + 17 * b, dev) for b in range(4)]
out = []
for depth in (1, 2, 3, 4):
    ctxs = [store.context() for _ in range(depth)]
    streams = [torch.cuda.Stream(device=dev) for _ in range(depth)]
    ids = [torch.empty((a.nq, 10), dtype=torch.int64, device=dev) for _ in range(depth)]
    dd = [torch.empty((a.nq, 10), dtype=torch.float32, device=dev) for _ in range(depth)]
    for c, s in zip(ctxs, streams):
        c.set_stream(s.cuda_stream)
    main = torch.cuda.Stream(device=dev)

    def run(n):
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        torch.cuda.synchronize()
        e0.record(main)
        for s in streams:
            s.wait_event(e0)
        for i in range(n):
            j = i % depth
            store.search_raw(qs[i % 4], a.nq, 10, a.L, route=route, flags=lmd.SEARCH_ASYNC, out_ids=ids[j], out_dist=dd[j],
                             ctx=ctxs[j])
        for s in streams:
            ev = torch.cuda.Event()
            ev.record(s)
            main.wait_event(ev)
        e1.record(main)
        torch.cuda.synchronize()
        return e0.elapsed_time(e1) / n
    run(8)
    ms = run(a.steps)
    out.append({"streams": depth, "ms_per_step": round(ms, 4), "qps": round(a.nq / ms * 1e3)})
    print(out[-1], file=sys.stderr, flush=True)
    for c in ctxs:
        c.close()
print(json.dumps({"probe": "batches in flight", "rows": a.rows, "dims": a.dims, "nq": a.nq, "L": a.L, "codec": a.codec,
                  "route": a.route, "tune": a.tune, "results": out}))
