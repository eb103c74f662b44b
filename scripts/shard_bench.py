#!/usr/bin/env python
"""Multi-GPU measurement driver for BASELINE.json configs C3 / C4 / C5 (not part of the product; bench.py keeps the
headline contract). One process per GPU under torchrun; data is generated ON the GPU (same clustered-latent recipe as
fixtures.make_dataset, torch generator) and the graph is built by lmd_build_graph, so that nothing large crosses PCIe.

  --mode replicated   C3 / C4: every rank holds the whole index, rank g searches its own query batch (no collective)
  --mode sharded      C5: rank s holds rows [lo_s, hi_s) with a shard-local graph; every rank searches the FULL batch,
                      one all-gather of [Q][k] (distance, row id) + lmd_merge_topk

    python -m torch.distributed.run --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29511 \
        scripts/shard_bench.py --mode sharded --rows 100000000 --dims 96 --nq 10000 --L 64

Timing: barrier + CUDA events on the searching stream, max over ranks; rank 0 prints one JSON line."""
import argparse
import importlib
import itertools
import json
import os
import sys
import time

import numpy as np
import torch
import torch.distributed as dist

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--mode", default="sharded", choices=["sharded", "replicated"])
    ap.add_argument("--rows", dest="n", type=int, default=1_000_000, help="TOTAL rows")
    ap.add_argument("--dims", type=int, default=96)
    ap.add_argument("--r", type=int, default=32)
    ap.add_argument("--nq", type=int, default=10_000, help="sharded: the batch; replicated: queries PER GPU")
    ap.add_argument("--k", type=int, default=10)
    ap.add_argument("--L", type=int, nargs="+", default=[64])
    ap.add_argument("--l-insert", type=int, default=64)
    ap.add_argument("--metric", default="L2")
    ap.add_argument("--route", default="exact", choices=["exact", "ternary", "int8"])
    ap.add_argument("--codec", action="store_true", help="enable the int8 edge codec (exact route: lower-bound pre-filter)")
    ap.add_argument("--flags", type=int, default=0, help="extra LMD_SEARCH_* flags, e.g. 32 = LMD_SEARCH_TENSOR_RERANK")
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--tag", default="")
    ap.add_argument("--sweep", nargs="*", default=[],
                    help="repeat every L over the product of these tuning values, e.g. latency_mode=0,2 filter_log2=12,13")
    ap.add_argument("--tune", nargs="*", default=[], help="lmd_set_tuning pairs, e.g. merge_min=3 tie_slots=32")
    a = ap.parse_args()
    rank, world, local = (int(os.environ.get(v, d)) for v, d in (("RANK", "0"), ("WORLD_SIZE", "1"), ("LOCAL_RANK", "0")))
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    if world > 1:
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        os.environ.setdefault("NCCL_DEBUG", "WARN")
        dist.init_process_group("nccl", device_id=dev)
    lmd = importlib.import_module("duckdb-lm-diskann_b200")
    sh = importlib.import_module("duckdb-lm-diskann_b200.sharded")
    wl = importlib.import_module("duckdb-lm-diskann_b200.workloads")
    gen, topk_scores, recall = wl.gen, wl.topk_scores, wl.recall
    cos = a.metric == "COSINE"
    route = {"exact": lmd.ROUTE_EXACT, "ternary": lmd.ROUTE_TERNARY, "int8": lmd.ROUTE_INT8}[a.route]

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    def log(*m):
        if rank == 0:
            print("[shard_bench]", *m, file=sys.stderr, flush=True)

    t0 = time.time()
    if a.mode == "sharded":
        lo, hi = sh.shard_rows(a.n, rank, world)
        x = gen(hi - lo, a.dims, 42 + rank, dev, cos)          # every shard its own stream of the same distribution
        row_ids = torch.arange(lo, hi, dtype=torch.int64, device=dev)
        q = gen(a.nq, a.dims, 43, dev, cos)                    # the same batch on every rank
    else:
        lo, hi = 0, a.n
        x = gen(a.n, a.dims, 42, dev, cos)                     # the same index on every rank
        row_ids = None
        q = gen(a.nq, a.dims, 43 + 1000 * rank, dev, cos)      # every rank its own batch
    torch.cuda.synchronize()
    t_gen = time.time() - t0
    store = lmd.Store(a.dims, a.r, metric=lmd.METRIC_BY_NAME[a.metric], capacity=hi - lo, device=local)
    t0 = time.time()
    store.upload_vectors(x, row_ids=row_ids)
    bst = store.build_graph(l_insert=a.l_insert, alpha=1.2)
    t_build = time.time() - t0
    log("rows/rank %d, gen %.1f s, upload+build %.1f s (build kernels %.0f ms, %d rounds), store %.1f GB" % (
        hi - lo, t_gen, t_build, bst.total_ms, bst.rounds, store.device_bytes / 1e9))

    # ground truth: exact top-k per shard (or of the whole replica), merged across shards
    t0 = time.time()
    S, I = topk_scores(x, q, a.k, a.metric, base_id=lo)
    if a.mode == "sharded" and world > 1:
        allS = torch.empty((world,) + S.shape, device=dev)
        allI = torch.empty((world,) + I.shape, dtype=torch.int64, device=dev)
        dist.all_gather_into_tensor(allS, S)
        dist.all_gather_into_tensor(allI, I)
        cs = allS.permute(1, 0, 2).reshape(a.nq, -1)
        ci = allI.permute(1, 0, 2).reshape(a.nq, -1)
        _, sel = torch.topk(cs, a.k, dim=1, largest=False)
        truth = torch.gather(ci, 1, sel)
    else:
        truth = I
    del S, I
    torch.cuda.synchronize()
    log("truth %.1f s" % (time.time() - t0))
    del x
    torch.cuda.empty_cache()

    if a.codec:
        t0 = time.time()
        store.enable_codec()
        log("int8 codec built in %.2f s, store %.1f GB" % (time.time() - t0, store.device_bytes / 1e9))
    for kv in a.tune:
        key, val = kv.split("=")
        store.set_tuning(key, int(val))
    # a real (non-default) stream: the default stream's handle is 0, which lmd_store_set_stream reads as "own stream"
    work = torch.cuda.Stream(device=dev)
    torch.cuda.set_stream(work)
    store.set_stream(work.cuda_stream)
    results = []
    sweep = [(kv.split("=")[0], [int(v) for v in kv.split("=")[1].split(",")]) for kv in a.sweep]
    points = [dict(zip([k for k, _ in sweep], vals)) for vals in itertools.product(*[v for _, v in sweep])]
    for L, point in [(L, pt) for L in a.L for pt in points]:
        for key, val in point.items():
            store.set_tuning(key, val)
        if a.mode == "sharded":
            searcher = sh.DatasetShardedSearcher(store, device=dev)

            def step():
                return searcher.search(q, a.k, L, route=route)
        else:
            ids = torch.empty((a.nq, a.k), dtype=torch.int64, device=dev)
            dd = torch.empty((a.nq, a.k), dtype=torch.float32, device=dev)

            def step():
                store.search_raw(q, a.nq, a.k, L, route=route, flags=lmd.SEARCH_ASYNC | a.flags, out_ids=ids, out_dist=dd)
                return ids, dd
        for _ in range(a.warmup):
            out = step()
        st = store.search_raw(q, a.nq, a.k, L, route=route, flags=a.flags, out_ids=torch.empty((a.nq, a.k), dtype=torch.int64, device=dev))
        barrier()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(a.steps):
            out = step()
        e1.record()
        torch.cuda.synchronize()
        ms = torch.tensor([e0.elapsed_time(e1) / a.steps], device=dev)
        rec = torch.tensor([recall(out[0], truth)], device=dev)
        kms = torch.tensor([st.traverse_ms + st.rerank_ms], device=dev)
        if world > 1:
            dist.all_reduce(ms, op=dist.ReduceOp.MAX)
            dist.all_reduce(kms, op=dist.ReduceOp.MAX)
            if a.mode == "replicated":
                dist.all_reduce(rec, op=dist.ReduceOp.SUM)
                rec /= world
        total_q = a.nq if a.mode == "sharded" else a.nq * world
        row = {"L": L, "tuning": point, "ms_per_step": round(float(ms), 4), "qps": round(total_q / (float(ms) * 1e-3), 1),
               "recall_at_%d" % a.k: round(float(rec), 4), "local_kernel_ms_max": round(float(kms), 4),
               "E_per_q": round(st.expansions / a.nq, 1), "V_per_q": round(st.distance_evals / a.nq, 1),
               "fetched_per_q": round(st.exact_fetches / a.nq, 1),
               "local_algorithmic_GBps": round(st.algorithmic_bytes / (st.traverse_ms + st.rerank_ms) / 1e6, 1),
               "traverse_ms": round(st.traverse_ms, 4), "rerank_ms": round(st.rerank_ms, 4),
               "tie_overflows": st.tie_overflows, "merge_fallbacks_per_q": round(st.merge_fallbacks / a.nq, 2),
               "filter_evictions_per_q": round(st.filter_evictions / a.nq, 2)}
        if a.mode == "sharded":
            row["allgather_bytes_per_rank"] = a.nq * a.k * 12
        results.append(row)
        log(row)
    if rank == 0:
        print(json.dumps({"driver": "scripts/shard_bench.py", "tag": a.tag, "mode": a.mode, "n_total": a.n,
                          "dims": a.dims, "r": a.r, "metric": a.metric, "route": a.route, "flags": a.flags, "tune": a.tune, "n_gpus": world,
                          "rows_per_gpu": hi - lo, "queries_per_step_total": a.nq if a.mode == "sharded" else a.nq * world,
                          "k": a.k, "l_insert": a.l_insert, "steps": a.steps, "warmup": a.warmup,
                          "gen_s": round(t_gen, 2), "upload_build_s": round(t_build, 2),
                          "build_kernels_ms": round(bst.total_ms, 1), "store_GB_per_gpu": round(store.device_bytes / 1e9, 2),
                          "data": "synthetic clustered-latent (torch CUDA generator)", "results": results}))
    store.close()
    if world > 1:
        dist.barrier()
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
