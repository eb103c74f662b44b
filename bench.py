#!/usr/bin/env python
"""bench.py - LM_DISKANN query hot path on B200: QPS at recall@10 >= 0.95 on synthetic 1M x 128 L2 (BASELINE.json
configs[1]), 10k-query batches, k = 10, graph replicated per GPU and queries sharded (one 10k batch per GPU).

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl ours|reference] [--L L] [--n N --nq Q] [--sweep]

A "step" is one pass of the hot path (traversal + exact top-k) over one 10k-query batch per GPU.
  value     whole-job QPS with queries / results resident in HBM, timed with CUDA events on the launching stream
  e2e       the same through the C ABI with HOST buffers (pinned): H2D of the queries and D2H of ids+distances inside
            the timed region
  roofline  traversal kernel: algorithmic bytes (SURVEY.md section 8d formula, from the kernel's own E / V counters)
            / CUDA-event duration of that kernel, against MEASURED_PEAKS.json hbm_gbs
  cpu_baseline  the reference's own Searcher.cpp (+ oracle/patches/searcher_fixed.patch) compiled into oracle/_ref,
            timed on this box's host cores on a bounded sample of the same queries
`--impl reference` times that CPU implementation alone on the same workload (all host threads).
"""
import argparse
import importlib
import json
import os
import subprocess
import sys
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)
CACHE = os.environ.get("LMD_BENCH_CACHE", "/tmp/lmd_bench_cache")


def log(*a):
    if int(os.environ.get("RANK", "0")) == 0:
        print("[bench]", *a, file=sys.stderr, flush=True)


def parse():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=100)
    ap.add_argument("--warmup", type=int, default=10)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--n", type=int, default=1_000_000)
    ap.add_argument("--dims", type=int, default=128)
    ap.add_argument("--r", type=int, default=32)
    ap.add_argument("--nq", type=int, default=10_000)
    ap.add_argument("--k", type=int, default=10)
    ap.add_argument("--L", type=int, default=0, help="0 = smallest L of the sweep reaching the recall target")
    ap.add_argument("--recall", type=float, default=0.95)
    ap.add_argument("--kind", default="clustered")
    ap.add_argument("--metric", default="L2", choices=["L2", "COSINE", "IP"])
    ap.add_argument("--route", default="exact", choices=["exact", "ternary"],
                    help="ternary = LM-DiskANN compressed-neighbour routing + exact rerank (COSINE / IP only)")
    ap.add_argument("--sweep", action="store_true", help="also print the whole L sweep (stderr)")
    ap.add_argument("--cpu-seconds", type=float, default=15.0)
    ap.add_argument("--no-cpu", action="store_true")
    return ap.parse_args()


L_SWEEP = [32, 40, 48, 56, 64, 80, 96, 128, 192, 256]
N_BATCHES = 4  # distinct query batches rotated through the steps


def load_or_build_index(a, fx, device):
    """x [n, dims] f32, nbrs [n, r] (slots, -1 empty), counts. Cached on local disk so that the reference arm and
    the other ranks of the same box reuse the graph (graph provenance is free: parity is same graph, same queries)."""
    tag = "n%d_d%d_r%d_%s%s_v2" % (a.n, a.dims, a.r, a.kind, "" if a.metric == "L2" else "_" + a.metric)
    path = os.path.join(CACHE, tag + ".npz")
    if os.path.exists(path):
        z = np.load(path)
        log("index cache hit", path)
        return z["x"], z["nbrs"].astype(np.int64), z["counts"].astype(np.int64)
    t0 = time.time()
    x = fx.make_dataset(a.n, a.dims, kind=a.kind, seed=42, normalize=(a.metric == "COSINE"))
    nbrs, counts = fx.build_graph(x, a.r, device=device, log=log)
    nbrs, counts = nbrs.numpy(), counts.numpy()
    log("graph built in %.1f s" % (time.time() - t0))
    try:
        os.makedirs(CACHE, exist_ok=True)
        tmp = path + ".%d.tmp.npz" % os.getpid()
        np.savez(tmp, x=x, nbrs=nbrs.astype(np.int32), counts=counts.astype(np.int32))
        os.replace(tmp, path)
    except OSError as e:
        log("cache write failed:", e)
    return x, nbrs, counts


def load_or_build_truth(a, fx, x, queries, device):
    tag = "truth_n%d_d%d_%s%s_q%d_k%d_v2" % (a.n, a.dims, a.kind, "" if a.metric == "L2" else "_" + a.metric,
                                             queries.shape[0], a.k)
    path = os.path.join(CACHE, tag + ".npy")
    if os.path.exists(path):
        return np.load(path)
    t = fx.brute_force_topk(x, queries, a.k, a.metric, device=device)
    try:
        os.makedirs(CACHE, exist_ok=True)
        np.save(path + ".%d.tmp.npy" % os.getpid(), t)
        os.replace(path + ".%d.tmp.npy" % os.getpid(), path)
    except OSError:
        pass
    return t


class ClockSampler:
    """nvidia-smi clocks + throttle reasons while the timed region runs (B200_PROFILING.md recipe)."""

    Q = ("clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,"
         "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,"
         "clocks_event_reasons.sw_power_cap")

    def __init__(self, index):
        self.index, self.rows, self.stop = index, [], threading.Event()
        self.th = threading.Thread(target=self.run, daemon=True)

    def run(self):
        while not self.stop.is_set():
            try:
                out = subprocess.run(["nvidia-smi", "-i", str(self.index), "--query-gpu=" + self.Q,
                                      "--format=csv,noheader,nounits"], capture_output=True, text=True, timeout=5)
                if out.returncode == 0 and out.stdout.strip():
                    self.rows.append([c.strip() for c in out.stdout.strip().split(",")])
            except Exception:
                pass
            self.stop.wait(0.05)

    def __enter__(self):
        self.th.start()
        return self

    def __exit__(self, *a):
        self.stop.set()
        self.th.join(timeout=6)

    def summary(self):
        if not self.rows:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["unavailable"], "samples": 0}
        sm = sorted(float(r[0]) for r in self.rows)
        reasons = []
        for i, name in enumerate(["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]):
            if any(r[3 + i].lower().startswith("active") for r in self.rows):
                reasons.append(name)
        return {"sm_mhz": sm[len(sm) // 2], "sm_max_mhz": float(self.rows[0][1]), "reasons": reasons,
                "samples": len(self.rows), "power_w_max": max(float(r[2]) for r in self.rows)}


def cpu_reference_qps(a, blocks_np, block_bytes, batches, L, seconds, nthreads=0):
    """Times the reference's CPU search (oracle/_ref fixed build when present, else the C port) on a bounded sample:
    the rotated query batches, over and over, until about `seconds` of CPU work are done. Returns the ids of batch 0
    (the parity / recall sample)."""
    sys.path.insert(0, os.path.join(ROOT, "oracle"))
    import pyoracle
    kind = "reference" if pyoracle.have_ref("fixed") and a.route == "exact" else "port"
    impl = "fixed" if kind == "reference" else "port"  # the reference never executes ternary routing
    metric = {"L2": 1, "COSINE": 2, "IP": 3}[a.metric]
    cores = os.cpu_count() or 1
    if nthreads <= 0:
        nthreads = cores
    done, dt, ids0, i = 0, 0.0, None, 0
    while dt < seconds:
        q = batches[i % len(batches)]
        t0 = time.perf_counter()
        ids, _, _, _ = pyoracle.search(blocks_np, block_bytes, a.dims, a.r, q, a.k, L, metric=metric, entry_row_id=0,
                                       route=1 if a.route == "ternary" else 0, nthreads=nthreads, impl=impl)
        dt += time.perf_counter() - t0
        done += q.shape[0]
        if ids0 is None:
            ids0 = ids
        i += 1
    return {"value": done / dt, "unit": "queries/s", "cores": nthreads, "kind": kind,
            "sample": "%d queries (%d passes over the %d-query batches) at L=%d, %d host threads, %.1f s" % (
                done, i, batches[0].shape[0], L, nthreads, dt)}, ids0, batches[0].shape[0]


def profiled_traffic(a, L):
    """DRAM bytes per traversal launch from the committed `ncu --set full` capture of this exact workload
    (profiles/traffic.json, dram__bytes_read.sum + dram__bytes_write.sum), or None."""
    try:
        t = json.load(open(os.path.join(ROOT, "profiles", "traffic.json")))
        key = "n%d_d%d_r%d_%s_q%d_k%d_L%d" % (a.n, a.dims, a.r, a.kind, a.nq, a.k, L)
        return t.get(key, {}).get("dram_bytes_per_launch")
    except Exception:
        return None


def main():
    a = parse()
    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if a.impl == "reference" and rank != 0:
        return 0
    import torch
    fx = importlib.import_module("duckdb-lm-diskann_b200.fixtures")
    have_cuda = torch.cuda.is_available()
    if a.impl == "ours" and not have_cuda:
        raise SystemExit("bench.py: no CUDA device - the product path has no CPU fallback")
    dev = "cuda:%d" % local if have_cuda else "cpu"
    if have_cuda:
        torch.cuda.set_device(local)
    dist_on = world > 1 and a.impl == "ours"
    if dist_on:
        import torch.distributed as dist
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        if os.environ.get("NCCL_DEBUG", "VERSION").upper() == "VERSION":
            os.environ["NCCL_DEBUG"] = "WARN"  # keep NCCL's version banner off stdout: rank 0 prints ONE JSON line
        dist.init_process_group("nccl", device_id=torch.device(dev))

    # ---- workload: one index (replicated), N_BATCHES query batches per rank -------------------------------------
    if rank == 0 or not dist_on:
        x, nbrs, counts = load_or_build_index(a, fx, dev)
    if dist_on:
        dist.barrier()
        if rank != 0:
            x, nbrs, counts = load_or_build_index(a, fx, dev)
    lay = fx.layout(a.dims, a.r)
    bb = lay["block_bytes"]
    batches = [fx.make_dataset(a.nq, a.dims, kind=a.kind, seed=43 + 1000 * rank + 17 * b,
                               normalize=(a.metric == "COSINE")) for b in range(N_BATCHES)]
    truth = load_or_build_truth(a, fx, x, batches[0], dev) if rank == 0 else None
    workload = "synthetic %dx%d float32 %s (%s, seed 42), R=%d, %d-query batch per GPU, k=%d" % (
        a.n, a.dims, a.metric, a.kind, a.r, a.nq, a.k)

    if a.impl == "reference":
        return run_reference(a, fx, x, nbrs, counts, batches, truth, workload, bb)

    lmd = importlib.import_module("duckdb-lm-diskann_b200")
    t0 = time.time()
    blocks_dev = fx.pack_blocks(x, nbrs, counts, a.r, device=dev)  # node blocks, reference layout, built in HBM
    store = lmd.Store(a.dims, a.r, metric=lmd.METRIC_BY_NAME[a.metric], capacity=a.n, device=local, l_search=100)
    route = lmd.ROUTE_TERNARY if a.route == "ternary" else lmd.ROUTE_EXACT
    store.upload(blocks_dev)
    store.set_entry(0)
    store.finalize()
    log("store ready in %.1f s: %d nodes, %.2f GB mirrored" % (time.time() - t0, store.node_count,
                                                                store.device_bytes / 1e9))
    keep_blocks_host = rank == 0 and not a.no_cpu
    blocks_np = blocks_dev.cpu().numpy() if keep_blocks_host else None
    del blocks_dev
    torch.cuda.empty_cache()

    # ---- L: smallest of the sweep reaching the recall target (rank 0 decides) ------------------------------------
    sweep = []
    L = a.L
    if rank == 0 and (a.L == 0 or a.sweep):
        for l in L_SWEEP:
            store.search(batches[1], a.k, l, route=route)  # warm-up launch of this L's kernel variant
            ids, _, _, st = store.search(batches[0], a.k, l, route=route)
            rec = fx.recall_at_k(ids, truth)
            gbps = st.algorithmic_bytes / (st.traverse_ms * 1e-3) / 1e9
            sweep.append({"L": l, "recall": round(rec, 4), "traverse_ms": round(st.traverse_ms, 4),
                          "kernel_qps": round(a.nq / (st.traverse_ms * 1e-3)), "algorithmic_GBps": round(gbps, 1),
                          "E_per_q": round(st.expansions / a.nq, 1), "V_per_q": round(st.distance_evals / a.nq, 1),
                          "filter_evictions_per_q": round(st.filter_evictions / a.nq, 2),
                          "merge_fallbacks": int(st.merge_fallbacks)})
            log("sweep", sweep[-1])
            if a.L == 0 and L == 0 and rec >= a.recall:
                L = l
                if not a.sweep:
                    break
        if L == 0:
            L = L_SWEEP[-1]
            log("recall target %.2f not reached by L <= %d; using L=%d" % (a.recall, L, L))
    if dist_on:
        t = torch.tensor([L], device=dev, dtype=torch.int64)
        dist.broadcast(t, 0)
        L = int(t.item())

    # ---- device-resident arm ("value") ---------------------------------------------------------------------------
    d_q = [torch.from_numpy(b).to(dev) for b in batches]
    d_ids = torch.empty((a.nq, a.k), dtype=torch.int64, device=dev)
    d_dist = torch.empty((a.nq, a.k), dtype=torch.float32, device=dev)
    d_cnt = torch.empty(a.nq, dtype=torch.int32, device=dev)
    stream = torch.cuda.Stream(device=dev)  # searches are enqueued on this stream; the events below bracket it
    torch.cuda.set_stream(stream)
    store.set_stream(stream.cuda_stream)

    def barrier():
        if dist_on:
            dist.barrier()
        torch.cuda.synchronize()

    for i in range(a.warmup):
        store.search_raw(d_q[i % N_BATCHES], a.nq, a.k, L, route=route, flags=lmd.SEARCH_ASYNC, out_ids=d_ids, out_dist=d_dist,
                         out_count=d_cnt)
    barrier()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    with ClockSampler(local) as clocks:
        e0.record(stream)
        for i in range(a.steps):
            store.search_raw(d_q[i % N_BATCHES], a.nq, a.k, L, route=route, flags=lmd.SEARCH_ASYNC, out_ids=d_ids, out_dist=d_dist,
                             out_count=d_cnt)
        e1.record(stream)
        barrier()
        dev_ms = e0.elapsed_time(e1)
        # per-kernel numbers for the roofline: same batches, synchronous calls with statistics
        kern_ms, rerank_ms, alg_bytes, launches, E, V = 0.0, 0.0, 0, 0, 0, 0
        for i in range(a.steps):
            st = store.search_raw(d_q[i % N_BATCHES], a.nq, a.k, L, route=route, out_ids=d_ids, out_dist=d_dist,
                                  out_count=d_cnt)
            kern_ms += st.traverse_ms
            rerank_ms += st.rerank_ms
            alg_bytes += st.algorithmic_bytes
            launches += st.launches
            E += st.expansions
            V += st.distance_evals
        barrier()
        # ---- host-buffer arm ("e2e"): pinned queries in, ids + distances out, through the C ABI --------------------
        h_q = [torch.from_numpy(b).pin_memory() for b in batches]
        h_ids = torch.empty((a.nq, a.k), dtype=torch.int64).pin_memory()
        h_dist = torch.empty((a.nq, a.k), dtype=torch.float32).pin_memory()
        # each call is synchronous: it returns when ids + distances of the step are in the caller's host buffers
        for i in range(max(a.warmup, 3)):
            store.search_raw(h_q[i % N_BATCHES], a.nq, a.k, L, route=route, out_ids=h_ids, out_dist=h_dist, stats=False)
        barrier()
        t0 = time.perf_counter()
        for i in range(a.steps):
            store.search_raw(h_q[i % N_BATCHES], a.nq, a.k, L, route=route, out_ids=h_ids, out_dist=h_dist, stats=False)
        torch.cuda.synchronize()
        e2e_s = time.perf_counter() - t0
        # what one `ORDER BY array_distance(...) LIMIT k` statement costs: a batch of ONE through the same call
        lat = []
        for i in range(220):
            t1 = time.perf_counter()
            store.search_raw(h_q[0][i:i + 1], 1, a.k, L, route=route, out_ids=h_ids[:1], out_dist=h_dist[:1], stats=False)
            lat.append(time.perf_counter() - t1)
        lat_us = sorted(lat[20:])[100] * 1e6
    if dist_on:
        t = torch.tensor([dev_ms, e2e_s], device=dev, dtype=torch.float64)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        dev_ms, e2e_s = float(t[0]), float(t[1])
    if rank != 0:
        if dist_on:
            dist.barrier()
            dist.destroy_process_group()
        return 0

    # ---- rank 0: recall at the chosen L, CPU baseline, report ------------------------------------------------------
    ids0, dist0, cnt0, st0 = store.search(batches[0], a.k, L, route=route)
    recall = fx.recall_at_k(ids0, truth)
    peaks = {}
    try:
        peaks = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))
    except Exception:
        pass
    peak = float(peaks.get("hbm_gbs", 6650.0))
    achieved = alg_bytes / (kern_ms * 1e-3) / 1e9
    cpu = None
    parity = None
    if not a.no_cpu:
        cpu, cpu_ids, nsamp = cpu_reference_qps(a, blocks_np, bb, batches, L, a.cpu_seconds)
        cpu["recall_at_10"] = round(fx.recall_at_k(cpu_ids, truth[:nsamp]), 4)
        parity = float((cpu_ids == ids0[:nsamp]).all(axis=1).mean())
    total_q = a.nq * world * a.steps
    out = {
        "metric": "QPS @ recall@10>=0.95, 1Mx128 L2 (LM_DISKANN beam search + exact top-k)",
        "value": total_q / (dev_ms * 1e-3), "unit": "queries/s", "n_gpus": world, "steps": a.steps,
        "warmup": a.warmup, "ms_per_step": dev_ms / a.steps, "higher_is_better": True, "scaling": "weak",
        "vs_baseline": None, "dtype": "f32", "data": "synthetic",
        "config": {"workload": workload, "L": L, "recall_at_10": round(recall, 4), "route": a.route,
                   "entry": "slot 0", "query_batches_rotated": N_BATCHES,
                   "single_query_latency_us": round(lat_us, 1),
                   "l2_note": ("node store %.1f GB >> 126 MB L2; every step gathers ~%.1f GB of it" % (
                       store.device_bytes / 1e9, alg_bytes / a.steps / 1e9)) if store.device_bytes > 4 * 126e6 else (
                       "node store %.0f MB FITS in the 126 MB L2 and is not flushed between steps: a parity / "
                       "small-index configuration, not a bandwidth measurement" % (store.device_bytes / 1e6)),
                   "parallelism": "queries sharded, graph replicated (%d x %d queries per step)" % (world, a.nq)},
        "e2e": {"value": total_q / e2e_s, "unit": "queries/s", "h2d_bytes_per_step": a.nq * a.dims * 4,
                "d2h_bytes_per_step": a.nq * a.k * 12, "ms_per_step": e2e_s * 1e3 / a.steps},
        "gpu_launches": launches,
        "roofline": {"bound": "hbm", "achieved": achieved, "peak": peak, "unit": "GB/s", "frac": achieved / peak,
                     "traffic": profiled_traffic(a, L), "kernel": "traverse_kernel", "kernel_ms": kern_ms / a.steps, "rerank_kernel_ms": rerank_ms / a.steps,
                     "algorithmic_bytes_per_launch": alg_bytes // a.steps,
                     "peak_source": "MEASURED_PEAKS.json hbm_gbs (measured copy)" if peaks else "fallback 6650",
                     "E_per_query": E / (a.nq * a.steps), "V_per_query": V / (a.nq * a.steps)},
        "cpu_baseline": cpu,
        "clocks": clocks.summary(),
    }
    if parity is not None:
        out["config"]["id_exact_queries_vs_reference_cpu"] = round(parity, 4)
    if sweep:
        out["config"]["L_sweep"] = sweep
    store.close()
    print(json.dumps(out), flush=True)
    if dist_on:
        dist.barrier()
        dist.destroy_process_group()
    return 0


def run_reference(a, fx, x, nbrs, counts, batches, truth, workload, bb):
    """The reference's own CPU implementation of the path (Searcher.cpp + H1-H4 compiled into oracle/_ref), all host
    threads, same graph / queries / k / L. Each step = a bounded sample of one 10k batch."""
    L = a.L or int(os.environ.get("LMD_REF_L", "0"))
    blocks_np = fx.pack_blocks(x, nbrs, counts, a.r, device="cpu").numpy()
    sys.path.insert(0, os.path.join(ROOT, "oracle"))
    import pyoracle
    impl = "fixed" if pyoracle.have_ref("fixed") and a.route == "exact" else "port"
    metric = {"L2": 1, "COSINE": 2, "IP": 3}[a.metric]

    def S(q, l):
        return pyoracle.search(blocks_np, bb, a.dims, a.r, q, a.k, l, metric=metric, entry_row_id=0,
                               route=1 if a.route == "ternary" else 0, impl=impl)
    cores = os.cpu_count() or 1
    if L == 0:  # same rule as our arm: smallest L of the sweep with recall >= target (on a 1000-query sample)
        for l in L_SWEEP:
            ids, _, _, _ = S(batches[0][:1000], l)
            if fx.recall_at_k(ids, truth[:1000]) >= a.recall:
                L = l
                break
        L = L or L_SWEEP[-1]
    # size the per-step sample so that warmup + steps stay within a few minutes
    t0 = time.perf_counter()
    S(batches[0][:512], L)
    per_q = (time.perf_counter() - t0) / 512
    budget = 120.0 / max(1, a.steps + a.warmup)
    nsamp = int(max(256, min(a.nq, budget / per_q)))
    for i in range(a.warmup):
        S(batches[i % N_BATCHES][:nsamp], L)
    t0 = time.perf_counter()
    for i in range(a.steps):
        ids, _, _, _ = S(batches[i % N_BATCHES][:nsamp], L)
    dt = time.perf_counter() - t0
    ids, _, _, _ = S(batches[0][:nsamp], L)
    qps = nsamp * a.steps / dt
    sample = "%d-query sample of each 10k batch, L=%d, %d host threads" % (nsamp, L, cores)
    out = {
        "impl": "reference",
        "metric": "QPS @ recall@10>=0.95, 1Mx128 L2 (LM_DISKANN beam search + exact top-k)",
        "value": qps, "unit": "queries/s", "n_gpus": a.gpus, "steps": a.steps, "warmup": a.warmup,
        "ms_per_step": dt * 1e3 / a.steps, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
        "dtype": "f32", "data": "synthetic",
        "config": {"workload": workload, "L": L, "recall_at_10": round(fx.recall_at_k(ids, truth[:nsamp]), 4),
                   "route": a.route, "entry": "slot 0", "sample": sample,
                   "code": "reference core/Searcher.cpp + oracle/patches/searcher_fixed.patch (oracle/_ref)"
                   if impl == "fixed" else "oracle/lmd_oracle.c port"},
        "cpu_baseline": {"value": qps, "unit": "queries/s", "cores": cores,
                         "kind": "reference" if impl == "fixed" else "port", "sample": sample},
        "e2e": {"value": qps, "unit": "queries/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
    }
    print(json.dumps(out), flush=True)
    return 0


if __name__ == "__main__":
    sys.exit(main())
