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

The headline keys keep the contract above; `config` additionally carries (all measured in this run, --extras none skips):
  parity        tie-aware comparison of the WHOLE batch with the reference CPU run (tests/parity.py rules) and the same for
                LMD_SEARCH_STRICT_FP (the reference's summation order: bit-equal distances)
  e2e_pageable  the e2e loop with pageable numpy buffers (what DuckDB's float* / std::vector hand the adapter)
  sustained     >= 2 s of back-to-back steps with the clock sampler's summary
  other_configs BASELINE.json configs on the GPUs of this run: c2_strong / c3_strong (ONE 10k batch split over the N
                ranks), c4 (10M x 96 replicated, 100k queries / N), c5 (dataset-sharded, NCCL all-gather + merge, N >= 2),
                c2_iid (SURVEY 8d hard case), c2_codec (int8 pre-filter / int8 route on the headline index)
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
    ap.add_argument("--route", default="exact", choices=["exact", "ternary", "int8"],
                    help="ternary = LM-DiskANN compressed-neighbour routing + exact rerank (COSINE / IP only); int8 = the "
                         "L2-capable int8 edge codec + exact rerank (PARITY UNPINNED: no reference code executes it)")
    ap.add_argument("--sweep", action="store_true", help="also print the whole L sweep (stderr)")
    ap.add_argument("--cpu-seconds", type=float, default=15.0)
    ap.add_argument("--no-cpu", action="store_true")
    ap.add_argument("--no-codec", action="store_true",
                    help="headline without the int8 code sidecar (the plain exact traversal instead of the pre-filtered one)")
    ap.add_argument("--depth", type=int, default=2, help="batches in flight (search contexts used in turn)")
    ap.add_argument("--single-process", action="store_true",
                    help="ONE process drives --gpus devices through lmd_group (query-sharded replicas) instead of one "
                         "torch.distributed rank per GPU; prints its own JSON line (not the driver's contract)")
    ap.add_argument("--extras", default="auto", choices=["auto", "none"],
                    help="auto = also run the other BASELINE configs / strong scaling / sustained records")
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
    the rotated query batches, over and over, until about `seconds` of CPU work are done. Returns the result of batch 0
    (the parity / recall sample)."""
    sys.path.insert(0, os.path.join(ROOT, "oracle"))
    import pyoracle
    if a.route == "int8":
        a = argparse.Namespace(**{**vars(a), "route": "exact"})  # the CPU arm of an int8-route run is the reference's exact search
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
        res = pyoracle.search(blocks_np, block_bytes, a.dims, a.r, q, a.k, L, metric=metric, entry_row_id=0,
                              route=1 if a.route == "ternary" else 0, nthreads=nthreads, impl=impl)
        dt += time.perf_counter() - t0
        done += q.shape[0]
        if ids0 is None:
            ids0 = res  # (ids, distances, counts, per-query [expansions, fetches]) of batch 0
        i += 1
    return {"value": done / dt, "unit": "queries/s", "cores": nthreads, "kind": kind,
            "sample": "%d queries (%d passes over the %d-query batches) at L=%d, %d host threads, %.1f s" % (
                done, i, batches[0].shape[0], L, nthreads, dt)}, ids0, batches[0].shape[0]


def profiled_traffic(a, L, prefiltered=False):
    """DRAM bytes per traversal launch from the committed `ncu --set full` capture of this exact workload
    (profiles/traffic.json, dram__bytes_read.sum + dram__bytes_write.sum), or None."""
    try:
        t = json.load(open(os.path.join(ROOT, "profiles", "traffic.json")))
        key = "n%d_d%d_r%d_%s_q%d_k%d_L%d%s" % (a.n, a.dims, a.r, a.kind, a.nq, a.k, L, "_pf" if prefiltered else "")
        return t.get(key, {}).get("dram_bytes_per_launch")
    except Exception:
        return None


# ------------------------------------------------------------------------------------------------------------------
# other_configs: BASELINE.json configs C2 (strong), C3, C4, C5 on the GPUs of this run. Every record is timed like the
# headline (barrier + CUDA events on the searching stream, max over ranks); local failures are reported in the record.
# ------------------------------------------------------------------------------------------------------------------
class Pipeline:
    """`depth` batches in flight: step i is enqueued on search context i % depth (lmd_ctx: own stream, counters, scratch)
    of the same store. The tail of a persistent-kernel launch - warps that find the query counter exhausted while others
    still work - is filled by the next batch's kernel (profiles/r2_10_pipe_*.json: 10k-query batches +3 % plain, +11 %
    pre-filtered, +23 % int8 route; 1 250-query batches +66 %). Timed on `main`, which brackets all streams by events."""

    def __init__(self, torch, store, dev, depth=2):
        self.torch, self.depth = torch, depth
        self.ctxs = [store.context() for _ in range(depth)]
        self.streams = [torch.cuda.Stream(device=dev) for _ in range(depth)]
        for c, st in zip(self.ctxs, self.streams):
            c.set_stream(st.cuda_stream)
        self.main = torch.cuda.Stream(device=dev)

    def begin(self):
        e0 = self.torch.cuda.Event(enable_timing=True)
        e0.record(self.main)
        for st in self.streams:
            st.wait_event(e0)
        return e0

    def end(self):
        for st in self.streams:
            ev = self.torch.cuda.Event()
            ev.record(st)
            self.main.wait_event(ev)
        e1 = self.torch.cuda.Event(enable_timing=True)
        e1.record(self.main)
        return e1

    def close(self):
        self.torch.cuda.synchronize()
        for c in self.ctxs:
            c.close()


class Job:
    def __init__(self, torch, dist, lmd, sh, wl, rank, world, local, dev, dist_on, steps, warmup):
        self.torch, self.dist, self.lmd, self.sh, self.wl = torch, dist, lmd, sh, wl
        self.rank, self.world, self.local, self.dev, self.dist_on = rank, world, local, dev, dist_on
        self.steps, self.warmup = max(6, min(steps, 20)), max(4, min(warmup, 6))
        self.stream = torch.cuda.current_stream(dev)
        self.depth = 2  # batches in flight (Pipeline)

    def barrier(self):
        if self.dist_on:
            self.dist.barrier()
        self.torch.cuda.synchronize()

    def reduce(self, vals, op="max"):
        t = self.torch.tensor(vals, device=self.dev, dtype=self.torch.float64)
        if self.dist_on:
            self.dist.all_reduce(t, op=self.dist.ReduceOp.MAX if op == "max" else self.dist.ReduceOp.SUM)
        return [float(v) for v in t]

    def agree(self, err):
        """Every rank reports whether its local (allocation-heavy) phase worked BEFORE the next collective, so that one
        rank's failure turns into an error record on all of them instead of a hang."""
        if self.reduce([1.0 if err else 0.0])[0] != 0.0:
            raise RuntimeError(err or "another rank failed in its local phase")

    def time_steps(self, step, pipe=None):
        """ms per step of `step(i)` (enqueues; on the pipeline's contexts when one is given, else on the current stream):
        warm-up, barrier, events, max over ranks."""
        torch = self.torch
        for i in range(self.warmup):
            step(i)
        self.barrier()
        if pipe is not None:
            e0 = pipe.begin()
            for i in range(self.steps):
                step(i)
            e1 = pipe.end()
        else:
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record(self.stream)
            for i in range(self.steps):
                step(i)
            e1.record(self.stream)
        self.barrier()
        return self.reduce([e0.elapsed_time(e1) / self.steps])[0]

    def query_sharded(self, store, q_local, truth_local, k, L, route, total_q, flags=0):
        """This rank searches q_local (its slice / its own batch); no collective on the path. -> record"""
        torch, lmd = self.torch, self.lmd
        nq = q_local.shape[0]
        pipe = Pipeline(torch, store, self.dev, self.depth)
        outs = [(torch.empty((max(nq, 1), k), dtype=torch.int64, device=self.dev),
                 torch.empty((max(nq, 1), k), dtype=torch.float32, device=self.dev)) for _ in range(self.depth)]
        ids, dd = outs[0]
        store.set_stream(self.stream.cuda_stream)

        def step(i):
            if nq:
                j = i % self.depth
                store.search_raw(q_local, nq, k, L, route=route, flags=lmd.SEARCH_ASYNC | flags, out_ids=outs[j][0],
                                 out_dist=outs[j][1], ctx=pipe.ctxs[j])
        ms = self.time_steps(step, pipe)
        pipe.close()
        st = store.search_raw(q_local, nq, k, L, route=route, flags=flags, out_ids=ids, out_dist=dd) if nq else None
        hits = self.wl.recall_hits(ids[:nq], truth_local) if nq and truth_local is not None else 0
        tot = truth_local.numel() if nq and truth_local is not None else 0
        kms, hits, tot, E, V, F, ab = (st.traverse_ms + st.rerank_ms if st else 0.0), hits, tot, (st.expansions if st else 0), \
            (st.distance_evals if st else 0), (st.exact_fetches if st else 0), (st.algorithmic_bytes if st else 0)
        kmax = self.reduce([kms])[0]
        hits, tot, E, V, F, ab, nqs = self.reduce([hits, tot, E, V, F, ab, nq], op="sum")
        return {"L": L, "queries_per_step": int(total_q), "queries_per_gpu": int(nq), "ms_per_step": round(ms, 4),
                "qps": round(total_q / (ms * 1e-3), 1), "recall_at_%d" % k: round(hits / max(tot, 1), 4),
                "kernel_ms_max_over_gpus": round(kmax, 4), "E_per_q": round(E / max(nqs, 1), 1),
                "V_per_q": round(V / max(nqs, 1), 1), "rows_fetched_per_q": round(F / max(nqs, 1), 1),
                "algorithmic_GBps_per_gpu": round(ab / max(self.world, 1) / max(kmax, 1e-9) / 1e6, 1),
                "batches_in_flight": self.depth}


def extra_c2_strong(J, store, batch0, truth0, k, L, route):
    """configs[2] as written: the headline index, ONE 10k-query batch split over the N ranks."""
    torch = J.torch
    lo, hi = J.sh.query_slice(batch0.shape[0], J.rank, J.world)
    q = torch.from_numpy(batch0[lo:hi]).to(J.dev)
    t = torch.from_numpy(truth0[lo:hi]).to(J.dev)
    r = J.query_sharded(store, q, t, k, L, route, batch0.shape[0])
    r["note"] = "one %d-query batch split %d ways (strong scaling); graph replicated" % (batch0.shape[0], J.world)
    return r


def extra_codec(J, store, batch_local, truth_local, k, L):
    """Row f4 on the headline index (its int8 code sidecar is already built): the plain exact traversal
    (LMD_SEARCH_NO_PREFILTER), exact routing behind the lower-bound pre-filter (= the headline; same ids and distance bits,
    fewer rows fetched) and int8 routing + exact rerank (parity unpinned). One full batch per GPU like the headline."""
    torch, lmd = J.torch, J.lmd
    q = torch.from_numpy(batch_local).to(J.dev)
    t = torch.from_numpy(truth_local).to(J.dev) if truth_local is not None else None
    tot = batch_local.shape[0] * J.world
    out = {"plain": J.query_sharded(store, q, t, k, L, lmd.ROUTE_EXACT, tot, flags=lmd.SEARCH_NO_PREFILTER),
           "prefilter": J.query_sharded(store, q, t, k, L, lmd.ROUTE_EXACT, tot)}
    rows = []
    for l in (56, 64, 80, 96):
        rows.append(J.query_sharded(store, q, t, k, l, lmd.ROUTE_INT8, tot))
    out["int8_route"] = rows
    out["note"] = ("exact + pre-filter returns the plain traversal's ids and distance bits (tests/test_gpu_round2.py); "
                   "int8 route: PARITY UNPINNED (restatement only); recall measured on rank 0's batch when N > 1")
    return out


def extra_built(J, tag, n, dims, metric, kind, nq_total, split, k, runs, r=32, l_insert=64, release=False, codec=False):
    """A GPU-generated, GPU-built index replicated on every rank. split=True: ONE batch of nq_total queries split over the
    ranks (strong); False: nq_total / N queries per rank, every rank its own. runs = [(label, route, L, flags)]."""
    torch, lmd, wl = J.torch, J.lmd, J.wl
    cos = metric == "COSINE"
    t0 = time.time()
    err = store = None
    try:
        x = wl.gen(n, dims, 42 if kind == "clustered" else 44, J.dev, cos, kind=kind)
        store, bst = wl.build_store(lmd, x, r, metric, J.local, l_insert=l_insert)
        t_build = time.time() - t0
        if split:
            qall = wl.gen(nq_total, dims, 43 if kind == "clustered" else 45, J.dev, cos, kind=kind)
            lo, hi = J.sh.query_slice(nq_total, J.rank, J.world)
            q = qall[lo:hi].contiguous()
        else:
            q = wl.gen(nq_total // J.world, dims, (43 if kind == "clustered" else 45) + 1000 * J.rank, J.dev, cos,
                       kind=kind)
        ns = min(q.shape[0], 5000)  # recall sample per rank
        _, truth = wl.topk_scores(x, q[:ns], k, metric)
        del x
        if release:
            full_gb = store.device_bytes / 1e9
            store.finalize()
            store.release_blocks()
        released_gb = store.device_bytes / 1e9
        if codec:  # int8 code sidecar: exact routing runs behind the lower-bound pre-filter (same results, fewer rows)
            store.enable_codec()
    except Exception as e:  # noqa: BLE001
        err = repr(e)
    torch.cuda.empty_cache()
    try:
        J.agree(err)
    except RuntimeError:
        if store is not None:
            store.close()
        raise
    rec = {"workload": "synthetic %dx%d float32 %s (%s), R=%d, graph built by lmd_build_graph (l_insert=%d)" % (
               n, dims, metric, kind, r, l_insert),
           "gen_build_s": round(t_build, 2), "build_kernels_ms": round(bst.total_ms, 1),
           "store_GB_per_gpu": round(full_gb if release else store.device_bytes / 1e9, 2),
           "mode": ("one %d-query batch split %d ways (strong)" if split else "%d queries per step over %d GPUs, replicated") % (
               nq_total, J.world)}
    if release:
        rec["store_GB_per_gpu_blocks_released"] = round(released_gb, 2)
    if codec:
        rec["store_GB_per_gpu_with_codes"] = round(store.device_bytes / 1e9, 2)
        rec["int8_code_sidecar"] = "on (runs without LMD_SEARCH_NO_PREFILTER use the pre-filtered exact kernel)"
    for label, route, L, flags in runs:
        # recall on the sample: time the whole slice, count hits on its first ns rows
        row = J.query_sharded(store, q, None, k, L, route, nq_total if split else (nq_total // J.world) * J.world, flags)
        ids = torch.empty((ns, k), dtype=torch.int64, device=J.dev)
        if ns:
            store.search_raw(q[:ns], ns, k, L, route=route, flags=flags, out_ids=ids, stats=False)
        hits, tot = J.reduce([wl.recall_hits(ids, truth) if ns else 0, truth.numel() if ns else 0], op="sum")
        row["recall_at_%d" % k] = round(hits / max(tot, 1), 4)
        row["recall_sample"] = "%d queries per GPU" % ns
        rec.setdefault("runs", {})[label] = row
    store.close()
    torch.cuda.empty_cache()
    return rec


def extra_c5(J, rows_per_gpu, dims, k, Ls, r=32, l_insert=64, codec=False):
    """configs[4]: dataset-sharded - rank s holds rows [lo_s, hi_s) with a shard-local graph; every rank searches the FULL
    batch, ONE NCCL all-gather of the [Q][k] (distance, row id) lists, k-way merge kernel."""
    torch, lmd, wl, sh = J.torch, J.lmd, J.wl, J.sh
    n = rows_per_gpu * J.world
    nq = 10000
    lo, hi = sh.shard_rows(n, J.rank, J.world)
    t0 = time.time()
    err = store = None
    try:
        x = wl.gen(hi - lo, dims, 42 + J.rank, J.dev)
        row_ids = torch.arange(lo, hi, dtype=torch.int64, device=J.dev)
        store, bst = wl.build_store(lmd, x, r, "L2", J.local, row_ids=row_ids, l_insert=l_insert)
        t_build = time.time() - t0
        q = wl.gen(nq, dims, 43, J.dev)
        S, I = wl.topk_scores(x, q, k, "L2", base_id=lo)
        del x, row_ids
        full_gb = store.device_bytes / 1e9
        store.finalize()
        store.release_blocks()
        released_gb = store.device_bytes / 1e9
        if codec:
            store.enable_codec()
    except Exception as e:  # noqa: BLE001
        err = repr(e)
    try:
        J.agree(err)
    except RuntimeError:
        if store is not None:
            store.close()
        raise
    allS = torch.empty((J.world,) + S.shape, device=J.dev)
    allI = torch.empty((J.world,) + I.shape, dtype=torch.int64, device=J.dev)
    J.dist.all_gather_into_tensor(allS, S)
    J.dist.all_gather_into_tensor(allI, I)
    _, sel = torch.topk(allS.permute(1, 0, 2).reshape(nq, -1), k, dim=1, largest=False)
    truth = torch.gather(allI.permute(1, 0, 2).reshape(nq, -1), 1, sel)
    del S, I, allS, allI
    torch.cuda.empty_cache()
    rec = {"workload": "synthetic %dx%d float32 L2 (clustered), %d shards of %d rows, shard-local graphs (lmd_build_graph, "
                       "l_insert=%d), global row ids" % (n, dims, J.world, hi - lo, l_insert),
           "gen_build_s": round(t_build, 2), "store_GB_per_gpu": round(full_gb, 2),
           "store_GB_per_gpu_blocks_released": round(released_gb, 2),
           "store_GB_per_gpu_searched": round(store.device_bytes / 1e9, 2),
           "nccl_ranks": J.dist.get_world_size(), "allgather_bytes_per_rank": nq * k * 12,
           "int8_code_sidecar": "on (pre-filtered exact kernel)" if codec else "off", "runs": []}
    store.set_stream(J.stream.cuda_stream)
    searcher = sh.DatasetShardedSearcher(store, device=torch.device(J.dev))
    ids = torch.empty((nq, k), dtype=torch.int64, device=J.dev)
    dd = torch.empty((nq, k), dtype=torch.float32, device=J.dev)
    for L in Ls:
        out = [None]

        def full(i):
            out[0] = searcher.search(q, k, L)

        def local_only(i):
            store.search_raw(q, nq, k, L, flags=lmd.SEARCH_ASYNC, out_ids=ids, out_dist=dd)
        ms_full = J.time_steps(full)
        ms_local = J.time_steps(local_only)
        st = store.search_raw(q, nq, k, L, out_ids=ids, out_dist=dd)
        kmax = J.reduce([st.traverse_ms + st.rerank_ms])[0]
        rec["runs"].append({"L": L, "ms_per_step": round(ms_full, 4), "qps": round(nq / (ms_full * 1e-3), 1),
                            "recall_at_%d" % k: round(wl.recall(out[0][0], truth), 4),
                            "local_search_ms": round(ms_local, 4), "local_kernel_ms_max_over_gpus": round(kmax, 4),
                            "allgather_plus_merge_ms": round(ms_full - ms_local, 4),
                            "E_per_q": round(st.expansions / nq, 1), "V_per_q": round(st.distance_evals / nq, 1)})
    store.close()
    torch.cuda.empty_cache()
    return rec


def guarded(J, name, fn, *a, **kw):
    t0 = time.time()
    try:
        r = fn(J, *a, **kw)
    except Exception as e:  # noqa: BLE001 - an extra must never take the headline line down
        log("extra %s failed: %r" % (name, e))
        r = {"error": repr(e)[:300]}
    if isinstance(r, dict):
        r["wall_s"] = round(time.time() - t0, 1)
    log("extra", name, json.dumps(r)[:600])
    return r


def run_single_process(a):
    """The headline workload through lmd_group: one host thread, a.gpus devices, one replica of the index on each, a batch
    of a.gpus x a.nq queries per step in pinned host memory, results written to pinned host memory (the e2e definition).
    Shows that the multi-GPU modes are reachable from a single-process host (DuckDB) and cost what the ranks cost."""
    import torch
    fx = importlib.import_module("duckdb-lm-diskann_b200.fixtures")
    lmd = importlib.import_module("duckdb-lm-diskann_b200")
    n_dev = min(a.gpus, torch.cuda.device_count())
    x, nbrs, counts = load_or_build_index(a, fx, "cuda:0")
    L = a.L or 56
    stores = []
    for d in range(n_dev):
        torch.cuda.set_device(d)
        blocks = fx.pack_blocks(x, nbrs, counts, a.r, device="cuda:%d" % d)
        st = lmd.Store(a.dims, a.r, metric=lmd.METRIC_BY_NAME[a.metric], capacity=a.n, device=d, l_search=100)
        st.upload(blocks)
        st.set_entry(0)
        st.finalize()
        if a.metric == "L2" and not a.no_codec:
            st.enable_codec()
        del blocks
        torch.cuda.empty_cache()
        stores.append(st)
    nq = a.nq * n_dev
    qs = [torch.from_numpy(fx.make_dataset(nq, a.dims, kind=a.kind, seed=43 + 17 * b)).pin_memory() for b in range(2)]
    ids = torch.empty((nq, a.k), dtype=torch.int64).pin_memory()
    dd = torch.empty((nq, a.k), dtype=torch.float32).pin_memory()
    truth = fx.brute_force_topk(x, qs[0][:2000].numpy(), a.k, a.metric, device="cuda:0")
    out = {}
    for mode_name, mode in (("query_sharded", lmd.GROUP_QUERY_SHARDED),):
        with lmd.Group(stores, mode) as g:
            for i in range(max(a.warmup, 3)):
                g.search_raw(qs[i % 2], nq, a.k, L, out_ids=ids, out_dist=dd, stats=False)
            t0 = time.perf_counter()
            for i in range(a.steps):
                g.search_raw(qs[i % 2], nq, a.k, L, out_ids=ids, out_dist=dd, stats=False)
            dt = time.perf_counter() - t0
            st = g.search_raw(qs[0], nq, a.k, L, out_ids=ids, out_dist=dd)
            rec = fx.recall_at_k(ids[:2000].numpy(), truth)
            out[mode_name] = {"qps": nq * a.steps / dt, "ms_per_step": dt * 1e3 / a.steps, "recall_at_10": round(rec, 4),
                              "search_ms_per_member": [round(st.search_ms[i], 4) for i in range(n_dev)],
                              "wall_ms_one_call": round(st.wall_ms, 4)}
    print(json.dumps({"mode": "single process, lmd_group", "n_gpus": n_dev, "steps": a.steps, "L": L,
                      "workload": "synthetic %dx%d %s, %d x %d queries per step, pinned host buffers" % (
                          a.n, a.dims, a.metric, n_dev, a.nq), "results": out}), flush=True)
    for st in stores:
        st.close()
    return 0


def main():
    a = parse()
    if a.single_process:
        return run_single_process(a)
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
    dist = None
    if dist_on:
        import datetime
        import torch.distributed as dist
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        if os.environ.get("NCCL_DEBUG", "VERSION").upper() == "VERSION":
            os.environ["NCCL_DEBUG"] = "WARN"  # keep NCCL's version banner off stdout: rank 0 prints ONE JSON line
        dist.init_process_group("nccl", device_id=torch.device(dev), timeout=datetime.timedelta(seconds=600))

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
    # rank 0's first batch doubles as THE batch of the strong-scaling record (identical on every rank)
    batch0 = batches[0] if rank == 0 else fx.make_dataset(a.nq, a.dims, kind=a.kind, seed=43,
                                                          normalize=(a.metric == "COSINE"))
    truth = load_or_build_truth(a, fx, x, batch0, dev) if rank == 0 else None
    workload = "synthetic %dx%d float32 %s (%s, seed 42), R=%d, %d-query batch per GPU, k=%d" % (
        a.n, a.dims, a.metric, a.kind, a.r, a.nq, a.k)

    if a.impl == "reference":
        return run_reference(a, fx, x, nbrs, counts, batches, truth, workload, bb)

    lmd = importlib.import_module("duckdb-lm-diskann_b200")
    sh = importlib.import_module("duckdb-lm-diskann_b200.sharded")
    wl = importlib.import_module("duckdb-lm-diskann_b200.workloads")
    t0 = time.time()
    blocks_dev = fx.pack_blocks(x, nbrs, counts, a.r, device=dev)  # node blocks, reference layout, built in HBM
    store = lmd.Store(a.dims, a.r, metric=lmd.METRIC_BY_NAME[a.metric], capacity=a.n, device=local, l_search=100)
    route = {"ternary": lmd.ROUTE_TERNARY, "int8": lmd.ROUTE_INT8}.get(a.route, lmd.ROUTE_EXACT)
    store.upload(blocks_dev)
    store.set_entry(0)
    store.finalize()
    # L2 + exact routing: the int8 code sidecar lets the traversal skip rows whose lower bound already exceeds the list's
    # worst distance. Ids, distance bits and expansion counts are those of the plain traversal (tests/test_gpu_round2.py
    # ::test_prefilter_changes_nothing_but_the_bytes; the parity report below is computed from THIS store's results).
    use_codec = a.metric == "L2" and a.route == "exact" and not a.no_codec
    if use_codec:
        store.enable_codec()
    log("store ready in %.1f s: %d nodes, %.2f GB mirrored%s" % (time.time() - t0, store.node_count,
                                                                  store.device_bytes / 1e9,
                                                                  " (incl. int8 code sidecar)" if use_codec else ""))
    keep_blocks_host = rank == 0 and not a.no_cpu
    blocks_np = blocks_dev.cpu().numpy() if keep_blocks_host else None
    del blocks_dev, x, nbrs, counts
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
        tt = torch.from_numpy(truth).to(dev) if rank == 0 else torch.empty((a.nq, a.k), dtype=torch.int64, device=dev)
        dist.broadcast(tt, 0)
        truth = tt.cpu().numpy()

    # ---- device-resident arm ("value"): a.depth batches in flight on as many search contexts ------------------------
    d_q = [torch.from_numpy(b).to(dev) for b in batches]
    depth = max(1, a.depth)
    d_out = [(torch.empty((a.nq, a.k), dtype=torch.int64, device=dev), torch.empty((a.nq, a.k), dtype=torch.float32, device=dev),
              torch.empty(a.nq, dtype=torch.int32, device=dev)) for _ in range(depth)]
    d_ids, d_dist, d_cnt = d_out[0]
    stream = torch.cuda.Stream(device=dev)  # the store's own (statistics) calls run here
    torch.cuda.set_stream(stream)
    store.set_stream(stream.cuda_stream)
    pipe = Pipeline(torch, store, dev, depth)

    def barrier():
        if dist_on:
            dist.barrier()
        torch.cuda.synchronize()

    def dev_step(i):
        j = i % depth
        store.search_raw(d_q[i % N_BATCHES], a.nq, a.k, L, route=route, flags=lmd.SEARCH_ASYNC, out_ids=d_out[j][0],
                         out_dist=d_out[j][1], out_count=d_out[j][2], ctx=pipe.ctxs[j])

    for i in range(a.warmup):
        dev_step(i)
    barrier()
    with ClockSampler(local) as clocks:
        e0 = pipe.begin()
        for i in range(a.steps):
            dev_step(i)
        e1 = pipe.end()
        barrier()
        dev_ms = e0.elapsed_time(e1)
        # per-kernel numbers for the roofline: same batches, one at a time, synchronous calls with statistics
        kern_ms, rerank_ms, alg_bytes, launches, E, V, F = 0.0, 0.0, 0, 0, 0, 0, 0
        for i in range(a.steps):
            st = store.search_raw(d_q[i % N_BATCHES], a.nq, a.k, L, route=route, out_ids=d_ids, out_dist=d_dist,
                                  out_count=d_cnt)
            kern_ms += st.traverse_ms
            rerank_ms += st.rerank_ms
            alg_bytes += st.algorithmic_bytes
            launches += st.launches
            E += st.expansions
            V += st.distance_evals
            F += st.exact_fetches
        # the plain exact traversal on the same batches (the kernel the pre-filter is measured against)
        plain_ms, plain_bytes = 0.0, 0
        if use_codec:
            for i in range(3):  # warm-up: the first launch of a kernel variant pays for its lazy load
                store.search_raw(d_q[i % N_BATCHES], a.nq, a.k, L, route=route, flags=lmd.SEARCH_NO_PREFILTER,
                                 out_ids=d_ids, out_dist=d_dist, out_count=d_cnt)
            for i in range(min(a.steps, 10)):
                st = store.search_raw(d_q[i % N_BATCHES], a.nq, a.k, L, route=route, flags=lmd.SEARCH_NO_PREFILTER,
                                      out_ids=d_ids, out_dist=d_dist, out_count=d_cnt)
                plain_ms += st.traverse_ms
                plain_bytes += st.algorithmic_bytes
        barrier()
        # ---- host-buffer arm ("e2e"): pinned queries in, ids + distances out, through the C ABI --------------------
        # The kernels read the pinned queries and write the pinned results in place (zero copy). Two calls are in flight
        # (one per context); before a context is used again its stream is synchronised and the step's first result row
        # is read by the host, so every step's ids and distances have landed in host memory inside the timed region.
        h_q = [torch.from_numpy(b).pin_memory() for b in batches]
        h_out = [(torch.empty((a.nq, a.k), dtype=torch.int64).pin_memory(),
                  torch.empty((a.nq, a.k), dtype=torch.float32).pin_memory()) for _ in range(depth)]
        seen = 0

        def e2e_step(i):
            j = i % depth
            pipe.streams[j].synchronize()
            s_ = int(h_out[j][0][0, 0])  # the host reads the previous result of this buffer pair
            store.search_raw(h_q[i % N_BATCHES], a.nq, a.k, L, route=route, flags=lmd.SEARCH_ASYNC, out_ids=h_out[j][0],
                             out_dist=h_out[j][1], ctx=pipe.ctxs[j])
            return s_
        for i in range(max(a.warmup, 4)):
            e2e_step(i)
        barrier()
        t0 = time.perf_counter()
        for i in range(a.steps):
            seen += e2e_step(i)
        for j in range(depth):
            pipe.streams[j].synchronize()
            seen += int(h_out[j][0][0, 0])
        e2e_s = time.perf_counter() - t0
    # ... and with PAGEABLE buffers (what DuckDB's float* / std::vector<row_t> give GpuSearcher::Search): staged copies
    p_ids = np.empty((a.nq, a.k), dtype=np.int64)
    p_dist = np.empty((a.nq, a.k), dtype=np.float32)
    for i in range(3):
        store.search_raw(batches[i % N_BATCHES], a.nq, a.k, L, route=route, out_ids=p_ids, out_dist=p_dist, stats=False)
    barrier()
    t0 = time.perf_counter()
    for i in range(a.steps):
        store.search_raw(batches[i % N_BATCHES], a.nq, a.k, L, route=route, out_ids=p_ids, out_dist=p_dist, stats=False)
    torch.cuda.synchronize()
    e2e_page_s = time.perf_counter() - t0
    # what one `ORDER BY array_distance(...) LIMIT k` statement costs: a batch of ONE (pageable) through the same call
    lat = []
    for i in range(220):
        t1 = time.perf_counter()
        store.search_raw(batches[0][i:i + 1], 1, a.k, L, route=route, out_ids=p_ids[:1], out_dist=p_dist[:1], stats=False)
        lat.append(time.perf_counter() - t1)
    lat_us = sorted(lat[20:])[100] * 1e6
    st1 = store.search_raw(batches[0][:1], 1, a.k, L, route=route, out_ids=p_ids[:1], out_dist=p_dist[:1])
    log("single query: median %.1f us of %d calls (min %.1f, max %.1f), kernel %.1f us, launches %d" % (
        lat_us, len(lat) - 20, min(lat[20:]) * 1e6, max(lat[20:]) * 1e6, st1.traverse_ms * 1e3, st1.launches))
    # ---- sustained: >= 2 s of back-to-back device-resident steps, clocks sampled throughout --------------------------
    sustained = None
    if a.extras != "none":
        n_sus = int(max(a.steps, min(20000, 2200.0 / max(dev_ms / a.steps, 1e-3))))
        barrier()
        with ClockSampler(local) as sus_clocks:
            s0 = pipe.begin()
            for i in range(n_sus):
                dev_step(i)
            s1 = pipe.end()
            barrier()
        sus_ms = s0.elapsed_time(s1)
        if dist_on:
            t = torch.tensor([sus_ms], device=dev, dtype=torch.float64)
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
            sus_ms = float(t[0])
        sustained = {"steps": n_sus, "seconds": round(sus_ms * 1e-3, 3), "ms_per_step": round(sus_ms / n_sus, 4),
                     "value": round(a.nq * world * n_sus / (sus_ms * 1e-3), 1), "unit": "queries/s",
                     "clocks": sus_clocks.summary()}
    if dist_on:
        t = torch.tensor([dev_ms, e2e_s, e2e_page_s], device=dev, dtype=torch.float64)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        dev_ms, e2e_s, e2e_page_s = float(t[0]), float(t[1]), float(t[2])

    # ---- rank 0, GPU side: results at the chosen L (default and strict fp) for recall / parity ----------------------
    ids0 = dist0 = cnt0 = strict_res = None
    if rank == 0:
        ids0, dist0, cnt0, st0 = store.search(batches[0], a.k, L, route=route)
        if not a.no_cpu:
            strict_res = store.search(batches[0], a.k, L, route=route, flags=lmd.SEARCH_STRICT_FP, qstats=True)
    store_bytes = store.device_bytes
    pipe.close()

    # ---- other configs (all ranks) -----------------------------------------------------------------------------------
    other = {}
    if a.extras != "none":
        J = Job(torch, dist, lmd, sh, wl, rank, world, local, dev, dist_on, a.steps, a.warmup)
        J.stream = stream
        J.depth = depth
        other["c2_strong"] = guarded(J, "c2_strong", extra_c2_strong, store, batch0, truth, a.k, L, route)
        if use_codec:
            other["c2_codec"] = guarded(J, "c2_codec", extra_codec, store, batches[0], truth if rank == 0 else None, a.k, L)
    store.close()
    del d_q
    torch.cuda.empty_cache()
    if a.extras != "none":
        other["c3_strong"] = guarded(J, "c3_strong", extra_built, "c3", 1_000_000, 768, "COSINE", "clustered", 10_000, True,
                                     a.k, [("exact_L32", lmd.ROUTE_EXACT, 32, 0),
                                           ("exact_L32_no_prefilter", lmd.ROUTE_EXACT, 32, lmd.SEARCH_NO_PREFILTER),
                                           ("ternary_L80", lmd.ROUTE_TERNARY, 80, 0)], codec=use_codec)
        other["c2_iid"] = guarded(J, "c2_iid", extra_built, "iid", 1_000_000, 128, "L2", "iid", 10_000 * world, False, a.k,
                                  [("exact_L64", lmd.ROUTE_EXACT, 64, 0), ("exact_L256", lmd.ROUTE_EXACT, 256, 0)])
        other["c4"] = guarded(J, "c4", extra_built, "c4", 10_000_000, 96, "L2", "clustered", 100_000, False, a.k,
                              [("exact_L96", lmd.ROUTE_EXACT, 96, 0),
                               ("exact_L96_no_prefilter", lmd.ROUTE_EXACT, 96, lmd.SEARCH_NO_PREFILTER)], release=True,
                              codec=use_codec)
        if dist_on:
            # 12.5 M rows per GPU: the 100 M x 96 configuration at N = 8, the same shard size below that
            other["c5"] = guarded(J, "c5", extra_c5, 12_500_000, 96, a.k, [48, 64], codec=use_codec)
        else:
            other["c5"] = {"skipped": "dataset sharding needs N >= 2 (its all-gather is the point)"}
    if rank != 0:
        if dist_on:
            dist.barrier()
            dist.destroy_process_group()
        return 0

    # ---- rank 0: recall at the chosen L, CPU baseline, parity, report ---------------------------------------------
    recall = fx.recall_at_k(ids0, truth)
    peaks = {}
    try:
        peaks = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))
    except Exception:
        pass
    peak = float(peaks.get("hbm_gbs", 6650.0))
    achieved = alg_bytes / ((kern_ms + rerank_ms) * 1e-3) / 1e9
    cpu = None
    parity = None
    if not a.no_cpu:
        sys.path.insert(0, os.path.join(ROOT, "tests"))
        from parity import parity_report
        cpu, cpu_res, nsamp = cpu_reference_qps(a, blocks_np, bb, batches, L, a.cpu_seconds)
        cpu["recall_at_10"] = round(fx.recall_at_k(cpu_res[0], truth[:nsamp]), 4)
        one, _, _ = cpu_reference_qps(a, blocks_np, bb, [b[:2000] for b in batches], L, min(4.0, a.cpu_seconds), nthreads=1)
        cpu["single_thread"] = {"value": one["value"], "unit": "queries/s", "cores": 1, "sample": one["sample"]}
        parity = {"against": "reference CPU run (%s) of the same %d queries" % (cpu["kind"], nsamp),
                  "default": parity_report(ids0[:nsamp], dist0[:nsamp], cpu_res[0], cpu_res[1], cpu_res[2])}
        if strict_res is not None:
            sp = parity_report(strict_res[0][:nsamp], strict_res[1][:nsamp], cpu_res[0], cpu_res[1], cpu_res[2])
            sp["dist_bits_identical"] = bool(np.array_equal(strict_res[1][:nsamp].view(np.uint32), cpu_res[1].view(np.uint32)))
            sp["expansion_counts_identical"] = bool(np.array_equal(strict_res[4][:nsamp, 0].astype(np.int64),
                                                                   cpu_res[3][:, 0].astype(np.int64)))
            sp["flag"] = "LMD_SEARCH_STRICT_FP (reference summation order; parity tool, not the timed path)"
            parity["strict_fp"] = sp
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
                       store_bytes / 1e9, alg_bytes / a.steps / 1e9)) if store_bytes > 4 * 126e6 else (
                       "node store %.0f MB FITS in the 126 MB L2 and is not flushed between steps: a parity / "
                       "small-index configuration, not a bandwidth measurement" % (store_bytes / 1e6)),
                   "parallelism": "queries sharded, graph replicated (%d x %d queries per step)" % (world, a.nq),
                   "batches_in_flight": depth,
                   "parity_note": ("int8 routing: PARITY UNPINNED - the parity block compares with the reference's EXACT search"
                                   if a.route == "int8" else None),
                   "traversal": ("exact routing behind the int8 lower-bound pre-filter (results bit-identical to the plain "
                                 "traversal; %.1f of %.1f evaluated rows fetched per query)" % (
                                     F / (a.nq * a.steps), V / (a.nq * a.steps))) if use_codec else "plain exact traversal"},
        "e2e": {"value": total_q / e2e_s, "unit": "queries/s", "h2d_bytes_per_step": a.nq * a.dims * 4,
                "d2h_bytes_per_step": a.nq * a.k * 12, "ms_per_step": e2e_s * 1e3 / a.steps,
                "buffers": "pinned host memory, read / written in place by the kernels; %d calls in flight, every step's "
                           "result read by the host before its buffers are reused" % depth},
        "e2e_pageable": {"value": total_q / e2e_page_s, "unit": "queries/s", "ms_per_step": e2e_page_s * 1e3 / a.steps,
                         "buffers": "pageable host memory (numpy), staged through cudaMemcpyAsync inside the call"},
        "gpu_launches": launches,
        "roofline": {"bound": "hbm", "achieved": achieved, "peak": peak, "unit": "GB/s", "frac": achieved / peak,
                     "traffic": profiled_traffic(a, L, use_codec),
                     "traffic_source": "profiles/traffic.json (ncu --set full capture of this workload by the builder, "
                                       "not measured in this run)",
                     "kernel": "traverse_kernel" + (" (exact routing, int8 pre-filter)" if use_codec else
                                                    " (int8 routing; the rerank kernel's bytes and time are included)"
                                                    if a.route == "int8" else ""),
                     "kernel_ms": kern_ms / a.steps, "rerank_kernel_ms": rerank_ms / a.steps,
                     "kernel_ms_note": "CUDA-event duration of one launch at a time (lmd_stats.traverse_ms); with %d batches "
                                       "in flight a step takes ms_per_step" % depth,
                     "achieved_in_flight": alg_bytes / a.steps / (dev_ms / a.steps * 1e-3) / 1e9 / world if world else None,
                     "algorithmic_bytes_per_launch": alg_bytes // a.steps,
                     "bytes_formula": ("E * code_block_bytes + F * 4D + Q * (4D + 12k)" if use_codec else
                                       "E * code_block_bytes + C * 4D + Q * (8D + 12k + 16L)" if a.route == "int8" else
                                       "E * 4R + V * 4D + Q * (4D + 12k)"),
                     "peak_source": "MEASURED_PEAKS.json hbm_gbs (measured copy)" if peaks else "fallback 6650",
                     "E_per_query": E / (a.nq * a.steps), "V_per_query": V / (a.nq * a.steps),
                     "rows_fetched_per_query": F / (a.nq * a.steps)},
        "cpu_baseline": cpu,
        "clocks": clocks.summary(),
    }
    if use_codec and plain_ms > 0:
        n_pl = min(a.steps, 10)
        pa = plain_bytes / (plain_ms * 1e-3) / 1e9
        out["roofline_plain_kernel"] = {"kernel": "traverse_kernel (plain exact traversal, LMD_SEARCH_NO_PREFILTER)",
                                        "kernel_ms": plain_ms / n_pl, "achieved": pa, "peak": peak, "frac": pa / peak,
                                        "algorithmic_bytes_per_launch": plain_bytes // n_pl,
                                        "bytes_formula": "E * 4R + V * 4D + Q * (4D + 12k)"}
    if sustained:
        out["sustained"] = sustained
    if parity is not None:
        out["config"]["parity"] = parity
        out["config"]["id_exact_queries_vs_reference_cpu"] = round(
            parity["default"]["id_identical_queries"] / max(parity["default"]["queries"], 1), 4)
        out["config"]["id_mismatch_outside_ties"] = parity["default"]["id_mismatch_outside_ties"]
        out["config"]["dist_max_rel_err"] = parity["default"]["dist_max_rel_err"]
    if sweep:
        out["config"]["L_sweep"] = sweep
    if other:
        out["config"]["other_configs"] = other
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
