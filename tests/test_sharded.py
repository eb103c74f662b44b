"""Multi-GPU host logic (duckdb-lm-diskann_b200/sharded.py) with world_size 2.

not-gpu: `gloo`, CPU tensors; the local search is the CPU oracle and the merge a numpy restatement, so these tests
         cover exactly the plumbing: query slicing, the gather layout, the merge order, -1 padding.
gpu:     the same two ranks with the REAL local search and merge kernel (both ranks share cuda:0 when the box has one
         GPU, so the collective stays on gloo there; with >= 2 GPUs it is NCCL, one rank per GPU).
"""
import importlib
import os
import socket
import sys

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


def _np_merge(all_ids, all_d, k):
    """[G][Q][k] -> [Q][k] smallest by (distance, id); ids < 0 are padding."""
    G, Q, _ = all_ids.shape
    out_i = np.full((Q, k), -1, dtype=np.int64)
    out_d = np.zeros((Q, k), dtype=np.float32)
    for q in range(Q):
        c = [(float(all_d[g, q, j]), int(all_ids[g, q, j])) for g in range(G) for j in range(k) if all_ids[g, q, j] >= 0]
        c.sort()
        for j, (d, i) in enumerate(c[:k]):
            out_i[q, j], out_d[q, j] = i, d
    return out_i, out_d


def _make_case(n=1600, dims=32, r=12, nq=37, world=2):
    """Base table, a replicated index over all of it, and per-shard indexes with GLOBAL row ids."""
    fx = importlib.import_module("duckdb-lm-diskann_b200.fixtures")
    sh = importlib.import_module("duckdb-lm-diskann_b200.sharded")
    x = fx.make_dataset(n, dims, kind="iid", seed=21)
    q = fx.make_dataset(nq, dims, kind="iid", seed=22)
    bb = fx.layout(dims, r)["block_bytes"]
    nb, cn = fx.build_graph(x, r, device="cpu")
    full = fx.pack_blocks(x, nb, cn, r).numpy()
    shards = []
    for s in range(world):
        lo, hi = sh.shard_rows(n, s, world)
        nbs, cns = fx.build_graph(x[lo:hi], r, device="cpu")
        rid = np.arange(lo, hi, dtype=np.int64)
        shards.append((fx.pack_blocks(x[lo:hi], nbs, cns, r, row_ids=rid).numpy(), rid))
    return dict(x=x, q=q, bb=bb, full=full, shards=shards, dims=dims, r=r, n=n)


class _OracleStore:
    """Stand-in with the Store.search signature, backed by the CPU oracle (tests only)."""
    device = 0

    def __init__(self, blocks, bb, dims, r, row_ids=None, entry=0):
        self.a = (blocks, bb, dims, r)
        self.row_ids, self.entry = row_ids, entry

    def search(self, queries, k, L=0, **kw):
        sys.path.insert(0, os.path.join(ROOT, "oracle"))
        import pyoracle
        blocks, bb, dims, r = self.a
        ids, d, c, st = pyoracle.search(blocks, bb, dims, r, queries, k, L, row_ids=self.row_ids,
                                        entry_row_id=self.entry, nthreads=1)
        return ids, d, c, st


def _worker(rank, world, port, mode, use_cuda, case, k, L, ret):
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    sys.path.insert(0, ROOT)
    sh = importlib.import_module("duckdb-lm-diskann_b200.sharded")
    ngpu = torch.cuda.device_count() if use_cuda else 0
    nccl = use_cuda and ngpu >= world
    dist.init_process_group("nccl" if nccl else "gloo", rank=rank, world_size=world)
    try:
        gather_dev = torch.device("cuda", rank) if nccl else torch.device("cpu")
        cuda_dev = (rank if nccl else 0) if use_cuda else None
        dims, r, bb = case["dims"], case["r"], case["bb"]
        if mode == "query":
            if use_cuda:
                lmd = importlib.import_module("duckdb-lm-diskann_b200")
                store = lmd.Store(dims, r, capacity=case["n"], device=cuda_dev)
                store.upload(case["full"])
                store.set_entry(0)
            else:
                store = _OracleStore(case["full"], bb, dims, r)
            s = sh.QueryShardedSearcher(store, device=gather_dev)
            ids, d, c, _ = s.search(case["q"], k, L, gather=True)
            lo, hi = sh.query_slice(case["q"].shape[0], rank, world)
            ret[rank] = (ids, d, c, lo, hi)
        else:
            blocks, rid = case["shards"][rank]
            if use_cuda:
                lmd = importlib.import_module("duckdb-lm-diskann_b200")
                store = lmd.Store(dims, r, capacity=blocks.shape[0], device=cuda_dev)
                store.upload(blocks, row_ids=rid)
                store.set_entry(int(rid[0]))
                if nccl:
                    s = sh.DatasetShardedSearcher(store, device=gather_dev)
                else:  # one GPU, two ranks: real kernels on cuda:0, gloo moves CPU copies
                    def local(st, qd, nq, kk, LL, route):
                        i, dd = sh._cuda_local_search(st, qd.to("cuda:0"), nq, kk, LL, route)
                        return i.cpu(), dd.cpu()

                    def merge(ai, ad, g, nq, kk, di):
                        i, dd = sh._cuda_merge(ai.to("cuda:0"), ad.to("cuda:0"), g, nq, kk, 0)
                        return i.cpu(), dd.cpu()
                    s = sh.DatasetShardedSearcher(store, device=gather_dev, local_search=local, merge=merge)
            else:
                ostore = _OracleStore(blocks, bb, dims, r, row_ids=rid, entry=int(rid[0]))

                def local(st, qd, nq, kk, LL, route):
                    i, dd, _, _ = st.search(qd.numpy(), kk, LL)
                    return torch.from_numpy(i), torch.from_numpy(dd)

                def merge(ai, ad, g, nq, kk, di):
                    i, dd = _np_merge(ai.numpy(), ad.numpy(), kk)
                    return torch.from_numpy(i), torch.from_numpy(dd)
                s = sh.DatasetShardedSearcher(ostore, device=gather_dev, local_search=local, merge=merge)
            ids, d = s.search(case["q"], k, L)
            if use_cuda:
                # again on a side stream: the store is bound to it, search and merge are only enqueued (no host sync)
                side = torch.cuda.Stream(device="cuda:%d" % cuda_dev)
                with torch.cuda.stream(side):
                    ids2, d2 = s.search(case["q"], k, L)
                    ids2, d2 = ids2.cpu(), d2.cpu()
                assert torch.equal(ids2, ids.cpu()) and torch.equal(d2, d.cpu())
            ret[rank] = (ids.cpu().numpy(), d.cpu().numpy(), s.last_gather_bytes)
        dist.barrier()
    finally:
        dist.destroy_process_group()


def _run(mode, use_cuda, case, k, L, world=2):
    mgr = mp.Manager()
    ret = mgr.dict()
    mp.spawn(_worker, args=(world, _free_port(), mode, use_cuda, case, k, L, ret), nprocs=world, join=True)
    return [ret[r] for r in range(world)]


def _expected(case, k, L):
    sys.path.insert(0, os.path.join(ROOT, "oracle"))
    import pyoracle
    full = pyoracle.search(case["full"], case["bb"], case["dims"], case["r"], case["q"], k, L, entry_row_id=0, nthreads=1)
    per = [pyoracle.search(b, case["bb"], case["dims"], case["r"], case["q"], k, L, row_ids=rid,
                           entry_row_id=int(rid[0]), nthreads=1) for b, rid in case["shards"]]
    merged = _np_merge(np.stack([p[0] for p in per]), np.stack([p[1] for p in per]), k)
    return full, merged


def test_slices_are_a_partition():
    sh = importlib.import_module("duckdb-lm-diskann_b200.sharded")
    for n in (0, 1, 7, 10, 10_000, 100_003):
        for w in (1, 2, 3, 8):
            cuts = [sh.query_slice(n, r, w) for r in range(w)]
            assert cuts[0][0] == 0 and cuts[-1][1] == n
            assert all(cuts[i][1] == cuts[i + 1][0] for i in range(w - 1))
            sizes = [b - a for a, b in cuts]
            assert max(sizes) - min(sizes) <= 1


def test_query_sharding_gloo_world2():
    case, k, L = _make_case(), 10, 32
    full, _ = _expected(case, k, L)
    out = _run("query", False, case, k, L)
    for ids, d, c, lo, hi in out:  # every rank ends with the full, identical result
        assert np.array_equal(ids, full[0]) and np.array_equal(c, full[2])
        assert np.array_equal(d.view(np.uint32), full[1].view(np.uint32))
    assert out[0][3:] == (0, 19) and out[1][3:] == (19, 37)  # 37 queries over 2 ranks


def test_dataset_sharding_gloo_world2():
    case, k, L = _make_case(), 10, 32
    _, merged = _expected(case, k, L)
    out = _run("dataset", False, case, k, L)
    for ids, d, nbytes in out:
        assert np.array_equal(ids, merged[0])
        assert np.array_equal(d.view(np.uint32), merged[1].view(np.uint32))
        assert nbytes == 37 * k * 12 * 2
    # the merged list draws from both shards and is ascending
    lo1 = case["shards"][1][1][0]
    assert (merged[0] >= lo1).any() and ((merged[0] < lo1) & (merged[0] >= 0)).any()
    assert np.all(np.diff(merged[1], axis=1) >= 0)


def test_both_modes_gloo_world3_ragged():
    """Three ranks, 1 601 rows and 37 queries: neither divides by 3, so the slices and shards are ragged."""
    case, k, L = _make_case(n=1601, world=3), 10, 32
    full, merged = _expected(case, k, L)
    out = _run("query", False, case, k, L, world=3)
    for ids, d, c, lo, hi in out:
        assert np.array_equal(ids, full[0]) and np.array_equal(c, full[2])
        assert np.array_equal(d.view(np.uint32), full[1].view(np.uint32))
    assert [o[3:] for o in out] == [(0, 13), (13, 25), (25, 37)]
    out = _run("dataset", False, case, k, L, world=3)
    for ids, d, nbytes in out:
        assert np.array_equal(ids, merged[0]) and np.array_equal(d.view(np.uint32), merged[1].view(np.uint32))
        assert nbytes == 37 * k * 12 * 3
    assert [b.shape[0] for b, _ in case["shards"]] == [534, 534, 533]


@pytest.mark.gpu
def test_query_sharding_cuda_world2():
    case, k, L = _make_case(n=3000, dims=128, r=32, nq=101), 10, 48
    full, _ = _expected(case, k, L)
    out = _run("query", True, case, k, L)
    for ids, d, c, lo, hi in out:
        assert np.array_equal(c, full[2])
        assert (ids == full[0]).all(axis=1).mean() >= 0.98
        assert np.allclose(d, full[1], rtol=1e-5, atol=1e-12)
    assert np.array_equal(out[0][0], out[1][0])


@pytest.mark.gpu
def test_dataset_sharding_cuda_world2():
    case, k, L = _make_case(n=3000, dims=128, r=32, nq=101), 10, 48
    _, merged = _expected(case, k, L)
    out = _run("dataset", True, case, k, L)
    for ids, d, _ in out:
        assert (ids == merged[0]).all(axis=1).mean() >= 0.98
        assert np.allclose(d, merged[1], rtol=1e-5, atol=1e-12)
    assert np.array_equal(out[0][0], out[1][0])
