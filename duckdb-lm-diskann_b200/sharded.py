"""Multi-GPU host logic of the LM_DISKANN query path. One process per GPU
(torch.distributed; NCCL on the GPUs, gloo in the CPU tests of the plumbing).

  QueryShardedSearcher    node store replicated in every GPU's HBM, rank g searches queries [lo_g, hi_g).
                          NO data-path collective; an optional all-gather hands every rank the full result.
  DatasetShardedSearcher  rank s holds shard s (shard-local graph + entry point, GLOBAL row ids in the blocks);
                          every rank searches the full batch on its shard, ONE all-gather of the per-shard
                          [Q][k] (distance, row id) lists, then the k-way merge kernel (lmd_merge_topk).

The local search and the merge default to the CUDA path (the C ABI); they are constructor arguments only so that the
CPU tests can exercise the slicing / gather layout with a stand-in. Nothing here falls back by itself.
"""
import numpy as np
import torch
import torch.distributed as dist


def query_slice(nq, rank, world):
    """Contiguous, balanced: the first nq % world ranks get one extra query."""
    base, extra = divmod(nq, world)
    lo = rank * base + min(rank, extra)
    return lo, lo + base + (1 if rank < extra else 0)


def shard_rows(n, rank, world):
    """Rows [lo, hi) of the base table held by shard `rank` (same rule as query_slice)."""
    return query_slice(n, rank, world)


def _world(group):
    if dist.is_available() and dist.is_initialized():
        return dist.get_rank(group), dist.get_world_size(group)
    return 0, 1


class QueryShardedSearcher:
    def __init__(self, store, group=None, device=None):
        self.store, self.group = store, group
        self.rank, self.world = _world(group)
        self.device = device if device is not None else (
            torch.device("cuda", store.device) if torch.cuda.is_available() else torch.device("cpu"))

    def search_local(self, queries, k, L=0, **kw):
        """This rank's slice of `queries` (the full [nq, dims] batch, identical on every rank) ->
        (lo, hi, ids [hi-lo, k], dist, count, stats)."""
        lo, hi = query_slice(queries.shape[0], self.rank, self.world)
        ids, d, cnt, st = self.store.search(queries[lo:hi], k, L, **kw)[:4]
        return lo, hi, ids, d, cnt, st

    def search(self, queries, k, L=0, gather=True, **kw):
        """Full-batch result on every rank (gather=True) or only this rank's rows filled in (gather=False; the caller
        owns a shared / pinned [nq][k] buffer in the reference's single-process setting)."""
        nq = queries.shape[0]
        lo, hi, ids, d, cnt, st = self.search_local(queries, k, L, **kw)
        out_ids = np.full((nq, k), -1, dtype=np.int64)
        out_d = np.zeros((nq, k), dtype=np.float32)
        out_c = np.zeros(nq, dtype=np.uint32)
        out_ids[lo:hi], out_d[lo:hi], out_c[lo:hi] = ids, d, cnt
        if gather and self.world > 1:
            # ranks hold slices of different length: pad to the longest, gather, cut
            m = -(-nq // self.world)
            pack = torch.zeros((m, 2 * k + 1), dtype=torch.int64)
            pack[: hi - lo, :k] = torch.from_numpy(ids)
            pack[: hi - lo, k:2 * k] = torch.from_numpy(d.view(np.int32).astype(np.int64))
            pack[: hi - lo, 2 * k] = torch.from_numpy(cnt.astype(np.int64))
            pack = pack.to(self.device)
            allp = torch.empty((self.world * m, 2 * k + 1), dtype=torch.int64, device=self.device)
            dist.all_gather_into_tensor(allp, pack, group=self.group)  # concatenated along dim 0 (gloo and NCCL)
            allp = allp.view(self.world, m, 2 * k + 1).cpu().numpy()
            for r in range(self.world):
                rl, rh = query_slice(nq, r, self.world)
                out_ids[rl:rh] = allp[r, : rh - rl, :k]
                out_d[rl:rh] = allp[r, : rh - rl, k:2 * k].astype(np.int32).view(np.float32)
                out_c[rl:rh] = allp[r, : rh - rl, 2 * k].astype(np.uint32)
        return out_ids, out_d, out_c, st


def _cuda_local_search(store, queries_dev, nq, k, L, route):
    """Device-resident search through the C ABI: returns (ids [nq,k] i64, dist [nq,k] f32) torch tensors on the GPU.
    On a non-default torch stream the store is bound to that stream and the search is only ENQUEUED (the all-gather and
    the merge follow on the same stream, no host round trip); on the default stream (handle 0, which the C ABI reads as
    "the store's own stream") the call synchronises before the collectives are issued."""
    from . import SEARCH_ASYNC
    dev = queries_dev.device
    ids = torch.empty((nq, k), dtype=torch.int64, device=dev)
    d = torch.empty((nq, k), dtype=torch.float32, device=dev)
    cur = torch.cuda.current_stream(dev).cuda_stream
    if not cur:
        # default stream: the store searches on its own non-blocking stream, which nothing orders against pending
        # default-stream work that may still be producing `queries_dev`
        torch.cuda.current_stream(dev).synchronize()
    if getattr(store, "_bound_stream", 0) != cur:
        store.set_stream(cur or None)
        store._bound_stream = cur
    store.search_raw(queries_dev, nq, k, L, route=route, flags=SEARCH_ASYNC if cur else 0, out_ids=ids, out_dist=d,
                     stats=False)
    return ids, d


def _cuda_merge(all_ids, all_d, shards, nq, k, device_index):
    from . import merge_topk_async  # the C ABI (lmd_merge_topk_async): same stream as the all-gather before it
    out_ids = torch.empty((nq, k), dtype=torch.int64, device=all_ids.device)
    out_d = torch.empty((nq, k), dtype=torch.float32, device=all_ids.device)
    merge_topk_async(all_ids, all_d, shards, nq, k, out_ids, out_d, device=device_index,
                     cuda_stream=torch.cuda.current_stream(all_ids.device).cuda_stream)
    return out_ids, out_d


class DatasetShardedSearcher:
    """store: this rank's shard. search() returns the merged global top-k on every rank."""

    def __init__(self, store, group=None, device=None, local_search=None, merge=None):
        self.store, self.group = store, group
        self.rank, self.world = _world(group)
        self.device = device if device is not None else (
            torch.device("cuda", store.device) if torch.cuda.is_available() else torch.device("cpu"))
        self._local_search = local_search or _cuda_local_search
        self._merge = merge or _cuda_merge
        self.last_gather_bytes = 0

    def search(self, queries, k, L=0, route=0):
        """queries: [nq, dims] float32 numpy array or torch tensor, identical on every rank.
        -> (ids [nq,k] int64, dist [nq,k] float32) torch tensors on self.device, ascending by (distance, id), -1 padded."""
        q = torch.as_tensor(queries, dtype=torch.float32).contiguous().to(self.device)
        nq = q.shape[0]
        ids, d = self._local_search(self.store, q, nq, k, L, route)
        if self.world == 1:
            return ids, d
        all_ids = torch.empty((self.world * nq, k), dtype=torch.int64, device=self.device)  # [G][nq][k], rank-major
        all_d = torch.empty((self.world * nq, k), dtype=torch.float32, device=self.device)
        # the one exchange step of the path: Q*k*12 bytes per rank over NVLink / NVSwitch
        dist.all_gather_into_tensor(all_ids, ids.contiguous(), group=self.group)
        dist.all_gather_into_tensor(all_d, d.contiguous(), group=self.group)
        self.last_gather_bytes = nq * k * 12 * self.world
        dev_index = self.device.index if self.device.type == "cuda" else 0
        return self._merge(all_ids.view(self.world, nq, k), all_d.view(self.world, nq, k), self.world, nq, k, dev_index)
