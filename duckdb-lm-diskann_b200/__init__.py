"""Host-side Python binding of the B200 LM_DISKANN query path (ctypes over include/lmd_b200.h).

The product is the C-ABI library ``liblmd_b200.so`` (CUDA, sm_100a). This module only marshals
pointers; it never computes a distance and has no CPU fallback: if the library is missing or no
CUDA device is present, calls fail with :class:`LmdError`.

The directory name contains a hyphen (it mirrors the reference's repository name), so import it with
``importlib.import_module("duckdb-lm-diskann_b200")`` with the repository root on ``sys.path``.
"""
import ctypes as C
import os

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.environ.get("LMD_B200_LIB") or os.path.join(HERE, "liblmd_b200.so")  # the override is for kernel-tuning experiments

# common/ann.hpp:16-32
METRIC_L2, METRIC_COSINE, METRIC_IP = 1, 2, 3
VEC_FLOAT32, VEC_INT8 = 1, 2
ROUTE_EXACT, ROUTE_TERNARY, ROUTE_INT8 = 0, 1, 2
CODEC_NONE, CODEC_INT8 = 0, 1
SEARCH_SEPARATE_RERANK = 1
SEARCH_PER_QUERY_STATS = 2
SEARCH_ASYNC = 4
SEARCH_STRICT_FP = 8
SEARCH_NO_PREFILTER = 16
SEARCH_TENSOR_RERANK = 32
GROUP_QUERY_SHARDED, GROUP_DATASET_SHARDED = 0, 1
EMPTY_ROW_ID = np.iinfo(np.int64).max  # core/GraphManager.cpp:154-156
OK, ERR_INVALID_ARG, ERR_CUDA, ERR_OOM, ERR_STATE, ERR_UNSUPPORTED = range(6)
METRIC_BY_NAME = {"L2": METRIC_L2, "COSINE": METRIC_COSINE, "IP": METRIC_IP}


class LmdError(RuntimeError):
    def __init__(self, status, message):
        super().__init__("lmd status %d: %s" % (status, message))
        self.status = status


class Layout(C.Structure):
    _fields_ = [(n, C.c_uint64) for n in (
        "count_off", "vector_off", "ids_off", "pos_off", "neg_off", "total_node_size", "ternary_edge_bytes")]


class Config(C.Structure):
    _fields_ = [("dims", C.c_uint32), ("r", C.c_uint32), ("metric", C.c_uint32), ("vec_type", C.c_uint32),
                ("l_search", C.c_uint32), ("reserved", C.c_uint32), ("block_bytes", C.c_uint64)]


class Stats(C.Structure):
    _fields_ = [("queries", C.c_uint64), ("expansions", C.c_uint64), ("distance_evals", C.c_uint64),
                ("reranked", C.c_uint64), ("filter_evictions", C.c_uint64), ("tie_overflows", C.c_uint64),
                ("merge_fallbacks", C.c_uint64), ("algorithmic_bytes", C.c_uint64), ("exact_fetches", C.c_uint64),
                ("traverse_ms", C.c_float), ("rerank_ms", C.c_float), ("h2d_ms", C.c_float), ("d2h_ms", C.c_float),
                ("total_ms", C.c_float), ("launches", C.c_uint32), ("l_used", C.c_uint32)]

    def as_dict(self):
        return {n: getattr(self, n) for n, _ in self._fields_}


class GroupStats(C.Structure):
    _fields_ = [("members", C.c_uint32), ("mode", C.c_uint32), ("peer_access", C.c_uint32), ("reserved", C.c_uint32),
                ("queries", C.c_uint64 * 16), ("search_ms", C.c_float * 16), ("search_ms_max", C.c_float),
                ("merge_ms", C.c_float), ("wall_ms", C.c_float), ("reserved2", C.c_uint32),
                ("merge_input_bytes", C.c_uint64)]


class BuildParams(C.Structure):
    _fields_ = [("l_insert", C.c_uint32), ("alpha", C.c_float), ("max_batch", C.c_uint32), ("reserved", C.c_uint32)]


class BuildStats(C.Structure):
    _fields_ = [("nodes", C.c_uint64), ("rounds", C.c_uint64), ("search_expansions", C.c_uint64),
                ("search_evals", C.c_uint64), ("prune_evals", C.c_uint64), ("backlink_prunes", C.c_uint64),
                ("inbox_overflows", C.c_uint64), ("search_ms", C.c_float), ("prune_ms", C.c_float),
                ("backlink_ms", C.c_float), ("finalize_ms", C.c_float), ("total_ms", C.c_float),
                ("launches", C.c_uint32), ("reserved", C.c_uint32)]

    def as_dict(self):
        return {n: getattr(self, n) for n, _ in self._fields_}


# every symbol include/lmd_b200.h declares: (name, restype, argtypes)
_P = C.c_void_p
SYMBOLS = [
    ("lmd_layout_calc", C.c_int32, [C.c_uint64, C.c_uint32, C.c_uint32, C.POINTER(Layout)]),
    ("lmd_block_bytes", C.c_uint64, [C.c_uint64]),
    ("lmd_store_create", C.c_int32, [C.POINTER(Config), C.c_int, C.c_uint64, C.POINTER(_P)]),
    ("lmd_store_destroy", C.c_int32, [_P]),
    ("lmd_store_upload", C.c_int32, [_P, _P, _P, C.c_uint64]),
    ("lmd_store_update", C.c_int32, [_P, C.c_int64, _P]),
    ("lmd_store_apply", C.c_int32, [_P, _P, _P, C.c_uint64, C.POINTER(C.c_uint64)]),
    ("lmd_store_remove", C.c_int32, [_P, _P, C.c_uint64, C.POINTER(C.c_uint64)]),
    ("lmd_store_set_entry", C.c_int32, [_P, C.c_int64]),
    ("lmd_store_finalize", C.c_int32, [_P]),
    ("lmd_store_download_block", C.c_int32, [_P, C.c_int64, _P]),
    ("lmd_store_set_stream", C.c_int32, [_P, _P]),
    ("lmd_store_node_count", C.c_uint64, [_P]),
    ("lmd_store_device_bytes", C.c_uint64, [_P]),
    ("lmd_search_batch", C.c_int32, [_P, _P, C.c_uint64, C.c_uint32, C.c_uint32, C.c_uint32, C.c_uint32, _P, _P, _P,
                                    _P, C.POINTER(Stats)]),
    ("lmd_store_enable_codec", C.c_int32, [_P, C.c_uint32]),
    ("lmd_store_get_codec", C.c_int32, [_P, _P, C.POINTER(C.c_float)]),
    ("lmd_store_release_blocks", C.c_int32, [_P]),
    ("lmd_ctx_create", C.c_int32, [_P, C.POINTER(_P)]),
    ("lmd_ctx_destroy", C.c_int32, [_P]),
    ("lmd_ctx_set_stream", C.c_int32, [_P, _P]),
    ("lmd_ctx_search_batch", C.c_int32, [_P, _P, _P, C.c_uint64, C.c_uint32, C.c_uint32, C.c_uint32, C.c_uint32, _P, _P,
                                        _P, _P, C.POINTER(Stats)]),
    ("lmd_filter_create", C.c_int32, [_P, _P, C.c_uint64, C.POINTER(_P)]),
    ("lmd_filter_destroy", C.c_int32, [_P]),
    ("lmd_filter_count", C.c_uint64, [_P]),
    ("lmd_search_batch_filtered", C.c_int32, [_P, _P, _P, C.c_uint64, C.c_uint32, C.c_uint32, C.c_uint32, C.c_uint32,
                                             _P, _P, _P, _P, C.POINTER(Stats)]),
    ("lmd_group_create", C.c_int32, [C.POINTER(_P), C.c_uint32, C.c_uint32, C.POINTER(_P)]),
    ("lmd_group_destroy", C.c_int32, [_P]),
    ("lmd_group_size", C.c_uint32, [_P]),
    ("lmd_group_search_batch", C.c_int32, [_P, _P, C.c_uint64, C.c_uint32, C.c_uint32, C.c_uint32, C.c_uint32, _P, _P, _P,
                                          C.POINTER(GroupStats)]),
    ("lmd_merge_topk", C.c_int32, [C.c_int, _P, _P, C.c_uint32, C.c_uint64, C.c_uint32, _P, _P]),
    ("lmd_merge_topk_async", C.c_int32, [C.c_int, _P, _P, C.c_uint32, C.c_uint64, C.c_uint32, _P, _P, _P]),
    ("lmd_store_upload_vectors", C.c_int32, [_P, _P, _P, C.c_uint64]),
    ("lmd_build_graph", C.c_int32, [_P, C.POINTER(BuildParams), C.POINTER(BuildStats)]),
    ("lmd_robust_prune_batch", C.c_int32, [_P, _P, _P, C.c_uint64, C.c_uint32, C.c_float, _P, _P]),
    ("lmd_store_download_blocks", C.c_int32, [_P, C.c_uint64, C.c_uint64, _P]),
    ("lmd_store_download_row_ids", C.c_int32, [_P, C.c_uint64, C.c_uint64, _P]),
    ("lmd_static_scratch_bytes", C.c_uint64, []),
    ("lmd_search_batch_static", C.c_int32, [_P, _P, C.c_uint64, C.c_uint32, C.c_uint32, C.c_uint32, C.c_uint32, _P, _P, _P, _P]),
    ("lmd_last_error", C.c_char_p, []),
    ("lmd_abi_version", C.c_uint32, []),
    ("lmd_set_tuning", C.c_int32, [_P, C.c_char_p, C.c_int64]),
]

_lib = None


def lib():
    """Loads liblmd_b200.so. Raises if it has not been built - there is nothing to fall back to."""
    global _lib
    if _lib is None:
        if not os.path.exists(LIB_PATH):
            raise LmdError(ERR_STATE, "%s is missing: run `python duckdb-lm-diskann_b200/build.py` "
                                      "(there is no CPU fallback)" % LIB_PATH)
        l = C.CDLL(LIB_PATH)
        for name, res, args in SYMBOLS:
            f = getattr(l, name)
            f.restype = res
            f.argtypes = args
        _lib = l
    return _lib


def _check(status):
    if status != OK:
        raise LmdError(status, lib().lmd_last_error().decode("utf-8", "replace"))


def _ptr(x):
    """Address of a numpy array (host) or of anything with .data_ptr() (torch tensor, host or device)."""
    if x is None:
        return None
    if isinstance(x, int):
        return x
    if hasattr(x, "data_ptr"):
        if hasattr(x, "is_contiguous") and not x.is_contiguous():
            raise ValueError("tensor must be contiguous")
        return x.data_ptr()
    if isinstance(x, np.ndarray):
        if not x.flags["C_CONTIGUOUS"]:
            raise ValueError("array must be C-contiguous")
        return x.ctypes.data
    raise TypeError("unsupported buffer type %r" % type(x))


def layout_calc(dims, vec_type=VEC_FLOAT32, r=64):
    """core/index_config.cpp:104-148 through the C ABI (host code; no device needed)."""
    out = Layout()
    _check(lib().lmd_layout_calc(dims, vec_type, r, C.byref(out)))
    return out


def block_bytes(total_node_size):
    return int(lib().lmd_block_bytes(total_node_size))


class Store:
    """Device node store + search. Mirrors the query-side surface of the reference:
    IStorageManager::GetNodeBlockData / IGraphManager::{TryGetNodePointer, GetEntryPoint*} are replaced by
    upload()/set_entry(), ISearcher::Search by search()."""

    def __init__(self, dims, r, metric=METRIC_L2, vec_type=VEC_FLOAT32, l_search=100, capacity=0, device=0,
                 block_bytes=0):
        if isinstance(metric, str):
            metric = METRIC_BY_NAME[metric.upper()]
        self.cfg = Config(dims, r, metric, vec_type, l_search, 0, block_bytes)
        self._h = _P()
        _check(lib().lmd_store_create(C.byref(self.cfg), device, capacity, C.byref(self._h)))
        self.layout = layout_calc(dims, vec_type, r)
        self.block_bytes = block_bytes or int(lib().lmd_block_bytes(self.layout.total_node_size))
        self.device = device

    def close(self):
        try:
            if self._h:
                lib().lmd_store_destroy(self._h)
                self._h = _P()
        except TypeError:  # interpreter shutdown: module globals are already gone
            pass

    __del__ = close

    def __enter__(self):
        return self

    def __exit__(self, *a):
        self.close()

    def upload(self, blocks, row_ids=None, n=None):
        if n is None:
            n = blocks.shape[0]
        if row_ids is not None and isinstance(row_ids, np.ndarray):
            row_ids = np.ascontiguousarray(row_ids, dtype=np.int64)
        _check(lib().lmd_store_upload(self._h, _ptr(row_ids), _ptr(blocks), n))

    def update(self, row_id, block):
        _check(lib().lmd_store_update(self._h, int(row_id), _ptr(block)))

    def apply(self, row_ids, blocks):
        """Batch upsert of dirty / newly allocated blocks (MarkBlockDirty + AllocateNodeBlock) -> nodes appended."""
        rid = np.ascontiguousarray(row_ids, dtype=np.int64)
        out = C.c_uint64(0)
        _check(lib().lmd_store_apply(self._h, _ptr(rid), _ptr(blocks), rid.shape[0], C.byref(out)))
        return out.value

    def remove(self, row_ids):
        """HandleNodeDeletion / FreeNode for a batch of row ids -> how many were present."""
        rid = np.ascontiguousarray(row_ids, dtype=np.int64)
        out = C.c_uint64(0)
        _check(lib().lmd_store_remove(self._h, _ptr(rid), rid.shape[0], C.byref(out)))
        return out.value

    def set_entry(self, row_id):
        _check(lib().lmd_store_set_entry(self._h, int(row_id)))

    def finalize(self):
        _check(lib().lmd_store_finalize(self._h))

    def download_block(self, row_id):
        out = np.empty(self.block_bytes, dtype=np.uint8)
        _check(lib().lmd_store_download_block(self._h, int(row_id), _ptr(out)))
        return out

    # ---- graph write side (core/GraphManager.cpp:65-203, :259-457) ----
    def upload_vectors(self, vectors, row_ids=None):
        """Appends nodes given only their fp32 vectors [n, dims] (blocks get neighbour count 0)."""
        v = vectors if hasattr(vectors, "data_ptr") else np.ascontiguousarray(vectors, dtype=np.float32)
        if row_ids is not None and isinstance(row_ids, np.ndarray):
            row_ids = np.ascontiguousarray(row_ids, dtype=np.int64)
        _check(lib().lmd_store_upload_vectors(self._h, _ptr(row_ids), _ptr(v), v.shape[0]))

    def build_graph(self, l_insert=128, alpha=1.2, max_batch=0):
        """Batch-parallel Vamana build over every node in the store -> BuildStats."""
        bp = BuildParams(l_insert, alpha, max_batch, 0)
        st = BuildStats()
        _check(lib().lmd_build_graph(self._h, C.byref(bp), C.byref(st)))
        return st

    def robust_prune(self, node_row_ids, cand_row_ids, alpha=1.2):
        """RobustPrune of m nodes: cand_row_ids [m, c] (INT64_MAX = none) -> (ids [m, r] INT64_MAX padded, count [m])."""
        nodes = np.ascontiguousarray(node_row_ids, dtype=np.int64)
        cands = np.ascontiguousarray(cand_row_ids, dtype=np.int64)
        m, c = cands.shape
        out = np.full((m, self.cfg.r), EMPTY_ROW_ID, dtype=np.int64)
        cnt = np.zeros(m, dtype=np.uint32)
        _check(lib().lmd_robust_prune_batch(self._h, _ptr(nodes), _ptr(cands), m, c, alpha, _ptr(out), _ptr(cnt)))
        return out, cnt

    def download_blocks(self, first_slot=0, n=None, out=None):
        n = self.node_count - first_slot if n is None else n
        if out is None:
            out = np.empty((n, self.block_bytes), dtype=np.uint8)
        _check(lib().lmd_store_download_blocks(self._h, first_slot, n, _ptr(out)))
        return out

    def download_row_ids(self, first_slot=0, n=None):
        n = self.node_count - first_slot if n is None else n
        out = np.empty(n, dtype=np.int64)
        _check(lib().lmd_store_download_row_ids(self._h, first_slot, n, _ptr(out)))
        return out

    def enable_codec(self, codec=CODEC_INT8):
        """Builds the int8 edge-codec sidecar (row f4): ROUTE_INT8 searches, and the exact route's lower-bound pre-filter."""
        _check(lib().lmd_store_enable_codec(self._h, codec))

    def get_codec(self):
        """-> (offsets [dims] float32, scale) of the codec grid."""
        off = np.zeros(self.cfg.dims, dtype=np.float32)
        sc = C.c_float(0)
        _check(lib().lmd_store_get_codec(self._h, _ptr(off), C.byref(sc)))
        return off, float(sc.value)

    def release_blocks(self):
        """Frees the raw block mirror (search-only store)."""
        _check(lib().lmd_store_release_blocks(self._h))

    def context(self):
        """A per-caller search context (own stream / counters / scratch): one per host thread."""
        return Context(self)

    def set_stream(self, cuda_stream):
        """cuda_stream: a raw cudaStream_t address (e.g. torch.cuda.current_stream().cuda_stream) or None."""
        _check(lib().lmd_store_set_stream(self._h, cuda_stream or None))

    def set_tuning(self, key, value):
        _check(lib().lmd_set_tuning(self._h, key.encode(), int(value)))

    @property
    def node_count(self):
        return int(lib().lmd_store_node_count(self._h))

    @property
    def device_bytes(self):
        return int(lib().lmd_store_device_bytes(self._h))

    def make_filter(self, allowed_row_ids):
        """allowed_ids set of a filtered search (row f4) -> Filter (close() it, or use `with`)."""
        return Filter(self, allowed_row_ids)

    def search_raw(self, queries, nq, k, L=0, route=ROUTE_EXACT, flags=0, out_ids=None, out_dist=None, out_count=None,
                   out_qstats=None, stats=True, filter=None, ctx=None):
        """Thin call: any mix of numpy arrays / torch tensors / raw addresses. stats=False passes a NULL lmd_stats
        (the call still synchronises unless SEARCH_ASYNC is set) and returns None."""
        if ctx is not None:
            def call(*a):
                return lib().lmd_ctx_search_batch(ctx._h, filter._h if filter is not None else None, *a)
        elif filter is not None:
            def call(*a):
                return lib().lmd_search_batch_filtered(self._h, filter._h, *a)
        else:
            def call(*a):
                return lib().lmd_search_batch(self._h, *a)
        if (flags & SEARCH_ASYNC) or not stats:
            _check(call(_ptr(queries), nq, k, L, route, flags, _ptr(out_ids), _ptr(out_dist), _ptr(out_count),
                        _ptr(out_qstats), None))
            return None
        st = Stats()
        _check(call(_ptr(queries), nq, k, L, route, flags, _ptr(out_ids), _ptr(out_dist), _ptr(out_count),
                    _ptr(out_qstats), C.byref(st)))
        return st

    def search_static(self, queries, nq, k, L, route, out_ids, out_dist, out_count, flags=0, cuda_stream=None):
        """lmd_search_batch_static: device / pinned buffers only, the kernels' scratch in the library's static device
        array (no allocation inside the call; see include/lmd_b200.h for the limit and the sharing rule)."""
        _check(lib().lmd_search_batch_static(self._h, _ptr(queries), nq, k, L, route, flags, _ptr(out_ids), _ptr(out_dist),
                                             _ptr(out_count), cuda_stream or None))

    def search(self, queries, k, L=0, route=ROUTE_EXACT, flags=0, qstats=False, filter=None, ctx=None):
        """Host-buffer search: queries [nq, dims] float32 numpy -> (ids [nq,k] int64, dist [nq,k] f32,
        count [nq] u32, Stats[, qstats [nq,2] u32])."""
        q = np.ascontiguousarray(queries, dtype=np.float32)
        if q.ndim != 2 or q.shape[1] != self.cfg.dims:
            # db/LmDiskannIndex.cpp:651-655
            raise LmdError(ERR_INVALID_ARG, "Query vector dimension (%d) does not match index dimension (%d)." %
                           (q.shape[-1] if q.ndim else 0, self.cfg.dims))
        nq = q.shape[0]
        ids = np.full((nq, k), -1, dtype=np.int64)
        dist = np.zeros((nq, k), dtype=np.float32)
        count = np.zeros(nq, dtype=np.uint32)
        qs = np.zeros((nq, 2), dtype=np.uint32) if qstats else None
        if qstats:
            flags |= SEARCH_PER_QUERY_STATS
        st = self.search_raw(q, nq, k, L, route, flags, ids, dist, count, qs, filter=filter, ctx=ctx)
        return (ids, dist, count, st, qs) if qstats else (ids, dist, count, st)


class Context:
    """lmd_ctx: searches through different contexts of one store run side by side (one per host thread)."""

    def __init__(self, store):
        self._h = _P()
        self.store = store
        _check(lib().lmd_ctx_create(store._h, C.byref(self._h)))

    def search(self, queries, k, L=0, **kw):
        return self.store.search(queries, k, L, ctx=self, **kw)

    def set_stream(self, cuda_stream):
        _check(lib().lmd_ctx_set_stream(self._h, cuda_stream or None))

    def close(self):
        try:
            if self._h:
                lib().lmd_ctx_destroy(self._h)
                self._h = _P()
        except TypeError:
            pass

    __del__ = close

    def __enter__(self):
        return self

    def __exit__(self, *a):
        self.close()


class Group:
    """lmd_group: several stores on several devices of this ONE process searched as one (query- or dataset-sharded)."""

    def __init__(self, stores, mode=GROUP_QUERY_SHARDED):
        self.stores = list(stores)
        self.mode = mode
        arr = (_P * len(self.stores))(*[s._h for s in self.stores])
        self._h = _P()
        _check(lib().lmd_group_create(arr, len(self.stores), mode, C.byref(self._h)))

    def search_raw(self, queries, nq, k, L=0, route=ROUTE_EXACT, flags=0, out_ids=None, out_dist=None, out_count=None,
                   stats=True):
        st = GroupStats() if stats else None
        _check(lib().lmd_group_search_batch(self._h, _ptr(queries), nq, k, L, route, flags, _ptr(out_ids), _ptr(out_dist),
                                            _ptr(out_count), C.byref(st) if stats else None))
        return st

    def search(self, queries, k, L=0, route=ROUTE_EXACT, flags=0):
        """queries [nq, dims] float32 numpy -> (ids [nq,k], dist [nq,k], count [nq], GroupStats)"""
        q = np.ascontiguousarray(queries, dtype=np.float32)
        nq = q.shape[0]
        ids = np.full((nq, k), -1, dtype=np.int64)
        dist = np.zeros((nq, k), dtype=np.float32)
        count = np.zeros(nq, dtype=np.uint32)
        st = self.search_raw(q, nq, k, L, route, flags, ids, dist, count)
        return ids, dist, count, st

    def close(self):
        try:
            if self._h:
                lib().lmd_group_destroy(self._h)
                self._h = _P()
        except TypeError:
            pass

    __del__ = close

    def __enter__(self):
        return self

    def __exit__(self, *a):
        self.close()


class Filter:
    """lmd_filter: an allowed-ids set resolved to the store's slots."""

    def __init__(self, store, allowed_row_ids):
        rid = allowed_row_ids if hasattr(allowed_row_ids, "data_ptr") else \
            np.ascontiguousarray(allowed_row_ids, dtype=np.int64)
        self._h = _P()
        self._keep = store
        _check(lib().lmd_filter_create(store._h, _ptr(rid), int(rid.shape[0]), C.byref(self._h)))

    @property
    def count(self):
        return int(lib().lmd_filter_count(self._h))

    def close(self):
        try:
            if self._h:
                lib().lmd_filter_destroy(self._h)
                self._h = _P()
        except TypeError:
            pass

    __del__ = close

    def __enter__(self):
        return self

    def __exit__(self, *a):
        self.close()


def merge_topk_async(ids, dist, shards, nq, k, out_ids, out_dist, device=0, cuda_stream=0):
    """Device buffers; enqueued on `cuda_stream` (a raw cudaStream_t address, 0 = default stream), no synchronisation."""
    _check(lib().lmd_merge_topk_async(device, _ptr(ids), _ptr(dist), shards, nq, k, _ptr(out_ids), _ptr(out_dist),
                                      cuda_stream or None))


def merge_topk(ids, dist, shards, nq, k, out_ids, out_dist, device=0):
    _check(lib().lmd_merge_topk(device, _ptr(ids), _ptr(dist), shards, nq, k, _ptr(out_ids), _ptr(out_dist)))
