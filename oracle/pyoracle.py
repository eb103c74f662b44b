"""TEST INFRASTRUCTURE ONLY (oracle/). ctypes loaders for the CPU checkers.

Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs import this module.
Nothing under duckdb-lm-diskann_b200/ does, and the product library never links or executes it.

  port()         oracle/liblmd_oracle.so        plain-C restatement (lmd_oracle.c)
  ref("fixed")   oracle/_ref/liblmd_ref_fixed.so   the reference's own sources + patches/searcher_fixed.patch
  ref("literal") oracle/_ref/liblmd_ref_literal.so the reference's own sources, untouched
"""
import ctypes as C
import os
import subprocess

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
PORT_PATH = os.path.join(HERE, "liblmd_oracle.so")
REF_PATHS = {"fixed": os.path.join(HERE, "_ref", "liblmd_ref_fixed.so"),
             "literal": os.path.join(HERE, "_ref", "liblmd_ref_literal.so")}

_P = C.c_void_p
_libs = {}


def build(ref=True):
    """make port (+ ref when /root/reference is present). Building the checker is not using it."""
    subprocess.run(["make", "-s", "-C", HERE, "port"], check=True)
    if ref:
        subprocess.run(["make", "-s", "-C", HERE, "ref"], check=True)


def _sig(fn, res, args):
    fn.restype = res
    fn.argtypes = args


def port():
    if "port" not in _libs:
        if not os.path.exists(PORT_PATH):
            build(ref=False)
        l = C.CDLL(PORT_PATH)
        _sig(l.lmdo_layout_calc, C.c_int, [C.c_uint64, C.c_int, C.c_uint32, _P])
        _sig(l.lmdo_block_bytes, C.c_uint64, [C.c_uint64])
        _sig(l.lmdo_distance, C.c_float, [_P, _P, C.c_uint64, C.c_int, _P])
        _sig(l.lmdo_convert_raw, C.c_int, [_P, _P, C.c_uint64, C.c_int])
        _sig(l.lmdo_words_per_plane, C.c_uint64, [C.c_uint64])
        _sig(l.lmdo_encode_ternary, None, [_P, C.c_uint64, _P, _P])
        _sig(l.lmdo_ternary_dot, C.c_int64, [_P, _P, _P, _P, C.c_uint64])
        _sig(l.lmdo_approx_distance, C.c_float, [_P, _P, C.c_uint64, C.c_int, _P])
        _sig(l.lmdo_search, C.c_int, [_P, C.c_uint64, C.c_uint64, _P, C.c_int64, C.c_int, C.c_int, C.c_uint64,
                                      C.c_uint32, C.c_uint32, _P, C.c_uint64, C.c_uint64, C.c_uint64, C.c_int, _P, _P,
                                      _P, _P, C.c_int])
        _sig(l.lmdo_search_filtered, C.c_int, [_P, C.c_uint64, C.c_uint64, _P, C.c_int64, C.c_int, C.c_int, C.c_uint64,
                                               C.c_uint32, C.c_uint32, _P, C.c_uint64, C.c_uint64, C.c_uint64, C.c_int,
                                               _P, _P, _P, _P, _P, C.c_int])
        _sig(l.lmdo_build_serial, C.c_int, [_P, C.c_uint64, C.c_uint64, C.c_int, C.c_int, C.c_uint32, C.c_uint32,
                                            C.c_float, _P, C.c_uint64, _P, _P])
        _sig(l.lmdo_robust_prune, C.c_int, [_P, C.c_uint64, C.c_uint64, C.c_int, C.c_int, C.c_uint32, C.c_float, _P,
                                            C.c_uint64, _P, _P, C.c_uint64, _P, _P])
        _sig(l.lmdo_set_codec, None, [_P, C.c_float])
        _libs["port"] = l
    return _libs["port"]


_codec_keep = None


def set_codec(offsets, scale):
    """The int8 codec grid the next route=2 searches of the port use (downloaded from the store under test)."""
    global _codec_keep
    _codec_keep = np.ascontiguousarray(offsets, dtype=np.float32)
    port().lmdo_set_codec(_codec_keep.ctypes.data, float(scale))


def have_ref(kind="fixed"):
    return os.path.exists(REF_PATHS[kind])


def ref(kind="fixed"):
    key = "ref_" + kind
    if key not in _libs:
        l = C.CDLL(REF_PATHS[kind])
        _sig(l.lmdref_is_fixed, C.c_int, [])
        _sig(l.lmdref_layout, C.c_int, [C.c_uint64, C.c_int, C.c_uint32, _P])
        _sig(l.lmdref_block_bytes, C.c_uint64, [C.c_uint64])
        _sig(l.lmdref_distance, C.c_float, [_P, _P, C.c_uint64, C.c_int, _P])
        _sig(l.lmdref_convert_raw, C.c_int, [_P, _P, C.c_uint64, C.c_int])
        _sig(l.lmdref_words_per_plane, C.c_uint64, [C.c_uint64])
        _sig(l.lmdref_encode_ternary, None, [_P, C.c_uint64, _P, _P])
        _sig(l.lmdref_ternary_dot, C.c_int64, [_P, _P, _P, _P, C.c_uint64])
        _sig(l.lmdref_ternary_dot_scalar, C.c_int64, [_P, _P, _P, _P, C.c_uint64])
        _sig(l.lmdref_compress_edge, C.c_int, [_P, _P, C.c_uint64])
        _sig(l.lmdref_approx_distance, C.c_float, [_P, _P, C.c_uint64, C.c_int, _P])
        _sig(l.lmdref_search, C.c_int, [_P, C.c_uint64, C.c_uint64, _P, C.c_int64, C.c_int, C.c_int, C.c_uint64,
                                        C.c_uint32, C.c_uint32, _P, C.c_uint64, C.c_uint64, C.c_uint64, _P, _P, _P, _P,
                                        C.c_int])
        _sig(l.lmdref_build_serial, C.c_int, [_P, C.c_uint64, C.c_uint64, C.c_int, C.c_int, C.c_uint32, C.c_uint32,
                                              C.c_float, _P, C.c_uint64, _P, _P])
        _sig(l.lmdref_robust_prune, C.c_int, [_P, C.c_uint64, C.c_uint64, C.c_int, C.c_int, C.c_uint32, C.c_float, _P,
                                              C.c_uint64, _P, _P, C.c_uint64, _P, _P])
        _libs[key] = l
    return _libs[key]


def _np(a, dt):
    return np.ascontiguousarray(a, dtype=dt)


def layout(dims, vec_type=1, r=64, impl="port"):
    out = np.zeros(7, dtype=np.uint64)
    if impl == "port":
        rc = port().lmdo_layout_calc(dims, vec_type, r, out.ctypes.data)
    else:
        rc = ref(impl).lmdref_layout(dims, vec_type, r, out.ctypes.data)
    if rc:
        raise ValueError("layout failed")
    names = ("count_off", "vector_off", "ids_off", "pos_off", "neg_off", "total_node_size", "ternary_edge_bytes")
    return dict(zip(names, (int(v) for v in out)))


def distance(a, b, metric, impl="port"):
    a = _np(a, np.float32)
    b = _np(b, np.float32)
    err = C.c_int(0)
    if impl == "port":
        d = port().lmdo_distance(a.ctypes.data, b.ctypes.data, a.size, metric, C.addressof(err))
    else:
        d = ref(impl).lmdref_distance(a.ctypes.data, b.ctypes.data, a.size, metric, C.addressof(err))
    return np.float32(d), err.value


def encode_ternary(v, impl="port"):
    v = _np(v, np.float32)
    words = (v.size + 63) // 64
    pos = np.zeros(words, dtype=np.uint64)
    neg = np.zeros(words, dtype=np.uint64)
    if impl == "port":
        port().lmdo_encode_ternary(v.ctypes.data, v.size, pos.ctypes.data, neg.ctypes.data)
    else:
        ref(impl).lmdref_encode_ternary(v.ctypes.data, v.size, pos.ctypes.data, neg.ctypes.data)
    return pos, neg


def ternary_dot(qp, qn, vp, vn, impl="port"):
    qp, qn, vp, vn = (_np(x, np.uint64) for x in (qp, qn, vp, vn))
    if impl == "port":
        return int(port().lmdo_ternary_dot(qp.ctypes.data, qn.ctypes.data, vp.ctypes.data, vn.ctypes.data, qp.size))
    return int(ref(impl).lmdref_ternary_dot(qp.ctypes.data, qn.ctypes.data, vp.ctypes.data, vn.ctypes.data, qp.size))


def approx_distance(q, compressed, dims, metric, impl="port"):
    q = _np(q, np.float32)
    c = _np(compressed, np.uint8)
    err = C.c_int(0)
    if impl == "port":
        d = port().lmdo_approx_distance(q.ctypes.data, c.ctypes.data, dims, metric, C.addressof(err))
    else:
        d = ref(impl).lmdref_approx_distance(q.ctypes.data, c.ctypes.data, dims, metric, C.addressof(err))
    return np.float32(d), err.value


def search(blocks, block_bytes, dims, r, queries, k, L=0, metric=1, vec_type=1, row_ids=None, entry_row_id=0,
           cfg_l_search=100, route=0, nthreads=0, impl="port", allowed=None):
    """-> ids [nq,k] int64 (-1 padded), dist [nq,k] f32, count [nq] u32, stats [nq,2] u64 {expansions, vector
    fetches}. impl: "port" (C restatement; route 0/1), "fixed" / "literal" (reference sources; route 0 only)."""
    b = _np(np.asarray(blocks).reshape(-1), np.uint8)
    n = b.size // block_bytes
    q = _np(queries, np.float32)
    nq = q.shape[0]
    rid = None if row_ids is None else _np(row_ids, np.int64)
    ids = np.full((nq, k), -1, dtype=np.int64)
    dist = np.zeros((nq, k), dtype=np.float32)
    cnt = np.zeros(nq, dtype=np.uint32)
    st = np.zeros((nq, 2), dtype=np.uint64)
    ridp = None if rid is None else rid.ctypes.data
    if allowed is not None:
        # filtered search (row f4; C restatement only - the reference has no such code): allowed [n] per block position
        if impl != "port":
            raise ValueError("the reference sources implement no filtered search")
        a = _np(np.asarray(allowed).astype(np.uint8), np.uint8)
        assert a.size == n
        rc = port().lmdo_search_filtered(b.ctypes.data, n, block_bytes, ridp, entry_row_id, metric, vec_type, dims, r,
                                         cfg_l_search, q.ctypes.data, nq, k, L, route, a.ctypes.data, ids.ctypes.data,
                                         dist.ctypes.data, cnt.ctypes.data, st.ctypes.data, nthreads)
    elif impl == "port":
        rc = port().lmdo_search(b.ctypes.data, n, block_bytes, ridp, entry_row_id, metric, vec_type, dims, r,
                                cfg_l_search, q.ctypes.data, nq, k, L, route, ids.ctypes.data, dist.ctypes.data,
                                cnt.ctypes.data, st.ctypes.data, nthreads)
    else:
        if route != 0:
            raise ValueError("the reference executes exact routing only (Searcher.cpp:168-171)")
        rc = ref(impl).lmdref_search(b.ctypes.data, n, block_bytes, ridp, entry_row_id, metric, vec_type, dims, r,
                                     cfg_l_search, q.ctypes.data, nq, k, L, ids.ctypes.data, dist.ctypes.data,
                                     cnt.ctypes.data, st.ctypes.data, nthreads)
    if rc:
        raise RuntimeError("oracle search failed (%d)" % rc)
    return ids, dist, cnt, st


INT64_MAX = np.iinfo(np.int64).max


def build_serial(x, r, l_insert=128, alpha=1.2, metric=1, vec_type=1, row_ids=None, impl="port"):
    """Serial Vamana build, one AddNode per row in order (core/GraphManager.cpp:65-203). x: [n, dims] float32 (for
    INT8 nodes the values are cast on write, as the reference does). -> (blocks [n, block_bytes] uint8, entry row id).
    impl: "port" (C restatement), "fixed" / "literal" (the reference's own GraphManager.cpp + Searcher.cpp)."""
    x = _np(x, np.float32)
    n, dims = x.shape
    lay = layout(dims, vec_type, r)
    bb = (lay["total_node_size"] + 4095) // 4096 * 4096
    blocks = np.zeros((n, bb), dtype=np.uint8)
    rid = None if row_ids is None else _np(row_ids, np.int64)
    entry = C.c_int64(INT64_MAX)
    args = (x.ctypes.data, n, dims, metric, vec_type, r, l_insert, alpha, None if rid is None else rid.ctypes.data, bb,
            blocks.ctypes.data, C.addressof(entry))
    rc = port().lmdo_build_serial(*args) if impl == "port" else ref(impl).lmdref_build_serial(*args)
    if rc:
        raise RuntimeError("oracle build failed (%d)" % rc)
    return blocks, int(entry.value)


def robust_prune(x, r, alpha, node_idx, cands, metric=1, vec_type=1, row_ids=None, impl="port"):
    """RobustPrune (core/GraphManager.cpp:259-457) of m nodes. node_idx [m] positions in x; cands [m, c] candidate ROW
    ids, INT64_MAX = none. -> (ids [m, r] INT64_MAX padded, count [m])."""
    x = _np(x, np.float32)
    n, dims = x.shape
    node_idx = _np(node_idx, np.uint64)
    cands = _np(cands, np.int64)
    m, c = cands.shape
    rid = None if row_ids is None else _np(row_ids, np.int64)
    out = np.full((m, r), INT64_MAX, dtype=np.int64)
    cnt = np.zeros(m, dtype=np.uint32)
    args = (x.ctypes.data, n, dims, metric, vec_type, r, alpha, None if rid is None else rid.ctypes.data, m,
            node_idx.ctypes.data, cands.ctypes.data, c, out.ctypes.data, cnt.ctypes.data)
    rc = port().lmdo_robust_prune(*args) if impl == "port" else ref(impl).lmdref_robust_prune(*args)
    if rc:
        raise RuntimeError("oracle robust_prune failed (%d)" % rc)
    return out, cnt
