#!/usr/bin/env python
"""Generates tests/golden/*.npz from the REFERENCE'S OWN SOURCES compiled here (oracle/_ref, built by
oracle/Makefile from /root/reference/src/lm_diskann where the files lie). Run in the build container only
(/root/reference is not on the GPU box); the outputs are committed.

    python tests/golden/make_golden.py

The reference ships no golden vectors for this path, so these are frozen outputs of
  liblmd_ref_literal.so  reference sources untouched           -> arithmetic KATs + the literal search results
  liblmd_ref_fixed.so    + oracle/patches/searcher_fixed.patch -> the search results the CUDA path is graded on
Inputs are stored next to the outputs (vectors, neighbour slots, queries), so nothing is re-derived from a seed
at test time except the byte-level block packing (fixtures.pack_blocks, itself checked against the layout).
"""
import importlib
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "oracle"))

import pyoracle  # noqa: E402

fx = importlib.import_module("duckdb-lm-diskann_b200.fixtures")

METRIC = {"L2": 1, "COSINE": 2, "IP": 3}

SEARCH_CASES = [
    # name, n, dims, r, metric, kind, vec_type, nq, k, L, row_id_mode
    ("l2_d32_r8", 600, 32, 8, "L2", "clustered", 1, 32, 10, 24, "identity"),
    ("l2_d128_r32", 1200, 128, 32, "L2", "clustered", 1, 32, 10, 64, "identity"),
    ("l2_d96_r16_iid", 800, 96, 16, "L2", "iid", 1, 24, 5, 40, "identity"),
    ("cos_d48_r12", 700, 48, 12, "COSINE", "iid", 1, 24, 10, 32, "identity"),
    ("ip_d64_r16", 700, 64, 16, "IP", "iid", 1, 24, 10, 32, "identity"),
    ("l2_d40_r8_int8", 500, 40, 8, "L2", "iid", 2, 16, 7, 20, "identity"),
    ("l2_d32_r8_rowids", 500, 32, 8, "L2", "clustered", 1, 16, 10, 0, "sparse"),  # L = 0 -> cfg.l_search
]


def make_search_case(name, n, dims, r, metric, kind, vec_type, nq, k, L, rid_mode):
    x = fx.make_dataset(n, dims, kind=kind, seed=42, normalize=(metric == "COSINE"))
    q = fx.make_dataset(nq, dims, kind=kind, seed=43, normalize=(metric == "COSINE"))
    if vec_type == 2:
        x = np.clip(np.round(x * 40), -127, 127).astype(np.int8)
        q = (q * 40).astype(np.float32)
    nbrs, counts = fx.build_graph(np.asarray(x, dtype=np.float32), r, device="cpu")
    nbrs, counts = nbrs.numpy(), counts.numpy()
    row_ids = None
    entry = 0
    if rid_mode == "sparse":
        rng = np.random.default_rng(11)
        row_ids = rng.choice(np.arange(5, 20 * n, dtype=np.int64), size=n, replace=False)
        entry = int(row_ids[3])
    blocks = fx.pack_blocks(x, nbrs, counts, r, row_ids=row_ids, vec_type=vec_type).numpy()
    lay = fx.layout(dims, r, vec_type)
    cfg_l = 30
    out = {"x": x, "nbrs": nbrs.astype(np.int32), "counts": counts.astype(np.int32), "queries": q,
           "meta": np.array([n, dims, r, METRIC[metric], vec_type, nq, k, L, cfg_l, entry], dtype=np.int64)}
    if row_ids is not None:
        out["row_ids"] = row_ids
    for impl in ("fixed", "literal"):
        ids, dist, cnt, st = pyoracle.search(blocks, lay["block_bytes"], dims, r, q, k, L, metric=METRIC[metric],
                                             vec_type=vec_type, row_ids=row_ids, entry_row_id=entry,
                                             cfg_l_search=cfg_l, nthreads=1, impl=impl)
        out[impl + "_ids"], out[impl + "_dist"], out[impl + "_count"], out[impl + "_stats"] = ids, dist, cnt, st
    np.savez_compressed(os.path.join(HERE, "search_" + name + ".npz"), **out)
    print("search_%s: fixed expansions/query %.1f" % (name, out["fixed_stats"][:, 0].mean()))


def make_arith():
    rng = np.random.default_rng(2024)
    out = {}
    # ComputeExactDistanceFloat (core/distance.hpp:50-105): random pairs + the special cases
    for dims in (1, 3, 33, 96, 128, 768):
        a = rng.standard_normal((16, dims)).astype(np.float32)
        b = rng.standard_normal((16, dims)).astype(np.float32)
        a[0] = 0.0           # zero norm -> cosine 1.0
        b[1] = a[1]          # identical -> L2 0, cosine ~0 (clamp)
        b[2] = -a[2]         # opposite
        a[3] *= 1e-20        # underflowing squared norm
        a[4] *= 1e18         # overflowing products (inf / nan paths)
        d = np.zeros((3, 16), dtype=np.float32)
        for m in (1, 2, 3):
            for i in range(16):
                d[m - 1, i], err = pyoracle.distance(a[i], b[i], m, impl="literal")
                assert err == 0
        out["dist_a_%d" % dims], out["dist_b_%d" % dims], out["dist_out_%d" % dims] = a, b, d
    # EncodeTernary / dot / CalculateApproxDistance (ternary_quantization.hpp:146-199,269-296; distance.hpp:212-243)
    for dims in (1, 63, 64, 65, 96, 128, 200, 768):
        v = rng.standard_normal((8, dims)).astype(np.float32)
        v[0, ::3] = 0.0
        v[1] = 0.0
        v[2, 0] = -0.0
        w = rng.standard_normal((8, dims)).astype(np.float32)
        words = (dims + 63) // 64
        pos = np.zeros((8, words), dtype=np.uint64)
        neg = np.zeros((8, words), dtype=np.uint64)
        wpos = np.zeros((8, words), dtype=np.uint64)
        wneg = np.zeros((8, words), dtype=np.uint64)
        dots = np.zeros(8, dtype=np.int64)
        approx = np.zeros((2, 8), dtype=np.float32)
        for i in range(8):
            pos[i], neg[i] = pyoracle.encode_ternary(v[i], impl="literal")
            wpos[i], wneg[i] = pyoracle.encode_ternary(w[i], impl="literal")
            dots[i] = pyoracle.ternary_dot(pos[i], neg[i], wpos[i], wneg[i], impl="literal")
            edge = np.concatenate([wpos[i], wneg[i]]).view(np.uint8)
            approx[0, i], e1 = pyoracle.approx_distance(v[i], edge, dims, 2, impl="literal")
            approx[1, i], e2 = pyoracle.approx_distance(v[i], edge, dims, 3, impl="literal")
            _, e3 = pyoracle.approx_distance(v[i], edge, dims, 1, impl="literal")
            assert e1 == 0 and e2 == 0 and e3 == 1  # L2 throws (distance.hpp:236-239)
        out["tern_v_%d" % dims], out["tern_w_%d" % dims] = v, w
        out["tern_pos_%d" % dims], out["tern_neg_%d" % dims] = pos, neg
        out["tern_dot_%d" % dims], out["tern_approx_%d" % dims] = dots, approx
    # CalculateLayoutInternal (core/index_config.cpp:104-148)
    shapes = [(128, 32, 1), (128, 64, 1), (768, 32, 1), (768, 64, 1), (96, 32, 1), (96, 64, 1), (100, 24, 1),
              (33, 16, 2), (1, 1, 1), (4096, 128, 2), (65, 7, 1)]
    lay = np.zeros((len(shapes), 8), dtype=np.int64)
    for i, (d, r, vt) in enumerate(shapes):
        l = pyoracle.layout(d, vt, r, impl="literal")
        lay[i] = [l["count_off"], l["vector_off"], l["ids_off"], l["pos_off"], l["neg_off"], l["total_node_size"],
                  l["ternary_edge_bytes"], pyoracle.ref("literal").lmdref_block_bytes(l["total_node_size"])]
    out["layout_shapes"] = np.array(shapes, dtype=np.int64)
    out["layout_out"] = lay
    np.savez_compressed(os.path.join(HERE, "arith.npz"), **out)
    print("arith.npz written")


BUILD_CASES = [
    # name, n, dims, r, l_insert, alpha, metric, kind, vec_type, sparse row ids
    ("l2_d24_r8", 400, 24, 8, 24, 1.2, "L2", "iid", 1, False),
    ("cos_d32_r12", 350, 32, 12, 40, 1.2, "COSINE", "iid", 1, True),
    ("l2_d16_r6_int8", 300, 16, 6, 20, 1.0, "L2", "iid", 2, False),
]


def make_build_case(name, n, dims, r, l_insert, alpha, metric, kind, vec_type, sparse):
    """GraphManager::AddNode x n (serial build) and GraphManager::RobustPrune on random candidate lists, by the
    reference's own GraphManager.cpp + Searcher.cpp (+ H1-H5). Stored: inputs, neighbour ids / counts of every block."""
    x = fx.make_dataset(n, dims, kind=kind, seed=52, normalize=(metric == "COSINE"))
    if vec_type == 2:
        x = np.round(x * 40).astype(np.float32)
    rid = None
    if sparse:
        rid = np.random.default_rng(13).choice(np.arange(5, 30 * n, dtype=np.int64), size=n, replace=False)
    lay = fx.layout(dims, r, vec_type)
    out = {"x": x, "meta": np.array([n, dims, r, l_insert, METRIC[metric], vec_type], dtype=np.int64),
           "alpha": np.float32(alpha)}
    if rid is not None:
        out["row_ids"] = rid
    for impl in ("fixed", "literal"):
        blocks, entry = pyoracle.build_serial(x, r, l_insert, alpha, metric=METRIC[metric], vec_type=vec_type,
                                              row_ids=rid, impl=impl)
        out[impl + "_counts"] = blocks[:, 0:2].copy().view(np.uint16)[:, 0]
        out[impl + "_ids"] = blocks[:, lay["ids_off"]:lay["ids_off"] + 8 * r].copy().view(np.int64)
        out[impl + "_entry"] = np.int64(entry)
    rng = np.random.default_rng(17)
    m, c = 64, 3 * r + 5
    node_idx = rng.integers(0, n, m).astype(np.uint64)
    ids = np.arange(n, dtype=np.int64) if rid is None else rid
    cands = ids[rng.integers(0, n, (m, c))]
    cands[:, ::7] = np.iinfo(np.int64).max      # holes
    cands[:, 3] = cands[:, 4]                   # duplicates
    cands[:, 5] = ids[node_idx.astype(np.int64)]  # the node itself
    cands[:, 6] = 10 ** 15                      # a row id that does not exist
    p_ids, p_cnt = pyoracle.robust_prune(x, r, alpha, node_idx, cands, metric=METRIC[metric], vec_type=vec_type,
                                         row_ids=rid, impl="fixed")
    out["prune_node_idx"], out["prune_cands"], out["prune_ids"], out["prune_count"] = node_idx, cands, p_ids, p_cnt
    np.savez_compressed(os.path.join(HERE, "build_" + name + ".npz"), **out)
    print("build_%s: mean degree %.2f (literal build: %d edges)" % (name, out["fixed_counts"].mean(),
                                                                     int(out["literal_counts"].sum())))


if __name__ == "__main__":
    assert pyoracle.have_ref("fixed") and pyoracle.have_ref("literal"), "run `make -C oracle ref` first"
    make_arith()
    for c in SEARCH_CASES:
        make_search_case(*c)
    for c in BUILD_CASES:
        make_build_case(*c)
