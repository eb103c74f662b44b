"""-m gpu: parity to the letter (LMD_SEARCH_STRICT_FP), the special values of core/distance.hpp:75-99, the routing
sidecars (ternary planes re-laid, int8 edge codec + the exact route's lower-bound pre-filter), per-caller search contexts
and the search-only store. All through the C ABI, against oracle/liblmd_oracle.so."""
import glob
import os
import threading

import numpy as np
import pytest

from parity import check_parity
from test_gpu_parity import CASES, METRIC, _store

pytestmark = pytest.mark.gpu

HERE = os.path.dirname(os.path.abspath(__file__))
INT64_MAX = np.iinfo(np.int64).max


# ---- strict fp: the reference's own summation order => the reference's own decisions, bit for bit ---------------------
@pytest.mark.parametrize("n,dims,r,metric,kind,k,L", CASES)
def test_strict_fp_is_bit_exact(lmd, oracle, make_index, n, dims, r, metric, kind, k, L):
    ix = make_index(n, dims, r, metric=metric, kind=kind, nq=96)
    with _store(lmd, ix, dims, r, metric) as s:
        s.upload(ix["blocks_np"])
        s.set_entry(0)
        ids, dist, cnt, st, qs = s.search(ix["queries"], k, L, qstats=True, flags=lmd.SEARCH_STRICT_FP)
        ids2, dist2, cnt2, _ = s.search(ix["queries"], k, L, flags=lmd.SEARCH_STRICT_FP | lmd.SEARCH_SEPARATE_RERANK)
    o_ids, o_dist, o_cnt, o_st = oracle.search(ix["blocks_np"], ix["block_bytes"], dims, r, ix["queries"], k, L,
                                                metric=METRIC[metric], entry_row_id=0, cfg_l_search=100)
    assert np.array_equal(cnt, o_cnt)
    assert np.array_equal(ids, o_ids), "%d of %d queries differ" % ((ids != o_ids).any(axis=1).sum(), ids.shape[0])
    assert np.array_equal(dist.view(np.uint32), o_dist.view(np.uint32))  # the same BITS
    assert np.array_equal(qs[:, 0].astype(np.int64), o_st[:, 0].astype(np.int64))  # expansions per query
    assert st.tie_overflows == 0
    # the oracle's second counter is vector fetches: traversal evaluations + the final pass over the list
    assert np.array_equal(ids2, o_ids) and np.array_equal(dist2.view(np.uint32), o_dist.view(np.uint32))


@pytest.mark.parametrize("metric,dims,r", [("COSINE", 128, 32), ("IP", 100, 24)])
def test_strict_fp_ternary_rerank(lmd, oracle, make_index, metric, dims, r):
    """Ternary routing is integer arithmetic (already bit-exact); with the strict rerank the returned distances are too."""
    ix = make_index(1500, dims, r, metric=metric, kind="iid", nq=64)
    with _store(lmd, ix, dims, r, metric) as s:
        s.upload(ix["blocks_np"])
        s.set_entry(0)
        s.set_tuning("tie_slots", 64)
        ids, dist, cnt, _ = s.search(ix["queries"], 10, 64, route=lmd.ROUTE_TERNARY, flags=lmd.SEARCH_STRICT_FP)
    o = oracle.search(ix["blocks_np"], ix["block_bytes"], dims, r, ix["queries"], 10, 64, metric=METRIC[metric],
                      entry_row_id=0, route=1)
    assert np.array_equal(ids, o[0]) and np.array_equal(dist.view(np.uint32), o[1].view(np.uint32))


def test_strict_fp_robust_prune_is_list_exact(lmd, oracle, fixtures_mod):
    """lmd_robust_prune_batch with "strict_fp": every alpha comparison of core/GraphManager.cpp:347-362 sees the
    reference's bits, so every list is the reference's list (the default mode is within fp32 summation noise of it)."""
    for metric, dims, r, kind in [("L2", 128, 32, "clustered"), ("COSINE", 96, 24, "iid"), ("IP", 100, 16, "iid")]:
        n, m, c = 1500, 300, 150
        x = fixtures_mod.make_dataset(n, dims, kind=kind, seed=8, normalize=(metric == "COSINE"))
        rng = np.random.default_rng(4)
        rid = rng.choice(np.arange(1, 20 * n, dtype=np.int64), size=n, replace=False)
        node_idx = rng.integers(0, n, m)
        d = ((x[node_idx, None, :16] - x[None, :, :16]) ** 2).sum(-1)
        near = np.argsort(d, axis=1)[:, :c - 20]
        cands = rid[np.concatenate([near, rng.integers(0, n, (m, 20))], axis=1)]
        cands[:, 7] = INT64_MAX
        o_ids, o_cnt = oracle.robust_prune(x, r, 1.2, node_idx.astype(np.uint64), cands, metric=METRIC[metric],
                                           row_ids=rid)
        with lmd.Store(dims, r, metric=METRIC[metric], capacity=n) as s:
            s.upload_vectors(x, row_ids=rid)
            s.set_tuning("strict_fp", 1)
            ids, cnt = s.robust_prune(rid[node_idx], cands, alpha=1.2)
        assert np.array_equal(cnt, o_cnt), metric
        assert np.array_equal(ids, o_ids), "%s: %d of %d lists differ" % (metric, (ids != o_ids).any(axis=1).sum(), m)


@pytest.mark.parametrize("path", sorted(glob.glob(os.path.join(HERE, "golden", "build_*.npz"))),
                         ids=lambda p: os.path.basename(p)[6:-4])
def test_strict_fp_robust_prune_golden(lmd, path):
    """... and on the frozen output of the reference's own RobustPrune."""
    z = np.load(path)
    n, dims, r, l_insert, metric, vec_type = (int(v) for v in z["meta"])
    rid = z["row_ids"] if "row_ids" in z.files else None
    with lmd.Store(dims, r, metric=metric, vec_type=vec_type, capacity=n) as s:
        s.upload_vectors(z["x"], row_ids=rid)
        s.set_tuning("strict_fp", 1)
        rows = np.arange(n, dtype=np.int64) if rid is None else rid
        ids, cnt = s.robust_prune(rows[z["prune_node_idx"].astype(np.int64)], z["prune_cands"], alpha=float(z["alpha"]))
    assert np.array_equal(cnt, z["prune_count"])
    assert np.array_equal(ids, z["prune_ids"])


# ---- special values (core/distance.hpp:56-99): zero norms, -0.0, +-Inf, NaN ---------------------------------------------
def _star_index(fixtures_mod, x, r):
    """Node 0 points at nodes 1..r, nobody else has neighbours: one expansion evaluates every row."""
    n = x.shape[0]
    nbrs = np.full((n, r), -1, dtype=np.int64)
    nbrs[0, : n - 1] = np.arange(1, n)
    counts = np.zeros(n, dtype=np.int64)
    counts[0] = n - 1
    return fixtures_mod.pack_blocks(x, nbrs, counts, r).numpy()


@pytest.mark.parametrize("strict", [False, True])
@pytest.mark.parametrize("metric", ["L2", "COSINE", "IP"])
@pytest.mark.parametrize("dims", [128, 100])
def test_special_values(lmd, oracle, fixtures_mod, metric, dims, strict):
    r, n = 32, 33
    rng = np.random.default_rng(11)
    x = rng.standard_normal((n, dims)).astype(np.float32)
    x[3] = 0.0                      # zero-norm row: COSINE -> 1.0 (distance.hpp:85-87)
    x[4] = -0.0
    x[5, ::2] = -0.0                # signed zeros inside an ordinary row
    x[6] = 1e-30                    # squared norm underflows to 0 in fp32 -> the zero-norm rule again
    x[7] = 1e19                     # squared terms overflow: L2 -> +Inf, COSINE norms -> Inf
    x[8, 5] = np.inf
    x[9, 7] = -np.inf
    x[10, 0] = np.nan
    x[11] = x[12]                   # an exact duplicate (a distance tie)
    q = rng.standard_normal((12, dims)).astype(np.float32)
    q[0] = 0.0                      # all-zero query
    q[1] = -0.0
    q[2] = 1e-30
    q[3] = x[12]                    # distance 0 to rows 11 and 12
    q[4, 3] = np.inf
    q[5, 3] = np.nan
    q[6] = 1e19
    blocks = _star_index(fixtures_mod, x, r)
    lay = fixtures_mod.layout(dims, r)
    k = L = 40
    with lmd.Store(dims, r, metric=METRIC[metric], capacity=n) as s:
        s.upload(blocks)
        s.set_entry(0)
        ids, dist, cnt, _ = s.search(q, k, L, flags=lmd.SEARCH_STRICT_FP if strict else 0)
    o_ids, o_dist, o_cnt, _ = oracle.search(blocks, lay["block_bytes"], dims, r, q, k, L, metric=METRIC[metric],
                                            entry_row_id=0)
    assert np.array_equal(cnt, o_cnt) and (cnt == n).all()
    for qi in range(q.shape[0]):
        got = dict(zip(ids[qi, :n].tolist(), dist[qi, :n].tolist()))
        want = dict(zip(o_ids[qi, :n].tolist(), o_dist[qi, :n].tolist()))
        assert set(got) == set(want) == set(range(n))
        for row in range(n):
            g, w = np.float32(got[row]), np.float32(want[row])
            if np.isnan(w) or np.isnan(g):
                assert np.isnan(w) and np.isnan(g), (metric, qi, row, g, w)
            elif np.isinf(w) or strict:
                assert g == w, (metric, qi, row, g, w)
            else:
                # 1e-5 relative (north_star); COSINE = 1 - c with c near 1 carries the rounding of c: a few ulps of 1.0
                # IP = -sum(q_i x_i): cancellation makes the result small against its terms, so the bound is relative to
                # |q| |x| (what the summation order can move), not to the result
                atol = 4e-7 if metric == "COSINE" else 1e-12
                if metric == "IP":
                    atol = 2e-6 * float(np.linalg.norm(q[qi].astype(np.float64)) * np.linalg.norm(x[row].astype(np.float64)))
                assert abs(float(g) - float(w)) <= 1e-5 * abs(float(w)) + atol, (metric, qi, row, g, w)
        # rows with ordinary (finite, non-NaN) distances come back in the oracle's order; NaN keys have no order in the
        # reference either (every comparison with NaN is false, core/Searcher.cpp:35-37)
        fin_g = [i for i in ids[qi, :n] if np.isfinite(got[i])]
        fin_o = [i for i in o_ids[qi, :n] if np.isfinite(want[i])]
        if strict and not any(np.isnan(v) for v in want.values()):
            assert fin_g == fin_o, (metric, qi)
    if metric == "COSINE":
        zero_rows = [3, 4, 6]
        for qi in range(q.shape[0]):
            got = dict(zip(ids[qi, :n].tolist(), dist[qi, :n].tolist()))
            for row in zero_rows:
                assert got[row] == 1.0
        for qi in (0, 1, 2):  # zero-norm query: every row at distance exactly 1.0
            assert (dist[qi, :n][~np.isnan(dist[qi, :n])] == 1.0).all()


# ---- int8 edge codec (row f4, parity unpinned: restatement only) ------------------------------------------------------
@pytest.mark.parametrize("n,dims,r,kind,L", [(3000, 128, 32, "clustered", 64), (2000, 96, 32, "iid", 100),
                                             (2000, 128, 64, "clustered", 128), (1500, 100, 24, "iid", 48),
                                             (1500, 768, 32, "clustered", 64)])
def test_int8_route_parity(lmd, oracle, make_index, fixtures_mod, n, dims, r, kind, L):
    ix = make_index(n, dims, r, metric="L2", kind=kind, nq=96)
    k = 10
    with _store(lmd, ix, dims, r, "L2") as s:
        s.upload(ix["blocks_np"])
        s.set_entry(0)
        s.set_tuning("tie_slots", 64)  # oracle-exact ties
        ids, dist, cnt, st, qs = s.search(ix["queries"], k, L, route=lmd.ROUTE_INT8, qstats=True)
        off, scale = s.get_codec()
        ex = s.search(ix["queries"], k, L, flags=lmd.SEARCH_NO_PREFILTER)
    assert scale > 0 and st.launches == 2 and st.reranked > 0
    oracle.set_codec(off, scale)
    o_ids, o_dist, o_cnt, o_st = oracle.search(ix["blocks_np"], ix["block_bytes"], dims, r, ix["queries"], k, L,
                                                metric=1, entry_row_id=0, route=2)
    # routing distances are integers under a square root: the traversal is bit-exact, only the rerank is floating point
    check_parity(ids, dist, cnt, o_ids, o_dist, o_cnt, what="int8 route D=%d" % dims)
    assert np.array_equal(qs[:, 0].astype(np.int64), o_st[:, 0].astype(np.int64))
    assert st.tie_overflows == 0
    # property: the codec finds (nearly) what exact routing finds
    truth = fixtures_mod.brute_force_topk(ix["x"], ix["queries"], k, "L2", device="cuda")
    r_int8, r_exact = fixtures_mod.recall_at_k(ids, truth), fixtures_mod.recall_at_k(ex[0], truth)
    assert r_int8 >= r_exact - 0.03, (r_int8, r_exact)


def test_int8_codec_metrics(lmd, make_index):
    """int8 ROUTING is L2-only; the code sidecar (pre-filter of the exact route) exists for all three metrics."""
    for metric in ("COSINE", "IP"):
        ix = make_index(1200, 33, 16, metric=metric, kind="clustered", nq=8)
        with _store(lmd, ix, 33, 16, metric) as s:
            s.upload(ix["blocks_np"])
            s.set_entry(0)
            with pytest.raises(lmd.LmdError) as e:
                s.search(ix["queries"], 5, 32, route=lmd.ROUTE_INT8)
            assert e.value.status == lmd.ERR_UNSUPPORTED
            s.enable_codec()
            with pytest.raises(lmd.LmdError) as e:
                s.search(ix["queries"], 5, 32, route=lmd.ROUTE_INT8)
            assert e.value.status == lmd.ERR_UNSUPPORTED


@pytest.mark.parametrize("n,dims,r,kind,L,metric", [
    (6000, 128, 32, "clustered", 64, "L2"), (4000, 96, 32, "clustered", 128, "L2"), (3000, 128, 64, "iid", 32, "L2"),
    (3000, 100, 24, "clustered", 256, "L2"), (2000, 768, 32, "clustered", 64, "L2"),
    (5000, 128, 32, "clustered", 64, "COSINE"), (2000, 768, 32, "clustered", 80, "COSINE"),
    (3000, 100, 24, "iid", 48, "COSINE"), (3000, 96, 32, "clustered", 128, "COSINE"),
    (5000, 128, 32, "clustered", 64, "IP"), (2000, 768, 32, "clustered", 48, "IP"), (3000, 100, 24, "iid", 64, "IP")])
def test_prefilter_changes_nothing_but_the_bytes(lmd, oracle, make_index, n, dims, r, kind, L, metric):
    """Exact routing on a codec store: the int8 lower bound only skips rows that the list would have turned away, so ids,
    distance BITS and per-query expansion / evaluation counts equal the plain traversal's - with fewer rows fetched.
    L2: bound from the grid distance; COSINE / IP: from the grid dot product (zero-offset grid; unit rows for COSINE)."""
    ix = make_index(n, dims, r, metric=metric, kind=kind, nq=128)
    k = 10
    with _store(lmd, ix, dims, r, metric) as s:
        s.upload(ix["blocks_np"])
        s.set_entry(0)
        plain = s.search(ix["queries"], k, L, qstats=True)
        s.enable_codec()
        auto = s.search(ix["queries"], k, L, qstats=True)  # 128 queries: a small batch keeps the plain latency kernels
        assert np.array_equal(auto[0], plain[0]) and np.array_equal(auto[1].view(np.uint32), plain[1].view(np.uint32))
        if dims in (96, 128):  # the shapes that have small-batch kernels
            assert auto[3].exact_fetches == plain[3].exact_fetches
        s.set_tuning("prefilter", 2)                       # ... unless the pre-filter is forced for every batch size
        pf = s.search(ix["queries"], k, L, qstats=True)
        off = s.search(ix["queries"], k, L, qstats=True, flags=lmd.SEARCH_NO_PREFILTER)
        # ... also with a visited filter that forgets, and with the one-by-one admission path
        s.set_tuning("filter_log2", 7)
        s.set_tuning("merge_min", 33)
        pf2 = s.search(ix["queries"], k, L)
        # the codec follows the store: after new blocks arrive the sidecar is rebuilt before the next search
        s.set_tuning("filter_log2", 0)
        s.set_tuning("merge_min", 0)
        s.apply(np.array([7], dtype=np.int64), ix["blocks_np"][7:8])
        pf3 = s.search(ix["queries"], k, L, qstats=True)
    for o in (pf, off, pf3):
        assert np.array_equal(o[0], plain[0]) and np.array_equal(o[1].view(np.uint32), plain[1].view(np.uint32))
        assert np.array_equal(o[2], plain[2]) and np.array_equal(o[4][:, 0], plain[4][:, 0])  # expansions per query
        # evaluations per query too, unless a visited filter forgot entries: the code-route kernels keep the query's int8
        # image in shared memory and may get a smaller filter than the plain ones (re-evaluations never change results)
        if o[3].filter_evictions == 0 and plain[3].filter_evictions == 0:
            assert np.array_equal(o[4], plain[4])
    assert np.array_equal(pf2[0], plain[0]) and np.array_equal(pf2[1].view(np.uint32), plain[1].view(np.uint32))
    assert plain[3].exact_fetches == plain[3].distance_evals == off[3].exact_fetches
    if pf[3].filter_evictions == 0 and plain[3].filter_evictions == 0:
        assert pf[3].distance_evals == plain[3].distance_evals
    assert pf[3].exact_fetches <= plain[3].exact_fetches
    if kind == "clustered":
        assert pf[3].exact_fetches < plain[3].exact_fetches
    assert pf[3].algorithmic_bytes != plain[3].algorithmic_bytes


@pytest.mark.parametrize("metric", ["COSINE", "IP"])
def test_cosine_prefilter_with_unnormalised_rows(lmd, fixtures_mod, metric):
    """The cosine lower bound divides by || q || || x ||: rows and queries of any length, a zero row and a zero query must
    come out exactly as without the pre-filter (core/distance.hpp:85-98: zero norm -> distance 1.0)."""
    n, dims, r, k, L = 4000, 128, 32, 10, 64
    ix = fixtures_mod.build_index(n, dims, r, metric=metric, kind="clustered", device="cpu")
    rng = np.random.default_rng(11)
    x = np.asarray(ix["x"], dtype=np.float32) * rng.uniform(0.05, 20.0, size=(n, 1)).astype(np.float32)
    x[17] = 0.0
    x[1234] *= 1e-20  # denormal-sized row
    blocks = fixtures_mod.pack_blocks(x, ix["nbrs"], ix["counts"], r).numpy()
    q = np.ascontiguousarray(np.asarray(fixtures_mod.make_dataset(128, dims, kind="clustered", seed=5, normalize=True),
                                        dtype=np.float32) * rng.uniform(0.01, 100.0, size=(128, 1)).astype(np.float32))
    q[3] = 0.0
    with lmd.Store(dims, r, metric=METRIC[metric], capacity=n) as s:
        s.upload(blocks)
        s.set_entry(0)
        plain = s.search(q, k, L, qstats=True)
        s.enable_codec()
        s.set_tuning("prefilter", 2)
        pf = s.search(q, k, L, qstats=True)
        off = s.search(q, k, L, qstats=True, flags=lmd.SEARCH_NO_PREFILTER)
    for o in (pf, off):
        assert np.array_equal(o[0], plain[0]) and np.array_equal(o[1].view(np.uint32), plain[1].view(np.uint32))
        assert np.array_equal(o[2], plain[2]) and np.array_equal(o[4][:, 0], plain[4][:, 0])
    assert pf[3].exact_fetches < plain[3].exact_fetches


# ---- routing sidecar follows the store ----------------------------------------------------------------------------------
def test_ternary_sidecar_follows_apply_and_remove(lmd, oracle, fixtures_mod):
    n, dims, r = 2000, 200, 24
    ix = fixtures_mod.build_index(n, dims, r, metric="COSINE", kind="iid", device="cpu")
    blocks = ix["blocks"].numpy()
    bb = ix["layout"]["block_bytes"]
    q = fixtures_mod.make_dataset(48, dims, kind="iid", seed=77, normalize=True)
    with lmd.Store(dims, r, metric=lmd.METRIC_COSINE, capacity=n) as s:
        s.upload(blocks[:1500])
        s.set_entry(0)
        a = s.search(q, 10, 64, route=lmd.ROUTE_TERNARY)
        o = oracle.search(blocks[:1500], bb, dims, r, q, 10, 64, metric=2, entry_row_id=0, route=1)
        check_parity(a[0], a[1], a[2], o[0], o[1], o[2], what="ternary, first 1500 rows")
        s.apply(np.arange(1500, n, dtype=np.int64), blocks[1500:])
        b = s.search(q, 10, 64, route=lmd.ROUTE_TERNARY)
        o = oracle.search(blocks, bb, dims, r, q, 10, 64, metric=2, entry_row_id=0, route=1)
        check_parity(b[0], b[1], b[2], o[0], o[1], o[2], what="ternary, after apply")
        gone = np.arange(100, 160, dtype=np.int64)
        s.remove(gone)
        c = s.search(q, 10, 64, route=lmd.ROUTE_TERNARY)
        assert not np.isin(c[0], gone).any()


def test_release_blocks(lmd, make_index):
    ix = make_index(3000, 128, 32, metric="COSINE", kind="clustered", nq=64)
    with _store(lmd, ix, 128, 32, "COSINE") as s:
        s.upload(ix["blocks_np"])
        s.set_entry(0)
        a = s.search(ix["queries"], 10, 64)
        t = s.search(ix["queries"], 10, 64, route=lmd.ROUTE_TERNARY)
        before = s.device_bytes
        s.release_blocks()
        assert s.device_bytes <= before - 3000 * s.block_bytes
        b = s.search(ix["queries"], 10, 64)
        t2 = s.search(ix["queries"], 10, 64, route=lmd.ROUTE_TERNARY)
        assert np.array_equal(a[0], b[0]) and np.array_equal(a[1], b[1])
        assert np.array_equal(t[0], t2[0]) and np.array_equal(t[1], t2[1])
        for call in (lambda: s.download_block(5), lambda: s.upload(ix["blocks_np"][:1]),
                     lambda: s.apply(np.array([1], dtype=np.int64), ix["blocks_np"][:1])):
            with pytest.raises(lmd.LmdError) as e:
                call()
            assert e.value.status == lmd.ERR_STATE


# ---- per-caller contexts ------------------------------------------------------------------------------------------------
def test_contexts_search_side_by_side(lmd, make_index):
    """16 host threads, one lmd_ctx each, point queries and small batches against one store, while another thread keeps
    applying (unchanged) blocks: every answer equals the single-threaded one."""
    ix = make_index(3000, 128, 32, metric="L2", kind="clustered", nq=96)
    q = ix["queries"]
    with _store(lmd, ix, 128, 32, "L2") as s:
        s.upload(ix["blocks_np"])
        s.set_entry(0)
        ref = s.search(q, 10, 64)
        errors = []
        stop = threading.Event()

        def worker(t):
            try:
                with s.context() as c:
                    for rep in range(30):
                        i = (t * 7 + rep) % 90
                        ids, dist, cnt, _ = c.search(q[i:i + 1], 10, 64)
                        assert np.array_equal(ids, ref[0][i:i + 1]) and np.array_equal(dist, ref[1][i:i + 1])
                        ids, dist, cnt, _ = c.search(q[i:i + 6], 10, 64, flags=lmd.SEARCH_SEPARATE_RERANK)
                        assert np.array_equal(ids, ref[0][i:i + 6])
            except Exception as e:  # noqa: BLE001
                errors.append(repr(e))

        def writer():
            while not stop.is_set():
                s.apply(np.array([11, 12], dtype=np.int64), ix["blocks_np"][11:13])

        th = [threading.Thread(target=worker, args=(t,)) for t in range(16)]
        w = threading.Thread(target=writer)
        w.start()
        for t in th:
            t.start()
        for t in th:
            t.join()
        stop.set()
        w.join()
        assert not errors, errors[:3]


def test_scratch_capacities_are_tracked_per_array(lmd, make_index):
    """a k = 10 batch followed by a LARGER batch with k = 1 and per-query statistics (2 nq counters whatever
    k is) used to overrun the statistics scratch; pageable buffers, so the staged path is the one under test."""
    ix = make_index(3000, 128, 32, metric="L2", kind="clustered", nq=96)
    rng = np.random.default_rng(1)
    with _store(lmd, ix, 128, 32, "L2") as s:
        s.upload(ix["blocks_np"])
        s.set_entry(0)
        small = ix["queries"][rng.integers(0, 96, 1000)]
        big = ix["queries"][rng.integers(0, 96, 8000)]
        a = s.search(small, 10, 32, qstats=True)
        b = s.search(big, 1, 32, qstats=True)
        c = s.search(big[:4000], 1, 32, qstats=True)
        assert int(b[4][:, 0].sum()) == b[3].expansions and int(b[4][:, 1].sum()) == b[3].distance_evals
        assert np.array_equal(b[4][:4000], c[4]) and np.array_equal(b[0][:4000], c[0])
        assert int(a[4][:, 0].sum()) == a[3].expansions


# ---- rerank as a tensor-core contraction (row a3): tcgen05 kind::tf32, split fp32, bulk-copy staging ---------------------
@pytest.mark.parametrize("metric,dims,r,L,kind", [("COSINE", 768, 32, 80, "clustered"), ("L2", 128, 32, 64, "iid"),
                                                  ("IP", 100, 24, 200, "iid"), ("L2", 768, 32, 256, "clustered"),
                                                  ("COSINE", 128, 32, 300, "iid"), ("L2", 256, 16, 20, "iid")])
def test_tensor_rerank_matches_the_fma_rerank(lmd, make_index, metric, dims, r, L, kind):
    """LMD_SEARCH_TENSOR_RERANK against rerank_kernel on the same candidate lists: same rows, distances within 1e-5
    relative (IP: relative to |q| |x|, the quantity its summation error scales with)."""
    ix = make_index(1500, dims, r, metric=metric, kind=kind, nq=200)
    k = 10
    with _store(lmd, ix, dims, r, metric) as s:
        s.upload(ix["blocks_np"])
        s.set_entry(0)
        fma = s.search(ix["queries"], k, L, flags=lmd.SEARCH_SEPARATE_RERANK)
        tc = s.search(ix["queries"], k, L, flags=lmd.SEARCH_TENSOR_RERANK)
        tc2 = s.search(ix["queries"][:3], k, L, flags=lmd.SEARCH_TENSOR_RERANK)       # fewer queries than SMs
        if metric != "L2":
            tt = s.search(ix["queries"], k, L, route=lmd.ROUTE_TERNARY, flags=lmd.SEARCH_TENSOR_RERANK)
            tf = s.search(ix["queries"], k, L, route=lmd.ROUTE_TERNARY)
    assert tc[3].launches == 2 and tc[3].reranked == fma[3].reranked
    assert np.array_equal(tc[2], fma[2])
    if metric == "IP":
        x = np.asarray(ix["x"])
        scale = np.linalg.norm(ix["queries"], axis=1)[:, None] * np.linalg.norm(x, axis=1).max()
        assert (np.abs(tc[1] - fma[1]) <= 1e-5 * scale).all()
        same = (tc[0] == fma[0]) | (np.abs(tc[1] - fma[1]) <= 1e-5 * scale)
        assert same.all()
    else:
        check_parity(tc[0], tc[1], tc[2], fma[0], fma[1], fma[2], what="tensor rerank %s D=%d L=%d" % (metric, dims, L))
        if metric != "L2":
            check_parity(tt[0], tt[1], tt[2], tf[0], tf[1], tf[2], what="ternary route + tensor rerank")
    assert np.array_equal(tc2[0], tc[0][:3])


def test_tensor_rerank_cancellation_bound(lmd, make_index):
    """SURVEY.md section 7 hard part 4: the contraction needs |q|^2 + |x|^2 - 2 q.x, which cancels for near-duplicate
    rows. The squared distance is off by at most ~2^-20 (|q|^2 + |x|^2) - fine on the test distributions (checked above),
    not relative to a distance that goes to 0. This pins the absolute bound (and is why the kernel is not the default)."""
    dims, r = 128, 32
    ix = make_index(1500, dims, r, metric="L2", kind="iid", nq=64)
    x = np.asarray(ix["x"])
    rng = np.random.default_rng(3)
    q = (x[rng.integers(0, 1500, 64)] + 1e-3 * rng.standard_normal((64, dims))).astype(np.float32)
    with _store(lmd, ix, dims, r, "L2") as s:
        s.upload(ix["blocks_np"])
        s.set_entry(0)
        fma = s.search(q, 10, 64, flags=lmd.SEARCH_SEPARATE_RERANK)
        tc = s.search(q, 10, 64, flags=lmd.SEARCH_TENSOR_RERANK)
    q2 = (q.astype(np.float64) ** 2).sum(1)
    shared = {}
    for qi in range(64):
        f = dict(zip(fma[0][qi].tolist(), fma[1][qi].tolist()))
        for i, d in zip(tc[0][qi].tolist(), tc[1][qi].tolist()):
            if i in f:
                assert abs(d * d - f[i] * f[i]) <= 2.0 ** -20 * (2 * q2[qi] + 1.0), (qi, i, d, f[i])
                shared[qi] = shared.get(qi, 0) + 1
    assert min(shared.values()) >= 8


def test_tensor_rerank_rejects_what_it_cannot_tile(lmd, make_index):
    ix = make_index(1200, 96, 32, metric="L2", kind="clustered", nq=8)
    with _store(lmd, ix, 96, 32, "L2") as s:
        s.upload(ix["blocks_np"])
        s.set_entry(0)
        with pytest.raises(lmd.LmdError) as e:
            s.search(ix["queries"], 5, 32, flags=lmd.SEARCH_TENSOR_RERANK)
        assert e.value.status == lmd.ERR_UNSUPPORTED
        with pytest.raises(lmd.LmdError):
            s.search(ix["queries"], 5, 32, flags=lmd.SEARCH_TENSOR_RERANK | lmd.SEARCH_STRICT_FP)


# ---- a neighbour row that names the same node twice (all visited-filter kinds, all routes) ----------------------------
@pytest.mark.parametrize("metric,route", [("L2", "exact"), ("L2", "prefilter"), ("L2", "int8"), ("COSINE", "ternary")])
def test_duplicate_neighbour_ids_in_a_row(lmd, oracle, fixtures_mod, metric, route):
    """The reference's visited set drops the second occurrence (core/Searcher.cpp:135-139). The atomic filters do the same;
    the direct-mapped filter (mode 2) has no atomics and needs the in-warp match: no id may come back twice, the expansion
    counts stay the oracle's."""
    n, dims, r = 1500, 128, 32
    ix = fixtures_mod.build_index(n, dims, r, metric=metric, device="cpu", kind="iid")
    blocks = ix["blocks"].numpy().copy()
    lay = ix["layout"]
    ids_view = blocks[:, lay["ids_off"]:lay["ids_off"] + 8 * r].view(np.int64)
    cnt_view = blocks[:, 0:2].view(np.uint16)
    rng = np.random.default_rng(6)
    for node in range(n):  # every row: two of its entries repeat another two
        c = int(cnt_view[node, 0])
        if c >= 6:
            a, b = rng.choice(c, 2, replace=False)
            ids_view[node, a] = ids_view[node, b]
            ids_view[node, (a + 1) % c] = ids_view[node, (b + 3) % c]
    if metric == "COSINE":  # the planes of a repeated entry must be the repeated node's (ternary routing reads them)
        blocks = fixtures_mod.pack_blocks(ix["x"], np.where(np.arange(r)[None, :] < cnt_view.astype(np.int64),
                                                            ids_view, -1), cnt_view[:, 0].astype(np.int64), r).numpy()
    q = fixtures_mod.make_dataset(64, dims, kind="iid", seed=46, normalize=(metric == "COSINE"))
    k, L = 10, 48
    kw = {"route": {"ternary": lmd.ROUTE_TERNARY, "int8": lmd.ROUTE_INT8}.get(route, lmd.ROUTE_EXACT)}
    o_route = {"ternary": 1, "int8": 2}.get(route, 0)
    with lmd.Store(dims, r, metric=METRIC[metric], capacity=n) as s:
        s.upload(blocks)
        s.set_entry(0)
        s.set_tuning("tie_slots", 64)
        if route in ("prefilter", "int8"):
            s.enable_codec()
            s.set_tuning("prefilter", 2)
            oracle.set_codec(*s.get_codec())
        o = oracle.search(blocks, lay["block_bytes"], dims, r, q, k, L, metric=METRIC[metric], entry_row_id=0, route=o_route)
        s.set_tuning("filter_entries", 1000)  # mode 3: modulo-mapped table, any size
        for mode in (0, 1, 2, 3):
            s.set_tuning("filter_mode", mode)
            ids, dist, cnt, st, qs = s.search(q, k, L, qstats=True, **kw)
            for row in ids:
                v = row[row >= 0]
                assert len(set(v.tolist())) == len(v), (mode, row)
            check_parity(ids, dist, cnt, o[0], o[1], o[2], what="duplicate neighbours, filter mode %d" % mode)
            assert np.array_equal(qs[:, 0].astype(np.int64), o[3][:, 0].astype(np.int64)), mode
