"""-m gpu: the CUDA path (through the C ABI, include/lmd_b200.h) against the CPU oracle on the same graph and
queries. Oracle = oracle/liblmd_oracle.so (C restatement, pinned against the reference's own sources by
tests/test_oracle.py and tests/golden/)."""
import numpy as np
import pytest

from parity import check_parity, rel_close

pytestmark = pytest.mark.gpu

CASES = [
    # n, dims, r, metric, kind, k, L
    (3000, 128, 32, "L2", "clustered", 10, 64),
    (3000, 128, 32, "L2", "iid", 10, 32),
    (2000, 96, 32, "L2", "clustered", 10, 100),
    (2000, 128, 64, "L2", "clustered", 10, 128),
    (1500, 768, 32, "COSINE", "clustered", 10, 64),
    (2000, 128, 32, "IP", "iid", 10, 64),
    (2000, 128, 32, "COSINE", "iid", 5, 256),
    (1500, 100, 24, "L2", "iid", 7, 48),      # generic (non-specialised) shape, odd R
    (1200, 33, 16, "COSINE", "clustered", 10, 40),
    (2000, 128, 32, "L2", "clustered", 100, 0),  # L = 0 -> cfg.l_search (100), k = L
    (2000, 128, 32, "L2", "clustered", 1, 10),
    (2500, 128, 32, "L2", "clustered", 10, 512),
]
METRIC = {"L2": 1, "COSINE": 2, "IP": 3}


def _store(lmd, ix, dims, r, metric, **kw):
    s = lmd.Store(dims, r, metric=METRIC[metric], capacity=ix["blocks_np"].shape[0], l_search=100, **kw)
    return s


@pytest.mark.parametrize("n,dims,r,metric,kind,k,L", CASES)
def test_exact_routing_parity(lmd, oracle, make_index, n, dims, r, metric, kind, k, L):
    ix = make_index(n, dims, r, metric=metric, kind=kind, nq=96)
    with _store(lmd, ix, dims, r, metric) as s:
        s.upload(ix["blocks_np"])
        s.set_entry(0)
        ids, dist, cnt, st, qs = s.search(ix["queries"], k, L, qstats=True)
    o_ids, o_dist, o_cnt, o_st = oracle.search(ix["blocks_np"], ix["block_bytes"], dims, r, ix["queries"], k, L,
                                                metric=METRIC[metric], entry_row_id=0, cfg_l_search=100)
    ties, total = check_parity(ids, dist, cnt, o_ids, o_dist, o_cnt, what="%s n=%d D=%d" % (metric, n, dims))
    assert ties <= max(2, total // 200)
    # expansions equal the oracle's per query (same traversal), up to rounding-induced divergence on a few queries
    same = (qs[:, 0].astype(np.int64) == o_st[:, 0].astype(np.int64)).mean()
    assert same >= 0.9, "expansion counts agree on only %.2f of the queries" % same
    assert st.launches >= 1 and st.expansions == int(qs[:, 0].sum())


@pytest.mark.parametrize("metric", ["COSINE", "IP"])
@pytest.mark.parametrize("dims,r", [(128, 32), (768, 32), (96, 64), (100, 24), (200, 40), (768, 64)])
def test_ternary_routing_parity(lmd, oracle, make_index, metric, dims, r):
    n, k, L = 1500, 10, 64
    ix = make_index(n, dims, r, metric=metric, kind="iid", nq=64)
    with _store(lmd, ix, dims, r, metric) as s:
        s.upload(ix["blocks_np"])
        s.set_entry(0)
        s.set_tuning("tie_slots", 64)  # oracle-exact ties: room for every tied frontier entry (the default keeps >= 8)
        ids, dist, cnt, st = s.search(ix["queries"], k, L, route=lmd.ROUTE_TERNARY)
        s.set_tuning("tie_slots", -1)
        d_ids, d_dist, d_cnt, d_st = s.search(ix["queries"], k, L, route=lmd.ROUTE_TERNARY)
    o_ids, o_dist, o_cnt, _ = oracle.search(ix["blocks_np"], ix["block_bytes"], dims, r, ix["queries"], k, L,
                                            metric=METRIC[metric], entry_row_id=0, route=1)
    # the default (short) tie region may drop a tied frontier entry, counted in tie_overflows; without a drop the
    # results are the same ones
    if d_st.tie_overflows == 0:
        assert np.array_equal(d_ids, ids) and np.array_equal(d_dist, dist)
    # routing distances are integers (popcounts): the traversal is bit-exact, only the rerank is floating point
    check_parity(ids, dist, cnt, o_ids, o_dist, o_cnt, what="ternary %s D=%d" % (metric, dims))
    assert st.reranked > 0 and st.launches == 2
    assert st.tie_overflows == 0  # the tie region held every tied frontier entry: oracle-exact


def test_ternary_l2_is_rejected(lmd, make_index):
    ix = make_index(1200, 33, 16, metric="L2", kind="iid", nq=8)
    with _store(lmd, ix, 33, 16, "L2") as s:
        s.upload(ix["blocks_np"])
        s.set_entry(0)
        with pytest.raises(lmd.LmdError) as e:  # core/distance.hpp:236-239
            s.search(ix["queries"], 5, 32, route=lmd.ROUTE_TERNARY)
        assert e.value.status == lmd.ERR_UNSUPPORTED


def test_separate_rerank_matches_fused(lmd, make_index):
    ix = make_index(3000, 128, 32, metric="L2", kind="clustered", nq=96)
    with _store(lmd, ix, 128, 32, "L2") as s:
        s.upload(ix["blocks_np"])
        s.set_entry(0)
        a = s.search(ix["queries"], 10, 64)
        b = s.search(ix["queries"], 10, 64, flags=lmd.SEARCH_SEPARATE_RERANK)
    assert np.array_equal(a[0], b[0]) and np.array_equal(a[2], b[2])
    assert np.array_equal(a[1], b[1])  # same arithmetic in both kernels
    assert b[3].launches == 2 and b[3].reranked > 0


def test_row_id_mapping(lmd, oracle, fixtures_mod):
    """Row ids that are not slot numbers (DuckDB row_t after deletes): GraphManager.cpp:523-531 semantics."""
    n, dims, r = 2000, 128, 32
    rng = np.random.default_rng(5)
    row_ids = np.sort(rng.choice(np.arange(10, 10 * n, dtype=np.int64), size=n, replace=False))
    rng.shuffle(row_ids)
    ix = fixtures_mod.build_index(n, dims, r, device="cpu", row_ids=row_ids)
    blocks = ix["blocks"].numpy()
    q = fixtures_mod.make_dataset(64, dims, seed=43)
    entry = int(row_ids[17])
    with lmd.Store(dims, r, capacity=n) as s:
        s.upload(blocks[:1200], row_ids=row_ids[:1200])   # two uploads: the sidecar resolves ids across both
        s.upload(blocks[1200:], row_ids=row_ids[1200:])
        s.set_entry(entry)
        ids, dist, cnt, _ = s.search(q, 10, 64)
        got = s.download_block(int(row_ids[1500]))
    assert np.array_equal(got, blocks[1500])
    o = oracle.search(blocks, ix["layout"]["block_bytes"], dims, r, q, 10, 64, row_ids=row_ids, entry_row_id=entry)
    check_parity(ids, dist, cnt, o[0], o[1], o[2], what="row-id map")
    assert set(np.unique(ids[ids >= 0])).issubset(set(row_ids.tolist()))


def test_skip_rules_and_edge_cases(lmd, oracle, fixtures_mod):
    """count < stored ids, INT64_MAX sentinels inside the count, dangling row ids, isolated entry, missing entry,
    k > reachable nodes (core/GraphManager.cpp:241-254, core/Searcher.cpp:74-83,120-122,132,146-148)."""
    n, dims, r = 600, 64, 16
    ix = fixtures_mod.build_index(n, dims, r, device="cpu", kind="iid")
    blocks = ix["blocks"].numpy().copy()
    lay = ix["layout"]
    ids_view = blocks[:, lay["ids_off"]:lay["ids_off"] + 8 * r].view(np.int64)
    cnt_view = blocks[:, 0:2].view(np.uint16)
    rng = np.random.default_rng(3)
    for node in rng.choice(n, 150, replace=False):
        c = int(cnt_view[node, 0])
        what = rng.integers(0, 3)
        if what == 0 and c > 2:
            cnt_view[node, 0] = c - 2            # ids beyond count must be ignored
        elif what == 1 and c > 1:
            ids_view[node, rng.integers(0, c)] = np.iinfo(np.int64).max  # sentinel inside the count
        elif c > 0:
            ids_view[node, rng.integers(0, c)] = 10 ** 9 + int(node)      # dangling row id
    q = fixtures_mod.make_dataset(48, dims, kind="iid", seed=45)
    with lmd.Store(dims, r, capacity=n) as s:
        s.upload(blocks)
        s.set_entry(0)
        g = s.search(q, 10, 32)
        o = oracle.search(blocks, lay["block_bytes"], dims, r, q, 10, 32, entry_row_id=0)
        check_parity(g[0], g[1], g[2], o[0], o[1], o[2], what="skip rules")
        # missing entry point -> empty results, no error (Searcher.cpp:74-77)
        s.set_entry(123456789)
        g = s.search(q, 10, 32)
        assert (g[0] == -1).all() and (g[2] == 0).all()
        s.set_entry(np.iinfo(np.int64).max)
        g = s.search(q, 10, 32)
        assert (g[0] == -1).all() and (g[2] == 0).all()
        # isolated entry: only the entry itself is returned
        iso = blocks[5].copy()
        iso[0:2] = 0
        s.update(5, iso)
        s.set_entry(5)
        g = s.search(q, 10, 32)
        assert (g[2] == 1).all() and (g[0][:, 0] == 5).all() and (g[0][:, 1:] == -1).all()
        b2 = blocks.copy()
        b2[5] = iso
        o = oracle.search(b2, lay["block_bytes"], dims, r, q, 10, 32, entry_row_id=5)
        check_parity(g[0], g[1], g[2], o[0], o[1], o[2], what="isolated entry")
        # k = 0 is an error (Coordinator.cpp:72-74); empty batch is not
        with pytest.raises(lmd.LmdError):
            s.search(q, 0, 32)
        g = s.search(q[:0], 10, 32)
        assert g[0].shape == (0, 10)
    # a tiny graph: k and L exceed the number of nodes
    tiny = fixtures_mod.build_index(8, dims, 4, device="cpu", kind="iid", n_random=1, n_cand=4)
    tb = tiny["blocks"].numpy()
    with lmd.Store(dims, 4, capacity=8) as s:
        s.upload(tb)
        s.set_entry(0)
        g = s.search(q, 20, 64)
    o = oracle.search(tb, tiny["layout"]["block_bytes"], dims, 4, q, 20, 64, entry_row_id=0)
    check_parity(g[0], g[1], g[2], o[0], o[1], o[2], what="tiny graph")


def test_int8_nodes(lmd, oracle, fixtures_mod):
    """node_vector_type INT8: plain static_cast<float> per element (common/ann.hpp:52-58)."""
    n, dims, r = 1500, 128, 32
    x = fixtures_mod.make_dataset(n, dims, kind="iid", seed=44)
    xi = np.clip(np.round(x * 40), -127, 127).astype(np.int8)
    nbrs, counts = fixtures_mod.build_graph(xi.astype(np.float32), r, device="cpu")
    blocks = fixtures_mod.pack_blocks(xi, nbrs, counts, r, vec_type=2).numpy()
    lay = fixtures_mod.layout(dims, r, 2)
    q = (fixtures_mod.make_dataset(64, dims, kind="iid", seed=45) * 40).astype(np.float32)
    with lmd.Store(dims, r, vec_type=lmd.VEC_INT8, capacity=n) as s:
        assert s.block_bytes == lay["block_bytes"]
        s.upload(blocks)
        s.set_entry(0)
        g = s.search(q, 10, 64)
    o = oracle.search(blocks, lay["block_bytes"], dims, r, q, 10, 64, vec_type=2, entry_row_id=0)
    check_parity(g[0], g[1], g[2], o[0], o[1], o[2], what="int8")


@pytest.mark.parametrize("metric,route", [("L2", 0), ("COSINE", 1), ("IP", 1)])
def test_filter_size_does_not_change_results(lmd, make_index, metric, route):
    """The visited filter may forget: results must not depend on its size or kind (u32 open addressing /
    two-choice 16-bit buckets), for exact and for ternary routing (where the entry point, whose key is an exact
    distance, must never be re-admitted with a routing distance)."""
    ix = make_index(3000, 128, 32, metric=metric, kind="iid", nq=96)
    outs = []
    with _store(lmd, ix, 128, 32, metric) as s:
        s.upload(ix["blocks_np"])
        s.set_entry(0)
        for mode in (0, 1):
            s.set_tuning("filter_mode", mode)
            for lg in (6, 8, 11, 13):
                s.set_tuning("filter_log2", lg)
                outs.append(s.search(ix["queries"], 10, 128, route=route))
        s.set_tuning("filter_mode", 3)  # direct-mapped by row id modulo the table size, any size
        for entries in (512, 1000, 3504):
            s.set_tuning("filter_entries", entries)
            outs.append(s.search(ix["queries"], 10, 128, route=route))
    for o in outs[1:]:
        assert np.array_equal(o[0], outs[0][0]) and np.array_equal(o[1], outs[0][1])
    assert outs[0][3].filter_evictions > 0 and outs[4][3].filter_evictions > 0  # the 64-entry filters really did forget
    assert outs[0][3].distance_evals >= outs[3][3].distance_evals
    assert outs[4][3].distance_evals >= outs[7][3].distance_evals
    assert outs[7][3].filter_evictions <= outs[3][3].filter_evictions


def test_batch_merge_equals_one_by_one(lmd, oracle, make_index):
    """list_merge_batch() (the default for >= 5 admitted candidates) and the one-by-one insert loop must build the
    same list: bit-identical ids, distances and per-query expansion / evaluation counts."""
    for (n, dims, r, metric, kind, L) in [(3000, 128, 32, "L2", "clustered", 64), (2000, 96, 64, "L2", "iid", 100),
                                          (1500, 768, 32, "COSINE", "clustered", 200), (1500, 100, 24, "IP", "iid", 33)]:
        ix = make_index(n, dims, r, metric=metric, kind=kind, nq=96)
        outs = []
        with _store(lmd, ix, dims, r, metric) as s:
            s.upload(ix["blocks_np"])
            s.set_entry(0)
            for mm in (1, 5, 33):  # always merge / default / never merge
                s.set_tuning("merge_min", mm)
                outs.append(s.search(ix["queries"], 10, L, qstats=True))
        for o in outs[1:]:
            assert np.array_equal(o[0], outs[0][0]) and np.array_equal(o[1], outs[0][1])
            assert np.array_equal(o[4], outs[0][4])
        assert outs[2][3].merge_fallbacks == 0


def test_duplicate_vectors_exact_ties(lmd, oracle, fixtures_mod):
    """A data set where every vector occurs four times: exact fp32 distance ties everywhere, including at the L-th
    place of the list. The batch merge must detect them and replay one by one, so that the results stay those of the
    oracle's (distance, row id) order."""
    n, dims, r = 1600, 64, 16
    base = fixtures_mod.make_dataset(n // 4, dims, kind="iid", seed=46)
    x = np.concatenate([base, base, base, base])
    rng = np.random.default_rng(9)
    x = x[rng.permutation(n)]
    nbrs, counts = fixtures_mod.build_graph(x, r, device="cpu")
    blocks = fixtures_mod.pack_blocks(x, nbrs, counts, r).numpy()
    lay = fixtures_mod.layout(dims, r)
    q = fixtures_mod.make_dataset(64, dims, kind="iid", seed=47)
    with lmd.Store(dims, r, capacity=n) as s:
        s.upload(blocks)
        s.set_entry(0)
        s.set_tuning("tie_slots", 32)
        g = s.search(q, 10, 24, qstats=True)
        o = oracle.search(blocks, lay["block_bytes"], dims, r, q, 10, 24, entry_row_id=0)
        check_parity(g[0], g[1], g[2], o[0], o[1], o[2], what="duplicate vectors")
        assert np.array_equal(g[0], o[0])  # identical even inside the ties: same (distance, id) order
        assert g[3].merge_fallbacks > 0 and g[3].tie_overflows == 0
        assert np.array_equal(g[4][:, 0].astype(np.int64), o[3][:, 0].astype(np.int64))  # same expansions per query


def test_device_resident_io(lmd, make_index):
    """queries / outputs already in HBM (torch tensors): no staging copies."""
    import torch
    ix = make_index(3000, 128, 32, metric="L2", kind="clustered", nq=96)
    nq, k = 96, 10
    with _store(lmd, ix, 128, 32, "L2") as s:
        s.upload(torch.from_numpy(ix["blocks_np"]).cuda())  # device-resident blocks
        s.set_entry(0)
        h = s.search(ix["queries"], k, 64)
        dq = torch.from_numpy(ix["queries"]).cuda()
        d_ids = torch.empty((nq, k), dtype=torch.int64, device="cuda")
        d_dist = torch.empty((nq, k), dtype=torch.float32, device="cuda")
        d_cnt = torch.empty(nq, dtype=torch.int32, device="cuda")
        st = s.search_raw(dq, nq, k, 64, out_ids=d_ids, out_dist=d_dist, out_count=d_cnt)
        torch.cuda.synchronize()
    assert np.array_equal(d_ids.cpu().numpy(), h[0]) and np.array_equal(d_dist.cpu().numpy(), h[1])
    assert st.h2d_ms == 0 and st.d2h_ms == 0


def test_pinned_host_io_is_zero_copy(lmd, make_index):
    """Pinned host buffers are read / written in place by the kernels (no staging copies), in every mix with device
    buffers, for both routes; pageable outputs next to pinned ones are rejected."""
    import torch
    ix = make_index(3000, 128, 32, metric="COSINE", kind="clustered", nq=96)
    nq, k = 96, 10
    with _store(lmd, ix, 128, 32, "COSINE") as s:
        s.upload(ix["blocks_np"])
        s.set_entry(0)
        for route in (0, 1):
            ref = s.search(ix["queries"], k, 64, route=route)  # pageable numpy arrays: staged
            assert ref[3].h2d_ms > 0 and ref[3].d2h_ms > 0
            hq = torch.from_numpy(ix["queries"]).pin_memory()
            h_ids = torch.full((nq, k), -7, dtype=torch.int64).pin_memory()
            h_dist = torch.zeros((nq, k), dtype=torch.float32).pin_memory()
            h_cnt = torch.zeros(nq, dtype=torch.int32).pin_memory()
            st = s.search_raw(hq, nq, k, 64, route=route, out_ids=h_ids, out_dist=h_dist, out_count=h_cnt)
            assert st.h2d_ms == 0 and st.d2h_ms == 0
            assert np.array_equal(h_ids.numpy(), ref[0]) and np.array_equal(h_dist.numpy(), ref[1])
            assert np.array_equal(h_cnt.numpy().astype(np.uint32), ref[2])
            # pinned queries, device outputs
            d_ids = torch.empty((nq, k), dtype=torch.int64, device="cuda")
            s.search_raw(hq, nq, k, 64, route=route, out_ids=d_ids)
            assert np.array_equal(d_ids.cpu().numpy(), ref[0])
            # pageable queries, pinned outputs
            h_ids.fill_(-7)
            st = s.search_raw(ix["queries"], nq, k, 64, route=route, out_ids=h_ids)
            assert st.h2d_ms > 0 and st.d2h_ms == 0 and np.array_equal(h_ids.numpy(), ref[0])
        with pytest.raises(lmd.LmdError) as e:
            s.search_raw(hq, nq, k, 64, out_ids=h_ids, out_dist=np.zeros((nq, k), dtype=np.float32))
        assert e.value.status == lmd.ERR_INVALID_ARG


def test_merge_topk(lmd):
    rng = np.random.default_rng(0)
    G, nq, k = 8, 257, 10
    dist = np.sort(rng.random((G, nq, k), dtype=np.float32), axis=2)
    dist[:, :, 3] = dist[:, :, 2]  # ties inside a shard
    dist[1] = dist[0]              # ties across shards -> id order decides
    ids = rng.permutation(G * nq * k).astype(np.int64).reshape(G, nq, k)
    # make each shard's list ascending by (dist, id)
    for g in range(G):
        for q in range(nq):
            o = np.lexsort((ids[g, q], dist[g, q]))
            ids[g, q], dist[g, q] = ids[g, q][o], dist[g, q][o]
    ids[2, :, 7:] = -1  # a shard that found fewer than k
    out_i = np.empty((nq, k), dtype=np.int64)
    out_d = np.empty((nq, k), dtype=np.float32)
    lmd.merge_topk(ids, dist, G, nq, k, out_i, out_d)
    for q in range(nq):
        cand = [(float(dist[g, q, j]), int(ids[g, q, j])) for g in range(G) for j in range(k) if ids[g, q, j] >= 0]
        cand.sort()
        assert [c[1] for c in cand[:k]] == out_i[q].tolist()
        assert np.array_equal(np.array([c[0] for c in cand[:k]], dtype=np.float32), out_d[q])
    # the stream-ordered form (device buffers, caller's stream, no synchronisation inside) gives the same lists
    import torch
    st = torch.cuda.Stream()
    with torch.cuda.stream(st):
        d_ids, d_dist = torch.from_numpy(ids).cuda(), torch.from_numpy(dist).cuda()
        o_i = torch.empty((nq, k), dtype=torch.int64, device="cuda")
        o_d = torch.empty((nq, k), dtype=torch.float32, device="cuda")
        lmd.merge_topk_async(d_ids, d_dist, G, nq, k, o_i, o_d, cuda_stream=st.cuda_stream)
    st.synchronize()
    assert np.array_equal(o_i.cpu().numpy(), out_i) and np.array_equal(o_d.cpu().numpy(), out_d)
    with pytest.raises(lmd.LmdError):  # host memory is refused rather than staged
        lmd.merge_topk_async(ids, dist, G, nq, k, out_i, out_d)


def test_recall_matches_oracle_10k(lmd, oracle, fixtures_mod):
    """BASELINE config 1 shape (10k x 128 L2, R=32, L=64, 1k queries, k=10): recall@10 equal to the oracle's."""
    import torch
    n, dims, r, nq, k, L = 10000, 128, 32, 1000, 10, 64
    ix = fixtures_mod.build_index(n, dims, r, device="cuda")
    blocks = ix["blocks"].numpy()
    q = fixtures_mod.make_dataset(nq, dims, seed=43)
    truth = fixtures_mod.brute_force_topk(ix["x"], q, k, "L2", device="cuda")
    with lmd.Store(dims, r, capacity=n, l_search=L) as s:
        s.upload(blocks)
        s.set_entry(0)
        g = s.search(q, k)  # L = 0 -> cfg.l_search
    o = oracle.search(blocks, ix["layout"]["block_bytes"], dims, r, q, k, 0, entry_row_id=0, cfg_l_search=L)
    check_parity(g[0], g[1], g[2], o[0], o[1], o[2], what="10k config")
    rg, ro = fixtures_mod.recall_at_k(g[0], truth), fixtures_mod.recall_at_k(o[0], truth)
    assert rg == ro, (rg, ro)
    assert rg >= 0.9, rg
    del torch


# ---- live mirror: lmd_store_apply / lmd_store_remove ---------------------------------
@pytest.mark.parametrize("sparse_ids", [False, True])
def test_apply_and_remove_track_the_host_blocks(lmd, oracle, make_index, fixtures_mod, sparse_ids):
    """The host's storage changes between searches (new rows, re-pruned neighbour lists, deletions); after
    apply()/remove() the mirror must search exactly like the CPU reference semantics on the host's CURRENT blocks."""
    n, dims, r, k, L = 2400, 64, 16, 10, 48
    rng = np.random.default_rng(5)
    rid = None
    if sparse_ids:
        rid = np.sort(rng.choice(np.arange(5, 20 * n, dtype=np.int64), size=n, replace=False))
        rid[0], rid[7] = rid[7], rid[0]  # not monotone either
    ix = make_index(n, dims, r, metric="L2", kind="clustered", nq=64, row_ids=rid)
    rid = np.arange(n, dtype=np.int64) if rid is None else rid
    blocks, bb, q, lay = ix["blocks_np"], ix["block_bytes"], ix["queries"], ix["layout"]
    n0 = 1500  # the index as first mirrored: later rows are unknown ids inside the early blocks -> skipped

    def check(s, live, blk, entry, what):
        ids, dist, cnt, _ = s.search(q, k, L)
        o = oracle.search(blk[live], bb, dims, r, q, k, L, row_ids=rid[live], entry_row_id=int(entry))
        check_parity(ids, dist, cnt, o[0], o[1], o[2], what=what)
        return ids

    with lmd.Store(dims, r, capacity=n + 8, l_search=100) as s:
        s.upload(blocks[:n0], row_ids=rid[:n0])
        s.set_entry(rid[0])
        live = np.arange(n0)
        check(s, live, blocks, rid[0], "initial")
        # 1. inserts: 900 new rows arrive in three batches (each batch = new blocks only; old blocks unchanged)
        for lo, hi in ((n0, 1800), (1800, 1801), (1801, n)):
            assert s.apply(rid[lo:hi], blocks[lo:hi]) == hi - lo
            live = np.arange(hi)
            check(s, live, blocks, rid[0], "after insert up to %d" % hi)
        assert s.node_count == n
        # 2. dirty blocks: re-wire 300 existing nodes (shuffled neighbour lists cut to a random count), one of them
        #    given twice in the batch with the LAST version the valid one; none are new
        cur = blocks.copy()
        touched = rng.choice(n, size=300, replace=False)
        for t in touched:
            ids_view = cur[t, lay["ids_off"]:lay["ids_off"] + 8 * r].view(np.int64)
            cnt_t = int(cur[t, 0:2].view(np.uint16)[0])
            perm = rng.permutation(cnt_t)
            ids_view[:cnt_t] = ids_view[:cnt_t][perm]
            cur[t, 0:2] = np.array([rng.integers(1, cnt_t + 1)], dtype=np.uint16).view(np.uint8)
        stale = blocks[touched[:1]].copy()
        stale[0, 0:2] = 0
        batch_ids = np.concatenate([rid[touched[:1]], rid[touched]])
        batch_blk = np.concatenate([stale, cur[touched]])
        assert s.apply(batch_ids, batch_blk) == 0
        check(s, live, cur, rid[0], "after dirty blocks")
        assert np.array_equal(s.download_block(rid[touched[0]]), cur[touched[0]])
        # 3. deletions (GraphManager.cpp:476-499): 200 rows leave; rows that still name them skip them
        gone = rng.choice(np.arange(1, n), size=200, replace=False)
        assert s.remove(np.concatenate([rid[gone], [-12345]])) == 200
        assert s.remove(rid[gone[:5]]) == 0  # already gone
        live = np.setdiff1d(np.arange(n), gone)
        got = check(s, live, cur, rid[0], "after deletions")
        assert not np.isin(got, rid[gone]).any()
        # 4. the entry point is deleted: no entry -> empty results (Searcher.cpp:74-77) until the host picks another
        assert s.remove(rid[:1]) == 1
        ids, dist, cnt, _ = s.search(q, k, L)
        assert (cnt == 0).all() and (ids == -1).all()
        live = live[live != 0]
        s.set_entry(rid[live[3]])
        check(s, live, cur, rid[live[3]], "new entry point")
        # 5. a deleted row id comes back as a NEW node with a different block
        back = gone[:3]
        newblk = cur[back].copy()
        assert s.apply(rid[back], newblk) == 3
        live = np.concatenate([live, back])  # appended at the end of the mirror, and of the oracle's block array
        check(s, live, cur, rid[live[3]], "re-inserted rows")
        # capacity is enforced before anything is written
        with pytest.raises(lmd.LmdError):
            s.apply(np.arange(10**9, 10**9 + 50, dtype=np.int64), np.zeros((50, bb), dtype=np.uint8))
        check(s, live, cur, rid[live[3]], "after the refused batch")


# ---- filtered search: allowed-ids set + dual heap ------------------------------------
@pytest.mark.parametrize("n,dims,r,metric,kind,route,L,frac", [
    (3000, 128, 32, "L2", "clustered", 0, 64, 0.25),
    (3000, 128, 32, "L2", "iid", 0, 48, 0.5),
    (2000, 96, 32, "L2", "clustered", 0, 100, 0.05),   # selective: most queries return fewer than k rows
    (1500, 768, 32, "COSINE", "clustered", 0, 64, 0.3),
    (1500, 100, 24, "IP", "iid", 0, 40, 0.2),          # generic shape
    (2000, 128, 32, "COSINE", "clustered", 1, 64, 0.3),  # ternary routing: RH feeds the rerank
    (2000, 128, 32, "IP", "iid", 1, 200, 0.1),
    (2500, 128, 32, "L2", "clustered", 0, 512, 0.02),
    (3000, 128, 32, "L2", "iid", 0, 16, 0.004),        # ~12 allowed rows: fewer than k are met, count says so
])
def test_filtered_search_parity(lmd, oracle, make_index, fixtures_mod, n, dims, r, metric, kind, route, L, frac):
    k = 10
    ix = make_index(n, dims, r, metric=metric, kind=kind, nq=96)
    rng = np.random.default_rng(n + dims + L)
    allowed = rng.random(n) < frac
    allowed_ids = np.nonzero(allowed)[0].astype(np.int64)
    with _store(lmd, ix, dims, r, metric) as s:
        s.upload(ix["blocks_np"])
        s.set_entry(0)
        # unknown and repeated row ids in the host's set are harmless
        with s.make_filter(np.concatenate([allowed_ids, allowed_ids[:7], [10**12, -5]])) as f:
            assert f.count == allowed_ids.size
            ids, dist, cnt, st, qs = s.search(ix["queries"], k, L, route=route, qstats=True, filter=f)
            plain = s.search(ix["queries"], k, L, route=route, qstats=True)
        with s.make_filter(np.arange(n, dtype=np.int64)) as everything:
            full = s.search(ix["queries"], k, L, route=route, filter=everything)
    o = oracle.search(ix["blocks_np"], ix["block_bytes"], dims, r, ix["queries"], k, L, metric=METRIC[metric],
                      entry_row_id=0, cfg_l_search=100, route=route, allowed=allowed)
    check_parity(ids, dist, cnt, o[0], o[1], o[2], what="filtered %s route %d" % (metric, route))
    assert allowed[ids[ids >= 0]].all()
    if frac < 0.01:
        assert (cnt < k).all()  # the host is told when fewer than k allowed rows were met (it warns or widens L)
    # the traversal itself is the unfiltered one (tunnelling through rows that fail the filter): same expansions
    # (distance evaluations may differ: RH shares the shared-memory budget with the visited filter, and a smaller
    # visited filter re-evaluates more forgotten nodes)
    assert np.array_equal(qs[:, 0], plain[4][:, 0])
    # a filter that allows every row changes nothing
    assert np.array_equal(full[0], plain[0]) and np.array_equal(full[2], plain[2])
    if frac > 0.05:
        truth = allowed_ids[fixtures_mod.brute_force_topk(ix["x"][allowed], ix["queries"], k, metric, device="cuda")]
        assert fixtures_mod.recall_at_k(ids, truth) == fixtures_mod.recall_at_k(o[0], truth)


def test_filter_goes_stale_when_rows_change(lmd, make_index):
    ix = make_index(1500, 64, 16, metric="L2", kind="iid", nq=16)
    b, q = ix["blocks_np"], ix["queries"]
    rid = np.arange(1500, dtype=np.int64) * 3 + 1  # sparse ids: the filter resolves them through the row-id map
    with lmd.Store(64, 16, capacity=1500) as s:
        s.upload(b[:1400], row_ids=rid[:1400])
        s.set_entry(rid[0])
        f = s.make_filter(rid[:700:2])
        assert f.count == 350
        ids = s.search(q, 10, 40, filter=f)[0]
        assert np.isin(ids[ids >= 0], rid[:700:2]).all()
        s.apply(rid[5:6], b[5:6])          # an overwritten block keeps the row-id -> slot mapping: still valid
        s.search(q, 10, 40, filter=f)
        s.apply(rid[1400:], b[1400:])      # new rows: the mapping changed
        with pytest.raises(lmd.LmdError) as e:
            s.search(q, 10, 40, filter=f)
        assert e.value.status == lmd.ERR_STATE
        f.close()
        with s.make_filter(rid[1400:]) as f2:
            assert f2.count == 100
            s.remove(rid[1400:1410])
            with pytest.raises(lmd.LmdError):
                s.search(q, 10, 40, filter=f2)
        with lmd.Store(64, 16, capacity=10) as other, s.make_filter(rid[:10]) as f3:
            other.upload(b[:10])
            other.set_entry(0)
            with pytest.raises(lmd.LmdError):
                other.search(q, 10, 40, filter=f3)


# ---- latency variants (small batches): same results, bit for bit --------------------------------------------------
@pytest.mark.parametrize("n,dims,r,metric,L", [(3000, 128, 32, "L2", 56), (2000, 96, 32, "L2", 100),
                                               (1500, 768, 32, "COSINE", 48), (2000, 128, 64, "IP", 200),
                                               (2500, 128, 32, "L2", 512)])
def test_latency_kernels_equal_throughput_kernels(lmd, oracle, make_index, n, dims, r, metric, L):
    ix = make_index(n, dims, r, metric=metric, kind="clustered", nq=96)
    q = ix["queries"]
    with _store(lmd, ix, dims, r, metric) as s:
        s.upload(ix["blocks_np"])
        s.set_entry(0)
        s.set_tuning("latency_mode", 0)
        a = s.search(q, 10, L, qstats=True)
        s.set_tuning("latency_mode", 1)
        b = s.search(q, 10, L, qstats=True)
        one = s.search(q[:1], 10, L)          # a single query: one warp per CTA
        s.set_tuning("latency_mode", -1)
        c = s.search(q, 10, L)                # 96 queries: picked automatically
    assert np.array_equal(a[0], b[0]) and np.array_equal(a[1].view(np.uint32), b[1].view(np.uint32))
    # same expansions per query; the evaluations differ only by what the visited filters forget (the small-batch kernels
    # run fewer warps per SM and get a larger filter)
    assert np.array_equal(a[4][:, 0], b[4][:, 0])
    va, vb = a[4][:, 1].astype(np.int64), b[4][:, 1].astype(np.int64)
    assert np.all(vb <= va + va // 50 + 2) and np.all(va <= vb + vb // 3)   # L = 512 on 2 500 rows: 20 % re-evaluated
    assert np.array_equal(a[0], c[0]) and np.array_equal(a[0][:1], one[0])
    o = oracle.search(ix["blocks_np"], ix["block_bytes"], dims, r, q, 10, L, metric=METRIC[metric], entry_row_id=0)
    check_parity(b[0], b[1], b[2], o[0], o[1], o[2], what="latency kernels %s" % metric)


def test_concurrent_host_threads_share_one_store(lmd, make_index):
    """include/lmd_b200.h: concurrent searches on one store are serialised internally. Four host threads, different
    query slices and list sizes, must each get what a lone caller gets; an error in one thread stays in that thread."""
    import threading
    ix = make_index(3000, 128, 32, metric="L2", kind="clustered", nq=96)
    q = ix["queries"]
    with _store(lmd, ix, 128, 32, "L2") as s:
        s.upload(ix["blocks_np"])
        s.set_entry(0)
        want = {L: s.search(q, 10, L)[:3] for L in (32, 64, 100, 200)}
        out, errs = {}, []

        def work(t, L):
            try:
                for rep in range(6):
                    lo = (t * 7 + rep * 5) % 40
                    ids, dist, cnt, _ = s.search(q[lo:lo + 50], 10, L)
                    assert np.array_equal(ids, want[L][0][lo:lo + 50]) and np.array_equal(cnt, want[L][2][lo:lo + 50])
                    assert np.array_equal(dist.view(np.uint32), want[L][1][lo:lo + 50].view(np.uint32))
                if t == 0:
                    with pytest.raises(lmd.LmdError):
                        s.search(q, 10, 4096)  # L beyond the build's limit: this thread's error only
                out[t] = True
            except BaseException as e:  # noqa: BLE001 - reported below
                errs.append((t, repr(e)))

        th = [threading.Thread(target=work, args=(t, L)) for t, L in enumerate((32, 64, 100, 200))]
        for t in th:
            t.start()
        for t in th:
            t.join()
        assert not errs, errs
        assert len(out) == 4
