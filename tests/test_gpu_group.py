"""-m gpu: lmd_group - one process, several devices, through the C ABI.
Members sit on different devices when the box has them and share cuda:0 otherwise (the code path is the same; only
the peer loads of the merge become local loads). Parity: query sharding = the single-store search; dataset sharding =
the oracle run per shard and merged by (distance, row id)."""
import numpy as np
import pytest
import torch

from parity import check_parity
from test_sharded import _make_case, _np_merge

pytestmark = pytest.mark.gpu


def _devices(n):
    have = torch.cuda.device_count()
    return [i % have for i in range(n)]


@pytest.mark.parametrize("members", [1, 2, 3])
def test_group_query_sharded_equals_single_store(lmd, members):
    case = _make_case(n=1600, dims=32, r=12, nq=37)
    stores = []
    try:
        for d in _devices(members):
            s = lmd.Store(case["dims"], case["r"], capacity=case["n"], device=d)
            s.upload(case["full"])
            s.set_entry(0)
            stores.append(s)
        ref = stores[0].search(case["q"], 10, 48)
        with lmd.Group(stores, lmd.GROUP_QUERY_SHARDED) as g:
            ids, dist, cnt, st = g.search(case["q"], 10, 48)
            assert st.members == members and sum(st.queries[:members]) == 37
            assert max(st.queries[:members]) - min(st.queries[:members]) <= 1  # balanced slices
            assert np.array_equal(ids, ref[0]) and np.array_equal(dist, ref[1]) and np.array_equal(cnt, ref[2])
            # fewer queries than members, pinned and device buffers
            ids1, dist1, cnt1, _ = g.search(case["q"][:2], 10, 48)
            assert np.array_equal(ids1, ref[0][:2])
            qd = torch.from_numpy(case["q"]).cuda()
            oi = torch.empty((37, 10), dtype=torch.int64, device="cuda")
            od = torch.empty((37, 10), dtype=torch.float32, device="cuda")
            g.search_raw(qd, 37, 10, 48, out_ids=oi, out_dist=od)
            assert np.array_equal(oi.cpu().numpy(), ref[0]) and np.array_equal(od.cpu().numpy(), ref[1])
            hq = torch.from_numpy(case["q"]).pin_memory()
            hi = torch.empty((37, 10), dtype=torch.int64).pin_memory()
            g.search_raw(hq, 37, 10, 48, out_ids=hi, flags=lmd.SEARCH_SEPARATE_RERANK)
            assert np.array_equal(hi.numpy(), ref[0])
            with pytest.raises(lmd.LmdError) as e:
                g.search(case["q"], 0, 48)
            assert e.value.status == lmd.ERR_INVALID_ARG
    finally:
        for s in stores:
            s.close()


@pytest.mark.parametrize("world", [2, 4])
def test_group_dataset_sharded_equals_oracle_merge(lmd, oracle, world):
    case = _make_case(n=2000, dims=32, r=12, nq=41, world=world)
    k, L = 10, 40
    stores = []
    try:
        per_i, per_d = [], []
        for (blocks, rid), d in zip(case["shards"], _devices(world)):
            s = lmd.Store(case["dims"], case["r"], capacity=blocks.shape[0], device=d)
            s.upload(blocks, row_ids=rid)
            s.set_entry(int(rid[0]))
            stores.append(s)
            o = oracle.search(blocks, case["bb"], case["dims"], case["r"], case["q"], k, L, row_ids=rid,
                              entry_row_id=int(rid[0]), nthreads=1)
            per_i.append(o[0])
            per_d.append(o[1])
        want_i, want_d = _np_merge(np.stack(per_i), np.stack(per_d), k)
        with lmd.Group(stores, lmd.GROUP_DATASET_SHARDED) as g:
            ids, dist, cnt, st = g.search(case["q"], k, L)
            assert st.members == world and st.merge_input_bytes == world * 41 * k * 12
            assert all(st.queries[i] == 41 for i in range(world))
            check_parity(ids, dist, cnt, want_i, want_d, np.full(41, k), what="group, %d shards" % world)
            # the merge of the per-shard GPU lists is exact: (distance, row id) order of what the members returned
            mine_i, mine_d = [], []
            for s in stores:
                r = s.search(case["q"], k, L)
                mine_i.append(r[0])
                mine_d.append(r[1])
            gi, gd = _np_merge(np.stack(mine_i), np.stack(mine_d), k)
            assert np.array_equal(ids, gi) and np.array_equal(dist, gd)
            # k larger than a shard's result: -1 padding survives the merge
            ids2, dist2, cnt2, _ = g.search(case["q"][:3], 64, 64)
            assert (cnt2 <= 64).all() and ((ids2 >= 0).sum(axis=1) == cnt2).all()
            for q in range(3):
                v = ids2[q, : cnt2[q]]
                assert len(set(v.tolist())) == len(v)
                assert (np.diff(dist2[q, : cnt2[q]]) >= 0).all()
    finally:
        for s in stores:
            s.close()


def test_group_rejects_mismatched_stores(lmd):
    a = lmd.Store(32, 12, capacity=16)
    b = lmd.Store(48, 12, capacity=16)
    try:
        with pytest.raises(lmd.LmdError) as e:
            lmd.Group([a, b])
        assert e.value.status == lmd.ERR_INVALID_ARG
        with pytest.raises(lmd.LmdError):
            lmd.Group([a, a])
    finally:
        a.close()
        b.close()
