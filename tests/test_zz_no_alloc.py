"""gpu: lmd_search_batch_static (include/lmd_b200.h) - the search whose kernel scratch lives in the library's static device
array, for callers that cannot allocate - against lmd_search_batch on the same store and queries, through the C ABI.
Both run the same deterministic kernels, so the bar is the one of the other parity tests: identical ids, identical
distance bits, identical counts. Collected last on purpose."""
import numpy as np
import pytest

METRIC = {"L2": 1, "COSINE": 2, "IP": 3}
pytestmark = pytest.mark.gpu

CASES = [("L2", "exact", 128), ("L2", "prefilter", 128), ("L2", "int8", 128), ("COSINE", "ternary", 96)]


def _device_search(torch, s, q_dev, k, L, route):
    nq = q_dev.shape[0]
    ids = torch.full((nq, k), -7, dtype=torch.int64, device="cuda")
    dist = torch.full((nq, k), -7.0, dtype=torch.float32, device="cuda")
    cnt = torch.full((nq,), 77, dtype=torch.int32, device="cuda")
    s.search_raw(q_dev, nq, k, L, route, 0, ids, dist, cnt, None)
    return ids.cpu().numpy(), dist.cpu().numpy(), cnt.cpu().numpy()


_WARM = []


def _static_search(torch, s, q_dev, k, L, route, flags=0):
    nq = q_dev.shape[0]
    if not _WARM:
        # CUDA loads modules lazily: the first touch of the file that holds the static array (cudaGetSymbolAddress) is when
        # the driver materialises its 8 MB, once per process and device. Not an allocation of the call - taken out of the
        # free-memory check below by one unmeasured call.
        w = [torch.empty((nq, k), dtype=torch.int64, device="cuda"), torch.empty((nq, k), dtype=torch.float32, device="cuda"),
             torch.empty((nq,), dtype=torch.int32, device="cuda")]
        s.search_static(q_dev, nq, k, L, route, *w)
        _WARM.append(1)
    ids = torch.full((nq, k), -7, dtype=torch.int64, device="cuda")
    dist = torch.full((nq, k), -7.0, dtype=torch.float32, device="cuda")
    cnt = torch.full((nq,), 77, dtype=torch.int32, device="cuda")
    torch.cuda.synchronize()
    free0 = torch.cuda.mem_get_info()[0]
    s.search_static(q_dev, nq, k, L, route, ids, dist, cnt, flags=flags)
    torch.cuda.synchronize()
    assert torch.cuda.mem_get_info()[0] >= free0, "the call took device memory"
    return ids.cpu().numpy(), dist.cpu().numpy(), cnt.cpu().numpy()


@pytest.mark.parametrize("metric,route,dims", CASES)
def test_static_scratch_search_equals_search_batch(lmd, make_index, metric, route, dims):
    import torch
    n, r, k = 3000, 32, 10
    ix = make_index(n, dims, r, metric=metric, kind="clustered", nq=96)
    kroute = {"ternary": lmd.ROUTE_TERNARY, "int8": lmd.ROUTE_INT8}.get(route, lmd.ROUTE_EXACT)
    with lmd.Store(dims, r, metric=METRIC[metric], capacity=n, l_search=100) as s:
        s.upload(ix["blocks_np"])
        s.set_entry(0)
        if route in ("prefilter", "int8"):
            s.enable_codec()
            s.set_tuning("prefilter", 2)
        s.finalize()
        q_all = torch.from_numpy(np.ascontiguousarray(ix["queries"], dtype=np.float32)).cuda()
        # two calls in a row on the shared array: a full batch at L = 64, then a shorter one at another L (whatever the first
        # call left in the scratch - counters, candidate keys of 96 queries - must not reach the second call's results)
        for q_dev, L in ((q_all, 64), (q_all[5:42].contiguous(), 40)):
            want = _device_search(torch, s, q_dev, k, L, kroute)
            got = _static_search(torch, s, q_dev, k, L, kroute)
            assert np.array_equal(got[0], want[0]), (metric, route, L)
            assert np.array_equal(got[1].view(np.uint32), want[1].view(np.uint32)), (metric, route, L)
            assert np.array_equal(got[2], want[2]), (metric, route, L)
            assert (got[2] == k).all() and (got[0] >= 0).all()
        # the enqueue-only form on the caller's stream
        got = _static_search(torch, s, q_all, k, 64, kroute, flags=lmd.SEARCH_ASYNC)
        want = _device_search(torch, s, q_all, k, 64, kroute)
        assert np.array_equal(got[0], want[0]) and np.array_equal(got[1].view(np.uint32), want[1].view(np.uint32))


def test_static_scratch_limits_and_refusals(lmd, make_index):
    import torch
    n, dims, r, k = 2000, 96, 32, 10
    ix = make_index(n, dims, r, metric="COSINE", kind="clustered", nq=96)
    cap = lmd.lib().lmd_static_scratch_bytes()
    assert cap == 8 << 20
    with lmd.Store(dims, r, metric=METRIC["COSINE"], capacity=n, l_search=100) as s:
        s.upload(ix["blocks_np"])
        s.set_entry(0)
        s.finalize()
        L = 64
        too_many = (cap - 256) // (8 * L) + 1
        q = torch.zeros((too_many, dims), dtype=torch.float32, device="cuda")
        ids = torch.empty((too_many, k), dtype=torch.int64, device="cuda")
        dist = torch.empty((too_many, k), dtype=torch.float32, device="cuda")
        cnt = torch.empty((too_many,), dtype=torch.int32, device="cuda")
        with pytest.raises(lmd.LmdError) as e:  # one query beyond what the static array holds on the ternary route
            s.search_static(q, too_many, k, L, lmd.ROUTE_TERNARY, ids, dist, cnt)
        assert e.value.status == lmd.ERR_INVALID_ARG and "static array holds" in str(e.value)
        q[:] = torch.from_numpy(ix["queries"][:1]).cuda()
        with pytest.raises(lmd.LmdError) as e:  # the ternary sidecar is not built yet, and building it allocates
            s.search_static(q, 8, k, L, lmd.ROUTE_TERNARY, ids, dist, cnt)
        assert e.value.status == lmd.ERR_STATE and "not prepared" in str(e.value)
        s.search(ix["queries"][:2], k, L, route=lmd.ROUTE_TERNARY)  # a regular search prepares the route
        s.search_static(q, too_many - 1, k, L, lmd.ROUTE_TERNARY, ids, dist, cnt)  # exactly at the limit: served
        s.search_static(q, too_many, k, L, lmd.ROUTE_EXACT, ids, dist, cnt)        # the exact route keeps no candidates
        torch.cuda.synchronize()
        assert (cnt.cpu().numpy() == k).all()
        host_q = np.ascontiguousarray(ix["queries"][:4], dtype=np.float32)
        with pytest.raises(lmd.LmdError) as e:  # pageable memory would need a staging allocation
            s.search_static(host_q, 4, k, L, lmd.ROUTE_EXACT, ids, dist, cnt)
        assert e.value.status == lmd.ERR_INVALID_ARG and "pinned" in str(e.value)
        with pytest.raises(lmd.LmdError) as e:
            s.search_static(q, 4, k, L, lmd.ROUTE_EXACT, ids, dist, cnt, flags=lmd.SEARCH_PER_QUERY_STATS)
        assert e.value.status == lmd.ERR_INVALID_ARG
