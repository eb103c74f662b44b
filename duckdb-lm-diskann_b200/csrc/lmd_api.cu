// lmd_search_batch / lmd_merge_topk: the batched replacement of core/ISearcher.hpp:30-34 (Search).
// Host orchestration only: argument checks in the reference's order, staging of host buffers through
// the store's stream, kernel launches, statistics.
#include "lmd_common.cuh"

#include <algorithm>
#include <cstring>

using namespace lmd;

namespace {

// Where a caller's buffer lives, and the address the kernels can use for it. Pinned (page-locked) host memory is
// mapped into the device's address space under UVA: the kernels read the queries and write the results straight
// through PCIe / NVLink-C2C ("zero copy", 512 B in and 120 B out per query at k = 10), so the host-buffer path needs
// no staging copy and no extra stream round trip. Pageable memory goes through the store's device scratch.
enum BufKind { BUF_DEVICE, BUF_PINNED, BUF_PAGEABLE };
BufKind classify(const void *p, void **dev_alias) {
	*dev_alias = const_cast<void *>(p);
	if (!p) {
		return BUF_DEVICE;
	}
	cudaPointerAttributes at;
	if (cudaPointerGetAttributes(&at, p) != cudaSuccess) {
		cudaGetLastError();
		return BUF_PAGEABLE;
	}
	if (at.type == cudaMemoryTypeDevice || at.type == cudaMemoryTypeManaged) {
		return BUF_DEVICE;
	}
	if (at.type == cudaMemoryTypeHost && at.devicePointer) {
		*dev_alias = at.devicePointer;
		return BUF_PINNED;
	}
	return BUF_PAGEABLE;
}

bool on_device(const void *p) {
	cudaPointerAttributes at;
	if (cudaPointerGetAttributes(&at, p) != cudaSuccess) {
		cudaGetLastError();
		return false;
	}
	return at.type == cudaMemoryTypeDevice || at.type == cudaMemoryTypeManaged;
}

template <typename T>
lmd_status grow(T *&ptr, uint64_t &have, uint64_t need_elems, lmd_store *s) {
	if (have >= need_elems && ptr) {
		return LMD_OK;
	}
	if (ptr) {
		cudaFree(ptr);
		ptr = nullptr;
	}
	cudaError_t e = cudaMalloc((void **)&ptr, need_elems * sizeof(T));
	if (e != cudaSuccess) {
		have = 0;
		set_error("cudaMalloc(%llu bytes) for search scratch failed: %s", (unsigned long long)(need_elems * sizeof(T)),
		          cudaGetErrorString(e));
		return LMD_ERR_OOM;
	}
	have = need_elems;
	(void)s;
	return LMD_OK;
}

} // namespace

static lmd_status search_impl(lmd_store *s, const lmd_filter *filter, const float *queries, uint64_t nq, uint32_t k,
                       uint32_t L, uint32_t route, uint32_t flags, int64_t *out_ids, float *out_dist,
                       uint32_t *out_count, uint32_t *out_qstats, lmd_stats *stats) {
	if (stats) {
		memset(stats, 0, sizeof(*stats));
	}
	if (!s) {
		set_error("lmd_search_batch: null store");
		return LMD_ERR_INVALID_ARG;
	}
	if (k == 0) { // core/Coordinator.cpp:72-74, db/LmDiskannIndex.cpp:656-658
		set_error("k_neighbors cannot be 0 for search");
		return LMD_ERR_INVALID_ARG;
	}
	if (nq == 0) {
		return LMD_OK;
	}
	if (!queries || !out_ids) {
		set_error("lmd_search_batch: null queries / out_ids");
		return LMD_ERR_INVALID_ARG;
	}
	if (nq > 0xFFFFFFF0ull) {
		set_error("lmd_search_batch: batch too large");
		return LMD_ERR_INVALID_ARG;
	}
	if (route != LMD_ROUTE_EXACT && route != LMD_ROUTE_TERNARY) {
		set_error("lmd_search_batch: unknown route %u", route);
		return LMD_ERR_INVALID_ARG;
	}
	if (route == LMD_ROUTE_TERNARY && s->cfg.metric == LMD_METRIC_L2) {
		// core/distance.hpp:236-239
		set_error("L2 metric is not directly compatible with the current ternary CalculateApproxDistance");
		return LMD_ERR_UNSUPPORTED;
	}
	// core/Searcher.cpp:261-265
	uint32_t l_used = L > 0 ? L : s->cfg.l_search;
	if (l_used == 0) {
		l_used = k > 20 ? k + 20 : 40;
	}
	l_used = std::max(l_used, k);
	if (l_used > 512) {
		set_error("search list size L = %u exceeds this build's limit of 512", l_used);
		return LMD_ERR_UNSUPPORTED;
	}
	if (s->dirty) {
		lmd_status st = lmd_store_finalize(s);
		if (st != LMD_OK) {
			return st;
		}
	}
	std::lock_guard<std::mutex> g(s->mu);
	if (filter && (filter->store != s || filter->epoch != s->epoch)) {
		set_error("lmd_search_batch_filtered: the filter was built for another store or before rows were added / "
		          "removed; create it again");
		return LMD_ERR_STATE;
	}
	LMD_CUDA_OK(cudaSetDevice(s->device));
	cudaStream_t st = s->stream;
	const uint32_t q32 = (uint32_t)nq;
	void *a_q, *a_ids, *a_dist, *a_count, *a_qstats;
	const BufKind kq = classify(queries, &a_q), ki = classify(out_ids, &a_ids), kd = classify(out_dist, &a_dist),
	              kc = classify(out_count, &a_count), ks = classify(out_qstats, &a_qstats);
	const bool stage_q = kq == BUF_PAGEABLE;
	const bool stage_out = ki == BUF_PAGEABLE; // outputs are staged together or not at all
	if (!stage_out && (kd == BUF_PAGEABLE || kc == BUF_PAGEABLE || ks == BUF_PAGEABLE)) {
		set_error("lmd_search_batch: out_ids is device / pinned memory but another output is pageable host memory");
		return LMD_ERR_INVALID_ARG;
	}
	const bool dev_io = !stage_q && !stage_out; // nothing to copy: device-resident or zero-copy pinned buffers
	const bool separate = route == LMD_ROUTE_TERNARY || (flags & LMD_SEARCH_SEPARATE_RERANK);
	const bool want_qstats = (flags & LMD_SEARCH_PER_QUERY_STATS) && out_qstats;
	const bool async = (flags & LMD_SEARCH_ASYNC) != 0;
	lmd_status rc;
	if (async && (!dev_io || stats)) {
		set_error("LMD_SEARCH_ASYNC needs device-resident or pinned buffers and stats == NULL");
		return LMD_ERR_INVALID_ARG;
	}

	if (!async) {
		LMD_CUDA_OK(cudaEventRecord(s->ev[0], st));
	}
	const float *d_q = static_cast<const float *>(a_q);
	int64_t *d_ids = static_cast<int64_t *>(a_ids);
	float *d_dist = static_cast<float *>(a_dist);
	uint32_t *d_count = static_cast<uint32_t *>(a_count);
	uint32_t *d_qstats = want_qstats ? static_cast<uint32_t *>(a_qstats) : nullptr;
	if (stage_q) {
		uint64_t need_q = nq * (uint64_t)s->cfg.dims;
		uint64_t have_q = s->scratch_q;
		if ((rc = grow(s->d_q, have_q, need_q, s)) != LMD_OK) {
			s->scratch_q = 0;
			return rc;
		}
		s->scratch_q = have_q;
		LMD_CUDA_OK(cudaMemcpyAsync(s->d_q, queries, need_q * sizeof(float), cudaMemcpyHostToDevice, st));
		d_q = s->d_q;
	}
	if (stage_out) {
		uint64_t need_k = nq * (uint64_t)k;
		if (s->scratch_k < need_k || !s->d_out_ids) {
			uint64_t a = 0, b = 0, c = 0, d = 0;
			if ((rc = grow(s->d_out_ids, a, need_k, s)) != LMD_OK || (rc = grow(s->d_out_dist, b, need_k, s)) != LMD_OK ||
			    (rc = grow(s->d_out_count, c, need_k, s)) != LMD_OK ||
			    (rc = grow(s->d_qstats, d, std::max<uint64_t>(need_k, 2 * nq), s)) != LMD_OK) {
				s->scratch_k = 0;
				return rc;
			}
			s->scratch_k = need_k;
		}
		d_ids = s->d_out_ids;
		d_dist = s->d_out_dist; // always computed on the device; copied back only when asked for
		d_count = s->d_out_count;
		d_qstats = want_qstats ? s->d_qstats : nullptr;
	}
	uint64_t *d_cand = nullptr;
	if (separate) {
		uint64_t need_c = nq * (uint64_t)l_used;
		uint64_t have_c = s->scratch_cand;
		if ((rc = grow(s->d_cand, have_c, need_c, s)) != LMD_OK) {
			s->scratch_cand = 0;
			return rc;
		}
		s->scratch_cand = have_c;
		d_cand = s->d_cand;
	}
	LMD_CUDA_OK(cudaMemsetAsync(s->d_counters, 0, 8 * sizeof(unsigned long long), st));

	TraverseArgs ta;
	ta.s = s;
	ta.d_queries = d_q;
	ta.nq = q32;
	ta.k = k;
	ta.L = l_used;
	ta.route = route;
	ta.d_out_ids = separate ? nullptr : d_ids;
	ta.d_out_dist = separate ? nullptr : d_dist;
	ta.d_out_count = separate ? nullptr : d_count;
	ta.d_qstats = d_qstats;
	ta.d_cand = d_cand;
	ta.d_allowed = filter ? filter->d_bits : nullptr;
	if (!async) {
		LMD_CUDA_OK(cudaEventRecord(s->ev[1], st));
	}
	if ((rc = launch_traverse(ta, st)) != LMD_OK) {
		return rc;
	}
	if (!async) {
		LMD_CUDA_OK(cudaEventRecord(s->ev[2], st));
	}
	uint32_t launches = 1;
	if (separate) {
		RerankArgs ra;
		ra.s = s;
		ra.d_queries = d_q;
		ra.nq = q32;
		ra.k = k;
		ra.L = l_used;
		ra.d_cand = d_cand;
		ra.d_out_ids = d_ids;
		ra.d_out_dist = d_dist;
		ra.d_out_count = d_count;
		if ((rc = launch_rerank(ra, st)) != LMD_OK) {
			return rc;
		}
		launches++;
	}
	if (async) {
		return LMD_OK;
	}
	LMD_CUDA_OK(cudaEventRecord(s->ev[3], st));
	if (stage_out) {
		LMD_CUDA_OK(cudaMemcpyAsync(out_ids, d_ids, nq * k * sizeof(int64_t), cudaMemcpyDeviceToHost, st));
		if (out_dist) {
			LMD_CUDA_OK(cudaMemcpyAsync(out_dist, d_dist, nq * k * sizeof(float), cudaMemcpyDeviceToHost, st));
		}
		if (out_count) {
			LMD_CUDA_OK(cudaMemcpyAsync(out_count, d_count, nq * sizeof(uint32_t), cudaMemcpyDeviceToHost, st));
		}
		if (want_qstats) {
			LMD_CUDA_OK(cudaMemcpyAsync(out_qstats, d_qstats, nq * 2 * sizeof(uint32_t), cudaMemcpyDeviceToHost, st));
		}
	}
	unsigned long long h_cnt[8] = {0, 0, 0, 0, 0, 0, 0, 0};
	if (stats) {
		LMD_CUDA_OK(cudaMemcpyAsync(h_cnt, s->d_counters, sizeof(h_cnt), cudaMemcpyDeviceToHost, st));
	}
	LMD_CUDA_OK(cudaEventRecord(s->ev[4], st));
	cudaError_t e = cudaStreamSynchronize(st);
	if (e != cudaSuccess) {
		set_error("search kernels failed: %s", cudaGetErrorString(e));
		return LMD_ERR_CUDA;
	}
	if (stats) {
		stats->queries = nq;
		stats->expansions = h_cnt[1];
		stats->distance_evals = h_cnt[2];
		stats->filter_evictions = h_cnt[3];
		stats->reranked = h_cnt[4];
		stats->tie_overflows = h_cnt[5];
		stats->merge_fallbacks = h_cnt[6];
		stats->launches = launches;
		stats->l_used = l_used;
		cudaEventElapsedTime(&stats->traverse_ms, s->ev[1], s->ev[2]);
		if (separate) {
			cudaEventElapsedTime(&stats->rerank_ms, s->ev[2], s->ev[3]);
		}
		cudaEventElapsedTime(&stats->total_ms, s->ev[0], s->ev[4]);
		if (stage_q) {
			cudaEventElapsedTime(&stats->h2d_ms, s->ev[0], s->ev[1]);
		}
		if (stage_out) {
			cudaEventElapsedTime(&stats->d2h_ms, s->ev[3], s->ev[4]);
		}
		// SURVEY.md section 8d. The sidecar holds uint32 slots, so an expansion reads 4R bytes of ids
		// (the reference's 8 + 8R becomes 4R); padding is never counted.
		const uint64_t vec = (uint64_t)s->cfg.dims * 4;
		uint64_t bytes = nq * ((uint64_t)s->cfg.dims * 4 + (uint64_t)k * 12);
		if (route == LMD_ROUTE_EXACT) {
			bytes += h_cnt[1] * 4ull * s->cfg.r + h_cnt[2] * vec;
		} else {
			bytes += h_cnt[1] * (4ull * s->cfg.r) + h_cnt[2] * (uint64_t)s->lay.ternary_edge_bytes + vec * nq;
		}
		bytes += h_cnt[4] * vec;
		if (separate) {
			bytes += 2ull * nq * l_used * 8; // candidate list written by the traversal, read by the rerank
		}
		stats->algorithmic_bytes = bytes;
	}
	return LMD_OK;
}

extern "C" {

lmd_status lmd_search_batch(lmd_store *s, const float *queries, uint64_t nq, uint32_t k, uint32_t L, uint32_t route,
                            uint32_t flags, int64_t *out_ids, float *out_dist, uint32_t *out_count,
                            uint32_t *out_qstats, lmd_stats *stats) {
	return search_impl(s, nullptr, queries, nq, k, L, route, flags, out_ids, out_dist, out_count, out_qstats, stats);
}

lmd_status lmd_search_batch_filtered(lmd_store *s, const lmd_filter *filter, const float *queries, uint64_t nq,
                                     uint32_t k, uint32_t L, uint32_t route, uint32_t flags, int64_t *out_ids,
                                     float *out_dist, uint32_t *out_count, uint32_t *out_qstats, lmd_stats *stats) {
	if (!filter) {
		set_error("lmd_search_batch_filtered: null filter");
		return LMD_ERR_INVALID_ARG;
	}
	return search_impl(s, filter, queries, nq, k, L, route, flags, out_ids, out_dist, out_count, out_qstats, stats);
}

lmd_status lmd_merge_topk_async(int device, const int64_t *ids, const float *dist, uint32_t shards, uint64_t nq,
                                uint32_t k, int64_t *out_ids, float *out_dist, void *cuda_stream) {
	if (!ids || !dist || !out_ids || !out_dist) {
		set_error("lmd_merge_topk_async: null argument");
		return LMD_ERR_INVALID_ARG;
	}
	if (k == 0 || shards == 0 || shards > 64) {
		set_error("lmd_merge_topk_async: k must be > 0 and shards in [1, 64]");
		return LMD_ERR_INVALID_ARG;
	}
	if (nq == 0) {
		return LMD_OK;
	}
	LMD_CUDA_OK(cudaSetDevice(device));
	if (!on_device(ids) || !on_device(dist) || !on_device(out_ids) || !on_device(out_dist)) {
		set_error("lmd_merge_topk_async: all four buffers must be device (or pinned) memory");
		return LMD_ERR_INVALID_ARG;
	}
	return launch_merge_topk(ids, dist, shards, nq, k, out_ids, out_dist, (cudaStream_t)cuda_stream);
}

lmd_status lmd_merge_topk(int device, const int64_t *ids, const float *dist, uint32_t shards, uint64_t nq, uint32_t k,
                          int64_t *out_ids, float *out_dist) {
	if (!ids || !dist || !out_ids || !out_dist) {
		set_error("lmd_merge_topk: null argument");
		return LMD_ERR_INVALID_ARG;
	}
	if (k == 0 || shards == 0 || shards > 64) {
		set_error("lmd_merge_topk: k must be > 0 and shards in [1, 64]");
		return LMD_ERR_INVALID_ARG;
	}
	if (nq == 0) {
		return LMD_OK;
	}
	LMD_CUDA_OK(cudaSetDevice(device));
	const bool dev_io = on_device(ids);
	const uint64_t in_elems = (uint64_t)shards * nq * k, out_elems = nq * k;
	if (dev_io) {
		lmd_status rc = launch_merge_topk(ids, dist, shards, nq, k, out_ids, out_dist, 0);
		if (rc != LMD_OK) {
			return rc;
		}
		LMD_CUDA_OK(cudaStreamSynchronize(0));
		return LMD_OK;
	}
	int64_t *d_ids = nullptr, *d_oids = nullptr;
	float *d_dist = nullptr, *d_odist = nullptr;
	lmd_status rc = LMD_OK;
	auto cleanup = [&]() {
		cudaFree(d_ids);
		cudaFree(d_oids);
		cudaFree(d_dist);
		cudaFree(d_odist);
	};
	if (cudaMalloc((void **)&d_ids, in_elems * 8) != cudaSuccess || cudaMalloc((void **)&d_dist, in_elems * 4) != cudaSuccess ||
	    cudaMalloc((void **)&d_oids, out_elems * 8) != cudaSuccess ||
	    cudaMalloc((void **)&d_odist, out_elems * 4) != cudaSuccess) {
		cleanup();
		set_error("lmd_merge_topk: cudaMalloc failed");
		return LMD_ERR_OOM;
	}
	cudaMemcpy(d_ids, ids, in_elems * 8, cudaMemcpyHostToDevice);
	cudaMemcpy(d_dist, dist, in_elems * 4, cudaMemcpyHostToDevice);
	rc = launch_merge_topk(d_ids, d_dist, shards, nq, k, d_oids, d_odist, 0);
	if (rc == LMD_OK) {
		cudaMemcpy(out_ids, d_oids, out_elems * 8, cudaMemcpyDeviceToHost);
		cudaError_t e = cudaMemcpy(out_dist, d_odist, out_elems * 4, cudaMemcpyDeviceToHost);
		if (e != cudaSuccess) {
			set_error("lmd_merge_topk: %s", cudaGetErrorString(e));
			rc = LMD_ERR_CUDA;
		}
	}
	cleanup();
	return rc;
}

} // extern "C"
