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

// Grows one scratch array of a SearchState; `have` is ITS capacity in elements (each array has its own: d_qstats needs
// 2 nq entries whatever k is, d_out_count nq, the id / distance arrays nq * k).
template <typename T>
static lmd_status grow_to(T *&ptr, uint64_t &have, uint64_t need_elems) {
	lmd_status rc = grow(ptr, have, need_elems, nullptr);
	if (rc != LMD_OK) {
		have = 0;
	}
	return rc;
}

static const size_t PIN_Q_BYTES = 48 * 1024, PIN_OUT_BYTES = 80 * 1024;

static lmd_status search_impl(lmd_store *s, SearchState &ss, const lmd_filter *filter, const float *queries, uint64_t nq,
                              uint32_t k, uint32_t L, uint32_t route, uint32_t flags, int64_t *out_ids, float *out_dist,
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
	if (route != LMD_ROUTE_EXACT && route != LMD_ROUTE_TERNARY && route != LMD_ROUTE_INT8) {
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
	// Searches only read the store: they hold its lock shared, so calls on different contexts run side by side; whatever
	// changes the store (upload, apply / remove, finalize, build, tuning) holds it exclusively.
	std::shared_lock<std::shared_mutex> rd(s->mu);
	if (route == LMD_ROUTE_INT8 && s->cfg.metric != LMD_METRIC_L2) {
		set_error("the int8 edge codec is defined for the L2 metric only");
		return LMD_ERR_UNSUPPORTED;
	}
	const uint32_t prep = route | ((flags & LMD_SEARCH_TENSOR_RERANK) ? LMD_PREP_NORMS : 0u);
	while (!route_ready(s, prep)) {
		rd.unlock();
		lmd_status st = store_prepare(s, prep);
		if (st != LMD_OK) {
			return st;
		}
		rd.lock();
	}
	std::lock_guard<std::mutex> g(ss.mu);
	if (filter && (filter->store != s || filter->epoch != s->epoch)) {
		set_error("lmd_search_batch_filtered: the filter was built for another store or before rows were added / "
		          "removed; create it again");
		return LMD_ERR_STATE;
	}
	if (filter && route == LMD_ROUTE_INT8) {
		set_error("lmd_search_batch_filtered: the int8 edge codec has no filtered kernels");
		return LMD_ERR_UNSUPPORTED;
	}
	LMD_CUDA_OK(cudaSetDevice(s->device));
	cudaStream_t st = ss.stream;
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
	const bool separate = route != LMD_ROUTE_EXACT || (flags & (LMD_SEARCH_SEPARATE_RERANK | LMD_SEARCH_TENSOR_RERANK));
	const bool want_qstats = (flags & LMD_SEARCH_PER_QUERY_STATS) && out_qstats;
	const bool async = (flags & LMD_SEARCH_ASYNC) != 0;
	lmd_status rc;
	if (async && (!dev_io || stats)) {
		set_error("LMD_SEARCH_ASYNC needs device-resident or pinned buffers and stats == NULL");
		return LMD_ERR_INVALID_ARG;
	}

	// the events only feed lmd_stats: a call without statistics (every point query of the adapter) skips five driver calls
	const bool timed = !async && stats != nullptr;
	if (timed) {
		LMD_CUDA_OK(cudaEventRecord(ss.ev[0], st));
	}
	const float *d_q = static_cast<const float *>(a_q);
	int64_t *d_ids = static_cast<int64_t *>(a_ids);
	float *d_dist = static_cast<float *>(a_dist);
	uint32_t *d_count = static_cast<uint32_t *>(a_count);
	uint32_t *d_qstats = want_qstats ? static_cast<uint32_t *>(a_qstats) : nullptr;
	// Pageable point queries and small batches (DuckDB hands Scan a pageable float* and a std::vector to fill,
	// core/ISearcher.hpp:30-34): the query is copied into this state's pinned, device-mapped block with a host memcpy and
	// the kernels read it and write the result there in place - no cudaMemcpy call on either side of the launch.
	const uint64_t out_bytes = nq * ((uint64_t)k * 12 + 12);
	const bool pin_q = stage_q && nq * (uint64_t)s->cfg.dims * 4 <= PIN_Q_BYTES && (!stage_out || out_bytes <= PIN_OUT_BYTES);
	const bool pin_out = stage_out && out_bytes <= PIN_OUT_BYTES && (!stage_q || pin_q);
	if ((pin_q || pin_out) && !ss.h_pin) {
		LMD_CUDA_OK(cudaHostAlloc((void **)&ss.h_pin, PIN_Q_BYTES + PIN_OUT_BYTES, cudaHostAllocMapped | cudaHostAllocPortable));
	}
	int64_t *p_ids = nullptr;
	float *p_dist = nullptr;
	uint32_t *p_count = nullptr, *p_qstats = nullptr;
	if (pin_q) {
		memcpy(ss.h_pin, queries, nq * (uint64_t)s->cfg.dims * 4);
		void *dp = nullptr;
		LMD_CUDA_OK(cudaHostGetDevicePointer(&dp, ss.h_pin, 0));
		d_q = static_cast<const float *>(dp);
	} else if (stage_q) {
		const uint64_t need_q = nq * (uint64_t)s->cfg.dims;
		if ((rc = grow_to(ss.d_q, ss.scratch_q, need_q)) != LMD_OK) {
			return rc;
		}
		LMD_CUDA_OK(cudaMemcpyAsync(ss.d_q, queries, need_q * sizeof(float), cudaMemcpyHostToDevice, st));
		d_q = ss.d_q;
	}
	if (pin_out) {
		uint8_t *o = ss.h_pin + PIN_Q_BYTES;
		p_ids = reinterpret_cast<int64_t *>(o);
		p_dist = reinterpret_cast<float *>(o + nq * k * 8);
		p_count = reinterpret_cast<uint32_t *>(o + nq * k * 12);
		p_qstats = reinterpret_cast<uint32_t *>(o + nq * k * 12 + nq * 4);
		void *dp = nullptr;
		LMD_CUDA_OK(cudaHostGetDevicePointer(&dp, o, 0));
		uint8_t *d = static_cast<uint8_t *>(dp);
		d_ids = reinterpret_cast<int64_t *>(d);
		d_dist = reinterpret_cast<float *>(d + nq * k * 8);
		d_count = reinterpret_cast<uint32_t *>(d + nq * k * 12);
		d_qstats = want_qstats ? reinterpret_cast<uint32_t *>(d + nq * k * 12 + nq * 4) : nullptr;
	} else if (stage_out) {
		const uint64_t need_k = nq * (uint64_t)k;
		if ((rc = grow_to(ss.d_out_ids, ss.scratch_k, need_k)) != LMD_OK) {
			return rc;
		}
		uint64_t have_d = ss.d_out_dist ? ss.scratch_k : 0; // ids and distances grow together
		if ((rc = grow_to(ss.d_out_dist, have_d, ss.scratch_k)) != LMD_OK ||
		    (rc = grow_to(ss.d_out_count, ss.scratch_cnt, nq)) != LMD_OK ||
		    (want_qstats && (rc = grow_to(ss.d_qstats, ss.scratch_qs, 2 * nq)) != LMD_OK)) {
			ss.scratch_k = 0;
			return rc;
		}
		d_ids = ss.d_out_ids;
		d_dist = ss.d_out_dist; // always computed on the device; copied back only when asked for
		d_count = ss.d_out_count;
		d_qstats = want_qstats ? ss.d_qstats : nullptr;
	}
	uint64_t *d_cand = nullptr;
	if (separate) {
		if ((rc = grow_to(ss.d_cand, ss.scratch_cand, nq * (uint64_t)l_used)) != LMD_OK) {
			return rc;
		}
		d_cand = ss.d_cand;
	}
	LMD_CUDA_OK(cudaMemsetAsync(ss.d_counters, 0, 8 * sizeof(unsigned long long), st));

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
	ta.strict = (flags & LMD_SEARCH_STRICT_FP) != 0;
	ta.no_prefilter = (flags & LMD_SEARCH_NO_PREFILTER) != 0;
	ta.d_counters = ss.d_counters;
	uint32_t kroute_used = route;
	ta.kroute_out = &kroute_used;
	if (timed) {
		LMD_CUDA_OK(cudaEventRecord(ss.ev[1], st));
	}
	if ((rc = launch_traverse(ta, st)) != LMD_OK) {
		return rc;
	}
	if (timed) {
		LMD_CUDA_OK(cudaEventRecord(ss.ev[2], st));
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
		ra.strict = (flags & LMD_SEARCH_STRICT_FP) != 0;
		ra.tensor = (flags & LMD_SEARCH_TENSOR_RERANK) != 0;
		ra.d_counters = ss.d_counters;
		if ((rc = launch_rerank(ra, st)) != LMD_OK) {
			return rc;
		}
		launches++;
	}
	if (async) {
		return LMD_OK;
	}
	if (timed) {
		LMD_CUDA_OK(cudaEventRecord(ss.ev[3], st));
	}
	if (stage_out && !pin_out) {
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
		LMD_CUDA_OK(cudaMemcpyAsync(h_cnt, ss.d_counters, sizeof(h_cnt), cudaMemcpyDeviceToHost, st));
	}
	if (timed) {
		LMD_CUDA_OK(cudaEventRecord(ss.ev[4], st));
	}
	cudaError_t e = cudaStreamSynchronize(st);
	if (e != cudaSuccess) {
		set_error("search kernels failed: %s", cudaGetErrorString(e));
		return LMD_ERR_CUDA;
	}
	if (pin_out) {
		memcpy(out_ids, p_ids, nq * k * sizeof(int64_t));
		if (out_dist) {
			memcpy(out_dist, p_dist, nq * k * sizeof(float));
		}
		if (out_count) {
			memcpy(out_count, p_count, nq * sizeof(uint32_t));
		}
		if (want_qstats) {
			memcpy(out_qstats, p_qstats, nq * 2 * sizeof(uint32_t));
		}
	}
	if (stats) {
		stats->queries = nq;
		stats->expansions = h_cnt[1];
		stats->distance_evals = h_cnt[2];
		stats->filter_evictions = h_cnt[3];
		stats->reranked = h_cnt[4];
		stats->tie_overflows = h_cnt[5];
		stats->merge_fallbacks = h_cnt[6];
		stats->exact_fetches = h_cnt[7];
		stats->launches = launches;
		stats->l_used = l_used;
		cudaEventElapsedTime(&stats->traverse_ms, ss.ev[1], ss.ev[2]);
		if (separate) {
			cudaEventElapsedTime(&stats->rerank_ms, ss.ev[2], ss.ev[3]);
		}
		cudaEventElapsedTime(&stats->total_ms, ss.ev[0], ss.ev[4]);
		if (stage_q) {
			cudaEventElapsedTime(&stats->h2d_ms, ss.ev[0], ss.ev[1]);
		}
		if (stage_out) {
			cudaEventElapsedTime(&stats->d2h_ms, ss.ev[3], ss.ev[4]);
		}
		// SURVEY.md section 8d. The sidecar holds uint32 slots, so an expansion reads 4R bytes of ids
		// (the reference's 8 + 8R becomes 4R); padding is never counted.
		const uint64_t vec = (uint64_t)s->cfg.dims * 4;
		uint64_t bytes = nq * ((uint64_t)s->cfg.dims * 4 + (uint64_t)k * 12);
		const bool prefiltered = kroute_used == 3; // LMD_KROUTE_EXACT_PF: launch_traverse decides (store, flags, batch size)
		if (route == LMD_ROUTE_INT8 || prefiltered) {
			// one code block per expansion (neighbour slots + code norms + R int8 codes), then an exact row per fetched node
			bytes += h_cnt[1] * (uint64_t)s->code_block_bytes + h_cnt[7] * vec + (route == LMD_ROUTE_INT8 ? vec * nq : 0);
		} else if (route == LMD_ROUTE_EXACT) {
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
	if (!s) {
		set_error("lmd_search_batch: null store");
		return LMD_ERR_INVALID_ARG;
	}
	return search_impl(s, s->def, nullptr, queries, nq, k, L, route, flags, out_ids, out_dist, out_count, out_qstats, stats);
}

lmd_status lmd_search_batch_filtered(lmd_store *s, const lmd_filter *filter, const float *queries, uint64_t nq,
                                     uint32_t k, uint32_t L, uint32_t route, uint32_t flags, int64_t *out_ids,
                                     float *out_dist, uint32_t *out_count, uint32_t *out_qstats, lmd_stats *stats) {
	if (!filter) {
		set_error("lmd_search_batch_filtered: null filter");
		return LMD_ERR_INVALID_ARG;
	}
	if (!s) {
		set_error("lmd_search_batch_filtered: null store");
		return LMD_ERR_INVALID_ARG;
	}
	return search_impl(s, s->def, filter, queries, nq, k, L, route, flags, out_ids, out_dist, out_count, out_qstats, stats);
}

// ---- per-caller search contexts ----------------------------------------------------------------------------------
lmd_status lmd_ctx_create(lmd_store *s, lmd_ctx **out) {
	if (!s || !out) {
		set_error("lmd_ctx_create: null argument");
		return LMD_ERR_INVALID_ARG;
	}
	*out = nullptr;
	LMD_CUDA_OK(cudaSetDevice(s->device));
	lmd_ctx *c = new (std::nothrow) lmd_ctx();
	if (!c) {
		set_error("host allocation failed");
		return LMD_ERR_OOM;
	}
	c->store = s;
	lmd_status rc = state_init(c->st);
	if (rc != LMD_OK) {
		state_free(c->st);
		delete c;
		return rc;
	}
	cudaStreamSynchronize(c->st.stream);
	*out = c;
	return LMD_OK;
}

lmd_status lmd_ctx_destroy(lmd_ctx *c) {
	if (!c) {
		return LMD_OK;
	}
	cudaSetDevice(c->store->device);
	{
		std::lock_guard<std::mutex> g(c->st.mu); // a search still running on another thread finishes first
	}
	state_free(c->st);
	delete c;
	return LMD_OK;
}

lmd_status lmd_ctx_set_stream(lmd_ctx *c, void *cuda_stream) {
	if (!c) {
		set_error("lmd_ctx_set_stream: null context");
		return LMD_ERR_INVALID_ARG;
	}
	std::lock_guard<std::mutex> g(c->st.mu);
	LMD_CUDA_OK(cudaSetDevice(c->store->device));
	LMD_CUDA_OK(cudaStreamSynchronize(c->st.stream));
	c->st.stream = cuda_stream ? (cudaStream_t)cuda_stream : c->st.own_stream;
	return LMD_OK;
}

lmd_status lmd_ctx_search_batch(lmd_ctx *c, const lmd_filter *filter, const float *queries, uint64_t nq, uint32_t k,
                                uint32_t L, uint32_t route, uint32_t flags, int64_t *out_ids, float *out_dist,
                                uint32_t *out_count, uint32_t *out_qstats, lmd_stats *stats) {
	if (!c) {
		set_error("lmd_ctx_search_batch: null context");
		return LMD_ERR_INVALID_ARG;
	}
	return search_impl(c->store, c->st, filter, queries, nq, k, L, route, flags, out_ids, out_dist, out_count, out_qstats,
	                   stats);
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
