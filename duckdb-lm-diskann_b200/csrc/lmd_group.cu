// lmd_group: several stores on several devices of ONE process searched as one (SURVEY.md section 8e behind the C ABI).
// The reference's host is a single multi-threaded process (DuckDB); core/ISearcher.hpp:30-37 is the seam this sits
// behind, so both multi-GPU modes must be reachable without torch.distributed or one process per GPU:
//
//   LMD_GROUP_QUERY_SHARDED    every member holds the whole index; member g searches queries [lo_g, hi_g) and its results
//                              land in that slice of the caller's [nq][k] buffers. No exchange step.
//   LMD_GROUP_DATASET_SHARDED  member s holds shard s (shard-local graph and entry point, GLOBAL row ids). Every member
//                              searches the full batch; then ONE kernel on the first member's device reads all G per-shard
//                              [nq][k] (distance, row id) lists IN PLACE - the other devices' lists through NVLink peer
//                              loads - and writes the k smallest per query: gather and merge fused, no staging copy and no
//                              collective library in between. Without peer access between two devices the lists are
//                              copied with cudaMemcpyPeerAsync first and the same kernel reads the copies.
//
// One host thread drives all devices: every member has its own lmd_ctx (stream, counters, scratch); all searches are
// enqueued before the first result is waited for, so the devices run side by side.
#include "lmd_common.cuh"

#include <algorithm>
#include <chrono>
#include <cstring>
#include <vector>

using namespace lmd;

#define LMD_GROUP_MAX 16

namespace {

struct MergeLists {
	const int64_t *ids[LMD_GROUP_MAX];
	const float *dist[LMD_GROUP_MAX];
};

// One warp per query. The G * k candidates (k <= 64, G <= 16) are loaded lane-strided from the members' lists (peer
// memory for the other devices), and every candidate finds its output rank by counting the candidates that sort before
// it by (distance, row id) - row ids are unique across shards, so ranks are too. -1 ids are padding.
__global__ void __launch_bounds__(128) merge_lists_kernel(const MergeLists in, uint32_t shards, uint64_t nq, uint32_t k,
                                                          int64_t *__restrict__ out_ids, float *__restrict__ out_dist,
                                                          uint32_t *__restrict__ out_count) {
	extern __shared__ __align__(16) uint8_t smem[];
	const int lane = threadIdx.x & 31, wib = threadIdx.x >> 5;
	const uint32_t total = shards * k;
	int64_t *s_id = reinterpret_cast<int64_t *>(smem) + (size_t)wib * total;
	float *s_d = reinterpret_cast<float *>(smem + (size_t)(blockDim.x >> 5) * total * 8) + (size_t)wib * total;
	const uint64_t warps = ((uint64_t)gridDim.x * blockDim.x) >> 5;
	for (uint64_t q = ((uint64_t)blockIdx.x * blockDim.x + threadIdx.x) >> 5; q < nq; q += warps) {
		for (uint32_t c = lane; c < total; c += 32) {
			const uint32_t s = c / k, j = c - s * k;
			const int64_t id = in.ids[s][q * k + j];
			s_id[c] = id;
			s_d[c] = id >= 0 ? in.dist[s][q * k + j] : 0.0f;
		}
		__syncwarp();
		uint32_t valid = 0;
		for (uint32_t c = lane; c < total; c += 32) {
			const int64_t id = s_id[c];
			if (id < 0) {
				continue;
			}
			valid++;
			const float d = s_d[c];
			uint32_t rank = 0;
			for (uint32_t o = 0; o < total; o++) {
				const int64_t oid = s_id[o];
				const float od = s_d[o];
				rank += (oid >= 0 && (od < d || (od == d && oid < id))) ? 1u : 0u;
			}
			if (rank < k) {
				out_ids[q * k + rank] = id;
				out_dist[q * k + rank] = d;
			}
		}
		valid = __reduce_add_sync(0xFFFFFFFFu, valid);
		for (uint32_t j = valid + lane; j < k; j += 32) { // fewer than k candidates in all shards together
			out_ids[q * k + j] = -1;
			out_dist[q * k + j] = 0.0f;
		}
		if (out_count && lane == 0) {
			out_count[q] = valid < k ? valid : k;
		}
		__syncwarp();
	}
}

template <typename T>
lmd_status grow_dev(T *&p, uint64_t &have, uint64_t need) {
	if (p && have >= need) {
		return LMD_OK;
	}
	if (p) {
		cudaFree(p);
		p = nullptr;
		have = 0;
	}
	cudaError_t e = cudaMalloc((void **)&p, need * sizeof(T));
	if (e != cudaSuccess) {
		cudaGetLastError();
		set_error("lmd_group: cudaMalloc(%llu bytes) failed: %s", (unsigned long long)(need * sizeof(T)), cudaGetErrorString(e));
		return LMD_ERR_OOM;
	}
	have = need;
	return LMD_OK;
}

struct Member {
	lmd_store *store = nullptr;
	lmd_ctx *ctx = nullptr;
	int device = 0;
	float *d_q = nullptr;
	int64_t *d_ids = nullptr;
	float *d_dist = nullptr;
	uint32_t *d_cnt = nullptr;
	uint64_t cap_q = 0, cap_ids = 0, cap_dist = 0, cap_cnt = 0;
	cudaEvent_t ev0 = nullptr, ev1 = nullptr;
	bool peer_of_first = true; // the first member's device can load from this member's memory
	// dataset sharding without peer access: copies of this member's lists on the first member's device
	int64_t *g_ids = nullptr;
	float *g_dist = nullptr;
	uint64_t cap_gids = 0, cap_gdist = 0;
};

} // namespace

struct lmd_group {
	uint32_t mode = 0;
	std::vector<Member> m;
	int64_t *d_out_ids = nullptr; // merged result on the first member's device
	float *d_out_dist = nullptr;
	uint32_t *d_out_cnt = nullptr;
	uint64_t cap_oids = 0, cap_odist = 0, cap_ocnt = 0;
	cudaEvent_t ev_m0 = nullptr, ev_m1 = nullptr;
	std::mutex mu;
};

extern "C" {

lmd_status lmd_group_create(lmd_store *const *stores, uint32_t n, uint32_t mode, lmd_group **out) {
	if (!stores || !out || n == 0 || n > LMD_GROUP_MAX) {
		set_error("lmd_group_create: need 1..%d stores", LMD_GROUP_MAX);
		return LMD_ERR_INVALID_ARG;
	}
	*out = nullptr;
	if (mode != LMD_GROUP_QUERY_SHARDED && mode != LMD_GROUP_DATASET_SHARDED) {
		set_error("lmd_group_create: unknown mode %u", mode);
		return LMD_ERR_INVALID_ARG;
	}
	for (uint32_t i = 0; i < n; i++) {
		if (!stores[i]) {
			set_error("lmd_group_create: store %u is null", i);
			return LMD_ERR_INVALID_ARG;
		}
		const lmd_config &a = stores[0]->cfg, &b = stores[i]->cfg;
		if (a.dims != b.dims || a.metric != b.metric) {
			set_error("lmd_group_create: store %u has another dimension / metric than store 0", i);
			return LMD_ERR_INVALID_ARG;
		}
		for (uint32_t j = 0; j < i; j++) {
			if (stores[j] == stores[i]) {
				set_error("lmd_group_create: store %u is listed twice", i);
				return LMD_ERR_INVALID_ARG;
			}
		}
	}
	lmd_group *g = new (std::nothrow) lmd_group();
	if (!g) {
		set_error("host allocation failed");
		return LMD_ERR_OOM;
	}
	g->mode = mode;
	g->m.resize(n);
	lmd_status rc = LMD_OK;
	for (uint32_t i = 0; i < n && rc == LMD_OK; i++) {
		Member &mb = g->m[i];
		mb.store = stores[i];
		mb.device = stores[i]->device;
		rc = lmd_ctx_create(stores[i], &mb.ctx);
		if (rc == LMD_OK && (cudaSetDevice(mb.device) != cudaSuccess || cudaEventCreate(&mb.ev0) != cudaSuccess ||
		                     cudaEventCreate(&mb.ev1) != cudaSuccess)) {
			set_error("lmd_group_create: event creation on device %d failed", mb.device);
			rc = LMD_ERR_CUDA;
		}
	}
	if (rc == LMD_OK) {
		const int d0 = g->m[0].device;
		cudaSetDevice(d0);
		if (cudaEventCreate(&g->ev_m0) != cudaSuccess || cudaEventCreate(&g->ev_m1) != cudaSuccess) {
			set_error("lmd_group_create: event creation failed");
			rc = LMD_ERR_CUDA;
		}
		for (uint32_t i = 1; i < n && mode == LMD_GROUP_DATASET_SHARDED; i++) {
			Member &mb = g->m[i];
			if (mb.device == d0) {
				continue;
			}
			int can = 0;
			cudaDeviceCanAccessPeer(&can, d0, mb.device);
			if (can) {
				cudaError_t e = cudaDeviceEnablePeerAccess(mb.device, 0);
				if (e != cudaSuccess && e != cudaErrorPeerAccessAlreadyEnabled) {
					can = 0;
				}
				cudaGetLastError();
			}
			mb.peer_of_first = can != 0;
		}
	}
	if (rc != LMD_OK) {
		lmd_group_destroy(g);
		return rc;
	}
	*out = g;
	return LMD_OK;
}

lmd_status lmd_group_destroy(lmd_group *g) {
	if (!g) {
		return LMD_OK;
	}
	{
		std::lock_guard<std::mutex> lk(g->mu);
	}
	for (Member &mb : g->m) {
		if (mb.ctx) {
			lmd_ctx_destroy(mb.ctx);
		}
		cudaSetDevice(mb.device);
		cudaFree(mb.d_q);
		cudaFree(mb.d_ids);
		cudaFree(mb.d_dist);
		cudaFree(mb.d_cnt);
		if (mb.ev0) {
			cudaEventDestroy(mb.ev0);
		}
		if (mb.ev1) {
			cudaEventDestroy(mb.ev1);
		}
	}
	if (!g->m.empty()) {
		cudaSetDevice(g->m[0].device);
		for (Member &mb : g->m) {
			cudaFree(mb.g_ids);
			cudaFree(mb.g_dist);
		}
		cudaFree(g->d_out_ids);
		cudaFree(g->d_out_dist);
		cudaFree(g->d_out_cnt);
		if (g->ev_m0) {
			cudaEventDestroy(g->ev_m0);
		}
		if (g->ev_m1) {
			cudaEventDestroy(g->ev_m1);
		}
	}
	cudaGetLastError();
	delete g;
	return LMD_OK;
}

uint32_t lmd_group_size(const lmd_group *g) {
	return g ? (uint32_t)g->m.size() : 0;
}

lmd_status lmd_group_search_batch(lmd_group *g, const float *queries, uint64_t nq, uint32_t k, uint32_t L, uint32_t route,
                                  uint32_t flags, int64_t *out_ids, float *out_dist, uint32_t *out_count,
                                  lmd_group_stats *stats) {
	if (stats) {
		memset(stats, 0, sizeof(*stats));
	}
	if (!g) {
		set_error("lmd_group_search_batch: null group");
		return LMD_ERR_INVALID_ARG;
	}
	if (k == 0) { // core/Coordinator.cpp:72-74
		set_error("k_neighbors cannot be 0 for search");
		return LMD_ERR_INVALID_ARG;
	}
	if (nq == 0) {
		return LMD_OK;
	}
	if (!queries || !out_ids) {
		set_error("lmd_group_search_batch: null queries / out_ids");
		return LMD_ERR_INVALID_ARG;
	}
	if (flags & (LMD_SEARCH_ASYNC | LMD_SEARCH_PER_QUERY_STATS)) {
		set_error("lmd_group_search_batch: LMD_SEARCH_ASYNC / LMD_SEARCH_PER_QUERY_STATS are not available on groups");
		return LMD_ERR_INVALID_ARG;
	}
	const uint32_t G = (uint32_t)g->m.size();
	if (g->mode == LMD_GROUP_DATASET_SHARDED && (uint64_t)G * k > 1024) {
		set_error("lmd_group_search_batch: members * k = %llu exceeds the merge kernel's 1024 candidates per query",
		          (unsigned long long)((uint64_t)G * k));
		return LMD_ERR_UNSUPPORTED;
	}
	std::lock_guard<std::mutex> lk(g->mu);
	const auto t_host0 = std::chrono::steady_clock::now();
	const uint32_t dims = g->m[0].store->cfg.dims;
	const bool sharded = g->mode == LMD_GROUP_DATASET_SHARDED;
	lmd_status rc = LMD_OK;
	uint64_t lo[LMD_GROUP_MAX + 1];
	for (uint32_t i = 0; i <= G; i++) { // contiguous, balanced: the first nq % G members get one query more
		lo[i] = sharded ? (i == 0 ? 0 : nq) : i * (nq / G) + std::min<uint64_t>(i, nq % G);
	}
	// 1. every member: queries in, search enqueued (nothing is waited for yet)
	uint32_t issued = 0;
	for (uint32_t i = 0; i < G && rc == LMD_OK; i++, issued++) {
		Member &mb = g->m[i];
		const uint64_t q0 = sharded ? 0 : lo[i], nqi = sharded ? nq : lo[i + 1] - lo[i];
		if (nqi == 0) {
			continue;
		}
		if (cudaSetDevice(mb.device) != cudaSuccess) {
			set_error("lmd_group_search_batch: cudaSetDevice(%d) failed", mb.device);
			rc = LMD_ERR_CUDA;
			break;
		}
		if ((rc = grow_dev(mb.d_q, mb.cap_q, nqi * dims)) != LMD_OK || (rc = grow_dev(mb.d_ids, mb.cap_ids, nqi * k)) != LMD_OK ||
		    (rc = grow_dev(mb.d_dist, mb.cap_dist, nqi * k)) != LMD_OK || (rc = grow_dev(mb.d_cnt, mb.cap_cnt, nqi)) != LMD_OK) {
			break;
		}
		cudaStream_t st = mb.ctx->st.stream;
		cudaEventRecord(mb.ev0, st);
		if (cudaMemcpyAsync(mb.d_q, queries + q0 * dims, nqi * dims * sizeof(float), cudaMemcpyDefault, st) != cudaSuccess) {
			set_error("lmd_group_search_batch: copying the queries to device %d failed: %s", mb.device,
			          cudaGetErrorString(cudaGetLastError()));
			rc = LMD_ERR_CUDA;
			break;
		}
		rc = lmd_ctx_search_batch(mb.ctx, nullptr, mb.d_q, nqi, k, L, route, flags | LMD_SEARCH_ASYNC, mb.d_ids, mb.d_dist,
		                          mb.d_cnt, nullptr, nullptr);
		cudaEventRecord(mb.ev1, st);
	}
	// 2. results out
	if (rc == LMD_OK && !sharded) {
		for (uint32_t i = 0; i < G; i++) {
			Member &mb = g->m[i];
			const uint64_t nqi = lo[i + 1] - lo[i];
			if (nqi == 0) {
				continue;
			}
			cudaSetDevice(mb.device);
			cudaStream_t st = mb.ctx->st.stream;
			cudaMemcpyAsync(out_ids + lo[i] * k, mb.d_ids, nqi * k * sizeof(int64_t), cudaMemcpyDefault, st);
			if (out_dist) {
				cudaMemcpyAsync(out_dist + lo[i] * k, mb.d_dist, nqi * k * sizeof(float), cudaMemcpyDefault, st);
			}
			if (out_count) {
				cudaMemcpyAsync(out_count + lo[i], mb.d_cnt, nqi * sizeof(uint32_t), cudaMemcpyDefault, st);
			}
		}
	} else if (rc == LMD_OK) {
		Member &m0 = g->m[0];
		cudaSetDevice(m0.device);
		cudaStream_t st0 = m0.ctx->st.stream;
		MergeLists ml;
		memset(&ml, 0, sizeof(ml));
		for (uint32_t i = 0; i < G && rc == LMD_OK; i++) {
			Member &mb = g->m[i];
			ml.ids[i] = mb.d_ids;
			ml.dist[i] = mb.d_dist;
			if (i == 0) {
				continue;
			}
			cudaStreamWaitEvent(st0, mb.ev1, 0); // the merge follows every member's search
			if (!mb.peer_of_first) {
				if ((rc = grow_dev(mb.g_ids, mb.cap_gids, nq * k)) != LMD_OK || (rc = grow_dev(mb.g_dist, mb.cap_gdist, nq * k)) != LMD_OK) {
					break;
				}
				cudaMemcpyPeerAsync(mb.g_ids, m0.device, mb.d_ids, mb.device, nq * k * sizeof(int64_t), st0);
				cudaMemcpyPeerAsync(mb.g_dist, m0.device, mb.d_dist, mb.device, nq * k * sizeof(float), st0);
				ml.ids[i] = mb.g_ids;
				ml.dist[i] = mb.g_dist;
			}
		}
		if (rc == LMD_OK && ((rc = grow_dev(g->d_out_ids, g->cap_oids, nq * k)) != LMD_OK ||
		                     (rc = grow_dev(g->d_out_dist, g->cap_odist, nq * k)) != LMD_OK ||
		                     (rc = grow_dev(g->d_out_cnt, g->cap_ocnt, nq)) != LMD_OK)) {
			// fall through to the synchronisation below
		}
		if (rc == LMD_OK) {
			const uint32_t warps = 4;
			const size_t smem = (size_t)warps * G * k * 12;
			const unsigned grid = (unsigned)std::min<uint64_t>((nq + warps - 1) / warps, (uint64_t)m0.store->sm_count * 8);
			cudaEventRecord(g->ev_m0, st0);
			merge_lists_kernel<<<grid, warps * 32, smem, st0>>>(ml, G, nq, k, g->d_out_ids, g->d_out_dist, g->d_out_cnt);
			cudaEventRecord(g->ev_m1, st0);
			if (cudaGetLastError() != cudaSuccess) {
				set_error("lmd_group_search_batch: merge kernel launch failed");
				rc = LMD_ERR_CUDA;
			} else {
				cudaMemcpyAsync(out_ids, g->d_out_ids, nq * k * sizeof(int64_t), cudaMemcpyDefault, st0);
				if (out_dist) {
					cudaMemcpyAsync(out_dist, g->d_out_dist, nq * k * sizeof(float), cudaMemcpyDefault, st0);
				}
				if (out_count) {
					cudaMemcpyAsync(out_count, g->d_out_cnt, nq * sizeof(uint32_t), cudaMemcpyDefault, st0);
				}
			}
		}
	}
	// 3. wait for every member that was given work (also after a failure: nothing may still be running on return)
	for (uint32_t i = 0; i < std::min(issued + 1, G); i++) {
		Member &mb = g->m[i];
		cudaSetDevice(mb.device);
		cudaError_t e = cudaStreamSynchronize(mb.ctx->st.stream);
		if (e != cudaSuccess && rc == LMD_OK) {
			set_error("lmd_group_search_batch: device %d: %s", mb.device, cudaGetErrorString(e));
			rc = LMD_ERR_CUDA;
		}
	}
	if (rc == LMD_OK && stats) {
		stats->members = G;
		stats->mode = g->mode;
		for (uint32_t i = 0; i < G; i++) {
			Member &mb = g->m[i];
			stats->queries[i] = sharded ? nq : lo[i + 1] - lo[i];
			stats->peer_access |= (mb.peer_of_first ? 1u : 0u) << i;
			if (stats->queries[i]) {
				cudaSetDevice(mb.device);
				cudaEventElapsedTime(&stats->search_ms[i], mb.ev0, mb.ev1);
				stats->search_ms_max = std::max(stats->search_ms_max, stats->search_ms[i]);
			}
		}
		if (sharded) {
			cudaSetDevice(g->m[0].device);
			cudaEventElapsedTime(&stats->merge_ms, g->ev_m0, g->ev_m1);
			stats->merge_input_bytes = (uint64_t)G * nq * k * 12;
		}
		stats->wall_ms = std::chrono::duration<float, std::milli>(std::chrono::steady_clock::now() - t_host0).count();
	}
	cudaGetLastError();
	return rc;
}

} // extern "C"
