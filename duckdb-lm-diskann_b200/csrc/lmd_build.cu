// Graph write side, host orchestration + the small kernels around prune_kernel (lmd_build_kernels.cuh).
//
// lmd_build_graph is a batch-parallel form of the reference's insertion loop (db/LmDiskannIndex.cpp:350-376 Append ->
// core/GraphManager.cpp:65-203 AddNode, one row at a time): round t inserts the next b_t nodes against the graph of all
// earlier rounds (b_t doubles up to max_batch). Per round, three launches:
//   1. traverse_kernel  L = l_insert candidates per new node (Searcher.cpp:275-294 SearchForInitialCandidates)
//   2. prune_kernel     RobustPrune of each new node over its candidates (:140), neighbour row written (:143-157),
//                       and the node pushed into the inbox of every neighbour it selected
//   3. prune_kernel     each touched neighbour re-pruned over its own row + its inbox (:171-200)
// and, once at the end, the blocks are brought up to date: count, row ids, INT64_MAX padding, and the ternary planes of
// every neighbour (core/ternary_quantization.hpp:146-199 EncodeTernary; the SoA plane arrays of index_config.cpp:136-143).
#include "lmd_build_kernels.cuh"

#include <algorithm>
#include <cstring>
#include <vector>

namespace lmd {

// Defined by lmd_kernel_inst.cu (-DLMD_INST_KIND=2), one object file per metric.
lmd_status dispatch_prune_m1(PruneParams &p, const lmd_store *s, cudaStream_t st);
lmd_status dispatch_prune_m2(PruneParams &p, const lmd_store *s, cudaStream_t st);
lmd_status dispatch_prune_m3(PruneParams &p, const lmd_store *s, cudaStream_t st);

static lmd_status dispatch_prune(PruneParams &p, const lmd_store *s, cudaStream_t st) {
	switch (s->cfg.metric) {
	case LMD_METRIC_L2:
		return dispatch_prune_m1(p, s, st);
	case LMD_METRIC_COSINE:
		return dispatch_prune_m2(p, s, st);
	case LMD_METRIC_IP:
		return dispatch_prune_m3(p, s, st);
	}
	set_error("prune: unsupported metric %u", s->cfg.metric);
	return LMD_ERR_UNSUPPORTED;
}

// One warp per node: a fresh block holding only the vector (GraphManager.cpp:81-108 + GraphNode.cpp:12-24):
// zeroed, count 0, vector stored as the node type (memcpy, or a plain cast to int8), ids INT64_MAX (:152-157).
__global__ void vectors_to_blocks_kernel(const float *__restrict__ vectors, uint64_t n, uint32_t dims, uint32_t vec_type,
                                         uint32_t r, uint64_t vector_off, uint64_t ids_off, uint64_t block_bytes,
                                         uint8_t *__restrict__ blocks) {
	const uint64_t node = (blockIdx.x * (uint64_t)blockDim.x + threadIdx.x) >> 5;
	const uint32_t lane = threadIdx.x & 31;
	if (node >= n) {
		return;
	}
	uint8_t *blk = blocks + node * block_bytes;
	uint4 *b4 = reinterpret_cast<uint4 *>(blk);
	for (uint64_t i = lane; i < block_bytes / 16; i += 32) {
		b4[i] = make_uint4(0u, 0u, 0u, 0u);
	}
	__syncwarp();
	const float *src = vectors + node * dims;
	for (uint32_t i = lane; i < dims; i += 32) {
		if (vec_type == LMD_VEC_FLOAT32) {
			reinterpret_cast<float *>(blk + vector_off)[i] = src[i];
		} else {
			reinterpret_cast<int8_t *>(blk + vector_off)[i] = (int8_t)src[i];
		}
	}
	for (uint32_t j = lane; j < r; j += 32) {
		reinterpret_cast<int64_t *>(blk + ids_off)[j] = INT64_MAX;
	}
}

// One warp per node: EncodeTernary (core/ternary_quantization.hpp:146-199) of the node's fp32 vector.
// planes[node] = [words of the positive plane | words of the negative plane]; bit d%64 of word d/64.
__global__ void encode_planes_kernel(const float *__restrict__ vecs, uint32_t row_floats, uint32_t dims, uint32_t words,
                                     uint64_t n, uint64_t *__restrict__ planes) {
	const uint64_t node = (blockIdx.x * (uint64_t)blockDim.x + threadIdx.x) >> 5;
	const uint32_t lane = threadIdx.x & 31;
	if (node >= n) {
		return;
	}
	const float *v = vecs + node * (uint64_t)row_floats;
	uint32_t *dst = reinterpret_cast<uint32_t *>(planes + node * 2ull * words);
	for (uint32_t g = 0; g < words * 2; g++) { // 32 dimensions per step = half a word
		const uint32_t d = g * 32 + lane;
		const float x = d < dims ? v[d] : 0.0f;
		const uint32_t bp = __ballot_sync(0xFFFFFFFFu, x > 0.0f);
		const uint32_t bn = __ballot_sync(0xFFFFFFFFu, x < 0.0f);
		if (lane == 0) {
			dst[g] = bp;             // little-endian halves of the u64 words
			dst[words * 2 + g] = bn;
		}
	}
}

// One thread per (node, j): brings slot j of the node's block up to date from the neighbour row.
__global__ void finalize_blocks_kernel(uint8_t *__restrict__ blocks, uint64_t block_bytes, uint64_t count_off,
                                       uint64_t ids_off, uint64_t pos_off, uint64_t neg_off, uint32_t r, uint32_t words,
                                       uint64_t n, const uint32_t *__restrict__ nbrs, const int64_t *__restrict__ row_ids,
                                       const uint64_t *__restrict__ planes) {
	const uint64_t t = blockIdx.x * (uint64_t)blockDim.x + threadIdx.x;
	const uint64_t node = t / r;
	const uint32_t j = (uint32_t)(t % r);
	if (node >= n) {
		return;
	}
	uint8_t *blk = blocks + node * block_bytes;
	const uint32_t nb = nbrs[node * r + j];
	uint64_t *pp = reinterpret_cast<uint64_t *>(blk + pos_off) + (uint64_t)j * words;
	uint64_t *pn = reinterpret_cast<uint64_t *>(blk + neg_off) + (uint64_t)j * words;
	if (nb != LMD_INVALID_SLOT) {
		reinterpret_cast<int64_t *>(blk + ids_off)[j] = row_ids[nb];
		const uint64_t *src = planes + (uint64_t)nb * 2ull * words;
		for (uint32_t w = 0; w < words; w++) {
			pp[w] = src[w];
			pn[w] = src[words + w];
		}
	} else {
		reinterpret_cast<int64_t *>(blk + ids_off)[j] = INT64_MAX; // GraphManager.cpp:449-452
		for (uint32_t w = 0; w < words; w++) {
			pp[w] = 0ull;
			pn[w] = 0ull;
		}
	}
	if (j == 0) {
		// neighbour rows are dense from the front (prune_kernel writes them that way)
		uint32_t cnt = 0;
		while (cnt < r && nbrs[node * r + cnt] != LMD_INVALID_SLOT) {
			cnt++;
		}
		*reinterpret_cast<uint16_t *>(blk + count_off) = (uint16_t)cnt;
	}
}

__device__ __forceinline__ uint64_t bmix64(uint64_t x) { // the row-id map's hash (lmd_store.cu mix64)
	x ^= x >> 33;
	x *= 0xff51afd7ed558ccdULL;
	x ^= x >> 33;
	x *= 0xc4ceb9fe1a85ec53ULL;
	x ^= x >> 33;
	return x;
}

// row ids -> slots through the store's map (core/GraphManager.cpp:523-531 TryGetNodePointer); unknown -> invalid.
__global__ void resolve_rows_kernel(const int64_t *__restrict__ rows, uint64_t m, int identity, uint64_t n,
                                    const long long *__restrict__ map_keys, const uint32_t *__restrict__ map_vals,
                                    uint64_t map_mask, uint32_t *__restrict__ slots) {
	const uint64_t i = blockIdx.x * (uint64_t)blockDim.x + threadIdx.x;
	if (i >= m) {
		return;
	}
	const int64_t rid = rows[i];
	uint32_t out = LMD_INVALID_SLOT;
	if (rid != INT64_MAX) {
		if (identity) {
			out = (rid >= 0 && (uint64_t)rid < n) ? (uint32_t)rid : LMD_INVALID_SLOT;
		} else {
			uint64_t h = bmix64((uint64_t)rid) & map_mask;
			for (;;) {
				const long long k = map_keys[h];
				if (k == (long long)rid) {
					out = map_vals[h];
					break;
				}
				if (k == (long long)0x8000000000000000LL) {
					break;
				}
				h = (h + 1) & map_mask;
			}
		}
	}
	slots[i] = out;
}

__global__ void slots_to_rows_kernel(const uint32_t *__restrict__ slots, uint64_t m, const int64_t *__restrict__ row_ids,
                                     int64_t *__restrict__ rows) {
	const uint64_t i = blockIdx.x * (uint64_t)blockDim.x + threadIdx.x;
	if (i < m) {
		rows[i] = slots[i] == LMD_INVALID_SLOT ? INT64_MAX : row_ids[slots[i]];
	}
}

static uint32_t prune_smem_bytes(const lmd_store *s, uint32_t list_cap) {
	const bool generic = s->tune.strict_fp || !(s->row_floats == 128 || s->row_floats == 96 || s->row_floats == 768);
	const uint32_t rpad = (s->cfg.r + 31) & ~31u;
	return (uint32_t)(((size_t)list_cap * 8 + 384 + (size_t)rpad * 4 + (generic ? (size_t)s->row_floats * 4 : 0) + 15) /
	                  16 * 16);
}

static PruneParams prune_base(const lmd_store *s, float alpha, uint32_t list_cap, unsigned long long *counters) {
	PruneParams p;
	memset(&p, 0, sizeof(p));
	p.kp.vecs = s->d_vecs;
	p.kp.row_floats = s->row_floats;
	p.kp.dims = s->cfg.dims;
	p.kp.metric = s->cfg.metric;
	p.kp.strict = s->tune.strict_fp ? 1u : 0u;
	p.alpha = alpha;
	p.r = s->cfg.r;
	p.nbrs = s->d_nbrs;
	p.counters = counters;
	p.list_cap = list_cap;
	p.warp_smem_bytes = prune_smem_bytes(s, list_cap);
	return p;
}

} // namespace lmd

using namespace lmd;

namespace {

struct DevBuf { // frees on scope exit
	void *p = nullptr;
	~DevBuf() {
		if (p) {
			cudaFree(p);
		}
	}
	lmd_status alloc(size_t bytes, const char *what) {
		cudaError_t e = cudaMalloc(&p, bytes ? bytes : 16);
		if (e != cudaSuccess) {
			p = nullptr;
			cudaGetLastError();
			set_error("cudaMalloc(%zu bytes) for %s failed: %s", bytes, what, cudaGetErrorString(e));
			return LMD_ERR_OOM;
		}
		return LMD_OK;
	}
};

template <typename T>
lmd_status ensure(T *&ptr, size_t elems, lmd_store *s, const char *what) {
	if (ptr) {
		return LMD_OK;
	}
	cudaError_t e = cudaMalloc((void **)&ptr, elems * sizeof(T));
	if (e != cudaSuccess) {
		ptr = nullptr;
		cudaGetLastError();
		set_error("cudaMalloc(%zu bytes) for %s failed: %s", elems * sizeof(T), what, cudaGetErrorString(e));
		return LMD_ERR_OOM;
	}
	s->device_bytes += elems * sizeof(T);
	return LMD_OK;
}

} // namespace

extern "C" {

lmd_status lmd_store_upload_vectors(lmd_store *s, const int64_t *row_ids, const float *vectors, uint64_t n) {
	if (!s || (!vectors && n)) {
		set_error("lmd_store_upload_vectors: null argument");
		return LMD_ERR_INVALID_ARG;
	}
	if (n == 0) {
		return LMD_OK;
	}
	std::unique_lock<std::shared_mutex> g(s->mu);
	if (s->blocks_released) {
		set_error("lmd_store_upload_vectors: the block mirror of this store was released (lmd_store_release_blocks)");
		return LMD_ERR_STATE;
	}
	if (s->n + n > s->capacity) {
		set_error("lmd_store_upload_vectors: %llu + %llu nodes exceed the store capacity %llu", (unsigned long long)s->n,
		          (unsigned long long)n, (unsigned long long)s->capacity);
		return LMD_ERR_INVALID_ARG;
	}
	LMD_CUDA_OK(cudaSetDevice(s->device));
	const uint64_t bb = s->cfg.block_bytes, dims = s->cfg.dims;
	const uint64_t chunk = std::max<uint64_t>(1, std::min<uint64_t>(n, (256ull << 20) / (dims * 4)));
	DevBuf tmp;
	lmd_status rc = tmp.alloc(chunk * dims * 4, "vector staging");
	if (rc != LMD_OK) {
		return rc;
	}
	const uint64_t base = s->n;
	for (uint64_t done = 0; done < n; done += chunk) {
		const uint64_t now = std::min(chunk, n - done);
		rc = store_copy_to_device(s, tmp.p, vectors + done * dims, now * dims * 4);
		if (rc != LMD_OK) {
			return rc;
		}
		vectors_to_blocks_kernel<<<(unsigned)((now * 32 + 255) / 256), 256, 0, s->stream>>>(
		    (const float *)tmp.p, now, (uint32_t)dims, s->cfg.vec_type, s->cfg.r, s->lay.vector_off, s->lay.ids_off, bb,
		    s->d_blocks + (base + done) * bb);
		LMD_CUDA_OK(cudaGetLastError());
		LMD_CUDA_OK(cudaStreamSynchronize(s->stream)); // the staging buffer is reused by the next chunk
	}
	return store_ingest(s, row_ids, base, n);
}

lmd_status lmd_store_download_blocks(lmd_store *s, uint64_t first_slot, uint64_t n, void *out_blocks) {
	if (!s || (!out_blocks && n)) {
		set_error("lmd_store_download_blocks: null argument");
		return LMD_ERR_INVALID_ARG;
	}
	std::unique_lock<std::shared_mutex> g(s->mu);
	if (s->blocks_released) {
		set_error("lmd_store_download_blocks: the block mirror of this store was released (lmd_store_release_blocks)");
		return LMD_ERR_STATE;
	}
	if (first_slot + n > s->n) {
		set_error("lmd_store_download_blocks: slots [%llu, %llu) are beyond the %llu mirrored nodes",
		          (unsigned long long)first_slot, (unsigned long long)(first_slot + n), (unsigned long long)s->n);
		return LMD_ERR_INVALID_ARG;
	}
	if (n == 0) {
		return LMD_OK;
	}
	LMD_CUDA_OK(cudaSetDevice(s->device));
	LMD_CUDA_OK(cudaMemcpyAsync(out_blocks, s->d_blocks + first_slot * s->cfg.block_bytes, n * s->cfg.block_bytes,
	                            cudaMemcpyDefault, s->stream));
	LMD_CUDA_OK(cudaStreamSynchronize(s->stream));
	return LMD_OK;
}

lmd_status lmd_store_download_row_ids(lmd_store *s, uint64_t first_slot, uint64_t n, int64_t *out_row_ids) {
	if (!s || (!out_row_ids && n)) {
		set_error("lmd_store_download_row_ids: null argument");
		return LMD_ERR_INVALID_ARG;
	}
	std::shared_lock<std::shared_mutex> g(s->mu);
	if (first_slot + n > s->n) {
		set_error("lmd_store_download_row_ids: slots [%llu, %llu) are beyond the %llu mirrored nodes",
		          (unsigned long long)first_slot, (unsigned long long)(first_slot + n), (unsigned long long)s->n);
		return LMD_ERR_INVALID_ARG;
	}
	if (n == 0) {
		return LMD_OK;
	}
	LMD_CUDA_OK(cudaSetDevice(s->device));
	LMD_CUDA_OK(cudaMemcpy(out_row_ids, s->d_row_ids + first_slot, n * sizeof(int64_t), cudaMemcpyDefault));
	return LMD_OK;
}

lmd_status lmd_robust_prune_batch(lmd_store *s, const int64_t *node_row_ids, const int64_t *cand_row_ids, uint64_t m,
                                  uint32_t c, float alpha, int64_t *out_ids, uint32_t *out_count) {
	if (!s || !node_row_ids || !cand_row_ids || !out_ids || !out_count) {
		set_error("lmd_robust_prune_batch: null argument");
		return LMD_ERR_INVALID_ARG;
	}
	if (m == 0) {
		return LMD_OK;
	}
	if (c == 0 || c > 512 || s->cfg.r > 512 || m > 0x7FFFFFFFull) {
		set_error("lmd_robust_prune_batch: candidate count must be in [1, 512], R <= 512");
		return LMD_ERR_UNSUPPORTED;
	}
	std::unique_lock<std::shared_mutex> g(s->mu);
	LMD_CUDA_OK(cudaSetDevice(s->device));
	cudaStream_t st = s->stream;
	const uint32_t r = s->cfg.r;
	DevBuf d_rows, d_slots, d_out, d_cnt, d_counters, d_outrows;
	lmd_status rc;
	if ((rc = d_rows.alloc((m + m * c) * 8, "prune row ids")) != LMD_OK ||
	    (rc = d_slots.alloc((m + m * c) * 4, "prune slots")) != LMD_OK ||
	    (rc = d_out.alloc(m * r * 4, "prune output")) != LMD_OK || (rc = d_cnt.alloc(m * 4, "prune counts")) != LMD_OK ||
	    (rc = d_outrows.alloc(m * r * 8, "prune output rows")) != LMD_OK ||
	    (rc = d_counters.alloc(64, "prune counters")) != LMD_OK) {
		return rc;
	}
	LMD_CUDA_OK(cudaMemcpyAsync(d_rows.p, node_row_ids, m * 8, cudaMemcpyDefault, st));
	LMD_CUDA_OK(cudaMemcpyAsync((int64_t *)d_rows.p + m, cand_row_ids, m * c * 8, cudaMemcpyDefault, st));
	LMD_CUDA_OK(cudaMemsetAsync(d_counters.p, 0, 64, st));
	const uint64_t tot = m + m * c;
	resolve_rows_kernel<<<(unsigned)((tot + 255) / 256), 256, 0, st>>>(
	    (const int64_t *)d_rows.p, tot, s->identity ? 1 : 0, s->n, (const long long *)s->d_map_keys, s->d_map_vals,
	    s->map_cap - 1, (uint32_t *)d_slots.p);
	LMD_CUDA_OK(cudaGetLastError());
	// a node row id that does not resolve cannot be pruned: report it
	std::vector<uint32_t> h_nodes(m);
	LMD_CUDA_OK(cudaMemcpyAsync(h_nodes.data(), d_slots.p, m * 4, cudaMemcpyDeviceToHost, st));
	LMD_CUDA_OK(cudaStreamSynchronize(st));
	for (uint64_t i = 0; i < m; i++) {
		if (h_nodes[i] == LMD_INVALID_SLOT) {
			set_error("lmd_robust_prune_batch: node row id %lld is not in the store", (long long)node_row_ids[i]);
			return LMD_ERR_INVALID_ARG;
		}
	}
	PruneParams p = prune_base(s, alpha, (c + 31) & ~31u, (unsigned long long *)d_counters.p);
	p.n_items = (uint32_t)m;
	p.item_node = (const uint32_t *)d_slots.p;
	p.cand_slots = (const uint32_t *)d_slots.p + m;
	p.cand_stride = c;
	p.out_slots = (uint32_t *)d_out.p;
	p.out_count = (uint32_t *)d_cnt.p;
	if ((rc = dispatch_prune(p, s, st)) != LMD_OK) {
		return rc;
	}
	slots_to_rows_kernel<<<(unsigned)((m * r + 255) / 256), 256, 0, st>>>((const uint32_t *)d_out.p, m * r, s->d_row_ids,
	                                                                      (int64_t *)d_outrows.p);
	LMD_CUDA_OK(cudaGetLastError());
	LMD_CUDA_OK(cudaMemcpyAsync(out_ids, d_outrows.p, m * r * 8, cudaMemcpyDefault, st));
	LMD_CUDA_OK(cudaMemcpyAsync(out_count, d_cnt.p, m * 4, cudaMemcpyDefault, st));
	cudaError_t e = cudaStreamSynchronize(st);
	if (e != cudaSuccess) {
		set_error("prune kernel failed: %s", cudaGetErrorString(e));
		return LMD_ERR_CUDA;
	}
	return LMD_OK;
}

lmd_status lmd_build_graph(lmd_store *s, const lmd_build_params *params, lmd_build_stats *stats) {
	if (stats) {
		memset(stats, 0, sizeof(*stats));
	}
	if (!s || !params) {
		set_error("lmd_build_graph: null argument");
		return LMD_ERR_INVALID_ARG;
	}
	const uint32_t r = s->cfg.r;
	uint32_t L = params->l_insert ? params->l_insert : 128; // core/index_config.hpp:33-38
	const float alpha = params->alpha > 0.0f ? params->alpha : 1.2f;
	if (L > 512 || r > 128) {
		set_error("lmd_build_graph: this build supports L_INSERT <= 512 and R <= 128 (got %u, %u)", L, r);
		return LMD_ERR_UNSUPPORTED;
	}
	std::unique_lock<std::shared_mutex> g(s->mu);
	if (s->blocks_released) {
		set_error("lmd_build_graph: the block mirror of this store was released (lmd_store_release_blocks)");
		return LMD_ERR_STATE;
	}
	const uint64_t n = s->n;
	if (n == 0) {
		return LMD_OK;
	}
	if (s->removed) {
		set_error("lmd_build_graph: %llu rows were removed from this store; rebuild into a fresh store",
		          (unsigned long long)s->removed);
		return LMD_ERR_UNSUPPORTED;
	}
	LMD_CUDA_OK(cudaSetDevice(s->device));
	cudaStream_t st = s->stream;
	const uint32_t inbox_cap = 2 * r;
	lmd_status rc;
	if (s->inbox_cap != inbox_cap && s->d_inbox) {
		cudaFree(s->d_inbox);
		s->d_inbox = nullptr;
	}
	s->inbox_cap = inbox_cap;
	if ((rc = ensure(s->d_inbox, s->capacity * inbox_cap, s, "build inbox")) != LMD_OK ||
	    (rc = ensure(s->d_inbox_cnt, s->capacity, s, "build inbox counters")) != LMD_OK ||
	    (rc = ensure(s->d_touched, s->capacity, s, "build touched list")) != LMD_OK ||
	    (rc = ensure(s->d_planes, s->capacity * 2ull * s->words, s, "ternary planes")) != LMD_OK) {
		return rc;
	}
	uint64_t max_batch = params->max_batch;
	if (max_batch == 0) {
		max_batch = std::min<uint64_t>(32768, std::max<uint64_t>(1024, n / 48));
	}
	DevBuf d_keys, d_bcnt;
	if ((rc = d_keys.alloc(max_batch * (uint64_t)L * 8, "build candidate lists")) != LMD_OK ||
	    (rc = d_bcnt.alloc(64, "build counters")) != LMD_OK) {
		return rc;
	}
	unsigned long long *bcnt = (unsigned long long *)d_bcnt.p;
	cudaEvent_t ev[4];
	for (auto &e : ev) {
		LMD_CUDA_OK(cudaEventCreate(&e));
	}
	auto drop_events = [&]() {
		for (auto &e : ev) {
			cudaEventDestroy(e);
		}
	};
	lmd_build_stats bs;
	memset(&bs, 0, sizeof(bs));
	bs.nodes = n;
	cudaEvent_t ev_all0 = s->ev[0], ev_all1 = s->ev[4];
	cudaEventRecord(ev_all0, st);

	// empty graph: no node has neighbours; the first node is the entry point (GraphManager.cpp:119-121)
	cudaMemsetAsync(s->d_nbrs, 0xFF, n * (uint64_t)r * sizeof(uint32_t), st);
	cudaMemsetAsync(s->d_inbox_cnt, 0, n * sizeof(uint32_t), st);
	int64_t row0 = 0;
	cudaMemcpyAsync(&row0, s->d_row_ids, sizeof(int64_t), cudaMemcpyDeviceToHost, st);
	if (cudaStreamSynchronize(st) != cudaSuccess) {
		drop_events();
		set_error("lmd_build_graph: initialisation failed: %s", cudaGetErrorString(cudaGetLastError()));
		return LMD_ERR_CUDA;
	}
	s->entry_row = row0;
	s->entry_slot = 0;
	s->has_entry = true;

	const uint32_t touched_cap = (uint32_t)(((uint64_t)3 * r + 31) & ~31ull); // own row (r) + inbox (2r)
	uint64_t done = 1;
	rc = LMD_OK;
	while (done < n && rc == LMD_OK) {
		const uint64_t b = std::min<uint64_t>(std::min<uint64_t>(done, max_batch), n - done);
		// 1. candidates: beam search of the new nodes' own vectors over the graph built so far
		cudaMemsetAsync(s->d_counters, 0, 8 * sizeof(unsigned long long), st);
		cudaMemsetAsync(bcnt, 0, 64, st);
		cudaEventRecord(ev[0], st);
		TraverseArgs ta;
		ta.s = s;
		ta.d_queries = s->d_vecs + done * (uint64_t)s->row_floats;
		ta.q_stride = s->row_floats;
		ta.nq = (uint32_t)b;
		ta.k = 1;
		ta.L = L;
		ta.route = LMD_ROUTE_EXACT;
		ta.d_out_ids = nullptr;
		ta.d_out_dist = nullptr;
		ta.d_out_count = nullptr;
		ta.d_qstats = nullptr;
		ta.d_cand = (uint64_t *)d_keys.p;
		if ((rc = launch_traverse(ta, st)) != LMD_OK) {
			break;
		}
		cudaEventRecord(ev[1], st);
		// 2. RobustPrune of the new nodes; each pushes itself into the inbox of the neighbours it chose
		PruneParams pn = prune_base(s, alpha, (L + 31) & ~31u, bcnt);
		pn.n_items = (uint32_t)b;
		pn.node_base = (uint32_t)done;
		pn.keys = (const uint64_t *)d_keys.p;
		pn.keys_stride = L;
		pn.scatter = 1;
		pn.inbox = nullptr; // new nodes have nothing in their own inbox ...
		pn.scatter_inbox = s->d_inbox; // ... they write into their neighbours'
		pn.inbox_cnt = s->d_inbox_cnt;
		pn.inbox_cap = inbox_cap;
		pn.touched = s->d_touched;
		if ((rc = dispatch_prune(pn, s, st)) != LMD_OK) {
			break;
		}
		cudaEventRecord(ev[2], st);
		unsigned long long h[8] = {0, 0, 0, 0, 0, 0, 0, 0}, hs[8] = {0, 0, 0, 0, 0, 0, 0, 0};
		cudaMemcpyAsync(h, bcnt, sizeof(h), cudaMemcpyDeviceToHost, st);
		cudaMemcpyAsync(hs, s->d_counters, sizeof(hs), cudaMemcpyDeviceToHost, st);
		if (cudaStreamSynchronize(st) != cudaSuccess) {
			set_error("lmd_build_graph: round at %llu nodes failed: %s", (unsigned long long)done,
			          cudaGetErrorString(cudaGetLastError()));
			rc = LMD_ERR_CUDA;
			break;
		}
		const uint64_t n_touched = h[1];
		bs.search_expansions += hs[1];
		bs.search_evals += hs[2];
		bs.prune_evals += h[2];
		bs.inbox_overflows += h[3];
		// 3. interconnect: every touched neighbour is re-pruned over its own row + the new nodes that chose it
		if (n_touched) {
			cudaMemsetAsync(bcnt, 0, 64, st);
			PruneParams pt = prune_base(s, alpha, touched_cap, bcnt);
			pt.n_items = (uint32_t)n_touched;
			pt.item_node = s->d_touched;
			pt.use_existing = 1;
			pt.inbox = s->d_inbox;
			pt.inbox_cnt = s->d_inbox_cnt;
			pt.inbox_cap = inbox_cap;
			if ((rc = dispatch_prune(pt, s, st)) != LMD_OK) {
				break;
			}
			bs.backlink_prunes += n_touched;
			bs.launches++;
		}
		cudaEventRecord(ev[3], st);
		unsigned long long h2[8] = {0, 0, 0, 0, 0, 0, 0, 0};
		cudaMemcpyAsync(h2, bcnt, sizeof(h2), cudaMemcpyDeviceToHost, st);
		if (cudaStreamSynchronize(st) != cudaSuccess) {
			set_error("lmd_build_graph: interconnect at %llu nodes failed: %s", (unsigned long long)done,
			          cudaGetErrorString(cudaGetLastError()));
			rc = LMD_ERR_CUDA;
			break;
		}
		if (n_touched) {
			bs.prune_evals += h2[2];
		}
		float ms = 0.0f;
		cudaEventElapsedTime(&ms, ev[0], ev[1]);
		bs.search_ms += ms;
		cudaEventElapsedTime(&ms, ev[1], ev[2]);
		bs.prune_ms += ms;
		cudaEventElapsedTime(&ms, ev[2], ev[3]);
		bs.backlink_ms += ms;
		bs.launches += 2;
		bs.rounds++;
		done += b;
	}
	if (rc == LMD_OK) {
		// blocks: count, row ids, INT64_MAX padding, ternary planes of every neighbour
		cudaEventRecord(ev[0], st);
		encode_planes_kernel<<<(unsigned)((n * 32 + 255) / 256), 256, 0, st>>>(s->d_vecs, s->row_floats, s->cfg.dims,
		                                                                       s->words, n, s->d_planes);
		const uint64_t total = n * (uint64_t)r;
		finalize_blocks_kernel<<<(unsigned)((total + 255) / 256), 256, 0, st>>>(
		    s->d_blocks, s->cfg.block_bytes, s->lay.count_off, s->lay.ids_off, s->lay.pos_off, s->lay.neg_off, r, s->words, n,
		    s->d_nbrs, s->d_row_ids, s->d_planes);
		cudaEventRecord(ev[1], st);
		cudaEventRecord(ev_all1, st);
		cudaError_t e = cudaStreamSynchronize(st);
		if (e != cudaSuccess || (e = cudaGetLastError()) != cudaSuccess) {
			set_error("lmd_build_graph: finalising the blocks failed: %s", cudaGetErrorString(e));
			rc = LMD_ERR_CUDA;
		} else {
			cudaEventElapsedTime(&bs.finalize_ms, ev[0], ev[1]);
			cudaEventElapsedTime(&bs.total_ms, ev_all0, ev_all1);
			bs.launches += 2;
			s->dirty = false; // d_nbrs IS the sidecar of the blocks just written; d_vecs is unchanged
			s->views_gen++;
		}
	}
	drop_events();
	if (stats) {
		*stats = bs;
	}
	return rc;
}

} // extern "C"
