// Device node store: HBM mirror of the LM_DISKANN node blocks + the derived search views.
//
// Replaces, on the query path, core/IStorageManager.hpp:156 (GetNodeBlockData) and the
// std::map<row_t, IndexPointer> behind core/GraphManager.cpp:523-531 (TryGetNodePointer).
//
// HBM layout (per store, N = capacity):
//   d_blocks  [N][block_bytes]  raw blocks, reference layout untouched (core/index_config.cpp:104-148)
//   d_vecs    [N][row_floats]   node vectors as fp32, rows padded with zeros to 128 B multiples so that every
//                               gather is 16-byte aligned (the in-block vector sits at offset 8)
//   d_nbrs    [N][r] u32        neighbour SLOTS: ids[j] of the block resolved once through the row-id map,
//                               LMD_INVALID_SLOT for j >= count, INT64_MAX sentinels and unknown row ids
//                               (the three skip rules of GraphManager.cpp:241-254 and Searcher.cpp:132,146)
//   d_row_ids [N] i64           slot -> row id for the result write-out
#include "lmd_common.cuh"

#include <cstring>
#include <new>
#include <unordered_map>
#include <vector>

namespace lmd {

static thread_local std::string g_err;

void set_error(const char *fmt, ...) {
	char buf[1024];
	va_list ap;
	va_start(ap, fmt);
	vsnprintf(buf, sizeof(buf), fmt, ap);
	va_end(ap);
	g_err = buf;
}

__device__ __forceinline__ uint64_t mix64(uint64_t x) {
	x ^= x >> 33;
	x *= 0xff51afd7ed558ccdULL;
	x ^= x >> 33;
	x *= 0xc4ceb9fe1a85ec53ULL;
	x ^= x >> 33;
	return x;
}

#define LMD_MAP_EMPTY ((long long)0x8000000000000000LL) // INT64_MIN is never a DuckDB row id
#define LMD_MAP_TOMB ((long long)0x8000000000000001LL)  // a removed row id: probes walk over it, nothing reuses it

// One thread per uploaded node: record row id, insert row id -> slot.
// A row id seen twice keeps the LATEST slot (an update of the mapping, like std::map::operator[]).
__global__ void ingest_rowids_kernel(const int64_t *row_ids, uint64_t base, uint64_t n,
                                     int64_t *slot_row, long long *map_keys, uint32_t *map_vals,
                                     uint64_t map_mask) {
	uint64_t i = blockIdx.x * (uint64_t)blockDim.x + threadIdx.x;
	if (i >= n) {
		return;
	}
	int64_t rid = row_ids ? row_ids[i] : (int64_t)(base + i);
	slot_row[base + i] = rid;
	uint64_t h = mix64((uint64_t)rid) & map_mask;
	for (;;) {
		long long old = (long long)atomicCAS((unsigned long long *)&map_keys[h], (unsigned long long)LMD_MAP_EMPTY,
		                                     (unsigned long long)rid);
		if (old == LMD_MAP_EMPTY || old == (long long)rid) {
			atomicMax(&map_vals[h], (uint32_t)(base + i)); // later slot wins deterministically
			return;
		}
		h = (h + 1) & map_mask;
	}
}

__device__ __forceinline__ uint32_t map_lookup(const long long *__restrict__ map_keys,
                                               const uint32_t *__restrict__ map_vals, uint64_t map_mask, int64_t rid) {
	uint64_t h = mix64((uint64_t)rid) & map_mask;
	for (;;) {
		long long k = map_keys[h];
		if (k == (long long)rid) {
			return map_vals[h];
		}
		if (k == LMD_MAP_EMPTY) {
			return LMD_INVALID_SLOT;
		}
		h = (h + 1) & map_mask;
	}
}

// row ids -> slots, one thread each (LMD_INVALID_SLOT = not in the store)
__global__ void lookup_many_kernel(const long long *map_keys, const uint32_t *map_vals, uint64_t map_mask,
                                   const int64_t *rids, uint64_t n, uint32_t *out) {
	uint64_t i = blockIdx.x * (uint64_t)blockDim.x + threadIdx.x;
	if (i < n) {
		out[i] = map_lookup(map_keys, map_vals, map_mask, rids[i]);
	}
}

// core/GraphManager.cpp:476-499 (HandleNodeDeletion / FreeNode): `rowid_to_node_ptr_map_.erase(row_id)`.
// One thread per row id; counts the ids that were present.
__global__ void remove_rowids_kernel(long long *map_keys, const uint32_t *map_vals, uint64_t map_mask,
                                     const int64_t *rids, uint64_t n, int64_t *slot_row,
                                     unsigned long long *removed) {
	uint64_t i = blockIdx.x * (uint64_t)blockDim.x + threadIdx.x;
	if (i >= n) {
		return;
	}
	const long long rid = (long long)rids[i];
	uint64_t h = mix64((uint64_t)rid) & map_mask;
	for (;;) {
		long long k = map_keys[h];
		if (k == rid) {
			// two threads may carry the same row id: only the one that swaps the key counts it
			if ((long long)atomicCAS((unsigned long long *)&map_keys[h], (unsigned long long)rid,
			                         (unsigned long long)LMD_MAP_TOMB) == rid) {
				slot_row[map_vals[h]] = -1; // the slot stays allocated but is unreachable
				atomicAdd(removed, 1ull);
			}
			return;
		}
		if (k == LMD_MAP_EMPTY) {
			return;
		}
		h = (h + 1) & map_mask;
	}
}

// block i of the staging buffer -> slot slots[i] of the mirror (LMD_INVALID_SLOT = superseded inside the batch).
// One CTA per block, 16-byte pieces (block_bytes is a multiple of 8; the tail goes as 8-byte words).
__global__ void scatter_blocks_kernel(const uint8_t *__restrict__ stage, const uint32_t *__restrict__ slots,
                                      uint64_t block_bytes, uint8_t *__restrict__ blocks) {
	const uint32_t slot = slots[blockIdx.x];
	if (slot == LMD_INVALID_SLOT) {
		return;
	}
	const uint8_t *src = stage + (uint64_t)blockIdx.x * block_bytes;
	uint8_t *dst = blocks + (uint64_t)slot * block_bytes;
	if ((block_bytes & 15) == 0) {
		for (uint64_t o = threadIdx.x * 16ull; o < block_bytes; o += blockDim.x * 16ull) {
			*reinterpret_cast<uint4 *>(dst + o) = *reinterpret_cast<const uint4 *>(src + o);
		}
	} else {
		for (uint64_t o = threadIdx.x * 8ull; o < block_bytes; o += blockDim.x * 8ull) {
			*reinterpret_cast<uint2 *>(dst + o) = *reinterpret_cast<const uint2 *>(src + o);
		}
	}
}

// Filtered search: allowed row ids -> bitmap over slots. Unknown row ids are ignored; `count` = bits newly set.
__global__ void filter_bits_kernel(const long long *map_keys, const uint32_t *map_vals, uint64_t map_mask, int identity,
                                   uint64_t n_nodes, const int64_t *rids, uint64_t n, uint32_t *bits,
                                   unsigned long long *count) {
	uint64_t i = blockIdx.x * (uint64_t)blockDim.x + threadIdx.x;
	if (i >= n) {
		return;
	}
	const int64_t rid = rids[i];
	uint32_t slot;
	if (identity) {
		slot = (rid >= 0 && (uint64_t)rid < n_nodes) ? (uint32_t)rid : LMD_INVALID_SLOT;
	} else {
		slot = map_lookup(map_keys, map_vals, map_mask, rid);
	}
	if (slot == LMD_INVALID_SLOT) {
		return;
	}
	const uint32_t bit = 1u << (slot & 31);
	if (!(atomicOr(&bits[slot >> 5], bit) & bit)) {
		atomicAdd(count, 1ull);
	}
}

__global__ void lookup_one_kernel(const long long *map_keys, const uint32_t *map_vals, uint64_t map_mask, int64_t rid,
                                  uint32_t *out) {
	*out = map_lookup(map_keys, map_vals, map_mask, rid);
}

// One warp per node: unpack the node vector of block `slot` into the packed fp32 row (common/ann.hpp:42-62:
// FLOAT32 is a byte copy, INT8 is a plain static_cast<float>), zero the padding.
__global__ void unpack_vectors_kernel(const uint8_t *__restrict__ blocks, uint64_t block_bytes, uint64_t vector_off,
                                      uint32_t dims, uint32_t vec_type, uint32_t row_floats, uint64_t base, uint64_t n,
                                      float *__restrict__ vecs) {
	uint64_t node = (blockIdx.x * (uint64_t)blockDim.x + threadIdx.x) >> 5;
	uint32_t lane = threadIdx.x & 31;
	if (node >= n) {
		return;
	}
	const uint8_t *src = blocks + (base + node) * block_bytes + vector_off;
	float *dst = vecs + (base + node) * (uint64_t)row_floats;
	for (uint32_t i = lane; i < row_floats; i += 32) {
		float v = 0.0f;
		if (i < dims) {
			if (vec_type == LMD_VEC_FLOAT32) {
				v = reinterpret_cast<const float *>(src)[i]; // vector_off is 8-aligned, block base 256-aligned
			} else {
				v = (float)reinterpret_cast<const int8_t *>(src)[i];
			}
		}
		dst[i] = v;
	}
}

// A row id listed twice among the first `count` entries of one block: the reference evaluates the first occurrence and
// skips the second (it is in `visited_nodes` by then, core/Searcher.cpp:135-139), for every query alike - so the derived
// neighbour-slot view keeps the first occurrence only. The blocks themselves are never changed. (The reference's own
// RobustPrune and lmd_build_graph never write such rows; a host may.)
__device__ __forceinline__ bool first_occurrence(const int64_t *ids, uint32_t j, int64_t rid) {
	for (uint32_t e = 0; e < j; e++) {
		if (ids[e] == rid) {
			return false;
		}
	}
	return true;
}

// One thread per (node, j): resolve neighbour j of the mirrored block to a slot.
// core/GraphManager.cpp:241-254: only the first `count` ids are neighbours, INT64_MAX entries are dropped;
// core/Searcher.cpp:146-148: a row id that does not resolve is skipped.
__global__ void build_sidecar_kernel(const uint8_t *__restrict__ blocks, uint64_t block_bytes, uint64_t count_off,
                                     uint64_t ids_off, uint32_t r, uint64_t n, int identity,
                                     const long long *__restrict__ map_keys, const uint32_t *__restrict__ map_vals,
                                     uint64_t map_mask, uint32_t *__restrict__ nbrs) {
	uint64_t t = blockIdx.x * (uint64_t)blockDim.x + threadIdx.x;
	uint64_t node = t / r;
	uint32_t j = (uint32_t)(t % r);
	if (node >= n) {
		return;
	}
	const uint8_t *blk = blocks + node * block_bytes;
	uint32_t count = *reinterpret_cast<const uint16_t *>(blk + count_off);
	uint32_t out = LMD_INVALID_SLOT;
	if (j < count) {
		int64_t rid = reinterpret_cast<const int64_t *>(blk + ids_off)[j];
		if (rid != INT64_MAX && first_occurrence(reinterpret_cast<const int64_t *>(blk + ids_off), j, rid)) {
			if (identity) {
				out = (rid >= 0 && (uint64_t)rid < n) ? (uint32_t)rid : LMD_INVALID_SLOT;
			} else {
				out = map_lookup(map_keys, map_vals, map_mask, rid);
			}
		}
	}
	nbrs[node * r + j] = out;
}

// The two view builders for a LIST of slots (the blocks an lmd_store_apply call overwrote). One warp per slot.
__global__ void refresh_slots_kernel(const uint8_t *__restrict__ blocks, uint64_t block_bytes, uint64_t vector_off,
                                     uint64_t count_off, uint64_t ids_off, uint32_t dims, uint32_t vec_type,
                                     uint32_t row_floats, uint32_t r, const uint32_t *__restrict__ slots, uint64_t m,
                                     uint64_t n_nodes, int identity, int with_sidecar,
                                     const long long *__restrict__ map_keys, const uint32_t *__restrict__ map_vals,
                                     uint64_t map_mask, float *__restrict__ vecs, uint32_t *__restrict__ nbrs) {
	uint64_t w = (blockIdx.x * (uint64_t)blockDim.x + threadIdx.x) >> 5;
	uint32_t lane = threadIdx.x & 31;
	if (w >= m) {
		return;
	}
	const uint32_t slot = slots[w];
	if (slot == LMD_INVALID_SLOT) {
		return;
	}
	const uint8_t *blk = blocks + (uint64_t)slot * block_bytes;
	float *dst = vecs + (uint64_t)slot * row_floats;
	for (uint32_t i = lane; i < row_floats; i += 32) {
		float v = 0.0f;
		if (i < dims) {
			v = vec_type == LMD_VEC_FLOAT32 ? reinterpret_cast<const float *>(blk + vector_off)[i]
			                                : (float)reinterpret_cast<const int8_t *>(blk + vector_off)[i];
		}
		dst[i] = v;
	}
	if (!with_sidecar) {
		return;
	}
	const uint32_t count = *reinterpret_cast<const uint16_t *>(blk + count_off);
	for (uint32_t j = lane; j < r; j += 32) {
		uint32_t out = LMD_INVALID_SLOT;
		if (j < count) {
			int64_t rid = reinterpret_cast<const int64_t *>(blk + ids_off)[j];
			if (rid != INT64_MAX && first_occurrence(reinterpret_cast<const int64_t *>(blk + ids_off), j, rid)) {
				out = identity ? ((rid >= 0 && (uint64_t)rid < n_nodes) ? (uint32_t)rid : LMD_INVALID_SLOT)
				               : map_lookup(map_keys, map_vals, map_mask, rid);
			}
		}
		nbrs[(uint64_t)slot * r + j] = out;
	}
}

__global__ void fill_i64_kernel(long long *p, uint64_t n, long long v) {
	uint64_t i = blockIdx.x * (uint64_t)blockDim.x + threadIdx.x;
	if (i < n) {
		p[i] = v;
	}
}

__global__ void check_identity_kernel(const int64_t *row_ids, uint64_t base, uint64_t n, int *not_identity) {
	uint64_t i = blockIdx.x * (uint64_t)blockDim.x + threadIdx.x;
	if (i < n && row_ids[i] != (int64_t)(base + i)) {
		*not_identity = 1;
	}
}

static bool is_device_ptr(const void *p) {
	cudaPointerAttributes at;
	if (cudaPointerGetAttributes(&at, p) != cudaSuccess) {
		cudaGetLastError();
		return false;
	}
	return at.type == cudaMemoryTypeDevice || at.type == cudaMemoryTypeManaged;
}

static lmd_status lookup_slot(lmd_store *s, int64_t row_id, uint32_t *slot) {
	if (s->identity) {
		*slot = (row_id >= 0 && (uint64_t)row_id < s->n) ? (uint32_t)row_id : LMD_INVALID_SLOT;
		return LMD_OK;
	}
	uint32_t *d_tmp = reinterpret_cast<uint32_t *>(s->d_counters + 9);
	lookup_one_kernel<<<1, 1, 0, s->stream>>>((const long long *)s->d_map_keys, s->d_map_vals, s->map_cap - 1, row_id,
	                                          d_tmp);
	LMD_CUDA_OK(cudaGetLastError());
	LMD_CUDA_OK(cudaMemcpyAsync(slot, d_tmp, sizeof(uint32_t), cudaMemcpyDeviceToHost, s->stream));
	LMD_CUDA_OK(cudaStreamSynchronize(s->stream));
	return LMD_OK;
}

} // namespace lmd

using namespace lmd;

namespace lmd {

lmd_status state_init(SearchState &st) {
	LMD_CUDA_OK(cudaMalloc((void **)&st.d_counters, 16 * sizeof(unsigned long long)));
	LMD_CUDA_OK(cudaStreamCreateWithFlags(&st.own_stream, cudaStreamNonBlocking));
	st.stream = st.own_stream;
	for (auto &e : st.ev) {
		LMD_CUDA_OK(cudaEventCreate(&e));
	}
	LMD_CUDA_OK(cudaMemsetAsync(st.d_counters, 0, 128, st.stream));
	return LMD_OK;
}

void state_free(SearchState &st) {
	if (st.stream) {
		cudaStreamSynchronize(st.stream);
	}
	cudaFree(st.d_counters);
	cudaFree(st.d_q);
	cudaFree(st.d_out_ids);
	cudaFree(st.d_out_dist);
	cudaFree(st.d_out_count);
	cudaFree(st.d_qstats);
	cudaFree(st.d_cand);
	if (st.h_pin) {
		cudaFreeHost(st.h_pin);
	}
	for (auto &e : st.ev) {
		if (e) {
			cudaEventDestroy(e);
		}
	}
	if (st.own_stream) {
		cudaStreamDestroy(st.own_stream);
	}
	st.stream = st.own_stream = nullptr;
	st.d_counters = nullptr;
	st.d_q = nullptr;
	st.d_out_ids = nullptr;
	st.d_out_dist = nullptr;
	st.d_out_count = nullptr;
	st.d_qstats = nullptr;
	st.d_cand = nullptr;
	st.h_pin = nullptr;
	for (auto &e : st.ev) {
		e = nullptr;
	}
	st.scratch_q = st.scratch_k = st.scratch_qs = st.scratch_cnt = st.scratch_cand = 0;
}

} // namespace lmd

extern "C" {

const char *lmd_last_error(void) {
	return g_err.c_str();
}

uint32_t lmd_abi_version(void) {
	return LMD_ABI_VERSION;
}

static uint64_t align_up64(uint64_t v, uint64_t a) {
	return (v + a - 1) / a * a;
}

// core/index_config.cpp:104-148 (CalculateLayoutInternal); element sizes :40-52, plane sizes :55-64,
// words per plane core/ternary_quantization.hpp:120-123.
lmd_status lmd_layout_calc(uint64_t dims, uint32_t vec_type, uint32_t r, lmd_layout *out) {
	if (!out) {
		set_error("lmd_layout_calc: out is null");
		return LMD_ERR_INVALID_ARG;
	}
	if (dims == 0 || (vec_type != LMD_VEC_FLOAT32 && vec_type != LMD_VEC_INT8)) {
		set_error("CalculateLayoutInternal requires dimensions and node_vector_type to be set in config");
		return LMD_ERR_INVALID_ARG;
	}
	const uint64_t elem = vec_type == LMD_VEC_FLOAT32 ? 4 : 1;
	const uint64_t plane = ((dims + 63) >> 6) * 8;
	uint64_t off = sizeof(uint16_t);
	out->count_off = 0;
	off = align_up64(off, 8);
	out->vector_off = off;
	off = align_up64(off + elem * dims, 8);
	out->ids_off = off;
	off = align_up64(off + (uint64_t)r * 8, 8);
	out->pos_off = off;
	off = align_up64(off + (uint64_t)r * plane, 8);
	out->neg_off = off;
	off += (uint64_t)r * plane;
	out->total_node_size = off;
	out->ternary_edge_bytes = 2 * plane;
	return LMD_OK;
}

// db/LmDiskannIndex.cpp:161-162 with Storage::SECTOR_SIZE = 4096
uint64_t lmd_block_bytes(uint64_t total_node_size) {
	return align_up64(total_node_size, 4096);
}

lmd_status lmd_store_create(const lmd_config *cfg, int device, uint64_t capacity_nodes, lmd_store **out) {
	if (!cfg || !out) {
		set_error("lmd_store_create: null argument");
		return LMD_ERR_INVALID_ARG;
	}
	*out = nullptr;
	// core/index_config.cpp:66-102 (ValidateParameters), the parts that concern the query path
	if (cfg->r == 0) {
		set_error("LM_DISKANN parameter R must be > 0");
		return LMD_ERR_INVALID_ARG;
	}
	if (cfg->metric != LMD_METRIC_L2 && cfg->metric != LMD_METRIC_COSINE && cfg->metric != LMD_METRIC_IP) {
		set_error("Unsupported metric type %u", cfg->metric);
		return LMD_ERR_INVALID_ARG;
	}
	if (capacity_nodes == 0 || capacity_nodes >= 0x7FFFFFFFull) {
		set_error("capacity_nodes must be in [1, 2^31-1)");
		return LMD_ERR_INVALID_ARG;
	}
	if (cfg->r > 1024) {
		set_error("R = %u exceeds this build's limit of 1024", cfg->r);
		return LMD_ERR_UNSUPPORTED;
	}
	lmd_layout lay;
	lmd_status st = lmd_layout_calc(cfg->dims, cfg->vec_type, cfg->r, &lay);
	if (st != LMD_OK) {
		return st;
	}
	uint64_t bb = cfg->block_bytes ? cfg->block_bytes : lmd_block_bytes(lay.total_node_size);
	if (bb < lay.total_node_size || (bb % 8) != 0) {
		set_error("block_bytes %llu smaller than the node layout (%llu) or not a multiple of 8",
		          (unsigned long long)bb, (unsigned long long)lay.total_node_size);
		return LMD_ERR_INVALID_ARG;
	}
	int ndev = 0;
	LMD_CUDA_OK(cudaGetDeviceCount(&ndev));
	if (device < 0 || device >= ndev) {
		set_error("device %d out of range (%d visible)", device, ndev);
		return LMD_ERR_INVALID_ARG;
	}
	LMD_CUDA_OK(cudaSetDevice(device));
	lmd_store *s = new (std::nothrow) lmd_store();
	if (!s) {
		set_error("host allocation failed");
		return LMD_ERR_OOM;
	}
	s->cfg = *cfg;
	s->cfg.block_bytes = bb;
	s->lay = lay;
	s->device = device;
	s->capacity = capacity_nodes;
	s->row_floats = (uint32_t)align_up64(cfg->dims, 32);
	s->words = (uint32_t)((cfg->dims + 63) >> 6);
	cudaDeviceGetAttribute(&s->sm_count, cudaDevAttrMultiProcessorCount, device);
	s->map_cap = 64;
	while (s->map_cap < capacity_nodes * 2) {
		s->map_cap <<= 1;
	}
	auto fail = [&](lmd_status code) {
		lmd_store_destroy(s);
		return code;
	};
#define LMD_ALLOC(ptr, bytes)                                                                                          \
	do {                                                                                                               \
		cudaError_t _e = cudaMalloc((void **)&(ptr), (bytes));                                                         \
		if (_e != cudaSuccess) {                                                                                       \
			set_error("cudaMalloc(%llu bytes) for %s failed: %s", (unsigned long long)(bytes), #ptr,                   \
			          cudaGetErrorString(_e));                                                                         \
			return fail(LMD_ERR_OOM);                                                                                  \
		}                                                                                                              \
		s->device_bytes += (bytes);                                                                                    \
	} while (0)
	LMD_ALLOC(s->d_blocks, capacity_nodes * bb);
	LMD_ALLOC(s->d_vecs, capacity_nodes * (uint64_t)s->row_floats * sizeof(float));
	LMD_ALLOC(s->d_nbrs, capacity_nodes * (uint64_t)cfg->r * sizeof(uint32_t));
	LMD_ALLOC(s->d_row_ids, capacity_nodes * sizeof(int64_t));
	LMD_ALLOC(s->d_map_keys, s->map_cap * sizeof(int64_t));
	LMD_ALLOC(s->d_map_vals, s->map_cap * sizeof(uint32_t));
#undef LMD_ALLOC
	if ((st = state_init(s->def)) != LMD_OK) {
		return fail(st);
	}
	if (cudaStreamCreateWithFlags(&s->copy_stream, cudaStreamNonBlocking) != cudaSuccess) {
		set_error("cudaStreamCreate failed");
		return fail(LMD_ERR_CUDA);
	}
	cudaGetLastError(); // start from a clean error state
	fill_i64_kernel<<<(unsigned)((s->map_cap + 255) / 256), 256, 0, s->stream>>>((long long *)s->d_map_keys, s->map_cap,
	                                                                              LMD_MAP_EMPTY);
	if (cudaGetLastError() != cudaSuccess || cudaMemsetAsync(s->d_map_vals, 0, s->map_cap * 4, s->stream) != cudaSuccess ||
	    cudaMemsetAsync(s->d_counters, 0, 128, s->stream) != cudaSuccess ||
	    cudaStreamSynchronize(s->stream) != cudaSuccess) {
		set_error("store initialisation failed: %s", cudaGetErrorString(cudaGetLastError()));
		return fail(LMD_ERR_CUDA);
	}
	*out = s;
	return LMD_OK;
}

lmd_status lmd_store_destroy(lmd_store *s) {
	if (!s) {
		return LMD_OK;
	}
	cudaSetDevice(s->device);
	if (s->stream) {
		cudaStreamSynchronize(s->stream);
	}
	cudaFree(s->d_blocks);
	cudaFree(s->d_norms);
	cudaFree(s->d_vecs);
	cudaFree(s->d_nbrs);
	cudaFree(s->d_row_ids);
	cudaFree(s->d_map_keys);
	cudaFree(s->d_map_vals);
	cudaFree(s->d_apply);
	cudaFree(s->d_apply_rids);
	cudaFree(s->d_apply_slots);
	cudaFree(s->d_inbox);
	cudaFree(s->d_inbox_cnt);
	cudaFree(s->d_touched);
	cudaFree(s->d_planes);
	cudaFree(s->d_side_t);
	cudaFree(s->d_codes);
	cudaFree(s->d_code_off);
	if (s->h_stage) {
		cudaFreeHost(s->h_stage);
	}
	state_free(s->def);
	if (s->copy_stream) {
		cudaStreamDestroy(s->copy_stream);
	}
	delete s;
	return LMD_OK;
}

lmd_status lmd_store_set_stream(lmd_store *s, void *cuda_stream) {
	if (!s) {
		set_error("lmd_store_set_stream: null store");
		return LMD_ERR_INVALID_ARG;
	}
	std::unique_lock<std::shared_mutex> g(s->mu);
	LMD_CUDA_OK(cudaSetDevice(s->device));
	LMD_CUDA_OK(cudaStreamSynchronize(s->stream)); // nothing of ours may still be running on the old stream
	s->stream = cuda_stream ? (cudaStream_t)cuda_stream : s->own_stream;
	return LMD_OK;
}

uint64_t lmd_store_node_count(const lmd_store *s) {
	return s ? s->n : 0;
}

uint64_t lmd_store_device_bytes(const lmd_store *s) {
	return s ? s->device_bytes : 0;
}


// Copies `bytes` from src (host or device) to device dst on the store's stream. Pageable host memory goes
// through the store's pinned staging buffer in 32 MiB pieces, double-buffered against the copy engine.
static lmd_status copy_to_device(lmd_store *s, void *dst, const void *src, size_t bytes) {
	if (bytes == 0) {
		return LMD_OK;
	}
	if (is_device_ptr(src)) {
		LMD_CUDA_OK(cudaMemcpyAsync(dst, src, bytes, cudaMemcpyDeviceToDevice, s->stream));
		return LMD_OK;
	}
	cudaPointerAttributes at;
	bool pinned = cudaPointerGetAttributes(&at, src) == cudaSuccess && at.type == cudaMemoryTypeHost;
	cudaGetLastError();
	if (pinned) {
		LMD_CUDA_OK(cudaMemcpyAsync(dst, src, bytes, cudaMemcpyHostToDevice, s->stream));
		return LMD_OK;
	}
	const size_t piece = 32u << 20;
	if (s->h_stage_bytes < 2 * piece) {
		if (s->h_stage) {
			cudaFreeHost(s->h_stage);
			s->h_stage = nullptr;
		}
		LMD_CUDA_OK(cudaMallocHost(&s->h_stage, 2 * piece));
		s->h_stage_bytes = 2 * piece;
	}
	size_t done = 0;
	int buf = 0;
	while (done < bytes) {
		size_t now = bytes - done < piece ? bytes - done : piece;
		uint8_t *stage = (uint8_t *)s->h_stage + (size_t)buf * piece;
		// the copy issued two pieces ago from this half must have drained
		LMD_CUDA_OK(cudaEventSynchronize(s->ev[6 + buf]));
		memcpy(stage, (const uint8_t *)src + done, now);
		LMD_CUDA_OK(cudaMemcpyAsync((uint8_t *)dst + done, stage, now, cudaMemcpyHostToDevice, s->stream));
		LMD_CUDA_OK(cudaEventRecord(s->ev[6 + buf], s->stream));
		done += now;
		buf ^= 1;
	}
	return LMD_OK;
}

} // extern "C"

namespace lmd {

lmd_status store_copy_to_device(lmd_store *s, void *dst, const void *src, size_t bytes) {
	return copy_to_device(s, dst, src, bytes);
}

// Everything lmd_store_upload does after the blocks [base, base + n) are in d_blocks: row ids, row-id map, packed
// vectors, entry-point resolution. The caller holds s->mu.
lmd_status store_ingest(lmd_store *s, const int64_t *row_ids, uint64_t base, uint64_t n) {
	const uint64_t bb = s->cfg.block_bytes;
	lmd_status st = LMD_OK;
	const int64_t *d_rid = nullptr;
	if (row_ids) {
		// stage the row ids in the slot->row table itself, then let the kernel read them from there
		st = copy_to_device(s, s->d_row_ids + base, row_ids, n * sizeof(int64_t));
		if (st != LMD_OK) {
			return st;
		}
		d_rid = s->d_row_ids + base;
		if (s->identity) {
			int *d_flag = reinterpret_cast<int *>(s->d_counters + 8);
			LMD_CUDA_OK(cudaMemsetAsync(d_flag, 0, sizeof(int), s->stream));
			check_identity_kernel<<<(unsigned)((n + 255) / 256), 256, 0, s->stream>>>(d_rid, base, n, d_flag);
			int h_flag = 0;
			LMD_CUDA_OK(cudaMemcpyAsync(&h_flag, d_flag, sizeof(int), cudaMemcpyDeviceToHost, s->stream));
			LMD_CUDA_OK(cudaStreamSynchronize(s->stream));
			if (h_flag) {
				s->identity = false;
			}
		}
	}
	ingest_rowids_kernel<<<(unsigned)((n + 255) / 256), 256, 0, s->stream>>>(
	    d_rid, base, n, s->d_row_ids, (long long *)s->d_map_keys, s->d_map_vals, s->map_cap - 1);
	LMD_CUDA_OK(cudaGetLastError());
	unpack_vectors_kernel<<<(unsigned)((n * 32 + 255) / 256), 256, 0, s->stream>>>(
	    s->d_blocks, bb, s->lay.vector_off, s->cfg.dims, s->cfg.vec_type, s->row_floats, base, n, s->d_vecs);
	LMD_CUDA_OK(cudaGetLastError());
	LMD_CUDA_OK(cudaStreamSynchronize(s->stream)); // caller may free / reuse its buffers on return
	s->n += n;
	s->epoch++;
	s->dirty = true;
	s->views_gen++;
	if (s->entry_row != INT64_MAX && !s->has_entry) {
		// an entry row id set before its block arrived resolves now
		uint32_t slot;
		st = lookup_slot(s, s->entry_row, &slot);
		if (st != LMD_OK) {
			return st;
		}
		s->entry_slot = slot;
		s->has_entry = slot != LMD_INVALID_SLOT;
	}
	return LMD_OK;
}

lmd_status store_lookup_slot(lmd_store *s, int64_t row_id, uint32_t *slot) {
	return lookup_slot(s, row_id, slot);
}

} // namespace lmd

extern "C" {

lmd_status lmd_store_upload(lmd_store *s, const int64_t *row_ids, const void *blocks, uint64_t n) {
	if (!s || (!blocks && n)) {
		set_error("lmd_store_upload: null argument");
		return LMD_ERR_INVALID_ARG;
	}
	if (n == 0) {
		return LMD_OK;
	}
	std::unique_lock<std::shared_mutex> g(s->mu);
	if (s->blocks_released) {
		set_error("lmd_store_upload: the block mirror of this store was released (lmd_store_release_blocks)");
		return LMD_ERR_STATE;
	}
	if (s->n + n > s->capacity) {
		set_error("lmd_store_upload: %llu + %llu nodes exceed the store capacity %llu", (unsigned long long)s->n,
		          (unsigned long long)n, (unsigned long long)s->capacity);
		return LMD_ERR_INVALID_ARG;
	}
	LMD_CUDA_OK(cudaSetDevice(s->device));
	const uint64_t bb = s->cfg.block_bytes;
	const uint64_t base = s->n;
	lmd_status st = copy_to_device(s, s->d_blocks + base * bb, blocks, n * bb);
	if (st != LMD_OK) {
		return st;
	}
	return lmd::store_ingest(s, row_ids, base, n);
}

lmd_status lmd_store_update(lmd_store *s, int64_t row_id, const void *block) {
	if (!s || !block) {
		set_error("lmd_store_update: null argument");
		return LMD_ERR_INVALID_ARG;
	}
	std::unique_lock<std::shared_mutex> g(s->mu);
	if (s->blocks_released) {
		set_error("lmd_store_update: the block mirror of this store was released (lmd_store_release_blocks)");
		return LMD_ERR_STATE;
	}
	LMD_CUDA_OK(cudaSetDevice(s->device));
	uint32_t slot;
	lmd_status st = lookup_slot(s, row_id, &slot);
	if (st != LMD_OK) {
		return st;
	}
	if (slot == LMD_INVALID_SLOT) {
		set_error("lmd_store_update: row id %lld is not in the store", (long long)row_id);
		return LMD_ERR_INVALID_ARG;
	}
	const uint64_t bb = s->cfg.block_bytes;
	st = copy_to_device(s, s->d_blocks + (uint64_t)slot * bb, block, bb);
	if (st != LMD_OK) {
		return st;
	}
	unpack_vectors_kernel<<<1, 32, 0, s->stream>>>(s->d_blocks, bb, s->lay.vector_off, s->cfg.dims, s->cfg.vec_type,
	                                               s->row_floats, slot, 1, s->d_vecs);
	LMD_CUDA_OK(cudaGetLastError());
	// only this node's neighbour row changes; the other rows keep pointing at valid slots
	build_sidecar_kernel<<<(s->cfg.r + 127) / 128, 128, 0, s->stream>>>(
	    s->d_blocks + (uint64_t)slot * bb, bb, s->lay.count_off, s->lay.ids_off, s->cfg.r, 1, 0,
	    (const long long *)s->d_map_keys, s->d_map_vals, s->map_cap - 1, s->d_nbrs + (uint64_t)slot * s->cfg.r);
	LMD_CUDA_OK(cudaGetLastError());
	LMD_CUDA_OK(cudaStreamSynchronize(s->stream));
	s->views_gen++;
	return LMD_OK;
}

// Batch form of MarkBlockDirty + AllocateNodeBlock (core/IStorageManager.hpp:142-177): the host hands over every block
// it changed or created since the last call. Known row ids overwrite their slot, new ones are appended in batch order;
// a row id that occurs twice keeps its LAST block (what a host that re-dirtied a block means).
lmd_status lmd_store_apply(lmd_store *s, const int64_t *row_ids, const void *blocks, uint64_t n,
                           uint64_t *out_appended) {
	if (out_appended) {
		*out_appended = 0;
	}
	if (!s || ((!blocks || !row_ids) && n)) {
		set_error("lmd_store_apply: null argument");
		return LMD_ERR_INVALID_ARG;
	}
	if (n == 0) {
		return LMD_OK;
	}
	if (is_device_ptr(row_ids)) {
		set_error("lmd_store_apply: row_ids must be host memory");
		return LMD_ERR_INVALID_ARG;
	}
	std::unique_lock<std::shared_mutex> g(s->mu);
	if (s->blocks_released) {
		set_error("lmd_store_apply: the block mirror of this store was released (lmd_store_release_blocks)");
		return LMD_ERR_STATE;
	}
	LMD_CUDA_OK(cudaSetDevice(s->device));
	const uint64_t bb = s->cfg.block_bytes;
	const uint64_t chunk = 16384; // blocks staged per pass (64 MiB at 4 KiB blocks)
	if (s->apply_cap < (n < chunk ? n : chunk)) {
		// grow in powers of two from 1024 blocks: a host that flushes a few dirty blocks per insert must not pay a
		// cudaFree + cudaMalloc round (tens of milliseconds) whenever a batch is a little larger than the last one
		uint64_t cap = 1024;
		while (cap < n && cap < chunk) {
			cap <<= 1;
		}
		if (cap > s->capacity && s->capacity >= 64) {
			cap = s->capacity;
		}
		cudaFree(s->d_apply);
		cudaFree(s->d_apply_rids);
		cudaFree(s->d_apply_slots);
		s->d_apply = nullptr;
		s->d_apply_rids = nullptr;
		s->d_apply_slots = nullptr;
		s->apply_cap = 0;
		LMD_CUDA_OK(cudaMalloc((void **)&s->d_apply, cap * bb));
		LMD_CUDA_OK(cudaMalloc((void **)&s->d_apply_rids, cap * sizeof(int64_t)));
		LMD_CUDA_OK(cudaMalloc((void **)&s->d_apply_slots, cap * sizeof(uint32_t)));
		s->apply_cap = cap;
	}
	// 1. resolve every row id (host side for identity stores, else one lookup kernel per chunk)
	std::vector<uint32_t> slots(n);
	if (s->identity) {
		for (uint64_t i = 0; i < n; i++) {
			slots[i] = (row_ids[i] >= 0 && (uint64_t)row_ids[i] < s->n) ? (uint32_t)row_ids[i] : LMD_INVALID_SLOT;
		}
	} else {
		for (uint64_t first = 0; first < n; first += s->apply_cap) {
			const uint64_t now = n - first < s->apply_cap ? n - first : s->apply_cap;
			LMD_CUDA_OK(cudaMemcpyAsync(s->d_apply_rids, row_ids + first, now * sizeof(int64_t), cudaMemcpyHostToDevice,
			                            s->stream));
			lookup_many_kernel<<<(unsigned)((now + 255) / 256), 256, 0, s->stream>>>(
			    (const long long *)s->d_map_keys, s->d_map_vals, s->map_cap - 1, s->d_apply_rids, now, s->d_apply_slots);
			LMD_CUDA_OK(cudaGetLastError());
			LMD_CUDA_OK(cudaMemcpyAsync(slots.data() + first, s->d_apply_slots, now * sizeof(uint32_t),
			                            cudaMemcpyDeviceToHost, s->stream));
			LMD_CUDA_OK(cudaStreamSynchronize(s->stream));
		}
	}
	// 2. new row ids get the next slots in batch order; an earlier block of a repeated row id is dropped
	std::vector<int64_t> new_rids;
	{
		std::unordered_map<int64_t, uint64_t> last; // row id -> index of its last occurrence
		last.reserve(n * 2);
		for (uint64_t i = 0; i < n; i++) {
			last[row_ids[i]] = i;
		}
		std::unordered_map<int64_t, uint32_t> fresh;
		for (uint64_t i = 0; i < n; i++) {
			if (row_ids[i] == INT64_MAX || row_ids[i] == INT64_MIN || row_ids[i] == INT64_MIN + 1) {
				set_error("lmd_store_apply: row id %lld is reserved", (long long)row_ids[i]);
				return LMD_ERR_INVALID_ARG;
			}
			if (last[row_ids[i]] != i) {
				slots[i] = LMD_INVALID_SLOT;
				continue;
			}
			if (slots[i] == LMD_INVALID_SLOT) {
				slots[i] = (uint32_t)(s->n + new_rids.size());
				new_rids.push_back(row_ids[i]);
			}
		}
	}
	if (s->n + new_rids.size() > s->capacity) {
		set_error("lmd_store_apply: %llu + %llu new nodes exceed the store capacity %llu", (unsigned long long)s->n,
		          (unsigned long long)new_rids.size(), (unsigned long long)s->capacity);
		return LMD_ERR_INVALID_ARG;
	}
	// 3. stage + scatter the blocks, refresh the views of the overwritten slots
	const bool rebuild_all = s->dirty || !new_rids.empty(); // a new row id may resolve ids other blocks already hold
	s->views_gen++;
	for (uint64_t first = 0; first < n; first += s->apply_cap) {
		const uint64_t now = n - first < s->apply_cap ? n - first : s->apply_cap;
		lmd_status st = copy_to_device(s, s->d_apply, (const uint8_t *)blocks + first * bb, now * bb);
		if (st != LMD_OK) {
			return st;
		}
		LMD_CUDA_OK(cudaMemcpyAsync(s->d_apply_slots, slots.data() + first, now * sizeof(uint32_t),
		                            cudaMemcpyHostToDevice, s->stream));
		scatter_blocks_kernel<<<(unsigned)now, 128, 0, s->stream>>>(s->d_apply, s->d_apply_slots, bb, s->d_blocks);
		LMD_CUDA_OK(cudaGetLastError());
		// slots >= s->n are new: store_ingest() below unpacks them; harmless to do it here as well
		refresh_slots_kernel<<<(unsigned)((now * 32 + 255) / 256), 256, 0, s->stream>>>(
		    s->d_blocks, bb, s->lay.vector_off, s->lay.count_off, s->lay.ids_off, s->cfg.dims, s->cfg.vec_type,
		    s->row_floats, s->cfg.r, s->d_apply_slots, now, s->n, s->identity ? 1 : 0, rebuild_all ? 0 : 1,
		    (const long long *)s->d_map_keys, s->d_map_vals, s->map_cap - 1, s->d_vecs, s->d_nbrs);
		LMD_CUDA_OK(cudaGetLastError());
		LMD_CUDA_OK(cudaStreamSynchronize(s->stream)); // `slots` / the staging buffer are reused by the next pass
	}
	// 4. register the new row ids (row-id map, slot -> row id, packed vectors, pending entry point)
	if (!new_rids.empty()) {
		lmd_status st = lmd::store_ingest(s, new_rids.data(), s->n, new_rids.size());
		if (st != LMD_OK) {
			return st;
		}
		if (out_appended) {
			*out_appended = new_rids.size();
		}
	}
	return LMD_OK;
}

// core/GraphManager.cpp:476-499 (HandleNodeDeletion + FreeNode): the row ids leave the row-id map, so every neighbour
// entry that names them is skipped from now on (Searcher.cpp:120-122,146-156), and a deleted entry point leaves the
// index without one (searches return nothing until the host sets another, Searcher.cpp:74-77).
lmd_status lmd_store_remove(lmd_store *s, const int64_t *row_ids, uint64_t n, uint64_t *out_removed) {
	if (out_removed) {
		*out_removed = 0;
	}
	if (!s || (!row_ids && n)) {
		set_error("lmd_store_remove: null argument");
		return LMD_ERR_INVALID_ARG;
	}
	if (n == 0) {
		return LMD_OK;
	}
	if (is_device_ptr(row_ids)) {
		set_error("lmd_store_remove: row_ids must be host memory");
		return LMD_ERR_INVALID_ARG;
	}
	std::unique_lock<std::shared_mutex> g(s->mu);
	LMD_CUDA_OK(cudaSetDevice(s->device));
	int64_t *d_rids = nullptr;
	LMD_CUDA_OK(cudaMalloc((void **)&d_rids, n * sizeof(int64_t)));
	unsigned long long *d_removed = s->d_counters + 8;
	cudaError_t e = cudaMemcpyAsync(d_rids, row_ids, n * sizeof(int64_t), cudaMemcpyHostToDevice, s->stream);
	if (e == cudaSuccess) {
		e = cudaMemsetAsync(d_removed, 0, sizeof(unsigned long long), s->stream);
	}
	unsigned long long removed = 0;
	if (e == cudaSuccess) {
		remove_rowids_kernel<<<(unsigned)((n + 255) / 256), 256, 0, s->stream>>>(
		    (long long *)s->d_map_keys, s->d_map_vals, s->map_cap - 1, d_rids, n, s->d_row_ids, d_removed);
		e = cudaGetLastError();
	}
	if (e == cudaSuccess) {
		e = cudaMemcpyAsync(&removed, d_removed, sizeof(removed), cudaMemcpyDeviceToHost, s->stream);
	}
	if (e == cudaSuccess) {
		e = cudaStreamSynchronize(s->stream);
	}
	cudaFree(d_rids);
	LMD_CUDA_OK(e);
	if (removed) {
		s->removed += removed;
		s->epoch++;
		s->identity = false; // slot numbers no longer stand for row ids: resolve through the map
		s->dirty = true;     // neighbour rows that point at the removed slots are rebuilt before the next search
		s->views_gen++;
	}
	for (uint64_t i = 0; i < n; i++) {
		if (row_ids[i] == s->entry_row) { // GraphManager.cpp:477-480
			s->entry_row = INT64_MAX;
			s->has_entry = false;
			s->entry_slot = LMD_INVALID_SLOT;
		}
	}
	if (out_removed) {
		*out_removed = removed;
	}
	return LMD_OK;
}

// Filtered search: the host's `allowed_ids` set, resolved once to a bitmap over the
// store's slots and reused by any number of searches.
lmd_status lmd_filter_create(lmd_store *s, const int64_t *allowed_row_ids, uint64_t n, lmd_filter **out) {
	if (!s || !out || (!allowed_row_ids && n)) {
		set_error("lmd_filter_create: null argument");
		return LMD_ERR_INVALID_ARG;
	}
	*out = nullptr;
	std::unique_lock<std::shared_mutex> g(s->mu);
	LMD_CUDA_OK(cudaSetDevice(s->device));
	lmd_filter *f = new (std::nothrow) lmd_filter();
	if (!f) {
		set_error("host allocation failed");
		return LMD_ERR_OOM;
	}
	f->store = s;
	f->epoch = s->epoch;
	f->words = (s->capacity + 31) / 32;
	int64_t *d_rids = nullptr;
	const int64_t *src = allowed_row_ids;
	cudaError_t e = cudaMalloc((void **)&f->d_bits, f->words * sizeof(uint32_t));
	if (e == cudaSuccess) {
		e = cudaMemsetAsync(f->d_bits, 0, f->words * sizeof(uint32_t), s->stream);
	}
	unsigned long long *d_count = s->d_counters + 8;
	unsigned long long h_count = 0;
	if (e == cudaSuccess && n) {
		if (!is_device_ptr(allowed_row_ids)) {
			e = cudaMalloc((void **)&d_rids, n * sizeof(int64_t));
			if (e == cudaSuccess) {
				e = cudaMemcpyAsync(d_rids, allowed_row_ids, n * sizeof(int64_t), cudaMemcpyHostToDevice, s->stream);
			}
			src = d_rids;
		}
		if (e == cudaSuccess) {
			e = cudaMemsetAsync(d_count, 0, sizeof(unsigned long long), s->stream);
		}
		if (e == cudaSuccess) {
			filter_bits_kernel<<<(unsigned)((n + 255) / 256), 256, 0, s->stream>>>(
			    (const long long *)s->d_map_keys, s->d_map_vals, s->map_cap - 1, s->identity ? 1 : 0, s->n, src, n,
			    f->d_bits, d_count);
			e = cudaGetLastError();
		}
		if (e == cudaSuccess) {
			e = cudaMemcpyAsync(&h_count, d_count, sizeof(h_count), cudaMemcpyDeviceToHost, s->stream);
		}
	}
	if (e == cudaSuccess) {
		e = cudaStreamSynchronize(s->stream);
	}
	cudaFree(d_rids);
	if (e != cudaSuccess) {
		cudaGetLastError();
		set_error("lmd_filter_create failed: %s", cudaGetErrorString(e));
		cudaFree(f->d_bits);
		delete f;
		return e == cudaErrorMemoryAllocation ? LMD_ERR_OOM : LMD_ERR_CUDA;
	}
	f->count = h_count;
	*out = f;
	return LMD_OK;
}

lmd_status lmd_filter_destroy(lmd_filter *f) {
	if (!f) {
		return LMD_OK;
	}
	if (f->store) {
		cudaSetDevice(f->store->device);
		cudaStreamSynchronize(f->store->stream);
	}
	cudaFree(f->d_bits);
	delete f;
	return LMD_OK;
}

uint64_t lmd_filter_count(const lmd_filter *f) {
	return f ? f->count : 0;
}

lmd_status lmd_store_set_entry(lmd_store *s, int64_t row_id) {
	if (!s) {
		set_error("lmd_store_set_entry: null store");
		return LMD_ERR_INVALID_ARG;
	}
	std::unique_lock<std::shared_mutex> g(s->mu);
	LMD_CUDA_OK(cudaSetDevice(s->device));
	s->entry_row = row_id;
	s->has_entry = false;
	s->entry_slot = LMD_INVALID_SLOT;
	if (row_id == INT64_MAX) {
		return LMD_OK; // "no entry point" (core/Searcher.cpp:74)
	}
	uint32_t slot;
	lmd_status st = lookup_slot(s, row_id, &slot);
	if (st != LMD_OK) {
		return st;
	}
	s->entry_slot = slot;
	s->has_entry = slot != LMD_INVALID_SLOT;
	return LMD_OK;
}

lmd_status lmd_store_finalize(lmd_store *s) {
	if (!s) {
		set_error("lmd_store_finalize: null store");
		return LMD_ERR_INVALID_ARG;
	}
	std::unique_lock<std::shared_mutex> g(s->mu);
	return lmd::finalize_locked(s);
}

} // extern "C"

lmd_status lmd::finalize_locked(lmd_store *s) {
	if (!s->dirty || s->n == 0) {
		s->dirty = false;
		return LMD_OK;
	}
	LMD_CUDA_OK(cudaSetDevice(s->device));
	uint64_t total = s->n * (uint64_t)s->cfg.r;
	build_sidecar_kernel<<<(unsigned)((total + 255) / 256), 256, 0, s->stream>>>(
	    s->d_blocks, s->cfg.block_bytes, s->lay.count_off, s->lay.ids_off, s->cfg.r, s->n, s->identity ? 1 : 0,
	    (const long long *)s->d_map_keys, s->d_map_vals, s->map_cap - 1, s->d_nbrs);
	LMD_CUDA_OK(cudaGetLastError());
	LMD_CUDA_OK(cudaStreamSynchronize(s->stream));
	s->dirty = false;
	s->views_gen++;
	return LMD_OK;
}

extern "C" {

lmd_status lmd_store_download_block(lmd_store *s, int64_t row_id, void *out_block) {
	if (!s || !out_block) {
		set_error("lmd_store_download_block: null argument");
		return LMD_ERR_INVALID_ARG;
	}
	std::unique_lock<std::shared_mutex> g(s->mu);
	if (s->blocks_released) {
		set_error("lmd_store_download_block: the block mirror of this store was released (lmd_store_release_blocks)");
		return LMD_ERR_STATE;
	}
	LMD_CUDA_OK(cudaSetDevice(s->device));
	uint32_t slot;
	lmd_status st = lookup_slot(s, row_id, &slot);
	if (st != LMD_OK) {
		return st;
	}
	if (slot == LMD_INVALID_SLOT) {
		set_error("lmd_store_download_block: row id %lld is not in the store", (long long)row_id);
		return LMD_ERR_INVALID_ARG;
	}
	LMD_CUDA_OK(cudaMemcpyAsync(out_block, s->d_blocks + (uint64_t)slot * s->cfg.block_bytes, s->cfg.block_bytes,
	                            cudaMemcpyDefault, s->stream));
	LMD_CUDA_OK(cudaStreamSynchronize(s->stream));
	return LMD_OK;
}

lmd_status lmd_set_tuning(lmd_store *s, const char *key, int64_t value) {
	if (!s || !key) {
		set_error("lmd_set_tuning: null argument");
		return LMD_ERR_INVALID_ARG;
	}
	std::string k(key);
	std::unique_lock<std::shared_mutex> g(s->mu); // searches read the tuning and the launch cache under the same lock
	s->launch_cache.clear();
	if (k == "filter_log2") {
		s->tune.filter_log2 = (int)value;
	} else if (k == "latency_mode") {
		s->tune.latency_mode = (int)value;
	} else if (k == "prefetch_planes") {
		s->tune.prefetch_planes = (int)value;
	} else if (k == "filter_entries") {
		s->tune.filter_entries = (int)value;
	} else if (k == "filter_mode") {
		s->tune.filter_mode = (int)value;
	} else if (k == "warps_per_cta") {
		s->tune.warps_per_cta = (int)value;
	} else if (k == "ctas_per_sm") {
		s->tune.ctas_per_sm = (int)value;
	} else if (k == "tie_slots") {
		s->tune.tie_slots = (int)value;
	} else if (k == "merge_min") {
		s->tune.merge_min = (int)value;
	} else if (k == "strict_fp") {
		s->tune.strict_fp = value ? 1 : 0;
	} else if (k == "code_sigmas") {
		s->tune.code_sigmas = (int)value;
		s->codes_gen = 0; // the grid changes: rebuild before the next search that uses it
	} else if (k == "plane_filter_cap") {
		s->tune.plane_filter_cap = value ? 1 : 0;
	} else if (k == "prefilter") {
		s->tune.prefilter = value == 2 ? 2 : (value ? 1 : 0);
	} else if (k == "nbr_prefetch") {
		s->tune.nbr_prefetch = (int)value;
	} else if (k == "row_prefetch") {
		s->tune.row_prefetch = value == 0 ? 0 : (value == 1 ? 1 : 2);
	} else if (k == "pf_g") {
		s->tune.pf_g = value == 1 ? 1 : 2;
	} else {
		set_error("lmd_set_tuning: unknown key %s", key);
		return LMD_ERR_INVALID_ARG;
	}
	return LMD_OK;
}

} // extern "C"
