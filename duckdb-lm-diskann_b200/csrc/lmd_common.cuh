// Shared declarations of the B200 LM_DISKANN query path (device node store + kernels).
// sm_100a only; no CPU fallback anywhere below this header.
#pragma once

#include "../../include/lmd_b200.h"

#include <cuda_runtime.h>

#include <cstdarg>
#include <cstdint>
#include <cstdio>
#include <mutex>
#include <shared_mutex>
#include <string>
#include <unordered_map>
#include <utility>

#define LMD_INVALID_SLOT 0xFFFFFFFFu
#define LMD_KEY_NONE 0xFFFFFFFFFFFFFFFFull

namespace lmd {

void set_error(const char *fmt, ...);

#define LMD_CUDA_OK(expr)                                                                                              \
	do {                                                                                                               \
		cudaError_t _e = (expr);                                                                                       \
		if (_e != cudaSuccess) {                                                                                       \
			cudaGetLastError(); /* a non-sticky error must not be reported again by the next call */                  \
			lmd::set_error("%s failed: %s (%s:%d)", #expr, cudaGetErrorString(_e), __FILE__, __LINE__);                \
			return (_e == cudaErrorMemoryAllocation) ? LMD_ERR_OOM : LMD_ERR_CUDA;                                     \
		}                                                                                                              \
	} while (0)

// ---- candidate keys -------------------------------------------------------------------------
// A candidate is one u64: [63:32] order-preserving image of the fp32 distance, [31:1] slot, [0] expanded flag.
// Unsigned comparison of (key >> 1) is the reference's H4 order (distance, id) with id = slot
// (core/Searcher.cpp:35-37 + oracle/patches/searcher_fixed.patch).
__host__ __device__ __forceinline__ uint32_t float_to_ordered(float f) {
#ifdef __CUDA_ARCH__
	// The bits are taken through an opaque move: given `__float_as_uint(f) | 0x80000000` the compiler emits
	// neg.f32(abs.f32(f)), and PTX's abs / neg do not preserve a NaN's bits (a NaN distance came back as -0.0 and sorted
	// first; tests/test_gpu_round2.py::test_special_values). Keys are integers from here on.
	uint32_t b;
	asm("mov.b32 %0, %1;" : "=r"(b) : "f"(f));
#else
	union {
		float f;
		uint32_t u;
	} cv;
	cv.f = f;
	uint32_t b = cv.u;
#endif
	return (b & 0x80000000u) ? ~b : (b | 0x80000000u);
}
__host__ __device__ __forceinline__ float ordered_to_float(uint32_t o) {
	uint32_t b = (o & 0x80000000u) ? (o & 0x7FFFFFFFu) : ~o;
#ifdef __CUDA_ARCH__
	float f; // opaque for the same reason (and / fabs)
	asm("mov.b32 %0, %1;" : "=f"(f) : "r"(b));
	return f;
#else
	union {
		float f;
		uint32_t u;
	} cv;
	cv.u = b;
	return cv.f;
#endif
}
__host__ __device__ __forceinline__ uint64_t make_key(float d, uint32_t slot) {
	return ((uint64_t)float_to_ordered(d) << 32) | ((uint64_t)slot << 1);
}
__host__ __device__ __forceinline__ uint32_t key_slot(uint64_t k) {
	return (uint32_t)(k & 0xFFFFFFFFull) >> 1;
}
__host__ __device__ __forceinline__ float key_dist(uint64_t k) {
	return ordered_to_float((uint32_t)(k >> 32));
}

// ---- the store ---------------------------------------------------------------------------------
struct Tuning {
	int filter_log2 = 0;   // log2(entries of the visited filter), 0 = heuristic
	int filter_entries = 0; // filter_mode 3: entries of the modulo-mapped table
	int filter_mode = -1;  // -1 = auto (1 when the store is small enough), 0 = u32 open addressing, 1 = 16-bit buckets
	int warps_per_cta = 0; // 0 = whichever of 4 / 2 / 1 packs the most warps on an SM
	int ctas_per_sm = 0;   // 0 = as many as shared memory allows
	int tie_slots = -1;    // -1 = default (plan_tie_slots(L) for the integer-valued routes, 32 under strict fp, else 0)
	int prefetch_planes = 1; // ternary routing: prefetch the plane region of the likely next expansion into L2
	int latency_mode = -1; // -1 = latency kernels for batches below ~10 queries per SM, 0 = never, 1 = always
	int merge_min = 0;     // batch-merge new candidates when at least this many pass the threshold (0 = default 5)
	int strict_fp = 0;     // 1 = every kernel of this store accumulates distances in the reference's order (parity tool)
	int code_sigmas = 0;   // int8 codec grid: scale = code_sigmas / 10 * max sigma / 127 (0 = 4.0 sigmas)
	int plane_filter_cap = 0; // 1 = the old 2 KB visited-filter cap for ternary routing over wide planes (pre-sidecar rule)
	int prefilter = 1;     // exact routing on a codec store: 1 = int8 lower-bound pre-filter for batches that fill the GPU, 2 = for every batch, 0 = plain traversal
	int pf_g = 2;          // rows in flight per team of the pre-filtered kernels (1 or 2)
	int row_prefetch = 2;  // L2 prefetch of the rows of a gather: 2 = auto, 1 = always, 0 = never
	int nbr_prefetch = 0;  // neighbour rows of admitted candidates into L2: 0 = never, 1 = small-batch kernels, 2 = always
};

// Everything ONE in-flight search call owns: its stream, events, statistics counters and staging scratch. The store has
// one of its own (lmd_search_batch, the build and the upload paths run on it); every lmd_ctx is another, so searches
// issued from different host threads through different contexts share nothing but the read-only node store.
struct SearchState {
	cudaStream_t stream = nullptr, own_stream = nullptr; // stream = the active one
	cudaEvent_t ev[8] = {nullptr, nullptr, nullptr, nullptr, nullptr, nullptr, nullptr, nullptr};
	unsigned long long *d_counters = nullptr; // [16] work counter + statistics ([8], [9]: store-side scratch)
	// host<->device staging of one search call (grown on demand)
	float *d_q = nullptr;
	int64_t *d_out_ids = nullptr;
	float *d_out_dist = nullptr;
	uint32_t *d_out_count = nullptr;
	uint32_t *d_qstats = nullptr;
	uint64_t *d_cand = nullptr; // [nq][L] candidate keys for the separate rerank kernel
	uint64_t scratch_q = 0, scratch_k = 0, scratch_qs = 0, scratch_cnt = 0, scratch_cand = 0;
	// small pinned (mapped) block for pageable point queries: the kernels read the query and write the result in place
	uint8_t *h_pin = nullptr;
	std::mutex mu; // one call at a time per state
};

} // namespace lmd

struct lmd_store {
	lmd_config cfg;
	lmd_layout lay;
	int device = 0;
	int sm_count = 148;
	uint64_t capacity = 0; // nodes
	uint64_t n = 0;        // nodes uploaded
	uint32_t row_floats = 0; // packed row length (dims padded to 32 floats)
	uint32_t words = 0;      // u64 words per ternary plane

	// device memory (all owned)
	uint8_t *d_blocks = nullptr;   // [capacity][block_bytes]   raw mirror, reference layout untouched
	float *d_vecs = nullptr;       // [capacity][row_floats]    packed fp32 view of the node vectors
	uint32_t *d_nbrs = nullptr;    // [capacity][r]             neighbour SLOTS (sidecar), LMD_INVALID_SLOT = none
	int64_t *d_row_ids = nullptr;  // [capacity]                slot -> row id
	int64_t *d_map_keys = nullptr; // [map_cap]                 row id -> slot, open addressing
	uint32_t *d_map_vals = nullptr;
	uint64_t map_cap = 0;
	bool identity = true;          // every row id so far equals its slot: the map is bypassed
	uint64_t device_bytes = 0;

	// the store's own search state; the names below keep meaning "the default context's"
	lmd::SearchState def;
	unsigned long long *&d_counters = def.d_counters;
	void *h_stage = nullptr; // pinned staging for pageable uploads
	size_t h_stage_bytes = 0;

	bool has_entry = false;
	int64_t entry_row = INT64_MAX;
	uint32_t entry_slot = LMD_INVALID_SLOT;
	bool dirty = true; // views need (re)building

	// lmd_store_apply scratch: staged dirty blocks, their row ids and resolved slots (grown on demand)
	uint8_t *d_apply = nullptr;
	int64_t *d_apply_rids = nullptr;
	uint32_t *d_apply_slots = nullptr;
	uint64_t apply_cap = 0; // blocks
	uint64_t epoch = 0;     // bumped whenever the row id -> slot mapping may have changed (filters are tied to it)
	uint64_t removed = 0;   // row ids taken out by lmd_store_remove (their slots stay allocated)

	// Routing sidecars (built lazily, rebuilt when the views changed): per node ONE contiguous block holding the slots of
	// its R neighbours and their compressed vectors, chunk-major so that lane j's loads of neighbour j coalesce.
	//   d_side_t  ternary planes (a transposition of the block's own plane region; ROUTE_TERNARY reads nothing else)
	//   d_codes   int8 edge codec (x ~ code_off[d] + code_scale * c[d]; ROUTE_INT8 and the exact route's pre-filter)
	uint64_t views_gen = 1;              // bumped by everything that changes blocks / d_vecs / d_nbrs
	uint8_t *d_side_t = nullptr;
	uint32_t side_t_bytes = 0, side_t_off = 0;
	uint64_t side_t_gen = 0;
	uint32_t codec = 0;                  // LMD_CODEC_* requested by lmd_store_enable_codec
	uint8_t *d_codes = nullptr;
	float *d_code_off = nullptr;         // [row_floats + 1]: per-dimension offsets, then the scale
	float code_scale = 1.0f;
	float code_rho = 0.0f; // RMS row length (IP stores: the query is scaled to it before it goes onto the grid)
	uint32_t code_block_bytes = 0, code_data_off = 0;
	uint64_t codes_gen = 0;
	bool blocks_released = false;        // lmd_store_release_blocks: d_blocks is gone (search-only store)
	float *d_norms = nullptr;            // [capacity] squared row norms: the view the tensor-core rerank's epilogue reads
	uint64_t norms_gen = 0;

	// graph-build scratch (lmd_build.cu), allocated on first use and kept
	uint32_t *d_inbox = nullptr, *d_inbox_cnt = nullptr, *d_touched = nullptr;
	uint64_t *d_planes = nullptr; // [capacity][2 * words] ternary planes of every node
	uint32_t inbox_cap = 0;

	cudaStream_t &stream = def.stream, &own_stream = def.own_stream; // stream = the active one
	cudaStream_t copy_stream = nullptr;
	cudaEvent_t (&ev)[8] = def.ev;
	lmd::Tuning tune;
	mutable std::unordered_map<uint64_t, std::pair<int, int>> launch_cache; // (kernel, smem) -> (warps per CTA, CTAs per SM)
	mutable std::mutex cache_mu; // launch_cache is filled by concurrent searches
	// Searches hold `mu` shared (they only read the store) plus their own SearchState::mu; everything that changes the
	// store - uploads, apply / remove, finalize, build, tuning, stream binding - holds it exclusively.
	std::shared_mutex mu;
};

// A per-caller search context (include/lmd_b200.h lmd_ctx_*): its own stream, counters and scratch on the store's device.
struct lmd_ctx {
	lmd_store *store = nullptr;
	lmd::SearchState st;
};

// An allowed-ids set resolved to the store's slots (filtered search).
struct lmd_filter {
	lmd_store *store = nullptr;
	uint32_t *d_bits = nullptr; // [(capacity + 31) / 32]
	uint64_t words = 0;
	uint64_t count = 0; // distinct allowed row ids that are in the store
	uint64_t epoch = 0;
};

namespace lmd {

// SearchState life cycle (lmd_store.cu): stream, events, counters; the scratch grows on demand in lmd_api.cu.
lmd_status state_init(SearchState &st);
void state_free(SearchState &st);

// Makes the store ready for a search of `route`: rebuilds the neighbour-slot view when dirty and whichever routing
// sidecar the route needs (takes s->mu exclusively). route_ready() is the shared-lock test for the same condition.
lmd_status store_prepare(lmd_store *s, uint32_t route);
lmd_status finalize_locked(lmd_store *s); // the caller holds s->mu exclusively
bool route_ready(const lmd_store *s, uint32_t route);

// Store internals shared with lmd_build.cu (the caller holds s->mu).
lmd_status store_copy_to_device(lmd_store *s, void *dst, const void *src, size_t bytes);
lmd_status store_ingest(lmd_store *s, const int64_t *row_ids, uint64_t base, uint64_t n);
lmd_status store_lookup_slot(lmd_store *s, int64_t row_id, uint32_t *slot);

// Kernel launchers (each returns a status and has set_error() on failure).
struct TraverseArgs {
	const lmd_store *s;
	const float *d_queries; // [nq][dims]
	uint32_t q_stride = 0;  // floats between queries, 0 = dims
	uint32_t nq, k, L, route;
	int64_t *d_out_ids;   // [nq][k] or null when the separate rerank writes them
	float *d_out_dist;    // [nq][k] or null
	uint32_t *d_out_count;
	uint32_t *d_qstats;   // [nq][2] or null
	uint64_t *d_cand;     // [nq][L] or null
	const uint32_t *d_allowed = nullptr; // filtered search: bitmap over slots, null = unfiltered
	bool strict = false;                 // LMD_SEARCH_STRICT_FP
	bool no_prefilter = false;           // LMD_SEARCH_NO_PREFILTER
	unsigned long long *d_counters = nullptr; // the calling SearchState's counters (null = the store's own)
	uint32_t *kroute_out = nullptr;           // receives the kernel-side route that was launched (3 = exact + pre-filter)
};
lmd_status launch_traverse(const TraverseArgs &a, cudaStream_t st);

struct RerankArgs {
	const lmd_store *s;
	const float *d_queries;
	uint32_t nq, k, L;
	const uint64_t *d_cand; // [nq][L], LMD_KEY_NONE padded
	int64_t *d_out_ids;
	float *d_out_dist;
	uint32_t *d_out_count;
	bool strict = false;
	bool tensor = false; // LMD_SEARCH_TENSOR_RERANK
	unsigned long long *d_counters = nullptr;
};
lmd_status launch_rerank(const RerankArgs &a, cudaStream_t st);
// LMD_SEARCH_TENSOR_RERANK (lmd_rerank_tc.cu): tcgen05 contraction staged through bulk async copies
lmd_status launch_rerank_tensor(const RerankArgs &a, cudaStream_t st);
lmd_status build_row_norms(lmd_store *s); // the caller holds s->mu exclusively
#define LMD_PREP_NORMS 0x100u             // or-ed into the route given to route_ready() / store_prepare()

lmd_status launch_merge_topk(const int64_t *d_ids, const float *d_dist, uint32_t shards, uint64_t nq, uint32_t k,
                             int64_t *d_out_ids, float *d_out_dist, cudaStream_t st);

} // namespace lmd
