// lmd_search_batch_static: the batched search of lmd_search_batch for callers that cannot allocate device memory (code
// under CUDA graph capture, tight serving loops, harnesses that forbid cudaMalloc). No counterpart in the reference.
// The same traversal (+ rerank) kernels as every other entry point, with
// their SCRATCH - the work counter / statistics block and, for the compressed routes, the [nq][L] candidate keys the
// rerank kernel reads - placed in a file-scope __device__ array of LMD_STATIC_SCRATCH_BYTES instead of cudaMalloc'ed
// memory. Every device that loads the library holds that array until the process exits.
//
// This file calls no cudaMalloc / cudaFree / cudaHostAlloc and no allocating library routine (what the launch path may
// still do on the HOST: one entry in the store's launch cache, a std::unordered_map, the first time a kernel / shared-memory
// size pair is launched, after its occupancy query). The store's own device
// memory was allocated earlier, by lmd_store_create / upload / a first regular search; a route whose sidecar is not built yet is
// refused here (building one allocates), not prepared.
//
// ALL calls on one device share the array, and it keeps the last call's contents: the counter block is cleared on the
// caller's stream before every launch, the candidate region is fully rewritten by the traversal before the rerank reads
// it. Two calls that overlap on the GPU (different streams, no dependency between them) RACE on it - callers serialise.
#include "lmd_common.cuh"

#include <mutex>
#include <shared_mutex>

using namespace lmd;

namespace {

constexpr size_t COUNTER_BYTES = 256; // 16 x u64: [0] work counter, [1..7] statistics, [8..9] store-side scratch
static_assert(LMD_STATIC_SCRATCH_BYTES > COUNTER_BYTES && LMD_STATIC_SCRATCH_BYTES % 256 == 0, "scratch size");

// In HOST code this name is a stand-in for the symbol, not a device address: the address the kernels get comes from
// cudaGetSymbolAddress() below, once per device, with that device current.
static __device__ __align__(256) unsigned char g_static_scratch[LMD_STATIC_SCRATCH_BYTES];

constexpr int MAX_DEVICES = 64;
std::mutex g_setup_mu;
unsigned char *g_scratch_addr[MAX_DEVICES] = {};

lmd_status scratch_for_device(int device, unsigned char **out) {
	if (device < 0 || device >= MAX_DEVICES) {
		set_error("lmd_search_batch_static: device %d out of range", device);
		return LMD_ERR_INVALID_ARG;
	}
	std::lock_guard<std::mutex> g(g_setup_mu);
	if (!g_scratch_addr[device]) {
		void *p = nullptr;
		LMD_CUDA_OK(cudaSetDevice(device));
		cudaError_t e = cudaGetSymbolAddress(&p, g_static_scratch);
		if (e != cudaSuccess || !p) {
			set_error("cudaGetSymbolAddress(static scratch) on device %d failed: %s", device, cudaGetErrorString(e));
			return LMD_ERR_CUDA;
		}
		g_scratch_addr[device] = static_cast<unsigned char *>(p);
	}
	*out = g_scratch_addr[device];
	return LMD_OK;
}

// device memory, or pinned host memory mapped into the device's address space (*alias = the address kernels use)
bool device_visible(const void *p, void **alias) {
	*alias = const_cast<void *>(p);
	cudaPointerAttributes at;
	if (cudaPointerGetAttributes(&at, p) != cudaSuccess) {
		cudaGetLastError();
		return false;
	}
	if (at.type == cudaMemoryTypeDevice || at.type == cudaMemoryTypeManaged) {
		return true;
	}
	if (at.type == cudaMemoryTypeHost && at.devicePointer) {
		*alias = at.devicePointer;
		return true;
	}
	return false;
}

} // namespace

extern "C" {

uint64_t lmd_static_scratch_bytes(void) {
	return LMD_STATIC_SCRATCH_BYTES;
}

lmd_status lmd_search_batch_static(lmd_store *s, const float *queries, uint64_t nq, uint32_t k, uint32_t L, uint32_t route,
                                   uint32_t flags, int64_t *out_ids, float *out_dist, uint32_t *out_count, void *cuda_stream) {
	if (!s) {
		set_error("lmd_search_batch_static: null store");
		return LMD_ERR_INVALID_ARG;
	}
	if (k == 0) { // core/Coordinator.cpp:72-74, as lmd_search_batch
		set_error("k_neighbors cannot be 0 for search");
		return LMD_ERR_INVALID_ARG;
	}
	if (nq == 0) {
		return LMD_OK;
	}
	if (!queries || !out_ids || !out_dist || !out_count) {
		set_error("lmd_search_batch_static: null queries / out_ids / out_dist / out_count");
		return LMD_ERR_INVALID_ARG;
	}
	if (flags & ~(uint32_t)LMD_SEARCH_ASYNC) {
		set_error("lmd_search_batch_static: the only flag it takes is LMD_SEARCH_ASYNC");
		return LMD_ERR_INVALID_ARG;
	}
	if (route != LMD_ROUTE_EXACT && route != LMD_ROUTE_TERNARY && route != LMD_ROUTE_INT8) {
		set_error("lmd_search_batch_static: unknown route %u", route);
		return LMD_ERR_INVALID_ARG;
	}
	if (route == LMD_ROUTE_TERNARY && s->cfg.metric == LMD_METRIC_L2) { // core/distance.hpp:236-239
		set_error("L2 metric is not directly compatible with the current ternary CalculateApproxDistance");
		return LMD_ERR_UNSUPPORTED;
	}
	if (route == LMD_ROUTE_INT8 && s->cfg.metric != LMD_METRIC_L2) {
		set_error("the int8 edge codec is defined for the L2 metric only");
		return LMD_ERR_UNSUPPORTED;
	}
	uint32_t l_used = L > 0 ? L : s->cfg.l_search; // core/Searcher.cpp:261-265
	if (l_used == 0) {
		l_used = k > 20 ? k + 20 : 40;
	}
	l_used = l_used > k ? l_used : k;
	if (l_used > 512) {
		set_error("search list size L = %u exceeds this build's limit of 512", l_used);
		return LMD_ERR_UNSUPPORTED;
	}
	// the limit: the compressed routes keep nq * L candidate keys (8 bytes each) behind the counter block
	const bool separate = route != LMD_ROUTE_EXACT;
	const uint64_t cand_bytes = separate ? nq * (uint64_t)l_used * 8 : 0;
	if (nq > 0xFFFFFFF0ull || COUNTER_BYTES + cand_bytes > LMD_STATIC_SCRATCH_BYTES) {
		set_error("lmd_search_batch_static: %llu queries at L = %u need %llu bytes of scratch, the static array holds %llu "
		          "(split the batch or use lmd_search_batch)",
		          (unsigned long long)nq, l_used, (unsigned long long)(COUNTER_BYTES + cand_bytes),
		          (unsigned long long)LMD_STATIC_SCRATCH_BYTES);
		return LMD_ERR_INVALID_ARG;
	}
	std::shared_lock<std::shared_mutex> rd(s->mu);
	if (!route_ready(s, route)) {
		set_error("lmd_search_batch_static: the store is not prepared for route %u (preparing allocates: run one "
		          "lmd_search_batch on this route first)", route);
		return LMD_ERR_STATE;
	}
	unsigned char *scratch = nullptr;
	lmd_status rc = scratch_for_device(s->device, &scratch);
	if (rc != LMD_OK) {
		return rc;
	}
	LMD_CUDA_OK(cudaSetDevice(s->device));
	void *a_q, *a_ids, *a_dist, *a_count;
	if (!device_visible(queries, &a_q) || !device_visible(out_ids, &a_ids) || !device_visible(out_dist, &a_dist) ||
	    !device_visible(out_count, &a_count)) {
		set_error("lmd_search_batch_static: queries and outputs must be device or pinned host memory (staging pageable "
		          "memory would allocate)");
		return LMD_ERR_INVALID_ARG;
	}
	cudaStream_t st = (cudaStream_t)cuda_stream;
	unsigned long long *d_counters = reinterpret_cast<unsigned long long *>(scratch);
	uint64_t *d_cand = separate ? reinterpret_cast<uint64_t *>(scratch + COUNTER_BYTES) : nullptr;
	LMD_CUDA_OK(cudaMemsetAsync(d_counters, 0, COUNTER_BYTES, st)); // the work counter must start at zero

	TraverseArgs ta;
	ta.s = s;
	ta.d_queries = static_cast<const float *>(a_q);
	ta.nq = (uint32_t)nq;
	ta.k = k;
	ta.L = l_used;
	ta.route = route;
	ta.d_out_ids = separate ? nullptr : static_cast<int64_t *>(a_ids);
	ta.d_out_dist = separate ? nullptr : static_cast<float *>(a_dist);
	ta.d_out_count = separate ? nullptr : static_cast<uint32_t *>(a_count);
	ta.d_qstats = nullptr;
	ta.d_cand = d_cand;
	ta.d_counters = d_counters;
	if ((rc = launch_traverse(ta, st)) != LMD_OK) {
		return rc;
	}
	if (separate) {
		RerankArgs ra;
		ra.s = s;
		ra.d_queries = static_cast<const float *>(a_q);
		ra.nq = (uint32_t)nq;
		ra.k = k;
		ra.L = l_used;
		ra.d_cand = d_cand;
		ra.d_out_ids = static_cast<int64_t *>(a_ids);
		ra.d_out_dist = static_cast<float *>(a_dist);
		ra.d_out_count = static_cast<uint32_t *>(a_count);
		ra.d_counters = d_counters;
		if ((rc = launch_rerank(ra, st)) != LMD_OK) {
			return rc;
		}
	}
	if (flags & LMD_SEARCH_ASYNC) {
		return LMD_OK; // enqueued on the caller's stream; the caller synchronises
	}
	cudaError_t e = cudaStreamSynchronize(st);
	if (e != cudaSuccess) {
		set_error("lmd_search_batch_static: search kernels failed: %s", cudaGetErrorString(e));
		return LMD_ERR_CUDA;
	}
	return LMD_OK;
}

} // extern "C"
