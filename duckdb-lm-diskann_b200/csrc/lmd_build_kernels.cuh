// Graph write side on the device: GraphManager::RobustPrune
// (core/GraphManager.cpp:259-457) for many nodes at once, and the pieces of GraphManager::AddNode (:65-203) that turn
// it into a batch-parallel Vamana build: candidates from the beam search (the traversal kernel with L = L_insert,
// Searcher.cpp:275-294), prune, write the list, then re-prune every selected neighbour over its own list plus the new
// nodes that chose it (:171-200).
//
// One WARP owns one node. Per node, in the warp's slice of shared memory:
//   s_list [cap] u64   candidates as (distance to the node, slot) keys, sorted ascending - the std::sort of :300; a
//                      duplicate (same slot, hence same distance) is dropped on insertion - the std::unique of :302-305
//   s_sel  [r]   u32   selected neighbours (final_selected_neighbors_rowids, :308)
//   s_nb / s_d / s_d2  the compute_batch() staging of lmd_kernels.cuh
// The alpha rule (:347-362) is evaluated for one candidate p at a time, in list order, against ALL selected r at
// once: p's vector sits in registers (it plays the "query"), the selected rows are gathered TEAM lanes per row exactly
// as the traversal gathers neighbours, and p is dropped when any lane sees alpha * d(p, r) < d(node, p).
#pragma once
#include "lmd_kernels.cuh"

namespace lmd {

struct PruneParams {
	KParams kp; // vecs, row_floats, dims, metric (compute_batch / load_vec read these)
	float alpha;
	uint32_t r;
	uint32_t n_items;
	const uint32_t *item_node; // [n_items] slot of the node to connect, or null: node = node_base + item
	uint32_t node_base;
	// candidate sources (any mix)
	const uint64_t *keys;       // [n_items][keys_stride] presorted (distance, slot) keys, LMD_KEY_NONE padded
	uint32_t keys_stride;
	const uint32_t *cand_slots; // [n_items][cand_stride] candidate slots, LMD_INVALID_SLOT = none
	uint32_t cand_stride;
	uint32_t use_existing;      // the node's current neighbour row (nbrs[node]) joins the candidates (:189-191)
	uint32_t *inbox;            // [N][inbox_cap] back-link candidates pushed by prune_scatter, consumed here
	uint32_t *inbox_cnt;        // [N]
	uint32_t inbox_cap;
	// outputs
	uint32_t *nbrs;       // [N][r] neighbour rows (written unless out_slots is set)
	uint32_t *out_slots;  // [n_items][r] (parity entry point) or null
	uint32_t *out_count;  // [n_items] or null
	// back-link scatter for freshly connected nodes: node -> inbox of each selected neighbour
	uint32_t scatter;
	uint32_t *scatter_inbox; // [N][inbox_cap]; counted by inbox_cnt (the node's own inbox is not read when `inbox` is null)
	uint32_t *touched; // nodes whose inbox went from empty to non-empty
	unsigned long long *counters; // [0] work queue, [1] touched, [2] distance evals, [3] inbox overflows, [4] pruned nodes
	uint32_t list_cap;
	uint32_t warp_smem_bytes;
};

// Distances from the vector in Q to the m (<= 32) nodes in s_nb, finished per the metric; lane j gets node j's.
template <int TEAM, int CH, int G, int METRIC>
__device__ __forceinline__ float batch_distances(const QueryRegs<TEAM, CH> &Q, const KParams &kp, const uint32_t *s_nb,
                                                 float *s_d, float *s_d2, int m, int lane) {
	compute_batch<TEAM, CH, G, METRIC>(Q, kp, s_nb, s_d, s_d2, m, lane);
	__syncwarp();
	const float d = finish_batch<METRIC>(s_d, s_d2, Q.na, m, lane);
	__syncwarp();
	return d;
}

template <int TEAM, int CH, int G, int METRIC>
__global__ void __launch_bounds__(128) prune_kernel(const PruneParams p) {
	extern __shared__ __align__(16) uint8_t smem[];
	const int lane = threadIdx.x & 31;
	uint8_t *wbase = smem + (threadIdx.x >> 5) * p.warp_smem_bytes;
	uint64_t *s_list = reinterpret_cast<uint64_t *>(wbase);
	uint32_t *s_nb = reinterpret_cast<uint32_t *>(s_list + p.list_cap);
	float *s_d = reinterpret_cast<float *>(s_nb + 32);
	float *s_d2 = s_d + 32;
	uint32_t *s_sel = reinterpret_cast<uint32_t *>(s_d2 + 32);
	float *s_q = reinterpret_cast<float *>(s_sel + ((p.r + 31) & ~31u)); // generic shape only
	const KParams &kp = p.kp;
	const int cap = (int)p.list_cap;

	for (;;) {
		uint32_t item = 0;
		if (lane == 0) {
			item = (uint32_t)atomicAdd(&p.counters[0], 1ull);
		}
		item = __shfl_sync(FULL, item, 0);
		if (item >= p.n_items) {
			break;
		}
		const uint32_t node = p.item_node ? p.item_node[item] : p.node_base + item;
		uint32_t n_eval = 0, ovf = 0;
		int n = 0;

		// ---- step 1 + 2 (:272-305): candidates with their distance to the node, sorted, duplicates dropped ----
		if (p.keys) {
			const uint64_t *src = p.keys + (uint64_t)item * p.keys_stride;
			for (uint32_t i = lane; i < p.keys_stride && i < (uint32_t)cap; i += 32) {
				const uint64_t k = src[i];
				// the list is dense from the front; the node itself is never a candidate (:280-282)
				s_list[i] = (k != LMD_KEY_NONE && key_slot(k) != node) ? (k & ~1ull) : LMD_KEY_NONE;
			}
			__syncwarp();
			// count valid entries; a self entry (possible only when the node is already in the graph) is squeezed out
			int cnt = 0, holes = 0;
			for (uint32_t base = 0; base < p.keys_stride && base < (uint32_t)cap; base += 32) {
				const uint32_t i = base + lane;
				const bool in = i < p.keys_stride && i < (uint32_t)cap;
				const uint64_t orig = in ? src[i] : LMD_KEY_NONE;
				cnt += __popc(__ballot_sync(FULL, in && orig != LMD_KEY_NONE));
				holes += __popc(__ballot_sync(FULL, in && orig != LMD_KEY_NONE && key_slot(orig) == node));
			}
			n = cnt;
			if (holes) { // rare: compact by re-inserting everything
				n = 0;
				uint32_t dummy = 0;
				for (int i = 0; i < cnt; i++) {
					const uint64_t k = src[i];
					if (key_slot(k) != node) {
						n = list_insert<16>(s_list, n, cap, cap, k & ~1ull, lane, dummy);
					}
				}
			}
		}
		const bool any_loose = p.cand_slots || p.use_existing || p.inbox;
		if (any_loose) {
			QueryRegs<TEAM, CH> Q;
			load_vec<TEAM, CH>(Q, kp.vecs + (uint64_t)node * kp.row_floats, kp.dims, kp.row_floats, lane, s_q,
			                   CH == 0 && kp.strict);
			uint32_t n_in = 0;
			if (p.inbox) {
				n_in = p.inbox_cnt[node];
				n_in = n_in < p.inbox_cap ? n_in : p.inbox_cap;
			}
			const uint32_t n_old = p.use_existing ? p.r : 0;
			const uint32_t n_ext = p.cand_slots ? p.cand_stride : 0;
			const uint32_t total = n_old + n_in + n_ext;
			for (uint32_t base = 0; base < total; base += 32) {
				const uint32_t t = base + lane;
				uint32_t c = LMD_INVALID_SLOT;
				if (t < n_old) {
					c = p.nbrs[(uint64_t)node * p.r + t];
				} else if (t < n_old + n_in) {
					c = p.inbox[(uint64_t)node * p.inbox_cap + (t - n_old)];
				} else if (t < total) {
					c = p.cand_slots[(uint64_t)item * p.cand_stride + (t - n_old - n_in)];
				}
				const bool ok = c != LMD_INVALID_SLOT && c != node;
				const uint32_t mask = __ballot_sync(FULL, ok);
				const int m = __popc(mask);
				if (m == 0) {
					continue;
				}
				if (ok) {
					s_nb[__popc(mask & ((1u << lane) - 1u))] = c;
				}
				__syncwarp();
				const float d = batch_distances<TEAM, CH, G, METRIC>(Q, kp, s_nb, s_d, s_d2, m, lane);
				const uint32_t mine = lane < m ? s_nb[lane] : 0u;
				__syncwarp();
				n_eval += m;
				for (int j = 0; j < m; j++) {
					const float dj = __shfl_sync(FULL, d, j);
					const uint32_t cj = __shfl_sync(FULL, mine, j);
					uint32_t dummy = 0;
					n = list_insert<16>(s_list, n, cap, cap, make_key(dj, cj), lane, dummy);
				}
			}
			if (p.inbox && lane == 0) {
				p.inbox_cnt[node] = 0;
			}
		}

		// ---- step 3 (:308-379): alpha pruning in list order ----
		uint32_t nsel = 0;
		for (int i = 0; i < n && nsel < p.r; i++) {
			const uint64_t key = s_list[i];
			const uint32_t cand = key_slot(key);
			const float d_np = key_dist(key);
			bool pruned = false;
			if (nsel > 0) {
				QueryRegs<TEAM, CH> Qp;
				load_vec<TEAM, CH>(Qp, kp.vecs + (uint64_t)cand * kp.row_floats, kp.dims, kp.row_floats, lane, s_q,
				                   CH == 0 && kp.strict);
				for (uint32_t b = 0; b < nsel && !pruned; b += 32) {
					const int m = (int)(nsel - b < 32u ? nsel - b : 32u);
					const float dpr = batch_distances<TEAM, CH, G, METRIC>(Qp, kp, s_sel + b, s_d, s_d2, m, lane);
					n_eval += m;
					pruned = __any_sync(FULL, lane < m && p.alpha * dpr < d_np); // :359
				}
			}
			if (!pruned) {
				if (lane == 0) {
					s_sel[nsel] = cand;
				}
				nsel++;
				__syncwarp();
			}
		}

		// ---- step 4 (:388-452): the node's neighbour row ----
		uint32_t *dst = p.out_slots ? p.out_slots + (uint64_t)item * p.r : p.nbrs + (uint64_t)node * p.r;
		for (uint32_t j = lane; j < p.r; j += 32) {
			dst[j] = j < nsel ? s_sel[j] : LMD_INVALID_SLOT;
		}
		if (p.out_count && lane == 0) {
			p.out_count[item] = nsel;
		}
		// ---- AddNode :171-200: every selected neighbour will be re-pruned with this node as a candidate ----
		if (p.scatter) {
			for (uint32_t j = lane; j < nsel; j += 32) {
				const uint32_t v = s_sel[j];
				const uint32_t pos = atomicAdd(&p.inbox_cnt[v], 1u);
				if (pos < p.inbox_cap) {
					p.scatter_inbox[(uint64_t)v * p.inbox_cap + pos] = node;
				} else {
					ovf++;
				}
				if (pos == 0) {
					p.touched[atomicAdd(&p.counters[1], 1ull)] = v;
				}
			}
		}
		ovf = __reduce_add_sync(FULL, ovf);
		if (lane == 0) {
			atomicAdd(&p.counters[2], (unsigned long long)n_eval);
			if (ovf) {
				atomicAdd(&p.counters[3], (unsigned long long)ovf);
			}
			atomicAdd(&p.counters[4], 1ull);
		}
		__syncwarp();
	}
}

template <typename K>
static lmd_status launch_prune(K kernel, PruneParams &p, const lmd_store *s, cudaStream_t st) {
	const size_t smem = (size_t)p.warp_smem_bytes * 4;
	if (smem > (size_t)227 * 1024) {
		set_error("prune kernel: %zu bytes of shared memory per CTA (candidate list too long)", smem);
		return LMD_ERR_UNSUPPORTED;
	}
	LMD_CUDA_OK(cudaFuncSetAttribute(kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, 227 * 1024));
	int per_sm = 0;
	LMD_CUDA_OK(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, kernel, 128, smem));
	if (per_sm < 1) {
		set_error("prune kernel does not fit on an SM");
		return LMD_ERR_UNSUPPORTED;
	}
	uint32_t grid = (p.n_items + 3) / 4;
	const uint32_t cap = (uint32_t)(per_sm * s->sm_count);
	grid = grid > cap ? cap : grid;
	kernel<<<grid ? grid : 1, 128, smem, st>>>(p);
	LMD_CUDA_OK(cudaGetLastError());
	return LMD_OK;
}

#define LMD_DISPATCH_PRUNE_SHAPE(METRIC)                                                                               \
	if (p.kp.strict) {                                                                                                 \
		return launch_prune(prune_kernel<32, 0, 1, METRIC>, p, s, st);                                                 \
	} else if (p.kp.row_floats == 128) {                                                                                      \
		return launch_prune(prune_kernel<8, 4, 2, METRIC>, p, s, st);                                                  \
	} else if (p.kp.row_floats == 96) {                                                                                \
		return launch_prune(prune_kernel<8, 3, 2, METRIC>, p, s, st);                                                  \
	} else if (p.kp.row_floats == 768) {                                                                               \
		return launch_prune(prune_kernel<32, 6, 6, METRIC>, p, s, st);                                                 \
	} else {                                                                                                           \
		return launch_prune(prune_kernel<32, 0, 1, METRIC>, p, s, st);                                                 \
	}

} // namespace lmd
