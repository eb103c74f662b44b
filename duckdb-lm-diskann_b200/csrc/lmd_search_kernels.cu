// Host side of the search kernels: launch configuration (list / filter sizing), dispatch to the per-variant
// translation units (lmd_kernel_inst.cu compiled once per kernel kind, metric and route), and the k-way merge kernel
// of the dataset-sharded mode.
#include "lmd_kernels.cuh"
#include "lmd_plan.hpp"

namespace lmd {

// k-way merge of per-shard results: one thread per query, shards*k <= a few hundred.
__global__ void merge_topk_kernel(const int64_t *__restrict__ ids, const float *__restrict__ dist, uint32_t shards,
                                  uint64_t nq, uint32_t k, int64_t *__restrict__ out_ids,
                                  float *__restrict__ out_dist) {
	uint64_t q = blockIdx.x * (uint64_t)blockDim.x + threadIdx.x;
	if (q >= nq) {
		return;
	}
	// every shard's list is ascending by (distance, id): classic k-way merge with one cursor per shard
	uint32_t cur[64];
	for (uint32_t s = 0; s < shards; s++) {
		cur[s] = 0;
	}
	for (uint32_t o = 0; o < k; o++) {
		int best_s = -1;
		float bd = 0.0f;
		int64_t bi = 0;
		for (uint32_t s = 0; s < shards; s++) {
			if (cur[s] >= k) {
				continue;
			}
			uint64_t at = ((uint64_t)s * nq + q) * k + cur[s];
			int64_t id = ids[at];
			if (id < 0) {
				cur[s] = k;
				continue;
			}
			float d = dist[at];
			if (best_s < 0 || d < bd || (d == bd && id < bi)) {
				best_s = (int)s;
				bd = d;
				bi = id;
			}
		}
		if (best_s < 0) {
			out_ids[q * k + o] = -1;
			out_dist[q * k + o] = 0.0f;
		} else {
			out_ids[q * k + o] = bi;
			out_dist[q * k + o] = bd;
			cur[best_s]++;
		}
	}
}

// Defined by lmd_kernel_inst.cu, one object file per (metric, route) / (metric).
#define LMD_DECL_TRAVERSE(M, R)                                                                                        \
	lmd_status dispatch_traverse_m##M##_r##R(KParams &p, const lmd_store *s, uint32_t nq, int lpl, cudaStream_t st);
#define LMD_DECL_RERANK(M) lmd_status dispatch_rerank_m##M(KParams &p, const lmd_store *s, uint32_t nq, int lpl, cudaStream_t st);
LMD_DECL_TRAVERSE(1, 0)
LMD_DECL_TRAVERSE(2, 0)
LMD_DECL_TRAVERSE(3, 0)
LMD_DECL_TRAVERSE(2, 1)
LMD_DECL_TRAVERSE(3, 1)
LMD_DECL_TRAVERSE(1, 2) // int8 routing (L2)
LMD_DECL_TRAVERSE(1, 3) // exact routing with the int8 pre-filter (L2)
LMD_DECL_TRAVERSE(2, 3) // ... and COSINE / IP (lower bound from the grid dot product)
LMD_DECL_TRAVERSE(3, 3)
#define LMD_DECL_TRAVERSE_FILT(M, R)                                                                                   \
	lmd_status dispatch_traverse_filt_m##M##_r##R(KParams &p, const lmd_store *s, uint32_t nq, int lpl, cudaStream_t st);
LMD_DECL_TRAVERSE_FILT(1, 0)
LMD_DECL_TRAVERSE_FILT(2, 0)
LMD_DECL_TRAVERSE_FILT(3, 0)
LMD_DECL_TRAVERSE_FILT(2, 1)
LMD_DECL_TRAVERSE_FILT(3, 1)
#define LMD_DECL_TRAVERSE_LAT(M)                                                                                        \
	lmd_status dispatch_traverse_lat_m##M(KParams &p, const lmd_store *s, uint32_t nq, int lpl, cudaStream_t st);
LMD_DECL_TRAVERSE_LAT(1)
LMD_DECL_TRAVERSE_LAT(2)
LMD_DECL_TRAVERSE_LAT(3)
#define LMD_DECL_TRAVERSE_MID(M)                                                                                        \
	lmd_status dispatch_traverse_mid_m##M(KParams &p, const lmd_store *s, uint32_t nq, int lpl, cudaStream_t st);
LMD_DECL_TRAVERSE_MID(1)
LMD_DECL_TRAVERSE_MID(2)
LMD_DECL_TRAVERSE_MID(3)
LMD_DECL_RERANK(1)
LMD_DECL_RERANK(2)
LMD_DECL_RERANK(3)

static KParams base_params(const lmd_store *s, const float *d_queries, uint32_t nq, uint32_t k, uint32_t L) {
	KParams p;
	memset(&p, 0, sizeof(p));
	p.vecs = s->d_vecs;
	p.nbrs = s->d_nbrs;
	p.row_ids = s->d_row_ids;
	p.blocks = s->d_blocks;
	p.queries = d_queries;
	p.counters = s->d_counters;
	p.block_bytes = s->cfg.block_bytes;
	p.pos_off = s->lay.pos_off;
	p.neg_off = s->lay.neg_off;
	p.nq = nq;
	p.q_stride = s->cfg.dims;
	p.dims = s->cfg.dims;
	p.row_floats = s->row_floats;
	p.r = s->cfg.r;
	p.words = s->words;
	p.L = L;
	p.k = k;
	p.metric = s->cfg.metric;
	p.entry_slot = s->entry_slot;
	p.has_entry = s->has_entry ? 1u : 0u;
	return p;
}

static lmd_status dispatch_traverse(KParams &p, const lmd_store *s, uint32_t nq, int lpl, uint32_t route,
                                    cudaStream_t st) {
	if (p.allowed) { // filtered search (row f4)
		if (route == LMD_ROUTE_EXACT) {
			switch (p.metric) {
			case LMD_METRIC_L2:
				return dispatch_traverse_filt_m1_r0(p, s, nq, lpl, st);
			case LMD_METRIC_COSINE:
				return dispatch_traverse_filt_m2_r0(p, s, nq, lpl, st);
			case LMD_METRIC_IP:
				return dispatch_traverse_filt_m3_r0(p, s, nq, lpl, st);
			}
		} else {
			switch (p.metric) {
			case LMD_METRIC_COSINE:
				return dispatch_traverse_filt_m2_r1(p, s, nq, lpl, st);
			case LMD_METRIC_IP:
				return dispatch_traverse_filt_m3_r1(p, s, nq, lpl, st);
			}
		}
		set_error("no filtered traversal kernel for this configuration (L=%u)", p.L);
		return LMD_ERR_UNSUPPORTED;
	}
	if (route == LMD_ROUTE_INT8) {
		return dispatch_traverse_m1_r2(p, s, nq, lpl, st);
	}
	if (route == LMD_KROUTE_EXACT_PF) {
		switch (p.metric) {
		case LMD_METRIC_COSINE:
			return dispatch_traverse_m2_r3(p, s, nq, lpl, st);
		case LMD_METRIC_IP:
			return dispatch_traverse_m3_r3(p, s, nq, lpl, st);
		default:
			return dispatch_traverse_m1_r3(p, s, nq, lpl, st);
		}
	}
	if (route == LMD_ROUTE_EXACT && p.latency == 2) {
		switch (p.metric) {
		case LMD_METRIC_L2:
			return dispatch_traverse_mid_m1(p, s, nq, lpl, st);
		case LMD_METRIC_COSINE:
			return dispatch_traverse_mid_m2(p, s, nq, lpl, st);
		case LMD_METRIC_IP:
			return dispatch_traverse_mid_m3(p, s, nq, lpl, st);
		}
	}
	if (route == LMD_ROUTE_EXACT && p.latency) {
		switch (p.metric) {
		case LMD_METRIC_L2:
			return dispatch_traverse_lat_m1(p, s, nq, lpl, st);
		case LMD_METRIC_COSINE:
			return dispatch_traverse_lat_m2(p, s, nq, lpl, st);
		case LMD_METRIC_IP:
			return dispatch_traverse_lat_m3(p, s, nq, lpl, st);
		}
	}
	if (route == LMD_ROUTE_EXACT) {
		switch (p.metric) {
		case LMD_METRIC_L2:
			return dispatch_traverse_m1_r0(p, s, nq, lpl, st);
		case LMD_METRIC_COSINE:
			return dispatch_traverse_m2_r0(p, s, nq, lpl, st);
		case LMD_METRIC_IP:
			return dispatch_traverse_m3_r0(p, s, nq, lpl, st);
		}
	} else {
		switch (p.metric) {
		case LMD_METRIC_COSINE:
			return dispatch_traverse_m2_r1(p, s, nq, lpl, st);
		case LMD_METRIC_IP:
			return dispatch_traverse_m3_r1(p, s, nq, lpl, st);
		}
	}
	set_error("no traversal kernel for this configuration (L=%u)", p.L);
	return LMD_ERR_UNSUPPORTED;
}

static lmd_status dispatch_rerank(KParams &p, const lmd_store *s, uint32_t nq, int lpl, cudaStream_t st) {
	switch (p.metric) {
	case LMD_METRIC_L2:
		return dispatch_rerank_m1(p, s, nq, lpl, st);
	case LMD_METRIC_COSINE:
		return dispatch_rerank_m2(p, s, nq, lpl, st);
	case LMD_METRIC_IP:
		return dispatch_rerank_m3(p, s, nq, lpl, st);
	}
	set_error("no rerank kernel for this configuration (L=%u)", p.L);
	return LMD_ERR_UNSUPPORTED;
}

lmd_status launch_traverse(const TraverseArgs &a, cudaStream_t st) {
	const lmd_store *s = a.s;
	KParams p = base_params(s, a.d_queries, a.nq, a.k, a.L);
	p.out_ids = a.d_out_ids;
	p.out_dist = a.d_out_dist;
	p.out_count = a.d_out_count;
	p.qstats = a.d_qstats;
	p.cand = a.d_cand;
	if (a.q_stride) {
		p.q_stride = a.q_stride;
	}
	p.strict = (a.strict || s->tune.strict_fp) ? 1u : 0u;
	if (a.d_counters) {
		p.counters = a.d_counters;
	}
	// room behind the top L for distance ties (see list_insert): integer ternary distances tie all the time, fp32
	// distances only for duplicate vectors
	// ... and, under LMD_SEARCH_STRICT_FP, whenever two different sums round to the same sqrtf (about once per thousand
	// queries at L = 64): the oracle expands such an entry, so strict parity of the expansion counts needs the region too
	uint32_t tie_slots = s->tune.tie_slots >= 0 ? (uint32_t)s->tune.tie_slots
	                                            : (a.route != LMD_ROUTE_EXACT ? plan_tie_slots(a.L) : (p.strict ? 32u : 0u));
	// exact routing on a store with the int8 codec: the codes decide which rows are fetched (same results, fewer bytes)
	// ... for batches that fill the GPU. Small batches (one query, index-join chunks) are latency-bound, not byte-bound: they
	// keep the plain kernels that gather a whole expansion in one pass (one query 230 us against 271 us pre-filtered;
	// "prefilter" = 2 forces the pre-filter for every batch size).
	uint32_t kroute = a.route;
	const bool specialised_shape = !p.strict && (p.row_floats == 128 || p.row_floats == 96);
	const bool small_batch = specialised_shape && plan_family(a.nq, (uint32_t)s->sm_count, s->tune.latency_mode) != 0;
	if (a.route == LMD_ROUTE_EXACT && s->codec == LMD_CODEC_INT8 && s->codes_gen == s->views_gen && !a.no_prefilter &&
	    !a.d_allowed && !p.strict && s->tune.prefilter &&
	    (!small_batch || s->tune.prefilter == 2)) {
		kroute = LMD_KROUTE_EXACT_PF;
	}
	if (a.kroute_out) {
		*a.kroute_out = kroute;
	}
	if (a.route == LMD_ROUTE_TERNARY) {
		p.side = s->d_side_t;
		p.side_bytes = s->side_t_bytes;
		p.side_data_off = s->side_t_off;
	} else if (kroute != LMD_ROUTE_EXACT) {
		p.side = s->d_codes;
		p.side_bytes = s->code_block_bytes;
		p.side_data_off = s->code_data_off;
		p.code_off = s->d_code_off;
		p.code_scale = s->code_scale;
		p.code_rho = s->code_rho;
	}
	int lpl = lpl_for(a.L + tie_slots);
	if (!lpl) {
		lpl = lpl_for(a.L);
	}
	if (!lpl) {
		set_error("search list size L = %u exceeds this build's limit of 512", a.L);
		return LMD_ERR_UNSUPPORTED;
	}
	// small batches: the latency variants (all of an expansion's gathers in flight at once), see lmd_kernels.cuh
	const bool specialised = !p.strict && (p.row_floats == 128 || p.row_floats == 96); // 768-d rows already run with G = 6
	const int lat = s->tune.latency_mode;
	if (kroute == LMD_ROUTE_EXACT && !a.d_allowed && specialised) {
		// latency_mode: -1 = by batch size, 0 = throughput kernels only, 1 = G = 8 kernels, 2 = G = 4 kernels
		p.latency = (uint32_t)plan_family(a.nq, (uint32_t)s->sm_count, lat);
		if (p.latency) {
			lpl = lpl_coarse(lpl);
		}
	}
	if (a.d_allowed) {
		lpl = lpl_coarse(lpl); // the filtered kernels exist for these list sizes only
		p.allowed = a.d_allowed;
	}
	p.nbr_prefetch = (s->tune.nbr_prefetch == 2 || (s->tune.nbr_prefetch == 1 && p.latency)) ? 1u : 0u;
	p.list_cap = (uint32_t)lpl * 32;
	p.merge_min = (uint32_t)(s->tune.merge_min > 0 ? s->tune.merge_min : 5);
	p.prefetch_planes = s->tune.prefetch_planes ? 1u : 0u;
	p.pf_g = s->tune.pf_g == 1 ? 1u : 2u;
	p.row_prefetch = (uint32_t)s->tune.row_prefetch;
	const bool generic = p.strict || !(p.row_floats == 128 || p.row_floats == 96 || p.row_floats == 768);
	size_t fixed = (size_t)lpl * 32 * 8 + 384 + (generic ? (size_t)p.row_floats * 4 : 0) +
	               (a.route == LMD_ROUTE_TERNARY ? (size_t)p.words * 16 : 0) +
	               (kroute == LMD_ROUTE_INT8 || kroute == LMD_KROUTE_EXACT_PF ? (size_t)p.row_floats : 0);
	const size_t res_bytes = a.d_allowed ? (size_t)lpl * 32 * 8 : 0; // the results list RH of a filtered search
	fixed += res_bytes;
	// The 128-d / 96-d exact kernels keep 2 rows in flight per team and need 24 warps on an SM to cover HBM latency: a
	// list of 128+ entries with the 8 KB filter leaves room for 20-22, and the 4 KB filter with its re-evaluations (3 % of
	// the rows at L = 128, 17 % at L = 256) is the faster trade: 1 M x 128, L = 128 / 160 / 256: 2.94 -> 2.47, 3.82 ->
	// 3.25, 6.20 -> 5.73 ms per 10 k queries (profiles/r44_long_lists_d*.json). Other kernels hold fewer warps anyway.
	// The filtered kernels (72 registers up to 128 entries) gain the same way up to L = 128 (1 M x 128, all rows allowed,
	// L = 64 / 128: 1.79 -> 1.52, 3.66 -> 3.15 ms); at L = 256 they hold 96 registers, the small filter adds 14 % of
	// re-evaluations and loses (8.76 -> 9.97 ms), so longer filtered lists keep the old rule (profiles/r47_filtered.json).
	// Small batches run 12 (G = 8) or 16 (G = 4) warps per SM at most, which leaves room for the filter that never forgets:
	// 1 M x 128, L = 128, 16 KB against 4 KB: 1 query 383 -> 348 us, 1 024 queries 564 -> 524 us
	// (profiles/r50_small_batch_L128.json).
	const bool exact = a.route == LMD_ROUTE_EXACT; // (with or without the pre-filter: rows are gathered either way)
	const uint32_t min_warps = plan_min_warps((int)p.latency, exact, specialised, a.d_allowed != nullptr, lpl, a.L);
	FilterPlan fp = plan_filter(s->n, s->cfg.r, s->words, a.L, fixed, s->tune.plane_filter_cap != 0 && !exact, min_warps,
	                            plan_filter_mode(s->tune.filter_mode, kroute != LMD_ROUTE_EXACT, a.d_allowed != nullptr),
	                            (int)s->tune.filter_log2, s->tune.filter_entries);
	if (specialised && !a.d_allowed && p.latency == 0 && s->tune.filter_mode < 0 && s->tune.filter_log2 == 0) {
		// a 7th CTA per SM for the 72-register exact kernels: the filter takes what 28 warps leave (lmd_plan.hpp)
		FilterPlan fm;
		fm.entries = 0;
		if (kroute == LMD_KROUTE_EXACT_PF) {
			fm = plan_filter_modulo(s->n, fixed, 28);
		} else if (kroute == LMD_ROUTE_EXACT && p.row_floats == 128 && lpl <= 2) {
			fm = plan_filter_modulo(s->n, fixed, 28, 196, s->cfg.r, a.L, 0.6);
		}
		// (pre-filtered kernel: also where the power-of-two rule had shrunk to 4 KB for a 128-entry list - 7 CTAs hold the
		// 6 KB table too, and it forgets less: 1 M x 128 L = 100 / 128 in flight 1.537 -> 1.471 / 1.894 -> 1.814 ms;
		// 10 M x 96 L = 96, 100 k queries: kernel 12.5 -> 12.3 ms; profiles/r2_41_pf_long_lists.json)
		if (fm.entries && (fm.bytes < fp.bytes || kroute == LMD_KROUTE_EXACT_PF)) {
			fp = fm;
			// two batches in flight: CTAs of one warp hand their SM share to the next kernel as soon as that warp runs out
			// of queries (pre-filtered L = 56: 0.814 -> 0.794 ms; the plain kernels: no difference; profiles/r2_35_cta_*)
			// ... which only matters while a batch is a few rounds of the resident warps: larger batches keep the CTAs of 4
			// (100 k queries over 10 M x 96 rows: 13.2 ms either way, profiles/r2_42_bench_n1.json, r2_43_bench_n1.json)
			p.cta_warps = kroute == LMD_KROUTE_EXACT_PF && a.nq <= 8u * 28u * (uint32_t)s->sm_count ? 1u : 0u;
		}
	}
	if (kroute == LMD_ROUTE_EXACT && !p.strict && plan_96d_prefers_g4(p.row_floats, exact, a.d_allowed != nullptr, (int)p.latency, lat, lpl, fp.bytes)) {
		// 384-byte rows, and shared memory holds the G = 2 kernel at 24 warps or fewer: 4 rows in flight per team win
		// (10 M x 96, L = 48 / 128: 1.46 -> 1.37 / 3.53 -> 3.34 ms per 12.5 k queries; 1 M x 96, L = 64: 1.21 -> 1.13;
		// profiles/r43_g_sweep_*.json; those kernels exist for lists of 64 and 128 entries, not 96). With the 4 KB filter
		// the G = 2 kernel runs 32 warps and stays ahead.
		p.latency = 2;
	}
	p.filter_mode = fp.mode;
	p.filter_bytes = fp.bytes;
	p.filter_entries = fp.entries;
	p.filter_magic = fp.entries ? (uint32_t)(((1ull << 40) + fp.entries - 1) / fp.entries) : 0u;
	p.filter_mask = (1u << fp.lg) - 1;
	p.filter_shift = 32 - fp.lg;
	p.filter_bmask = (1u << (fp.lg - 3)) - 1;
	p.filter_hmask = fp.mode == 2 ? (uint32_t)((1ull << (fp.lg + 15)) - 1) : (uint32_t)((1ull << (fp.lg - 3 + 15)) - 1);
	p.res_off = (uint32_t)((fixed - res_bytes + fp.bytes + 15) / 16 * 16);
	p.warp_smem_bytes = (uint32_t)(p.res_off + (res_bytes + 15) / 16 * 16);
	return dispatch_traverse(p, s, a.nq, lpl, kroute, st);
}

lmd_status launch_rerank(const RerankArgs &a, cudaStream_t st) {
	if (a.tensor) {
		return launch_rerank_tensor(a, st);
	}
	const lmd_store *s = a.s;
	KParams p = base_params(s, a.d_queries, a.nq, a.k, a.L);
	p.out_ids = a.d_out_ids;
	p.out_dist = a.d_out_dist;
	p.out_count = a.d_out_count;
	p.cand = const_cast<uint64_t *>(a.d_cand);
	p.strict = (a.strict || s->tune.strict_fp) ? 1u : 0u;
	if (a.d_counters) {
		p.counters = a.d_counters;
	}
	const int lpl = lpl_for(a.L);
	if (!lpl) {
		set_error("search list size L = %u exceeds this build's limit of 512", a.L);
		return LMD_ERR_UNSUPPORTED;
	}
	const bool generic = p.strict || !(p.row_floats == 128 || p.row_floats == 96 || p.row_floats == 768);
	p.warp_smem_bytes = (uint32_t)(384 + (generic ? (size_t)p.row_floats * 4 : 0));
	return dispatch_rerank(p, s, a.nq, lpl, st);
}

lmd_status launch_merge_topk(const int64_t *d_ids, const float *d_dist, uint32_t shards, uint64_t nq, uint32_t k,
                             int64_t *d_out_ids, float *d_out_dist, cudaStream_t st) {
	if (shards == 0 || shards > 64) {
		set_error("lmd_merge_topk: shards must be in [1, 64]");
		return LMD_ERR_INVALID_ARG;
	}
	if (nq == 0) {
		return LMD_OK;
	}
	merge_topk_kernel<<<(unsigned)((nq + 127) / 128), 128, 0, st>>>(d_ids, d_dist, shards, nq, k, d_out_ids,
	                                                                d_out_dist);
	LMD_CUDA_OK(cudaGetLastError());
	return LMD_OK;
}

} // namespace lmd
