// Traversal + rerank kernels of the LM_DISKANN query path for sm_100a.
//
// Replaces core/Searcher.cpp:45-246 (PerformSearch) with the semantics of the patched oracle
// (oracle/patches/searcher_fixed.patch, hunks H1-H4). One WARP owns one query from start to finish;
// a persistent grid pulls queries from a global counter. Per query, in the warp's slice of shared memory:
//
//   s_list   [LPL*32] u64  the top-L list (core/Searcher.cpp:64-65 `top_candidates` and :67 `candidate_beam`
//                          in one structure): sorted keys (distance, slot, expanded-bit). The closest entry
//                          whose expanded bit is clear is the beam's top(); the search ends when none is left,
//                          which is the H2 `break` (the beam only holds list members and evicted, farther nodes).
//   s_filter [2^f]   u32   visited filter (core/Searcher.cpp:63 `visited_nodes`), open addressing with at most 8
//                          probes; when a probe window is full the home slot is overwritten. A forgotten node can
//                          only be re-evaluated, never re-admitted: it is either still in the list (dropped as a
//                          duplicate at insert) or was rejected/evicted with distance >= the list's worst, which
//                          only shrinks. Results are therefore independent of the filter size.
//   s_nb/s_d/s_d2 [32]     compacted new neighbour slots of the current expansion and their raw distance sums.
//
// Memory traffic per expansion (exact routing): one 4*R byte row of the slot sidecar, then one row_floats*4 byte
// row of the packed vector table per NEW neighbour, fetched with 128-bit loads by TEAM lanes per row, G rows per
// team in flight at once.
//
// This header holds the device code and the launch helpers; it is compiled once per (kernel kind, metric, route)
// by lmd_kernel_inst.cu so that the template variants build in parallel.
#pragma once
#include "lmd_common.cuh"

#include <algorithm>

namespace lmd {

struct KParams {
	const float *__restrict__ vecs;
	const uint32_t *__restrict__ nbrs;
	const int64_t *__restrict__ row_ids;
	const uint8_t *__restrict__ blocks;
	const float *__restrict__ queries;
	int64_t *out_ids;
	float *out_dist;
	uint32_t *out_count;
	uint32_t *qstats;
	uint64_t *cand;
	unsigned long long *counters; // [0] next query, [1] E, [2] V, [3] filter evictions, [4] reranked, [5] tie overflows, [6] merge fallbacks
	uint64_t block_bytes, pos_off, neg_off;
	uint32_t nq, dims, row_floats, r, words;
	uint32_t q_stride; // floats between consecutive queries (dims for a caller's batch, row_floats for store rows)
	uint32_t L, k, metric, list_cap, merge_min;
	uint32_t entry_slot, has_entry;
	uint32_t filter_mask, filter_shift; // filter_mode 0: slots - 1, 32 - log2(slots)
	uint32_t filter_mode;               // 0 = u32 open addressing, 1 = two-choice buckets of 8 x 16-bit entries
	uint32_t filter_bmask, filter_hmask; // mode 1: buckets - 1, 2^(log2(buckets) + 15) - 1
	uint32_t filter_bytes;              // bytes of the table (either mode)
	uint32_t filter_entries, filter_magic; // mode 3: entries S (any multiple of 8), ceil(2^40 / S)
	uint32_t prefetch_planes;           // ternary routing: also prefetch the runner-up's plane region
	uint32_t warp_smem_bytes;
	// filtered search: bit `slot` set = the row passes the host's filter
	const uint32_t *__restrict__ allowed;
	uint32_t res_off; // byte offset of the results list RH inside the warp's shared memory
	uint32_t latency; // host side only: dispatch to the latency variants
	uint32_t cta_warps; // host side only: warps per CTA the plan asks for (0 = the packing that fits most warps)
	// LMD_SEARCH_STRICT_FP: distances accumulated exactly as core/distance.hpp:56-99 does - one lane per row, element by
	// element, separate multiply and add (no FMA) - so that every comparison of the traversal sees the reference's own
	// bits. Served by the generic-shape kernels only (the query sits in shared memory); a parity tool, never timed.
	uint32_t strict;
	// Per-node routing sidecar (ternary planes or int8 codes of the node's R neighbours, laid out so that lane j's loads
	// of neighbour j coalesce across the warp - see lmd_store.cu build_plane_side_kernel / build_code_side_kernel):
	//   [u32 slot[R]] [int8 codec only: u32 csq[R], f32 err[R]] pad->16 [data: chunk-major, 16 bytes per (chunk, neighbour)]
	const uint8_t *__restrict__ side;
	uint32_t side_bytes;    // bytes per node
	uint32_t side_data_off; // offset of the chunk-major data inside a node's sidecar block
	// int8 edge codec: x ~ code_off[d] + code_scale * c[d], c in [-127, 127]
	const float *__restrict__ code_off; // [row_floats]
	float code_scale;
	float code_rho; // IP pre-filter: RMS row length
	uint32_t prefilter;     // exact routing: 1 = the int8 lower bound decides which neighbours' rows are fetched at all
	uint32_t pf_g;          // host side only: rows in flight per team of the pre-filtered kernels (1 or 2)
	uint32_t row_prefetch;  // 2 = auto (see the gather), 1 = always prefetch the rows of a gather into L2, 0 = never
	uint32_t nbr_prefetch;  // 1 = prefetch the neighbour rows of admitted candidates (small-batch kernels)
};

#define FULL 0xFFFFFFFFu
// kernel-side route: the public LMD_ROUTE_* values plus exact routing with the int8 lower-bound pre-filter
#define LMD_KROUTE_EXACT_PF 3
#define FILTER_EMPTY 0xFFFFFFFFu

__device__ __forceinline__ float4 ldg4(const float *p) {
	return __ldg(reinterpret_cast<const float4 *>(p));
}

// core/distance.hpp:50-105 (ComputeExactDistanceFloat), partial sums of one lane over one float4
template <int METRIC>
__device__ __forceinline__ void accum(float4 q, float4 x, float &a, float &b) {
	if (METRIC == LMD_METRIC_L2) {
		float d0 = q.x - x.x, d1 = q.y - x.y, d2 = q.z - x.z, d3 = q.w - x.w;
		a = fmaf(d0, d0, a);
		a = fmaf(d1, d1, a);
		a = fmaf(d2, d2, a);
		a = fmaf(d3, d3, a);
	} else {
		a = fmaf(q.x, x.x, a);
		a = fmaf(q.y, x.y, a);
		a = fmaf(q.z, x.z, a);
		a = fmaf(q.w, x.w, a);
		if (METRIC == LMD_METRIC_COSINE) {
			b = fmaf(x.x, x.x, b);
			b = fmaf(x.y, x.y, b);
			b = fmaf(x.z, x.z, b);
			b = fmaf(x.w, x.w, b);
		}
	}
}

// the tail of each metric (distance.hpp:62-67, :74, :85-98); na = squared norm of the query
template <int METRIC>
__device__ __forceinline__ float finish(float a, float b, float na) {
	float d;
	if (METRIC == LMD_METRIC_L2) {
		d = sqrtf(a < 0.0f ? 0.0f : a);
	} else if (METRIC == LMD_METRIC_IP) {
		d = -a;
	} else {
		if (na <= 0.0f || b <= 0.0f) {
			return 1.0f;
		}
		float ra = sqrtf(na), rb = sqrtf(b);
		if (ra == 0.0f || rb == 0.0f) {
			return 1.0f;
		}
		float c = a / (ra * rb);
		c = fmaxf(-1.0f, fminf(1.0f, c));
		d = 1.0f - c;
	}
	return d + 0.0f; // -0.0 -> +0.0 so that equal distances have equal keys
}

template <int TEAM>
__device__ __forceinline__ float team_sum(float v) {
#pragma unroll
	for (int o = TEAM / 2; o > 0; o >>= 1) {
		v += __shfl_xor_sync(FULL, v, o);
	}
	return v;
}

// Per-warp view of the query. CH > 0: each lane keeps its CH float4 pieces in registers (piece c of lane t
// inside a team covers floats [(c*TEAM + t)*4, +4)). CH == 0: generic shape, the query sits in shared memory.
template <int TEAM, int CH>
struct QueryRegs {
	float4 q[CH > 0 ? CH : 1];
	const float *sq; // generic only
	float na;
};

// Loads one vector (dims floats at src; row_floats = dims padded to a multiple of 32) into the warp's query view.
template <int TEAM, int CH>
__device__ __forceinline__ void load_vec(QueryRegs<TEAM, CH> &Q, const float *src, uint32_t dims, uint32_t row_floats,
                                         int lane, float *s_q, bool strict = false) {
	float na = 0.0f;
	if (CH > 0) {
		const int tl = lane % TEAM;
#pragma unroll
		for (int c = 0; c < (CH > 0 ? CH : 1); c++) {
			uint32_t f = (uint32_t)(c * TEAM + tl) * 4;
			float4 v;
			v.x = f + 0 < dims ? __ldg(src + f + 0) : 0.0f;
			v.y = f + 1 < dims ? __ldg(src + f + 1) : 0.0f;
			v.z = f + 2 < dims ? __ldg(src + f + 2) : 0.0f;
			v.w = f + 3 < dims ? __ldg(src + f + 3) : 0.0f;
			Q.q[c] = v;
			na = fmaf(v.x, v.x, fmaf(v.y, v.y, fmaf(v.z, v.z, fmaf(v.w, v.w, na))));
		}
		na = team_sum<TEAM>(na);
	} else {
		__syncwarp();
		for (uint32_t f = lane; f < row_floats; f += 32) {
			float v = f < dims ? __ldg(src + f) : 0.0f;
			s_q[f] = v;
			na = fmaf(v, v, na);
		}
		na = team_sum<32>(na);
		Q.sq = s_q;
		__syncwarp();
		if (strict) { // norm_a_sq of distance.hpp:77-83: element by element, multiply then add
			na = 0.0f;
			for (uint32_t f = 0; f < dims; f++) {
				const float v = s_q[f];
				na = __fadd_rn(na, __fmul_rn(v, v));
			}
		}
	}
	Q.na = na;
}

template <int TEAM, int CH, int METRIC>
__device__ __forceinline__ void load_query(QueryRegs<TEAM, CH> &Q, const KParams &p, uint32_t qi, int lane,
                                           float *s_q) {
	load_vec<TEAM, CH>(Q, p.queries + (uint64_t)qi * p.q_stride, p.dims, p.row_floats, lane, s_q, CH == 0 && p.strict);
}

// Raw partial results of the exact distance from the query to the m (1..32) nodes whose slots are in s_nb[0..m):
// s_a[j] = sum of squared differences (L2) or dot product (IP / COSINE), s_b2[j] = squared norm of the node (COSINE).
// TEAM lanes share one row; a warp works on (32/TEAM)*G rows at a time with all their loads issued first.
// The metric's tail (sqrt / clamp / ...) is applied once per batch by finish_batch(), one lane per row, instead of
// once per pass by every lane.
template <int TEAM, int CH, int G, int METRIC>
__device__ __forceinline__ void compute_batch(const QueryRegs<TEAM, CH> &Q, const KParams &p, const uint32_t *s_nb,
                                              float *s_a, float *s_b2, int m, int lane) {
	if (CH > 0) {
		constexpr int NT = 32 / TEAM;
		constexpr int CHN = CH > 0 ? CH : 1;
		const int team = lane / TEAM, tl = lane % TEAM;
		for (int j0 = 0; j0 < m; j0 += NT * G) {
			float4 x[G][CHN];
#pragma unroll
			for (int g = 0; g < G; g++) {
				int j = j0 + g * NT + team;
				uint32_t id = s_nb[j < m ? j : m - 1];
				const float *row = p.vecs + (uint64_t)id * p.row_floats + tl * 4;
#pragma unroll
				for (int c = 0; c < CHN; c++) {
					x[g][c] = ldg4(row + c * TEAM * 4);
				}
			}
#pragma unroll
			for (int g = 0; g < G; g++) {
				float a = 0.0f, b = 0.0f;
#pragma unroll
				for (int c = 0; c < CHN; c++) {
					accum<METRIC>(Q.q[c], x[g][c], a, b);
				}
				a = team_sum<TEAM>(a);
				if (METRIC == LMD_METRIC_COSINE) {
					b = team_sum<TEAM>(b);
				}
				int j = j0 + g * NT + team;
				if (tl == 0 && j < m) {
					s_a[j] = a;
					if (METRIC == LMD_METRIC_COSINE) {
						s_b2[j] = b;
					}
				}
			}
		}
	} else if (p.strict) {
		// the reference's own summation (core/distance.hpp:56-61, :69-73, :76-83): lane j walks row j front to back
		if (lane < m) {
			const float *row = p.vecs + (uint64_t)s_nb[lane] * p.row_floats;
			float a = 0.0f, b = 0.0f;
			for (uint32_t i = 0; i < p.dims; i++) {
				const float x = __ldg(row + i), q = Q.sq[i];
				if (METRIC == LMD_METRIC_L2) {
					const float d = __fsub_rn(q, x);
					a = __fadd_rn(a, __fmul_rn(d, d));
				} else {
					a = __fadd_rn(a, __fmul_rn(q, x));
					if (METRIC == LMD_METRIC_COSINE) {
						b = __fadd_rn(b, __fmul_rn(x, x));
					}
				}
			}
			s_a[lane] = a;
			if (METRIC == LMD_METRIC_COSINE) {
				s_b2[lane] = b;
			}
		}
	} else {
		const uint32_t nch = p.row_floats / 4;
		for (int j = 0; j < m; j++) {
			const float *row = p.vecs + (uint64_t)s_nb[j] * p.row_floats;
			float a = 0.0f, b = 0.0f;
			for (uint32_t c = lane; c < nch; c += 32) {
				float4 x = ldg4(row + c * 4);
				float4 q = *reinterpret_cast<const float4 *>(Q.sq + c * 4);
				accum<METRIC>(q, x, a, b);
			}
			a = team_sum<32>(a);
			if (METRIC == LMD_METRIC_COSINE) {
				b = team_sum<32>(b);
			}
			if (lane == 0) {
				s_a[j] = a;
				if (METRIC == LMD_METRIC_COSINE) {
					s_b2[j] = b;
				}
			}
		}
	}
}

// Lane j < m turns the raw sums of row j into the reference's distance (must follow a __syncwarp()).
template <int METRIC>
__device__ __forceinline__ float finish_batch(const float *s_a, const float *s_b2, float na, int m, int lane) {
	if (lane >= m) {
		return 0.0f;
	}
	return finish<METRIC>(s_a[lane], METRIC == LMD_METRIC_COSINE ? s_b2[lane] : 0.0f, na);
}

// Pulls the m rows that compute_batch() is about to read into L2, every 128-byte line of every row at once, so that
// the passes of compute_batch() wait for L2 instead of one HBM round trip each. LPR = 128-byte lines per row when the
// shape is specialised (0 = read it from the parameters).
template <int LPR>
__device__ __forceinline__ void prefetch_rows(const KParams &p, const uint32_t *s_nb, int m, int lane) {
	const uint32_t lines_per_row = LPR > 0 ? (uint32_t)LPR : p.row_floats / 32; // row_floats is a multiple of 32 floats
	const uint32_t total = (uint32_t)m * lines_per_row;
	for (uint32_t t = lane; t < total; t += 32) {
		const uint32_t j = t / lines_per_row, l = t - j * lines_per_row;
		const float *a = p.vecs + (uint64_t)s_nb[j] * (LPR > 0 ? (uint32_t)LPR * 32u : p.row_floats) + l * 32;
		asm volatile("prefetch.global.L2 [%0];" ::"l"(a));
	}
}

// core/Searcher.cpp:135-139: `if (visited.count(id)) continue; visited.insert(id);` - true when id is new.
__device__ __forceinline__ bool filter_test_and_set(uint32_t *tab, uint32_t mask, uint32_t shift, uint32_t id,
                                                    uint32_t &evictions) {
	const uint32_t home = (id * 2654435761u) >> shift;
	uint32_t h = home;
#pragma unroll 1
	for (int probe = 0; probe < 8; probe++) {
		uint32_t old = atomicCAS(&tab[h], FILTER_EMPTY, id);
		if (old == FILTER_EMPTY) {
			return true;
		}
		if (old == id) {
			return false;
		}
		h = (h + 1) & mask;
	}
	tab[home] = id; // window full: forget whoever lived at the home slot
	evictions++;
	return true;
}

// ---- visited filter, mode 1: two-choice buckets of eight 16-bit entries -----------------------------------------
// Same contract as filter_test_and_set() at half the bytes per remembered node and a much higher usable load.
// slot -> h = xorshift(slot * odd mod 2^m), m = log2(buckets) + 15: a BIJECTION on [0, 2^m), so (home bucket =
// h >> 15, tag = h & 0x7FFF) identifies the node exactly for every store with fewer than 2^m nodes (the host checks).
// An entry is (tag << 1 | c): c = 0 in the node's home bucket, c = 1 in its alternate bucket home ^ x(tag); a node
// goes to the alternate bucket only when its home bucket is full, and buckets never lose entries except by being
// overwritten while full, so "home has a free entry" proves the node is absent. Entry 0 = free; nodes whose tag is 0
// (1 in 32768) are simply never remembered. When both buckets are full one entry of the home bucket is forgotten -
// which, as in mode 0, can only cost a re-evaluation.
__device__ __forceinline__ bool has16(uint4 v, uint32_t e) {
	const uint32_t t = e * 0x10001u;
	return (__vcmpeq2(v.x, t) | __vcmpeq2(v.y, t) | __vcmpeq2(v.z, t) | __vcmpeq2(v.w, t)) != 0u;
}
__device__ __forceinline__ int first_free16(uint4 v, uint32_t &word) {
	uint32_t z;
	if ((z = __vcmpeq2(v.x, 0u)) != 0u) {
		word = v.x;
		return (z & 0xFFFFu) ? 0 : 1;
	}
	if ((z = __vcmpeq2(v.y, 0u)) != 0u) {
		word = v.y;
		return (z & 0xFFFFu) ? 2 : 3;
	}
	if ((z = __vcmpeq2(v.z, 0u)) != 0u) {
		word = v.z;
		return (z & 0xFFFFu) ? 4 : 5;
	}
	if ((z = __vcmpeq2(v.w, 0u)) != 0u) {
		word = v.w;
		return (z & 0xFFFFu) ? 6 : 7;
	}
	return -1;
}
__device__ __forceinline__ uint4 lds_bucket(const uint32_t *tab, uint32_t b) {
	uint4 v;
	const uint32_t a = (uint32_t)__cvta_generic_to_shared(tab + b * 4);
	asm volatile("ld.volatile.shared.v4.u32 {%0, %1, %2, %3}, [%4];" : "=r"(v.x), "=r"(v.y), "=r"(v.z), "=r"(v.w) : "r"(a));
	return v;
}
__device__ __forceinline__ bool bucket_test_and_set(uint32_t *tab, uint32_t bmask, uint32_t hmask, uint32_t id,
                                                    uint32_t &evictions) {
	uint32_t h = (id * 0x9E3779B1u) & hmask;
	h ^= h >> 15;
	const uint32_t tag = h & 0x7FFFu;
	if (tag == 0u) {
		return true;
	}
	const uint32_t home = h >> 15;
	uint32_t x = (tag * 0x2545u) & bmask;
	x = x ? x : 1u;
	const uint32_t alt = home ^ x;
	const uint32_t e0 = tag << 1, e1 = e0 | 1u;
#pragma unroll 1
	for (int attempt = 0; attempt < 12; attempt++) {
		const uint4 v = lds_bucket(tab, home);
		if (has16(v, e0)) {
			return false;
		}
		uint32_t old, entry = e0, bucket = home;
		int way = first_free16(v, old);
		if (way < 0) {
			const uint4 u = lds_bucket(tab, alt);
			if (has16(u, e1)) {
				return false;
			}
			way = first_free16(u, old);
			if (way < 0) {
				reinterpret_cast<uint16_t *>(tab + home * 4)[(id + evictions) & 7u] = (uint16_t)e0;
				evictions++;
				return true;
			}
			entry = e1;
			bucket = alt;
		}
		uint32_t *w = tab + bucket * 4 + (way >> 1);
		if (atomicCAS(w, old, old | (entry << ((way & 1) * 16))) == old) {
			return true;
		}
	}
	evictions++; // pathological contention inside one bucket: evaluate the node without remembering it
	return true;
}

// ---- visited filter, mode 2: direct-mapped 16-bit entries ----------------------------------------------------------
// One load, one compare, one store - a quarter of the bucket filter's instructions - for the kernels that are bound by
// instruction issue and for which a forgotten node is cheap (the compressed routes compute every neighbour's routing
// distance anyway; behind the int8 pre-filter a re-met node that was turned away before is turned away again without a row
// fetch). Same bijective hash as mode 1: entry index = h >> 15, entry = 15-bit tag | 0x8000, so (index, tag) identifies the
// node exactly below 2^(log2(entries) + 15) nodes. A colliding node simply replaces the resident one (same contract as
// the other modes: forgetting costs a re-evaluation, never a result). Lanes of one warp racing on an entry: either write
// may win. Two lanes can never hold the SAME node: the neighbour-slot view drops the second occurrence of a row id inside
// one row when it is built (lmd_store.cu first_occurrence(); the reference's visited set skips it just the same,
// Searcher.cpp:135-139) - an in-warp __match_any_sync here cost the ternary kernel 30 %.
__device__ __forceinline__ bool direct_test_and_set(uint32_t *tab, uint32_t hmask, uint32_t id, uint32_t &evictions) {
	uint32_t h = (id * 0x9E3779B1u) & hmask;
	h ^= h >> 15;
	const uint32_t entry = (h & 0x7FFFu) | 0x8000u;
	uint16_t *slot = reinterpret_cast<uint16_t *>(tab) + (h >> 15);
	const uint32_t old = *slot;
	if (old == entry) {
		return false;
	}
	*slot = (uint16_t)entry;
	evictions += old != 0u ? 1u : 0u;
	return true;
}

// Mode 3: the direct-mapped filter over ANY number of entries S, so the table can take exactly the shared memory that is
// left at the CTA count the kernel wants (a power of two wastes up to half of it). slot = id mod S, tag = id div S + 1:
// the pair identifies the node exactly for id < min(65535 S, 2^40 / S); the quotient comes from one wide multiply by
// ceil(2^40 / S). No mixing: row ids next to each other never collide, and a collision only forgets.
__device__ __forceinline__ bool modulo_test_and_set(uint32_t *tab, uint32_t entries, uint32_t magic, uint32_t id,
                                                    uint32_t &evictions) {
	const uint32_t q = (uint32_t)(((uint64_t)id * magic) >> 40);
	const uint32_t entry = q + 1u;
	uint16_t *slot = reinterpret_cast<uint16_t *>(tab) + (id - q * entries);
	const uint32_t old = *slot;
	if (old == entry) {
		return false;
	}
	*slot = (uint16_t)entry;
	evictions += old != 0u ? 1u : 0u;
	return true;
}

__device__ __forceinline__ bool visited_test_and_set(const KParams &p, uint32_t *tab, uint32_t id, uint32_t &evictions) {
	if (p.filter_mode == 3) {
		return modulo_test_and_set(tab, p.filter_entries, p.filter_magic, id, evictions);
	}
	if (p.filter_mode == 2) {
		return direct_test_and_set(tab, p.filter_hmask, id, evictions);
	}
	if (p.filter_mode) {
		return bucket_test_and_set(tab, p.filter_bmask, p.filter_hmask, id, evictions);
	}
	return filter_test_and_set(tab, p.filter_mask, p.filter_shift, id, evictions);
}

// Sorted insert of one key (warp-uniform) into s_list[0..n). The list holds the top-L entries in [0, L) and, behind
// them, the TIE REGION: entries pushed out of the top L whose distance still EQUALS the L-th distance. The oracle
// keeps such entries in its frontier heap and expands them, because its cut-off is `distance > worst`
// (core/Searcher.cpp:112-114 + H2), not `>=`; they sort after every top-L entry, so "first unexpanded entry of the
// whole array" is exactly the oracle's pop order. The region is bounded by cap (= LPL*32 or L + tie_slots); a tied,
// unexpanded entry that does not fit is dropped and counted in `overflow` (statistics: results may then differ from
// the oracle inside that distance tie only). Duplicates of a listed (distance, slot) pair are dropped.
// Returns the new length.
template <int LPL>
__device__ __forceinline__ int list_insert(uint64_t *s_list, int n, int L, int cap, uint64_t key, int lane,
                                           uint32_t &overflow) {
	// pass 1 (reads only): insert position and duplicate test
	int pos = 0;
	bool dup = false;
	const uint64_t k1 = key >> 1;
#pragma unroll
	for (int c = 0; c < LPL; c++) {
		if (c * 32 < n) {
			const int i = c * 32 + lane;
			const uint64_t e1 = (i < n ? s_list[i] : LMD_KEY_NONE) >> 1;
			pos += __popc(__ballot_sync(FULL, e1 < k1));
			dup |= __any_sync(FULL, e1 == k1);
		}
	}
	if (dup || pos >= cap) {
		return n;
	}
	// entry that falls off the end of the array (only when n == cap)
	const uint64_t last = n == cap ? s_list[cap - 1] : LMD_KEY_NONE;
	__syncwarp();
	// pass 2: shift [pos, n) right by one, top chunk first so that one chunk at a time lives in registers
#pragma unroll
	for (int c = LPL - 1; c >= 0; c--) {
		if (c * 32 < n && (c + 1) * 32 > pos) {
			const int i = c * 32 + lane;
			const uint64_t e = i < n ? s_list[i] : LMD_KEY_NONE;
			__syncwarp();
			if (i < n && i >= pos && i + 1 < cap) {
				s_list[i + 1] = e;
			}
			__syncwarp();
		}
	}
	if (lane == 0) {
		s_list[pos] = key;
	}
	__syncwarp();
	int nn = n < cap ? n + 1 : cap;
	if (nn > L) {
		// trim the tie region: keep entries behind the top L only while they tie with the L-th distance
		const uint32_t wd = (uint32_t)(s_list[L - 1] >> 32);
		int keep = L;
#pragma unroll
		for (int c = 0; c < LPL; c++) {
			if ((c + 1) * 32 > L && c * 32 < nn) {
				const int i = c * 32 + lane;
				const bool tie = i >= L && i < nn && (uint32_t)(s_list[i] >> 32) == wd;
				keep += __popc(__ballot_sync(FULL, tie));
			}
		}
		if (last != LMD_KEY_NONE && keep == cap && (uint32_t)(last >> 32) == wd && !(last & 1ull)) {
			overflow++;
		}
		nn = keep;
	}
	return nn;
}

// Ascending bitonic sort of one u64 per lane across the warp (LMD_KEY_NONE sorts last).
__device__ __forceinline__ uint64_t warp_sort_keys(uint64_t key, int lane) {
#pragma unroll
	for (int k = 2; k <= 32; k <<= 1) {
#pragma unroll
		for (int j = k >> 1; j > 0; j >>= 1) {
			const uint64_t other = __shfl_xor_sync(FULL, key, j);
			const bool keep_min = ((lane & j) == 0) == ((lane & k) == 0);
			const bool lt = key < other;
			key = (lt == keep_min) ? key : other;
		}
	}
	return key;
}

// Batch form of list_insert(): merges the warp's sorted batch (lane j holds the j-th smallest new key, cnt valid
// keys, LMD_KEY_NONE behind them) into s_list[0..n) in one pass. Every list entry and every batch key finds its
// merged position by a binary search in the other sequence (s_bk = the batch in shared memory), then all of them are
// written at once. The result equals cnt calls of list_insert() in neighbour order (core/Searcher.cpp:175-185)
// unless a batch key ties with the distance that ends up L-th (see the rule inside): otherwise the final list is the
// L smallest of the union by (distance, id), however the threshold moved in between. If such a tie (or a duplicate
// of a listed node) exists, nothing is written and `fallback` is set: the caller replays the batch one by one.
template <int LPL>
__device__ __forceinline__ int list_merge_batch(uint64_t *s_list, uint64_t *s_bk, int n, int L, int cap, uint64_t bkey,
                                                int cnt, int lane, uint32_t &overflow, bool &fallback) {
	constexpr int CAPP2 = LPL <= 1 ? 32 : LPL <= 2 ? 64 : LPL <= 4 ? 128 : LPL <= 8 ? 256 : 512;
	s_bk[lane] = bkey;
	__syncwarp();
	// batch keys: how many list entries sort before each (reads only), and the tie test
	const uint64_t b1 = bkey >> 1;
	const uint32_t bhi = (uint32_t)(bkey >> 32);
	int lo = 0;
#pragma unroll
	for (int step = CAPP2 / 2; step > 0; step >>= 1) {
		const int t = lo + step - 1;
		if (t < n && (s_list[t] >> 1) < b1) {
			lo += step;
		}
	}
	if (lo < n && (s_list[lo] >> 1) < b1) {
		lo++;
	}
	// When does one merge equal the one-by-one admission of core/Searcher.cpp:175-185? Let w be the distance at merged
	// position L - 1. Everything closer than w ends up listed either way, everything farther ends up outside either way;
	// only elements AT distance w depend on the order of arrival (a candidate that ties with the current worst is turned
	// away, one that arrived while the worst was still larger is not). Old list entries at w behave the same in both
	// (evictions take the largest (distance, id) first), so the merge is exact unless a BATCH key sits at distance w -
	// or repeats a listed node (a node the visited filter forgot). While the list is not full after the merge there is
	// no threshold at all. Ties anywhere else - the normal case with integer ternary distances - are harmless.
	bool tie = false;
	if (lane < cnt) {
		tie = lo < n && (s_list[lo] >> 1) == b1; // duplicate of a listed node
	}
	if (n + cnt > L) {
		const int cb = __popc(__ballot_sync(FULL, lane < cnt && lane + lo <= L - 1)); // batch keys among the first L
		uint32_t whi = 0;
		if (cb > 0) {
			whi = (uint32_t)(s_bk[cb - 1] >> 32);
		}
		if (L - cb > 0) {
			const uint32_t lhi = (uint32_t)(s_list[L - cb - 1] >> 32);
			whi = lhi > whi ? lhi : whi;
		}
		// ... and a batch key at distance w only matters when something else is at distance w too
		const uint32_t prev_hi = (uint32_t)(__shfl_up_sync(FULL, bkey, 1) >> 32);
		const uint32_t next_hi = (uint32_t)(__shfl_down_sync(FULL, bkey, 1) >> 32);
		if (lane < cnt && bhi == whi) {
			tie = tie || (lo < n && (uint32_t)(s_list[lo] >> 32) == bhi) ||
			      (lo > 0 && (uint32_t)(s_list[lo - 1] >> 32) == bhi) || (lane > 0 && prev_hi == bhi) ||
			      (lane + 1 < cnt && next_hi == bhi);
		}
	}
	if (__any_sync(FULL, tie)) {
		fallback = true;
		return n;
	}
	__syncwarp(); // every read of the old list by the batch side is done
	// list entries, top chunk first: each moves up by the number of batch keys that sort before it (<= 32), so a chunk
	// only ever writes into itself and the chunk above, whose entries have already moved. One chunk in registers.
	uint64_t e_push = LMD_KEY_NONE; // the entry of this lane pushed past `cap` (top chunk only: cap is a multiple of 32)
#pragma unroll
	for (int c = LPL - 1; c >= 0; c--) {
		if (c * 32 < n) {
			const int i = c * 32 + lane;
			const uint64_t e = i < n ? s_list[i] : LMD_KEY_NONE;
			const uint64_t e1 = e >> 1;
			int sh = 0;
#pragma unroll
			for (int step = 16; step > 0; step >>= 1) {
				if ((s_bk[sh + step - 1] >> 1) < e1) {
					sh += step;
				}
			}
			if ((s_bk[sh] >> 1) < e1) {
				sh++;
			}
			__syncwarp();
			if (i < n) {
				if (i + sh < cap) {
					s_list[i + sh] = e;
				} else {
					e_push = e;
				}
			}
			__syncwarp();
		}
	}
	if (lane < cnt && lane + lo < cap) {
		s_list[lane + lo] = bkey;
	}
	__syncwarp();
	int nn = n + cnt < cap ? n + cnt : cap;
	if (nn > L) {
		// trim the tie region exactly as list_insert() does; entries pushed past `cap` that would have stayed count
		// as overflow
		const uint32_t wd = (uint32_t)(s_list[L - 1] >> 32);
		int keep = L;
#pragma unroll
		for (int c = 0; c < LPL; c++) {
			if ((c + 1) * 32 > L && c * 32 < nn) {
				const int i = c * 32 + lane;
				const bool tie_kept = i >= L && i < nn && (uint32_t)(s_list[i] >> 32) == wd;
				keep += __popc(__ballot_sync(FULL, tie_kept));
			}
		}
		const bool lost = e_push != LMD_KEY_NONE && (uint32_t)(e_push >> 32) == wd && !(e_push & 1ull);
		overflow += __popc(__ballot_sync(FULL, lost));
		nn = keep;
	}
	return nn;
}

template <int LPL>
__device__ __forceinline__ int find_unexpanded(const uint64_t *s_list, int n, int lane) {
#pragma unroll
	for (int c = 0; c < LPL; c++) {
		int i = c * 32 + lane;
		bool un = i < n && !(s_list[i] & 1ull);
		uint32_t b = __ballot_sync(FULL, un);
		if (b) {
			return c * 32 + __ffs(b) - 1;
		}
		if ((c + 1) * 32 >= n) {
			break;
		}
	}
	return -1;
}

// Ternary planes of the query (core/ternary_quantization.hpp:146-199 EncodeTernary), one 16-byte entry per 64-bit word w:
// s_q4[w] = {pos bits [64w, 64w+32), pos bits [64w+32, 64w+64), neg likewise} - the layout of a sidecar entry, so that a
// word of the query is ONE broadcast 128-bit shared-memory load. (s_qn is unused; kept in the signature of the callers.)
__device__ __forceinline__ void encode_query_planes(const KParams &p, uint32_t qi, int lane, uint32_t *s_qp,
                                                    uint32_t *s_qn) {
	(void)s_qn;
	const float *src = p.queries + (uint64_t)qi * p.q_stride;
	for (uint32_t g = 0; g < p.words * 2; g++) {
		uint32_t f = g * 32 + lane;
		float v = f < p.dims ? __ldg(src + f) : 0.0f;
		uint32_t bp = __ballot_sync(FULL, v > 0.0f);
		uint32_t bn = __ballot_sync(FULL, v < 0.0f);
		if (lane == 0) {
			s_qp[(g >> 1) * 4 + (g & 1)] = bp;
			s_qp[(g >> 1) * 4 + 2 + (g & 1)] = bn;
		}
	}
	__syncwarp();
}

// One 64-bit word of core/ternary_quantization.hpp:269-296 (Dot_scalar): pop(q+ & v+) - pop(q+ & v-) - pop(q- & v+) +
// pop(q- & v-). A vector's positive and negative planes are disjoint (a dimension is > 0 or < 0, EncodeTernary :146-199;
// the sidecar zeroes dead neighbours), so (q+ & v+) and (q- & v-) are disjoint too and one popcount of their union counts
// both - half the popcounts (the slow instruction here) for the same integer.
__device__ __forceinline__ int ternary_word(const uint4 q, const uint4 v) {
	const uint32_t agree0 = (q.x & v.x) | (q.z & v.z), agree1 = (q.y & v.y) | (q.w & v.w);
	const uint32_t differ0 = (q.x & v.z) | (q.z & v.x), differ1 = (q.y & v.w) | (q.w & v.y);
	return __popc(agree0) + __popc(agree1) - __popc(differ0) - __popc(differ1);
}

// core/ternary_quantization.hpp:269-296 (Dot_scalar) + core/distance.hpp:229-235 for neighbour j of the expanded node.
// The planes come from the node's routing sidecar: word w of neighbour j is ONE 16-byte entry {pos.lo, pos.hi, neg.lo,
// neg.hi} at data[(w * R + j) * 16], so the warp's load of word w is 512 contiguous bytes (the block's own SoA planes put
// lane j's 8-byte words 8 * words bytes apart: 23 sectors per request at D = 768, profiles/r19_ternary768_ncu_details.csv).
// WORDS = u64 words per plane when the shape is specialised (every load issued before the first popcount), 0 = runtime.
template <int WORDS>
__device__ __forceinline__ float ternary_distance(const KParams &p, const uint8_t *side_blk, uint32_t j,
                                                  const uint32_t *s_qp, const uint32_t *s_qn) {
	(void)s_qn;
	const uint32_t words = WORDS > 0 ? (uint32_t)WORDS : p.words;
	const uint4 *pl = reinterpret_cast<const uint4 *>(side_blk + p.side_data_off) + j;
	const uint4 *q4 = reinterpret_cast<const uint4 *>(s_qp);
	int score = 0;
	if (WORDS > 0) {
		uint4 v[WORDS > 0 ? WORDS : 1];
#pragma unroll
		for (int w = 0; w < WORDS; w++) {
			v[w] = __ldg(pl + (uint32_t)w * p.r);
		}
#pragma unroll
		for (int w = 0; w < WORDS; w++) {
			score += ternary_word(q4[w], v[w]);
		}
	} else {
		for (uint32_t w = 0; w < words; w++) {
			score += ternary_word(q4[w], __ldg(pl + w * p.r));
		}
	}
	float sim = (float)score;
	float d = p.metric == LMD_METRIC_IP ? -sim : 1.0f - sim;
	return d + 0.0f;
}

// ---- int8 edge codec (the L2-capable codec; libsql/lm_diskann.hpp:55-62 INT8) -------------
// x ~ off[d] + scale * c[d]. The query is brought onto the same grid once per query: qi[d] = clamp(rint((q[d] - off[d]) /
// scale)), kept as int8 in shared memory; a_sq = sum qi^2 (exact integer); e_q = || (q - off) - scale * qi || (fp32).
// Then for a neighbour with codes c:  || q^ - x^ ||^2 = scale^2 * (a_sq - 2 * dot(qi, c) + csq)  - all integer, dp4a.
// `qmul`: COSINE stores keep UNIT rows on the grid, and the query goes there as q / || q || (qmul = 1 / || q ||; 0 for a
// zero query, whose e_q becomes +inf so that no bound ever excludes a row).
__device__ __forceinline__ void encode_query_int8(const KParams &p, uint32_t qi, int lane, int8_t *s_qi8, int &a_sq,
                                                  float &e_q, float qmul = 1.0f) {
	const float *src = p.queries + (uint64_t)qi * p.q_stride;
	const float inv = 1.0f / p.code_scale;
	int a = 0;
	float e = 0.0f;
	for (uint32_t f = lane; f < p.row_floats; f += 32) { // row_floats is a multiple of 32 >= dims: the tail encodes 0
		int c = 0;
		if (f < p.dims) {
			const float v = __ldg(src + f) * qmul - __ldg(p.code_off + f);
			c = __float2int_rn(v * inv);
			c = c < -127 ? -127 : (c > 127 ? 127 : c);
			const float r = v - p.code_scale * (float)c;
			e = fmaf(r, r, e);
		}
		s_qi8[f] = (int8_t)c;
		a += c * c;
	}
#pragma unroll
	for (int o = 16; o > 0; o >>= 1) {
		a += __shfl_xor_sync(FULL, a, o);
		e += __shfl_xor_sync(FULL, e, o);
	}
	a_sq = a;
	e_q = qmul > 0.0f ? sqrtf(e) * 1.0001f : __int_as_float(0x7f800000);
	__syncwarp();
}

// dot(qi, c_j) over all dimensions for neighbour j of the expanded node: chunk w (16 dimensions) of neighbour j is the
// 16-byte entry data[(w * R + j) * 16]; NCH = row_floats / 16 when the shape is specialised, 0 = runtime.
template <int NCH>
__device__ __forceinline__ int int8_dot(const KParams &p, const uint8_t *side_blk, uint32_t j, const int8_t *s_qi8) {
	const uint4 *cp = reinterpret_cast<const uint4 *>(side_blk + p.side_data_off) + j;
	const uint4 *qp = reinterpret_cast<const uint4 *>(s_qi8);
	int dot = 0;
	constexpr int GRP = 8; // 128 bytes of codes in flight per lane
	const uint32_t nch = NCH > 0 ? (uint32_t)NCH : p.row_floats / 16;
	for (uint32_t w0 = 0; w0 < nch; w0 += GRP) {
		uint4 c[GRP];
#pragma unroll
		for (int g = 0; g < GRP; g++) {
			c[g] = (w0 + g < nch) ? __ldg(cp + (w0 + g) * p.r) : make_uint4(0u, 0u, 0u, 0u);
		}
#pragma unroll
		for (int g = 0; g < GRP; g++) {
			if (w0 + g < nch) {
				const uint4 q = qp[w0 + g];
				dot = __dp4a((int)c[g].x, (int)q.x, dot);
				dot = __dp4a((int)c[g].y, (int)q.y, dot);
				dot = __dp4a((int)c[g].z, (int)q.z, dot);
				dot = __dp4a((int)c[g].w, (int)q.w, dot);
			}
		}
	}
	return dot;
}

__device__ __forceinline__ bool slot_allowed(const KParams &p, uint32_t slot) {
	return (__ldg(p.allowed + (slot >> 5)) >> (slot & 31)) & 1u;
}

// FILT = 1: filtered search, the dual-heap design of the reference's plan ("Consolidated Diskann Proposal v1", section
// "Filtered Searches"). The exploration heap EH is s_list, untouched: the traversal tunnels through rows that fail the
// filter exactly as an unfiltered search would. The results heap RH is a second sorted list (s_res, L entries) that is
// offered every node whose routing distance gets evaluated and admits only allowed rows; results and the rerank
// candidates come from RH alone.
template <int TEAM, int CH, int G, int LPL, int METRIC, int ROUTE, int FILT = 0>
__global__ void __launch_bounds__(128, (ROUTE == LMD_KROUTE_EXACT_PF && G == 1) ? 8 : 0) traverse_kernel(const KParams p) {
	extern __shared__ __align__(16) uint8_t smem[];
	const int lane = threadIdx.x & 31;
	uint8_t *wbase = smem + (threadIdx.x >> 5) * p.warp_smem_bytes;
	uint64_t *s_list = reinterpret_cast<uint64_t *>(wbase);
	uint32_t *s_nb = reinterpret_cast<uint32_t *>(s_list + LPL * 32);
	float *s_d = reinterpret_cast<float *>(s_nb + 32);
	uint64_t *s_bk = reinterpret_cast<uint64_t *>(s_nb); // the sorted batch of list_merge_batch() reuses s_nb + s_d
	float *s_d2 = s_d + 32;
	uint32_t *s_filter = reinterpret_cast<uint32_t *>(s_d2 + 32);
	uint8_t *extra = reinterpret_cast<uint8_t *>(s_filter) + p.filter_bytes;
	float *s_q = reinterpret_cast<float *>(extra); // generic shape only
	uint32_t *s_qp = reinterpret_cast<uint32_t *>(extra + (CH == 0 ? p.row_floats * 4 : 0));
	uint32_t *s_qn = s_qp + p.words * 2;
	int8_t *s_qi8 = reinterpret_cast<int8_t *>(s_qp); // int8 codec: the query on the code grid (row_floats bytes)
	constexpr bool SIDE = ROUTE != LMD_ROUTE_EXACT;   // the expansion reads the node's routing sidecar block
	constexpr bool CODES = ROUTE == LMD_ROUTE_INT8 || ROUTE == LMD_KROUTE_EXACT_PF;
	constexpr bool GATHER = ROUTE == LMD_ROUTE_EXACT || ROUTE == LMD_KROUTE_EXACT_PF; // exact rows of the new neighbours
	uint64_t *s_res = FILT ? reinterpret_cast<uint64_t *>(wbase + p.res_off) : s_list;
	const int L = (int)p.L;
	const int cap = (int)p.list_cap;
	const uint32_t nbr_lines = (p.r * 4 + 127) / 128;
	const uint32_t entry_guard = p.has_entry ? p.entry_slot : LMD_INVALID_SLOT;

	for (;;) {
		uint32_t qi = 0;
		if (lane == 0) {
			qi = (uint32_t)atomicAdd(&p.counters[0], 1ull);
		}
		qi = __shfl_sync(FULL, qi, 0);
		if (qi >= p.nq) {
			break;
		}
		uint32_t n_exp = 0, n_eval = 0, n_evict = 0, n_tie_ovf = 0, n_fallback = 0, n_fetch = 0;
		int n = 0;
		int nres = 0;
		int q_asq = 0;
		float q_err = 0.0f;
		{
			const uint32_t fill = p.filter_mode ? 0u : FILTER_EMPTY;
			uint4 e4 = make_uint4(fill, fill, fill, fill);
			uint4 *f4 = reinterpret_cast<uint4 *>(s_filter);
			for (uint32_t i = lane; i < p.filter_bytes / 16; i += 32) {
				f4[i] = e4;
			}
		}
		QueryRegs<TEAM, CH> Q;
		load_query<TEAM, CH, METRIC>(Q, p, qi, lane, s_q);
		if (ROUTE == LMD_ROUTE_TERNARY) {
			encode_query_planes(p, qi, lane, s_qp, s_qn);
		}
		if (CODES) {
			float qmul = 1.0f;
			if (METRIC == LMD_METRIC_COSINE) {
				qmul = Q.na > 0.0f ? 1.0f / sqrtf(Q.na) : 0.0f;
				qmul = isfinite(qmul) ? qmul : 0.0f;
			} else if (METRIC == LMD_METRIC_IP) { // the query at the rows' RMS length, so that it lands inside the grid
				qmul = Q.na > 0.0f ? p.code_rho / sqrtf(Q.na) : 0.0f;
				qmul = isfinite(qmul) ? qmul : 0.0f;
			}
			encode_query_int8(p, qi, lane, s_qi8, q_asq, q_err, qmul);
		}
		__syncwarp();

		if (p.has_entry) { // core/Searcher.cpp:70-103
			if (lane == 0) {
				s_nb[0] = p.entry_slot;
				visited_test_and_set(p, s_filter, p.entry_slot, n_evict);
			}
			__syncwarp();
			compute_batch<TEAM, CH, (GATHER ? G : 1), METRIC>(Q, p, s_nb, s_d, s_d2, 1, lane);
			__syncwarp();
			const float d0 = finish_batch<METRIC>(s_d, s_d2, Q.na, 1, lane);
			if (lane == 0) {
				s_list[0] = make_key(d0, p.entry_slot);
			}
			n = 1;
			n_eval = 1;
			n_fetch = 1;
			if (FILT && slot_allowed(p, p.entry_slot)) {
				if (lane == 0) {
					s_res[0] = make_key(d0, p.entry_slot);
				}
				nres = 1;
			}
			__syncwarp();
		}

		while (n > 0) { // core/Searcher.cpp:108-187
			int idx = find_unexpanded<LPL>(s_list, n, lane);
			if (idx < 0) {
				break;
			}
			const uint64_t cur_key = s_list[idx];
			const uint32_t cur = key_slot(cur_key);
			__syncwarp();
			if (lane == 0) {
				s_list[idx] = cur_key | 1ull;
			}
			__syncwarp();
			n_exp++;
			// SIDE: slots, (code norms,) planes / codes of the R neighbours sit in ONE contiguous sidecar block of the node
			const uint8_t *side_blk = SIDE ? p.side + (uint64_t)cur * p.side_bytes : nullptr;
			const uint32_t *nrow = SIDE ? reinterpret_cast<const uint32_t *>(side_blk) : p.nbrs + (uint64_t)cur * p.r;
			uint32_t id_next = lane < p.r ? __ldg(nrow + lane) : LMD_INVALID_SLOT;
			{
				// the runner-up of the frontier is the most likely next expansion: pull its neighbour row into L2 now
				const int idx2 = find_unexpanded<LPL>(s_list, n, lane);
				if (!SIDE && idx2 >= 0 && (uint32_t)lane < nbr_lines) {
					const uint32_t *nx = p.nbrs + (uint64_t)key_slot(s_list[idx2]) * p.r + lane * 32;
					asm volatile("prefetch.global.L2 [%0];" ::"l"(nx));
				}
				if (SIDE && idx2 >= 0) {
					// ... or its whole sidecar block (p.prefetch_planes = 0: only the slot row at its head)
					const uint8_t *b2 = p.side + (uint64_t)key_slot(s_list[idx2]) * p.side_bytes;
					const uint32_t span = p.prefetch_planes ? p.side_bytes : p.r * 4;
					for (uint32_t o = lane * 128; o < span; o += 32 * 128) {
						asm volatile("prefetch.global.L2 [%0];" ::"l"(b2 + o));
					}
				}
			}
			for (uint32_t base = 0; base < p.r; base += 32) {
				const uint32_t j = base + lane;
				const uint32_t id = id_next;
				if (base + 32 < p.r) {
					id_next = j + 32 < p.r ? __ldg(nrow + j + 32) : LMD_INVALID_SLOT;
				}
				// int8 codec: the code loads of neighbour j go out BEFORE the filter test (nothing depends on its outcome:
				// the bytes are contiguous with those of the other lanes), the integer dot product is ready when it ends
				int c_dot = 0;
				uint32_t c_sq = 0;
				float c_err = 0.0f;
				if (CODES && id != LMD_INVALID_SLOT) {
					c_sq = __ldg(nrow + p.r + j);
					c_err = __uint_as_float(__ldg(nrow + 2 * p.r + j));
					c_dot = int8_dot<(CH > 0 ? CH * TEAM / 4 : 0)>(p, side_blk, j, s_qi8);
				}
				bool isnew = false;
				// The entry point is the one node whose list key does not come from this loop (its distance is always
				// exact, core/Searcher.cpp:90-99, while ternary routing would give it a different one here), so it must
				// never be "forgotten" by the filter: it is tested by value.
				if (id != LMD_INVALID_SLOT && id != entry_guard) {
					isnew = visited_test_and_set(p, s_filter, id, n_evict);
				}

				const uint32_t mask = __ballot_sync(FULL, isnew);
				const int m = __popc(mask);
				if (m == 0) {
					continue;
				}
				n_eval += m;
				float d = 0.0f;
				uint32_t myid = id;
				bool have = isnew;
				uint32_t gmask = mask; // neighbours whose exact row is gathered
				int gm = m;
				if (CODES && ROUTE == LMD_KROUTE_EXACT_PF && METRIC == LMD_METRIC_COSINE) {
					// COSINE stores: the zero-offset grid holds the UNIT rows (x^ = scale * c ~ x / || x ||) and the unit query.
					// With u = q / || q ||, w = x / || x ||:  u.w = u^.w^ + (u - u^).w^ + u.(w - w^)
					//   <= scale^2 qi.c + || u - u^ || (1 + || w - w^ ||) + || w - w^ ||,
					// so 1 - that is a lower bound of the distance of core/distance.hpp:85-98 (whose clamp to [-1, 1] can only
					// raise the distance above the bound). fp32 rounding of either side stays below 5e-5 in these units
					// (D * 2^-24); the margin is 2e-4. Zero rows / a zero query carry err = +inf: the bound is -inf, they are
					// fetched and get the reference's 1.0.
					const float ub = p.code_scale * p.code_scale * (float)c_dot + q_err * (1.0f + c_err) + c_err;
					const float lb = 1.0f - ub;
					const float worst0 = n >= L ? key_dist(s_list[L - 1]) : 0.0f;
					const bool skip = n >= L && lb > worst0 + 2e-4f;
					gmask = __ballot_sync(FULL, isnew && !skip);
					gm = __popc(gmask);
					if (gm == 0) {
						continue;
					}
					isnew = isnew && !skip;
				} else if (CODES && ROUTE == LMD_KROUTE_EXACT_PF && METRIC == LMD_METRIC_IP) {
					// IP stores: raw rows on a zero-offset grid, the query scaled to the rows' RMS length rho (u = q rho / || q ||):
					// q.x = (|| q || / rho) u.x,  u.x <= scale^2 qi.c + || u - u^ || || x^ || + rho || x - x^ ||,  || x^ || = scale sqrt(csq);
					// the distance of core/distance.hpp:74 is -q.x. Margin: 1e-4 || q || (|| x^ || + || x - x^ ||) >= 1e-4 || q || || x ||,
					// against D * 2^-24 relative rounding of the exact dot product. A zero query (or an all-zero store) has
					// err = +inf: 0 * inf is a NaN, which never compares greater - nothing is skipped.
					const float qn = sqrtf(Q.na);
					const float xh = p.code_scale * sqrtf((float)c_sq);
					const float ub = (qn / p.code_rho) * (p.code_scale * p.code_scale * (float)c_dot + q_err * xh + p.code_rho * c_err);
					const float worst0 = n >= L ? key_dist(s_list[L - 1]) : 0.0f;
					const bool skip = n >= L && -ub > worst0 + 1e-4f * qn * (xh + c_err);
					gmask = __ballot_sync(FULL, isnew && !skip);
					gm = __popc(gmask);
					if (gm == 0) {
						continue;
					}
					isnew = isnew && !skip;
				} else if (CODES) {
					// || q^ - x^ || on the code grid: integers all the way to the square root
					const int i2 = q_asq - 2 * c_dot + (int)c_sq;
					d = p.code_scale * sqrtf((float)(i2 > 0 ? i2 : 0)) + 0.0f;
					if (ROUTE == LMD_KROUTE_EXACT_PF) {
						// d(q, x) >= || q^ - x^ || - || q - q^ || - || x - x^ ||: a neighbour whose lower bound is already
						// beyond the list's worst would be turned away by core/Searcher.cpp:175 whatever its exact
						// distance is - its row need not be read. The 5e-4 margin covers every fp32 rounding involved
						// (the kernel's own distance sum is within D * 2^-24 relative of the true one); the threshold can
						// only shrink while the batch is admitted, so testing against its value now is conservative.
						const float worst0 = n >= L ? key_dist(s_list[L - 1]) : 0.0f;
						const bool skip = n >= L && d > (worst0 + q_err + c_err) * 1.0005f;
						gmask = __ballot_sync(FULL, isnew && !skip);
						gm = __popc(gmask);
						if (gm == 0) {
							continue;
						}
						isnew = isnew && !skip;
					}
				}
				if (GATHER) {
					const int m = gm; // shadows the discovered count: rows actually fetched
					const uint32_t mask = gmask;
					n_fetch += m;
					const int pos = __popc(mask & ((1u << lane) - 1u));
					if (isnew) {
						s_nb[pos] = id;
					}
					__syncwarp();
					// Pull every line of every row into L2 first - unless ONE pass of a throughput kernel (G <= 2: at most
					// 8 rows, 4 loads per lane) gathers them all anyway. The small-batch kernels (G = 4 / 8) keep it: a lane
					// cannot have 32 loads in flight, the prefetches can.
					if (p.row_prefetch == 1 || (p.row_prefetch != 0 && (CH == 0 || G > 2 || m > (32 / TEAM) * G))) {
						prefetch_rows<(CH > 0 ? CH * TEAM / 8 : 0)>(p, s_nb, m, lane);
					}
					compute_batch<TEAM, CH, G, METRIC>(Q, p, s_nb, s_d, s_d2, m, lane);
					__syncwarp();
					have = lane < m;
					d = finish_batch<METRIC>(s_d, s_d2, Q.na, m, lane);
					if (have) {
						myid = s_nb[lane];
					}
					__syncwarp();
				} else if (ROUTE == LMD_ROUTE_TERNARY) {
					if (isnew) {
						d = ternary_distance<(CH > 0 && ROUTE == LMD_ROUTE_TERNARY ? (CH * TEAM * 4 + 63) / 64 : 0)>(p, side_blk, j, s_qp, s_qn);
					}
				}
				if (FILT) {
					// RH: allowed rows only, bounded at L, same (distance, slot) order and strict threshold as EH
					float rworst = nres >= L ? key_dist(s_res[L - 1]) : 0.0f;
					const bool ok = have && slot_allowed(p, myid) && (nres < L || d < rworst);
					uint32_t rmask = __ballot_sync(FULL, ok);
					uint32_t dummy = 0;
					if (__popc(rmask) >= (int)p.merge_min) { // same batch merge as EH below (s_bk is free until then)
						const uint64_t rkey = warp_sort_keys(ok ? make_key(d, myid) : LMD_KEY_NONE, lane);
						bool rfall = false;
						nres = list_merge_batch<LPL>(s_res, s_bk, nres, L, L, rkey, __popc(rmask), lane, dummy, rfall);
						__syncwarp();
						if (!rfall) {
							rmask = 0;
						}
					}
					while (rmask) {
						const int src = __ffs(rmask) - 1;
						rmask &= rmask - 1;
						const float dj = __shfl_sync(FULL, d, src);
						const uint32_t idj = __shfl_sync(FULL, myid, src);
						if (nres >= L && !(dj < rworst)) {
							continue;
						}
						nres = list_insert<LPL>(s_res, nres, L, L, make_key(dj, idj), lane, dummy);
						if (nres >= L) {
							rworst = key_dist(s_res[L - 1]);
						}
					}
				}
				// core/Searcher.cpp:175-185, candidates taken in neighbour order with a moving threshold
				float worst = n >= L ? key_dist(s_list[L - 1]) : 0.0f;
				const bool pass = have && (n < L || d < worst);
				uint32_t amask = __ballot_sync(FULL, pass);
				if (!SIDE && p.nbr_prefetch && pass) {
					// small batches: the next expansion is often one of the candidates admitted right now - pull their
					// neighbour rows into L2 while the list is being merged (the runner-up's row is already there)
					const uint32_t *nx = p.nbrs + (uint64_t)myid * p.r;
					for (uint32_t l = 0; l < nbr_lines; l++) {
						asm volatile("prefetch.global.L2 [%0];" ::"l"(nx + l * 32));
					}
				}
				if (__popc(amask) >= (int)p.merge_min) {
					// same result as the one-by-one loop below, in one merge (see list_merge_batch)
					const uint64_t bkey = warp_sort_keys(pass ? make_key(d, myid) : LMD_KEY_NONE, lane);
					bool fallback = false;
					n = list_merge_batch<LPL>(s_list, s_bk, n, L, cap, bkey, __popc(amask), lane, n_tie_ovf, fallback);
					if (!fallback) {
						amask = 0;
					} else {
						n_fallback++;
					}
				}
				while (amask) {
					const int src = __ffs(amask) - 1;
					amask &= amask - 1;
					const float dj = __shfl_sync(FULL, d, src);
					const uint32_t idj = __shfl_sync(FULL, myid, src);
					if (n >= L && !(dj < worst)) {
						continue;
					}
					n = list_insert<LPL>(s_list, n, L, cap, make_key(dj, idj), lane, n_tie_ovf);
					if (n >= L) {
						worst = key_dist(s_list[L - 1]);
					}
				}
			}
		}

		// results: the top-L part only (the oracle's final pass drains `top_candidates`, core/Searcher.cpp:193-196)
		if (n > L) {
			n = L;
		}
		if (FILT) { // only RH is returned
			n = nres;
		}
		if (p.cand) {
			uint64_t *dst = p.cand + (uint64_t)qi * p.L;
			for (int i = lane; i < L; i += 32) {
				dst[i] = i < n ? s_res[i] : LMD_KEY_NONE;
			}
		}
		if (p.out_ids) { // traversal distances are exact: the k best of the list are the answer (Searcher.cpp:189-245)
			for (uint32_t i = lane; i < p.k; i += 32) {
				int64_t rid = -1;
				float dd = 0.0f;
				if ((int)i < n) {
					uint64_t key = s_res[i];
					rid = p.row_ids[key_slot(key)];
					dd = key_dist(key);
				}
				p.out_ids[(uint64_t)qi * p.k + i] = rid;
				if (p.out_dist) {
					p.out_dist[(uint64_t)qi * p.k + i] = dd;
				}
			}
			if (p.out_count && lane == 0) {
				p.out_count[qi] = (uint32_t)n < p.k ? (uint32_t)n : p.k;
			}
		}
		n_evict = __reduce_add_sync(FULL, n_evict);
		if (lane == 0) {
			if (p.qstats) {
				p.qstats[2 * qi] = n_exp;
				p.qstats[2 * qi + 1] = n_eval;
			}
			atomicAdd(&p.counters[1], (unsigned long long)n_exp);
			atomicAdd(&p.counters[2], (unsigned long long)n_eval);
			atomicAdd(&p.counters[7], (unsigned long long)n_fetch);
			if (n_evict) {
				atomicAdd(&p.counters[3], (unsigned long long)n_evict);
			}
			if (n_tie_ovf) {
				atomicAdd(&p.counters[5], (unsigned long long)n_tie_ovf);
			}
			if (n_fallback) {
				atomicAdd(&p.counters[6], (unsigned long long)n_fallback);
			}
		}
		__syncwarp();
	}
}

// Rerank (core/Searcher.cpp:189-245 + H3): exact distance of every listed candidate, keep the k smallest by
// (distance, slot), ascending. One warp per query; candidates come from the traversal kernel's list dump.
template <int TEAM, int CH, int G, int LPL, int METRIC>
__global__ void __launch_bounds__(128) rerank_kernel(const KParams p) {
	extern __shared__ __align__(16) uint8_t smem[];
	const int lane = threadIdx.x & 31;
	uint8_t *wbase = smem + (threadIdx.x >> 5) * p.warp_smem_bytes;
	uint32_t *s_nb = reinterpret_cast<uint32_t *>(wbase);
	float *s_d = reinterpret_cast<float *>(s_nb + 32);
	float *s_d2 = s_d + 32;
	float *s_q = s_d2 + 32;
	const uint32_t warps = (gridDim.x * blockDim.x) >> 5;
	for (uint32_t qi = (blockIdx.x * blockDim.x + threadIdx.x) >> 5; qi < p.nq; qi += warps) {
		QueryRegs<TEAM, CH> Q;
		load_query<TEAM, CH, METRIC>(Q, p, qi, lane, s_q);
		uint64_t keys[LPL];
		uint32_t total = 0;
#pragma unroll
		for (int c = 0; c < LPL; c++) {
			const uint32_t i = c * 32 + lane;
			uint64_t ck = i < p.L ? p.cand[(uint64_t)qi * p.L + i] : LMD_KEY_NONE;
			const bool valid = ck != LMD_KEY_NONE;
			const uint32_t mask = __ballot_sync(FULL, valid);
			const int m = __popc(mask);
			keys[c] = LMD_KEY_NONE;
			if (m == 0) {
				continue;
			}
			// the list is dense from the front, so valid lanes are 0..m-1
			__syncwarp();
			if (valid) {
				s_nb[lane] = key_slot(ck);
			}
			__syncwarp();
			prefetch_rows<(CH > 0 ? CH * TEAM / 8 : 0)>(p, s_nb, m, lane);
			compute_batch<TEAM, CH, G, METRIC>(Q, p, s_nb, s_d, s_d2, m, lane);
			__syncwarp();
			const float dd = finish_batch<METRIC>(s_d, s_d2, Q.na, m, lane);
			if (valid) {
				keys[c] = make_key(dd, key_slot(ck));
			}
			total += m;
		}
		// k rounds of warp-wide min extraction
		const uint32_t kk = total < p.k ? total : p.k;
		for (uint32_t out = 0; out < p.k; out++) {
			uint64_t best = LMD_KEY_NONE;
#pragma unroll
			for (int c = 0; c < LPL; c++) {
				best = keys[c] < best ? keys[c] : best;
			}
			uint64_t wmin = best;
#pragma unroll
			for (int o = 16; o > 0; o >>= 1) {
				uint64_t other = __shfl_xor_sync(FULL, wmin, o);
				wmin = other < wmin ? other : wmin;
			}
#pragma unroll
			for (int c = 0; c < LPL; c++) {
				if (keys[c] == wmin) {
					keys[c] = LMD_KEY_NONE; // keys are unique per slot
				}
			}
			if (lane == 0) {
				const bool ok = out < kk && wmin != LMD_KEY_NONE;
				p.out_ids[(uint64_t)qi * p.k + out] = ok ? p.row_ids[key_slot(wmin)] : -1;
				if (p.out_dist) {
					p.out_dist[(uint64_t)qi * p.k + out] = ok ? key_dist(wmin) : 0.0f;
				}
			}
		}
		if (lane == 0) {
			if (p.out_count) {
				p.out_count[qi] = kk;
			}
			atomicAdd(&p.counters[4], (unsigned long long)total);
		}
		__syncwarp();
	}
}

// ------------------------------------------------------------------------------------------------
// launch helpers shared by the per-variant translation units
// ------------------------------------------------------------------------------------------------
template <typename K>
static lmd_status launch_persistent(K kernel, KParams &p, const lmd_store *s, uint32_t nq, cudaStream_t st,
                                    const char *what) {
	// Warps are independent (one query each), so the CTA size is only a packing choice: take the warps-per-CTA in
	// {4, 2, 1} that puts the most warps on an SM (shared memory is granted per CTA; smaller CTAs waste less of it).
	if ((size_t)p.warp_smem_bytes > (size_t)227 * 1024) {
		set_error("%s: %u bytes of shared memory per query exceed the 227 KB an sm_100a CTA can have (list size / "
		          "filter size too large)", what, p.warp_smem_bytes);
		return LMD_ERR_UNSUPPORTED;
	}
	// the occupancy queries cost several microseconds each: remember the answer per (kernel, shared memory) pair
	// a batch too small to give every SM a CTA of 4 warps runs one warp per CTA: the warps then spread over the SMs
	// instead of sharing a few SMs' load / shared-memory pipes (small index-join chunks, single queries)
	const bool spread = s->tune.warps_per_cta == 0 && nq <= (uint32_t)s->sm_count * 4;
	const uint64_t cache_key = (uint64_t)(uintptr_t)kernel * 0x9E3779B97F4A7C15ull + p.warp_smem_bytes * 2ull + (spread ? 1 : 0) + ((uint64_t)p.cta_warps << 40);
	int best_w = 0, best_ctas = 0;
	std::lock_guard<std::mutex> cache_guard(s->cache_mu);
	auto hit = s->launch_cache.find(cache_key);
	if (hit != s->launch_cache.end()) {
		best_w = hit->second.first;
		best_ctas = hit->second.second;
	} else {
		// opt in to the full 227 KB once: the same kernel is launched with different list / filter sizes
		LMD_CUDA_OK(cudaFuncSetAttribute(kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, 227 * 1024));
		for (int w = 4; w >= 1; w >>= 1) {
			const int forced = spread ? 1 : (s->tune.warps_per_cta ? s->tune.warps_per_cta : (int)p.cta_warps);
			if ((forced == 1 || forced == 2 || forced == 4) && w != forced) {
				continue;
			}
			const size_t smem = (size_t)p.warp_smem_bytes * w;
			if (smem > (size_t)227 * 1024) {
				continue;
			}
			int per_sm = 0;
			LMD_CUDA_OK(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, kernel, w * 32, smem));
			if (per_sm * w > best_w * best_ctas) {
				best_w = w;
				best_ctas = per_sm;
			}
		}
		if (best_ctas < 1) {
			set_error("%s: kernel does not fit on an SM with %u bytes of shared memory per warp", what, p.warp_smem_bytes);
			return LMD_ERR_UNSUPPORTED;
		}
		s->launch_cache[cache_key] = std::make_pair(best_w, best_ctas);
	}
	int warps_per_sm = best_w * best_ctas;
	if (s->tune.ctas_per_sm > 0 && s->tune.ctas_per_sm * 4 < warps_per_sm) {
		best_ctas = std::max(1, s->tune.ctas_per_sm * 4 / best_w); // the knob keeps its meaning: CTAs of 4 warps
	}
	uint32_t grid = (nq + best_w - 1) / best_w;
	const uint32_t cap = (uint32_t)(best_ctas * s->sm_count);
	if (grid > cap) {
		grid = cap;
	}
	kernel<<<grid, best_w * 32, (size_t)p.warp_smem_bytes * best_w, st>>>(p);
	LMD_CUDA_OK(cudaGetLastError());
	return LMD_OK;
}

#define LMD_DISPATCH_LPL(KERNEL, TEAM, CH, G, METRIC, ...)                                                             \
	switch (lpl) {                                                                                                     \
	case 1:                                                                                                            \
		return launch_persistent(KERNEL<TEAM, CH, G, 1, METRIC, ##__VA_ARGS__>, p, s, nq, st, #KERNEL);                \
	case 2:                                                                                                            \
		return launch_persistent(KERNEL<TEAM, CH, G, 2, METRIC, ##__VA_ARGS__>, p, s, nq, st, #KERNEL);                \
	case 3:                                                                                                            \
		return launch_persistent(KERNEL<TEAM, CH, G, 3, METRIC, ##__VA_ARGS__>, p, s, nq, st, #KERNEL);                \
	case 4:                                                                                                            \
		return launch_persistent(KERNEL<TEAM, CH, G, 4, METRIC, ##__VA_ARGS__>, p, s, nq, st, #KERNEL);                \
	case 6:                                                                                                            \
		return launch_persistent(KERNEL<TEAM, CH, G, 6, METRIC, ##__VA_ARGS__>, p, s, nq, st, #KERNEL);                \
	case 8:                                                                                                            \
		return launch_persistent(KERNEL<TEAM, CH, G, 8, METRIC, ##__VA_ARGS__>, p, s, nq, st, #KERNEL);                \
	case 12:                                                                                                           \
		return launch_persistent(KERNEL<TEAM, CH, G, 12, METRIC, ##__VA_ARGS__>, p, s, nq, st, #KERNEL);               \
	case 16:                                                                                                           \
		return launch_persistent(KERNEL<TEAM, CH, G, 16, METRIC, ##__VA_ARGS__>, p, s, nq, st, #KERNEL);               \
	default:                                                                                                           \
		break;                                                                                                         \
	}

// filtered traversal: half the list-size variants (lists of 64 / 128 / 256 / 512 entries)
#define LMD_DISPATCH_LPL_COARSE(KERNEL, TEAM, CH, G, METRIC, ...)                                                      \
	switch (lpl) {                                                                                                     \
	case 2:                                                                                                            \
		return launch_persistent(KERNEL<TEAM, CH, G, 2, METRIC, ##__VA_ARGS__>, p, s, nq, st, #KERNEL);                \
	case 4:                                                                                                            \
		return launch_persistent(KERNEL<TEAM, CH, G, 4, METRIC, ##__VA_ARGS__>, p, s, nq, st, #KERNEL);                \
	case 8:                                                                                                            \
		return launch_persistent(KERNEL<TEAM, CH, G, 8, METRIC, ##__VA_ARGS__>, p, s, nq, st, #KERNEL);                \
	case 16:                                                                                                           \
		return launch_persistent(KERNEL<TEAM, CH, G, 16, METRIC, ##__VA_ARGS__>, p, s, nq, st, #KERNEL);               \
	default:                                                                                                           \
		break;                                                                                                         \
	}

#define LMD_DISPATCH_SHAPE_COARSE(KERNEL, METRIC, ...)                                                                 \
	if (p.strict) {                                                                                                    \
		LMD_DISPATCH_LPL_COARSE(KERNEL, 32, 0, 1, METRIC, ##__VA_ARGS__)                                               \
	} else if (p.row_floats == 128) {                                                                                         \
		LMD_DISPATCH_LPL_COARSE(KERNEL, 8, 4, 2, METRIC, ##__VA_ARGS__)                                                \
	} else if (p.row_floats == 96) {                                                                                   \
		LMD_DISPATCH_LPL_COARSE(KERNEL, 8, 3, 2, METRIC, ##__VA_ARGS__)                                                \
	} else if (p.row_floats == 768) {                                                                                  \
		LMD_DISPATCH_LPL_COARSE(KERNEL, 32, 6, 6, METRIC, ##__VA_ARGS__)                                               \
	} else {                                                                                                           \
		LMD_DISPATCH_LPL_COARSE(KERNEL, 32, 0, 1, METRIC, ##__VA_ARGS__)                                               \
	}

// Rows in flight per warp (G): 128-d rows are 512 bytes and the kernel wants many warps (G = 2, 72 registers, 24 warps
// per SM; 10 k queries, L = 56: G = 2 1.21 ms, 4 1.35, 8 1.41); 96-d rows are 384 bytes and G = 4 wins (L = 64: G = 2
// 1.21 ms, 4 1.12, 8 1.35; profiles/r40_mode_sweep_d*.json); a 768-d row is 3 KB and one row per pass leaves the warp waiting on L2 thirteen times per expansion,
// so G = 6 (18 KB in flight per warp, 242 registers, 8 warps per SM) - measured on 1 M x 768 cosine, L = 48, 10 k
// queries: G = 2 8.72 ms, 3 7.90, 4 6.92, 5 6.99, 6 6.01 (profiles/r17_mode_sweep_d768_g*.json).
//
// Latency variants for small batches (fewer queries than the GPU has warp slots): the same kernel with G large enough
// that ALL new neighbours of an expansion (<= 32) are gathered in one pass - every load of the expansion is in flight at
// once, one HBM round trip instead of one per pass. Costs ~3x the registers, which a half-empty GPU does not miss.
// Results are bit-identical to the throughput variants (G only changes how many rows are in flight).
#define LMD_DISPATCH_SHAPE_LATENCY(KERNEL, METRIC, ...)                                                                \
	if (p.row_floats == 128) {                                                                                         \
		LMD_DISPATCH_LPL_COARSE(KERNEL, 8, 4, 8, METRIC, ##__VA_ARGS__)                                                \
	} else if (p.row_floats == 96) {                                                                                   \
		LMD_DISPATCH_LPL_COARSE(KERNEL, 8, 3, 8, METRIC, ##__VA_ARGS__)                                                \
	}

// ... and the step in between (G = 4, ~100 registers, 20 warps per SM) for batches that no longer fit the G = 8
// kernels' 12 warps per SM in one wave but still leave the throughput kernels' 24 warps per SM half empty - the index
// join's 2 048-row chunks sit exactly there.
#define LMD_DISPATCH_SHAPE_MID(KERNEL, METRIC, ...)                                                                    \
	if (p.row_floats == 128) {                                                                                         \
		LMD_DISPATCH_LPL_COARSE(KERNEL, 8, 4, 4, METRIC, ##__VA_ARGS__)                                                \
	} else if (p.row_floats == 96) {                                                                                   \
		LMD_DISPATCH_LPL_COARSE(KERNEL, 8, 3, 4, METRIC, ##__VA_ARGS__)                                                \
	}

// Exact routing behind the int8 pre-filter gathers ~6 rows per expansion instead of ~26: one row in flight per team
// (G = 1, 64 registers, 32 warps per SM) against the plain kernels' two ("pf_g" tuning key; measurements in DESIGN.md).
#define LMD_DISPATCH_SHAPE_PF(KERNEL, METRIC, ...)                                                                     \
	if (p.row_floats == 128 && p.pf_g == 1) {                                                                          \
		LMD_DISPATCH_LPL(KERNEL, 8, 4, 1, METRIC, ##__VA_ARGS__)                                                       \
	} else if (p.row_floats == 96 && p.pf_g == 1) {                                                                    \
		LMD_DISPATCH_LPL(KERNEL, 8, 3, 1, METRIC, ##__VA_ARGS__)                                                       \
	}

#define LMD_DISPATCH_SHAPE(KERNEL, METRIC, ...)                                                                        \
	if (p.strict) {                                                                                                    \
		LMD_DISPATCH_LPL(KERNEL, 32, 0, 1, METRIC, ##__VA_ARGS__)                                                      \
	} else if (p.row_floats == 128) {                                                                                         \
		LMD_DISPATCH_LPL(KERNEL, 8, 4, 2, METRIC, ##__VA_ARGS__)                                                       \
	} else if (p.row_floats == 96) {                                                                                   \
		LMD_DISPATCH_LPL(KERNEL, 8, 3, 2, METRIC, ##__VA_ARGS__)                                                       \
	} else if (p.row_floats == 768) {                                                                                  \
		LMD_DISPATCH_LPL(KERNEL, 32, 6, 6, METRIC, ##__VA_ARGS__)                                                      \
	} else {                                                                                                           \
		LMD_DISPATCH_LPL(KERNEL, 32, 0, 1, METRIC, ##__VA_ARGS__)                                                      \
	}

} // namespace lmd
