/* TEST INFRASTRUCTURE ONLY (oracle/) - see lmd_oracle.h for the rules and the parity-pinning note.
 * Plain-C restatement of the reference's query hot path. Citations are relative to
 * /root/reference/src/lm_diskann/. Built with -ffp-contract=off so that fp32 results equal the
 * reference's own default (non-FMA) build bit for bit. */
#include "lmd_oracle.h"

#include <math.h>
#include <pthread.h>
#include <stdlib.h>
#include <string.h>
#include <unistd.h>

static uint64_t align_up(uint64_t v, uint64_t a) {
	return (v + a - 1) / a * a;
}

/* core/ternary_quantization.hpp:120-123 */
uint64_t lmdo_words_per_plane(uint64_t dims) {
	return (dims + 63) >> 6;
}

/* core/index_config.cpp:104-148 (CalculateLayoutInternal), :40-64 (element / plane sizes) */
int lmdo_layout_calc(uint64_t dims, int vec_type, uint32_t r, lmdo_layout *out) {
	uint64_t elem;
	if (dims == 0) {
		return 1;
	}
	if (vec_type == LMDO_VEC_FLOAT32) {
		elem = 4;
	} else if (vec_type == LMDO_VEC_INT8) {
		elem = 1;
	} else {
		return 1;
	}
	uint64_t plane = lmdo_words_per_plane(dims) * 8;
	uint64_t off = 2; /* uint16 neighbour count at 0 */
	out->count_off = 0;
	off = align_up(off, 8);
	out->vector_off = off;
	off += elem * dims;
	off = align_up(off, 8);
	out->ids_off = off;
	off += (uint64_t)r * 8;
	out->ternary_edge_bytes = 2 * plane;
	off = align_up(off, 8);
	out->pos_off = off;
	off += (uint64_t)r * plane;
	off = align_up(off, 8);
	out->neg_off = off;
	off += (uint64_t)r * plane;
	out->total_node_size = off;
	return 0;
}

/* db/LmDiskannIndex.cpp:161-162: block = AlignValue<Storage::SECTOR_SIZE>(total) with SECTOR_SIZE 4096 */
uint64_t lmdo_block_bytes(uint64_t total_node_size) {
	return align_up(total_node_size, 4096);
}

/* core/distance.hpp:50-105 (ComputeExactDistanceFloat): sequential fp32 accumulation */
float lmdo_distance(const float *a, const float *b, uint64_t dims, int metric, int *err) {
	if (err) {
		*err = 0;
	}
	if (metric == LMDO_METRIC_L2) {
		float s = 0.0f;
		for (uint64_t i = 0; i < dims; i++) {
			float d = a[i] - b[i];
			s += d * d;
		}
		if (s < 0.0f) {
			s = 0.0f;
		}
		return sqrtf(s);
	}
	if (metric == LMDO_METRIC_IP) {
		float dot = 0.0f;
		for (uint64_t i = 0; i < dims; i++) {
			dot += a[i] * b[i];
		}
		return -dot;
	}
	if (metric == LMDO_METRIC_COSINE) {
		float dot = 0.0f, na = 0.0f, nb = 0.0f;
		for (uint64_t i = 0; i < dims; i++) {
			dot += a[i] * b[i];
			na += a[i] * a[i];
			nb += b[i] * b[i];
		}
		if (na <= 0.0f || nb <= 0.0f) {
			return 1.0f;
		}
		float ra = sqrtf(na), rb = sqrtf(nb);
		if (ra == 0.0f || rb == 0.0f) {
			return 1.0f;
		}
		float c = dot / (ra * rb);
		/* std::max(-1.0f, std::min(1.0f, c)) - written out so NaN handling matches */
		float m = (c < 1.0f) ? c : 1.0f;
		c = (-1.0f < m) ? m : -1.0f;
		return 1.0f - c;
	}
	if (err) {
		*err = 1; /* reference throws INVALID_INPUT (distance.hpp:100-103) */
	}
	return 0.0f;
}

/* common/ann.hpp:42-62 (ConvertRawVectorToFloat) */
int lmdo_convert_raw(const uint8_t *raw, float *out, uint64_t dims, int vec_type) {
	if (!raw || !out) {
		return 1;
	}
	if (vec_type == LMDO_VEC_FLOAT32) {
		memcpy(out, raw, dims * sizeof(float));
		return 0;
	}
	if (vec_type == LMDO_VEC_INT8) {
		const int8_t *p = (const int8_t *)raw;
		for (uint64_t i = 0; i < dims; i++) {
			out[i] = (float)p[i];
		}
		return 0;
	}
	return 1;
}

/* core/ternary_quantization.hpp:146-199 (EncodeTernary) */
void lmdo_encode_ternary(const float *v, uint64_t dims, uint64_t *pos, uint64_t *neg) {
	uint64_t words = lmdo_words_per_plane(dims);
	memset(pos, 0, words * 8);
	memset(neg, 0, words * 8);
	for (uint64_t d = 0; d < dims; d++) {
		uint64_t m = 1ULL << (d & 63);
		if (v[d] > 0.0f) {
			pos[d >> 6] |= m;
		} else if (v[d] < 0.0f) {
			neg[d >> 6] |= m;
		}
	}
}

/* core/ternary_quantization.hpp:269-296 (Dot_scalar) */
int64_t lmdo_ternary_dot(const uint64_t *qp, const uint64_t *qn, const uint64_t *vp, const uint64_t *vn,
                         uint64_t words) {
	int64_t acc = 0;
	for (uint64_t i = 0; i < words; i++) {
		acc += __builtin_popcountll(qp[i] & vp[i]);
		acc -= __builtin_popcountll(qp[i] & vn[i]);
		acc -= __builtin_popcountll(qn[i] & vp[i]);
		acc += __builtin_popcountll(qn[i] & vn[i]);
	}
	return acc;
}

/* Shared tail of CalculateApproxDistance (core/distance.hpp:229-242): similarity -> distance */
static float approx_from_score(int64_t score, int metric, int *err) {
	float sim = (float)score; /* distance.hpp:152 static_cast<float>(raw_score) */
	if (metric == LMDO_METRIC_IP) {
		return -sim;
	}
	if (metric == LMDO_METRIC_COSINE) {
		return 1.0f - sim;
	}
	if (err) {
		*err = 1; /* L2 and others throw (distance.hpp:236-241) */
	}
	return 0.0f;
}

/* core/distance.hpp:212-243 (CalculateApproxDistance) over a [pos||neg] contiguous edge, which
 * re-encodes the query on every call (:137-140) */
float lmdo_approx_distance(const float *q, const uint8_t *compressed, uint64_t dims, int metric, int *err) {
	if (err) {
		*err = 0;
	}
	uint64_t words = lmdo_words_per_plane(dims);
	uint64_t *qp = (uint64_t *)malloc(words * 16);
	uint64_t *qn = qp + words;
	uint64_t *vp = (uint64_t *)malloc(words * 16);
	uint64_t *vn = vp + words;
	memcpy(vp, compressed, words * 16); /* planes may be 8-byte aligned only */
	lmdo_encode_ternary(q, dims, qp, qn);
	int64_t s = lmdo_ternary_dot(qp, qn, vp, vn, words);
	free(qp);
	free(vp);
	return approx_from_score(s, metric, err);
}

/* ------------------------------------------------------------------------------------------------
 * Search: core/Searcher.cpp:45-246 (PerformSearch) + :256-272 (Searcher::Search) with the four
 * hunks of patches/searcher_fixed.patch applied (H1 top-L is a max-heap, H2 `break`, H3 rerank keeps
 * the k smallest ascending, H4 (distance,row_id) order).
 * ---------------------------------------------------------------------------------------------- */

typedef struct {
	float d;
	int64_t id;
} cand_t; /* Searcher.cpp:31-38 NodeCandidate */

/* H4: operator> orders by (distance,row_id) */
static int cand_gt(cand_t a, cand_t b) {
	return a.d > b.d || (a.d == b.d && a.id > b.id);
}

/* std::priority_queue stand-in. is_max=0: std::greater (top = smallest); is_max=1: std::less
 * (top = largest). The order is total over distinct ids, so top() is unique however the heap is
 * arranged internally. */
typedef struct {
	cand_t *a;
	size_t n, cap;
	int is_max;
} heap_t;

static int heap_before(const heap_t *h, cand_t x, cand_t y) { /* x should be nearer the top than y */
	return h->is_max ? cand_gt(x, y) : cand_gt(y, x);
}
static void heap_push(heap_t *h, cand_t c) {
	if (h->n == h->cap) {
		h->cap = h->cap ? h->cap * 2 : 64;
		h->a = (cand_t *)realloc(h->a, h->cap * sizeof(cand_t));
	}
	size_t i = h->n++;
	while (i > 0) {
		size_t p = (i - 1) / 2;
		if (!heap_before(h, c, h->a[p])) {
			break;
		}
		h->a[i] = h->a[p];
		i = p;
	}
	h->a[i] = c;
}
static cand_t heap_pop(heap_t *h) {
	cand_t top = h->a[0];
	cand_t last = h->a[--h->n];
	size_t i = 0;
	for (;;) {
		size_t l = 2 * i + 1, r = l + 1, m;
		if (l >= h->n) {
			break;
		}
		m = (r < h->n && heap_before(h, h->a[r], h->a[l])) ? r : l;
		if (!heap_before(h, h->a[m], last)) {
			break;
		}
		h->a[i] = h->a[m];
		i = m;
	}
	if (h->n) {
		h->a[i] = last;
	}
	return top;
}

/* std::set<row_t> / std::map<row_t,IndexPointer> stand-in: open-addressing int64 -> uint64 table */
typedef struct {
	int64_t *keys;
	uint64_t *vals;
	uint8_t *used;
	size_t cap, n;
} i64map_t;

static uint64_t mix64(uint64_t x) {
	x ^= x >> 33;
	x *= 0xff51afd7ed558ccdULL;
	x ^= x >> 33;
	x *= 0xc4ceb9fe1a85ec53ULL;
	x ^= x >> 33;
	return x;
}
static void map_init(i64map_t *m, size_t cap_pow2) {
	m->cap = cap_pow2;
	m->n = 0;
	m->keys = (int64_t *)malloc(cap_pow2 * sizeof(int64_t));
	m->vals = (uint64_t *)malloc(cap_pow2 * sizeof(uint64_t));
	m->used = (uint8_t *)calloc(cap_pow2, 1);
}
static void map_free(i64map_t *m) {
	free(m->keys);
	free(m->vals);
	free(m->used);
}
static int map_find(const i64map_t *m, int64_t k, uint64_t *v) {
	size_t i = mix64((uint64_t)k) & (m->cap - 1);
	while (m->used[i]) {
		if (m->keys[i] == k) {
			if (v) {
				*v = m->vals[i];
			}
			return 1;
		}
		i = (i + 1) & (m->cap - 1);
	}
	return 0;
}
static void map_put(i64map_t *m, int64_t k, uint64_t v);
static void map_grow(i64map_t *m) {
	i64map_t b;
	map_init(&b, m->cap * 2);
	for (size_t i = 0; i < m->cap; i++) {
		if (m->used[i]) {
			map_put(&b, m->keys[i], m->vals[i]);
		}
	}
	map_free(m);
	*m = b;
}
static void map_put(i64map_t *m, int64_t k, uint64_t v) {
	if ((m->n + 1) * 2 > m->cap) {
		map_grow(m);
	}
	size_t i = mix64((uint64_t)k) & (m->cap - 1);
	while (m->used[i]) {
		if (m->keys[i] == k) {
			return; /* first mapping wins, like std::map::insert / unordered_map[] fill order */
		}
		i = (i + 1) & (m->cap - 1);
	}
	m->used[i] = 1;
	m->keys[i] = k;
	m->vals[i] = v;
	m->n++;
}
static void map_clear(i64map_t *m) {
	memset(m->used, 0, m->cap);
	m->n = 0;
}

typedef struct {
	const uint8_t *blocks;
	uint64_t n, block_bytes;
	const i64map_t *rowmap; /* NULL: row_id == slot */
	int64_t entry_row;
	uint64_t entry_slot;
	int entry_ok;
	int metric, vec_type, route;
	uint64_t dims;
	uint32_t r, cfg_l_search;
	lmdo_layout lay;
	const uint8_t *allowed; /* filtered search (row f4): one byte per block position, NULL = unfiltered */
	const float *code_off;  /* int8 edge codec (row f4, PARITY UNPINNED): x ~ code_off[d] + code_scale * c[d] */
	float code_scale;
} index_t;

/* The int8 edge codec's grid (per-dimension offsets + one scale). The CUDA path derives it from the data
 * (csrc/lmd_sidecar.cu code_params_kernel); the tests download it with lmd_store_get_codec and hand it over here so
 * that both sides quantise on the same grid. Process-global: set before lmdo_search(route = LMDO_ROUTE_INT8). */
static const float *g_code_off = NULL;
static float g_code_scale = 1.0f;
void lmdo_set_codec(const float *off, float scale) {
	g_code_off = off;
	g_code_scale = scale;
}

static int code_of(float x, float off, float inv) {
	long c = lrintf((x - off) * inv); /* round to nearest even, as __float2int_rn */
	return c < -127 ? -127 : (c > 127 ? 127 : (int)c);
}

/* GraphManager.cpp:523-531 (TryGetNodePointer) over the harness' row_id -> slot mapping */
static int try_get_slot(const index_t *ix, int64_t row, uint64_t *slot) {
	if (!ix->rowmap) {
		if (row < 0 || (uint64_t)row >= ix->n) {
			return 0;
		}
		*slot = (uint64_t)row;
		return 1;
	}
	return map_find(ix->rowmap, row, slot);
}

typedef struct {
	heap_t top, beam, fin, res;
	i64map_t visited;
	float *tmp;
	uint64_t *qplanes;
	int8_t *qcodes; /* int8 codec: the query on the code grid */
	int q_asq;      /* sum of qcodes^2 */
	uint64_t expansions, fetches;
} scratch_t;

static const uint8_t *block_of(const index_t *ix, uint64_t slot) {
	return ix->blocks + slot * ix->block_bytes;
}

static float exact_to_slot(const index_t *ix, scratch_t *s, const float *q, uint64_t slot) {
	const uint8_t *blk = block_of(ix, slot);
	s->fetches++;
	lmdo_convert_raw(blk + ix->lay.vector_off, s->tmp, ix->dims, ix->vec_type); /* Searcher.cpp:161-166 */
	return lmdo_distance(q, s->tmp, ix->dims, ix->metric, NULL);            /* :172-173 */
}

/* Returns number of ids written to res (<= k; <= L when final_pass == 0). */
static uint64_t search_core(const index_t *ix, scratch_t *s, const float *q, uint64_t k, uint64_t l_param,
                            int64_t *res, int final_pass) {
	/* Searcher.cpp:261-265 */
	uint64_t L = l_param > 0 ? l_param : ix->cfg_l_search;
	if (L == 0) {
		L = k > 20 ? k + 20 : 40;
	}
	if (L < k) {
		L = k;
	}
	s->top.n = s->beam.n = s->fin.n = s->res.n = 0;
	s->res.is_max = 1;  /* filtered search: the results heap RH, bounded like `top` */
	s->top.is_max = 1;  /* H1 */
	s->beam.is_max = 0; /* :67 */
	s->fin.is_max = 1;  /* H3 */
	map_clear(&s->visited);
	s->expansions = s->fetches = 0;

	/* :70-83 */
	if (ix->entry_row == LMDO_EMPTY_ID || !ix->entry_ok) {
		return 0;
	}
	float d_entry = exact_to_slot(ix, s, q, ix->entry_slot); /* :85-99 */
	cand_t e = {d_entry, ix->entry_row};
	heap_push(&s->beam, e); /* :101-103 */
	map_put(&s->visited, ix->entry_row, 0);
	heap_push(&s->top, e);
	float worst = d_entry; /* :106 lowest_dist_in_results */
	/* Filtered search ("Consolidated Diskann Proposal v1", section "Filtered Searches": dual heap). The exploration
	 * heap EH is the unfiltered search's own state (`top` + `beam`): the traversal tunnels through nodes that fail the
	 * filter exactly as if there were none. The results heap RH (`res`) is offered every node whose distance gets
	 * evaluated and keeps the L best ALLOWED ones; only RH is returned. */
	if (ix->allowed && ix->allowed[ix->entry_slot]) {
		heap_push(&s->res, e);
	}

	uint64_t words = lmdo_words_per_plane(ix->dims);
	if (ix->route == LMDO_ROUTE_TERNARY) {
		/* the reference re-encodes the query per neighbour (distance.hpp:137-140); same bits, once */
		lmdo_encode_ternary(q, ix->dims, s->qplanes, s->qplanes + words);
	}
	if (ix->route == LMDO_ROUTE_INT8) {
		const float inv = 1.0f / ix->code_scale;
		s->q_asq = 0;
		for (uint64_t i = 0; i < ix->dims; i++) {
			int c = code_of(q[i], ix->code_off[i], inv);
			s->qcodes[i] = (int8_t)c;
			s->q_asq += c * c;
		}
	}

	while (s->beam.n) { /* :108 */
		cand_t cur = heap_pop(&s->beam);
		if (cur.d > worst && s->top.n >= L) {
			break; /* :112-114 + H2 */
		}
		uint64_t cur_slot;
		if (!try_get_slot(ix, cur.id, &cur_slot)) {
			continue; /* :120-122 */
		}
		/* GraphManager.cpp:230-257 (GetNeighbors): first `count` ids, INT64_MAX dropped */
		const uint8_t *blk = block_of(ix, cur_slot);
		uint16_t count;
		memcpy(&count, blk + ix->lay.count_off, 2);
		const uint8_t *idp = blk + ix->lay.ids_off;
		s->expansions++;
		for (uint16_t i = 0; i < count; i++) {
			int64_t nb;
			memcpy(&nb, idp + (size_t)i * 8, 8);
			if (nb == LMDO_EMPTY_ID) {
				continue; /* :132-133 */
			}
			if (map_find(&s->visited, nb, NULL)) {
				continue; /* :135-137 */
			}
			map_put(&s->visited, nb, 0); /* :139 visited at discovery */
			uint64_t nb_slot;
			if (!try_get_slot(ix, nb, &nb_slot)) {
				continue; /* :146-148 */
			}
			float d;
			if (ix->route == LMDO_ROUTE_EXACT) {
				d = exact_to_slot(ix, s, q, nb_slot); /* :153-173 */
			} else if (ix->route == LMDO_ROUTE_INT8) {
				/* libsql/lm_diskann.cpp:1596-1625 with LmDiskannEdgeType INT8: the compressed copy of neighbour i is its
				 * vector on the codec grid; routing distance = || q^ - x^ ||, integers up to the square root */
				const float inv = 1.0f / ix->code_scale;
				lmdo_convert_raw(block_of(ix, nb_slot) + ix->lay.vector_off, s->tmp, ix->dims, ix->vec_type);
				int dot = 0, csq = 0;
				for (uint64_t t = 0; t < ix->dims; t++) {
					int c = code_of(s->tmp[t], ix->code_off[t], inv);
					dot += c * (int)s->qcodes[t];
					csq += c * c;
				}
				int i2 = s->q_asq - 2 * dot + csq;
				d = ix->code_scale * sqrtf((float)(i2 > 0 ? i2 : 0)) + 0.0f;
			} else {
				/* the TODO at :168-171: compressed copy of neighbour i held in the EXPANDED block,
				 * SoA planes of index_config.cpp:136-143; distance.hpp:229-242 for the mapping */
				uint64_t vp[words], vn[words];
				memcpy(vp, blk + ix->lay.pos_off + (size_t)i * words * 8, words * 8);
				memcpy(vn, blk + ix->lay.neg_off + (size_t)i * words * 8, words * 8);
				d = approx_from_score(lmdo_ternary_dot(s->qplanes, s->qplanes + words, vp, vn, words),
				                      ix->metric, NULL);
			}
			if (ix->allowed && ix->allowed[nb_slot] && (s->res.n < L || d < s->res.a[0].d)) {
				cand_t c = {d, nb};
				heap_push(&s->res, c);
				if (s->res.n > L) {
					heap_pop(&s->res);
				}
			}
			if (s->top.n < L || d < worst) { /* :175 */
				cand_t c = {d, nb};
				heap_push(&s->beam, c);
				heap_push(&s->top, c);
				if (s->top.n > L) {
					heap_pop(&s->top); /* :179-181 */
				}
				if (s->top.n) {
					worst = s->top.a[0].d; /* :182-184 */
				}
			}
		}
	}

	if (ix->allowed) { /* only RH is returned: it takes the place of `top` below */
		heap_t t = s->top;
		s->top = s->res;
		s->res = t;
	}
	if (!final_pass) {
		/* Searcher.cpp:189 + H5: SearchForInitialCandidates gets the top list itself, worst first (heap pop order) */
		uint64_t cnt = 0;
		while (s->top.n) {
			res[cnt++] = heap_pop(&s->top).id;
		}
		return cnt;
	}
	/* :189-245 final pass: exact distance of every listed candidate, keep k smallest, ascending */
	while (s->top.n) {
		cand_t c = heap_pop(&s->top);
		uint64_t slot;
		if (!try_get_slot(ix, c.id, &slot)) {
			continue;
		}
		cand_t x = {exact_to_slot(ix, s, q, slot), c.id};
		heap_push(&s->fin, x);
		if (s->fin.n > k) {
			heap_pop(&s->fin);
		}
	}
	uint64_t cnt = s->fin.n;
	for (uint64_t j = cnt; j-- > 0;) {
		res[j] = heap_pop(&s->fin).id; /* max-heap pops descending; fill from the back (H3) */
	}
	return cnt;
}

static uint64_t search_one(const index_t *ix, scratch_t *s, const float *q, uint64_t k, uint64_t l_param,
                           int64_t *res) {
	return search_core(ix, s, q, k, l_param, res, 1);
}

typedef struct {
	const index_t *ix;
	const float *queries;
	uint64_t nq, k, l_param;
	int64_t *out_ids;
	float *out_dist;
	uint32_t *out_count;
	uint64_t *out_stats;
	volatile int64_t *next;
} job_t;

static void *worker(void *arg) {
	job_t *j = (job_t *)arg;
	const index_t *ix = j->ix;
	scratch_t s;
	memset(&s, 0, sizeof(s));
	map_init(&s.visited, 4096);
	s.tmp = (float *)malloc(ix->dims * sizeof(float));
	s.qplanes = (uint64_t *)malloc(lmdo_words_per_plane(ix->dims) * 16);
	s.qcodes = (int8_t *)malloc(ix->dims + 16);
	int64_t *res = (int64_t *)malloc((j->k ? j->k : 1) * sizeof(int64_t));
	for (;;) {
		int64_t q0 = __sync_fetch_and_add(j->next, 8);
		if (q0 >= (int64_t)j->nq) {
			break;
		}
		int64_t q1 = q0 + 8 < (int64_t)j->nq ? q0 + 8 : (int64_t)j->nq;
		for (int64_t qi = q0; qi < q1; qi++) {
			const float *q = j->queries + (size_t)qi * ix->dims;
			uint64_t cnt = search_one(ix, &s, q, j->k, j->l_param, res);
			if (j->out_count) {
				j->out_count[qi] = (uint32_t)cnt;
			}
			if (j->out_stats) {
				j->out_stats[qi * 2] = s.expansions;
				j->out_stats[qi * 2 + 1] = s.fetches;
			}
			for (uint64_t t = 0; t < j->k; t++) {
				int64_t id = t < cnt ? res[t] : -1;
				j->out_ids[qi * j->k + t] = id;
				if (j->out_dist) {
					uint64_t slot;
					float d = 0.0f;
					if (id >= 0 && try_get_slot(ix, id, &slot)) {
						d = exact_to_slot(ix, &s, q, slot);
					}
					j->out_dist[qi * j->k + t] = d;
				}
			}
		}
	}
	free(res);
	free(s.qplanes);
	free(s.qcodes);
	free(s.tmp);
	map_free(&s.visited);
	free(s.top.a);
	free(s.beam.a);
	free(s.fin.a);
	free(s.res.a);
	return NULL;
}

static int search_impl(const uint8_t *blocks, uint64_t n, uint64_t block_bytes, const int64_t *row_ids,
                       int64_t entry_row_id, int metric, int vec_type, uint64_t dims, uint32_t r, uint32_t cfg_l_search,
                       const float *queries, uint64_t nq, uint64_t k, uint64_t l_param, int route,
                       const uint8_t *allowed, int64_t *out_ids, float *out_dist, uint32_t *out_count,
                       uint64_t *out_stats, int nthreads) {
	index_t ix;
	memset(&ix, 0, sizeof(ix));
	ix.allowed = allowed;
	if (lmdo_layout_calc(dims, vec_type, r, &ix.lay)) {
		return 2;
	}
	if (metric != LMDO_METRIC_L2 && metric != LMDO_METRIC_COSINE && metric != LMDO_METRIC_IP) {
		return 2;
	}
	if (route == LMDO_ROUTE_TERNARY && metric == LMDO_METRIC_L2) {
		return 2; /* distance.hpp:236-239 throws */
	}
	if (route == LMDO_ROUTE_INT8 && (metric != LMDO_METRIC_L2 || !g_code_off)) {
		return 2;
	}
	ix.code_off = g_code_off;
	ix.code_scale = g_code_scale;
	ix.blocks = blocks;
	ix.n = n;
	ix.block_bytes = block_bytes;
	ix.metric = metric;
	ix.vec_type = vec_type;
	ix.route = route;
	ix.dims = dims;
	ix.r = r;
	ix.cfg_l_search = cfg_l_search;
	i64map_t rowmap;
	if (row_ids) {
		size_t cap = 64;
		while (cap < n * 2 + 2) {
			cap *= 2;
		}
		map_init(&rowmap, cap);
		for (uint64_t i = 0; i < n; i++) {
			/* ref_harness.cpp fills map_[row_ids[i]] = i in order: the LAST duplicate wins there */
			uint64_t old;
			if (map_find(&rowmap, row_ids[i], &old)) {
				size_t p = mix64((uint64_t)row_ids[i]) & (rowmap.cap - 1);
				while (rowmap.keys[p] != row_ids[i] || !rowmap.used[p]) {
					p = (p + 1) & (rowmap.cap - 1);
				}
				rowmap.vals[p] = i;
			} else {
				map_put(&rowmap, row_ids[i], i);
			}
		}
		ix.rowmap = &rowmap;
	}
	ix.entry_row = entry_row_id;
	ix.entry_ok = try_get_slot(&ix, entry_row_id, &ix.entry_slot);
	if (!ix.entry_ok) {
		ix.entry_row = LMDO_EMPTY_ID; /* ref_harness.cpp FakeGraph ctor */
	}
	if (nthreads <= 0) {
		long nc = sysconf(_SC_NPROCESSORS_ONLN);
		nthreads = nc > 0 ? (int)nc : 1;
	}
	volatile int64_t next = 0;
	job_t job = {&ix, queries, nq, k, l_param, out_ids, out_dist, out_count, out_stats, &next};
	if (nthreads == 1) {
		worker(&job);
	} else {
		pthread_t *th = (pthread_t *)malloc(sizeof(pthread_t) * (size_t)nthreads);
		for (int t = 0; t < nthreads; t++) {
			pthread_create(&th[t], NULL, worker, &job);
		}
		for (int t = 0; t < nthreads; t++) {
			pthread_join(th[t], NULL);
		}
		free(th);
	}
	if (row_ids) {
		map_free(&rowmap);
	}
	return 0;
}

int lmdo_search(const uint8_t *blocks, uint64_t n, uint64_t block_bytes, const int64_t *row_ids, int64_t entry_row_id,
                int metric, int vec_type, uint64_t dims, uint32_t r, uint32_t cfg_l_search, const float *queries,
                uint64_t nq, uint64_t k, uint64_t l_param, int route, int64_t *out_ids, float *out_dist,
                uint32_t *out_count, uint64_t *out_stats, int nthreads) {
	return search_impl(blocks, n, block_bytes, row_ids, entry_row_id, metric, vec_type, dims, r, cfg_l_search, queries,
	                   nq, k, l_param, route, NULL, out_ids, out_dist, out_count, out_stats, nthreads);
}

/* Row f4 (designed in the reference's plan, implemented nowhere in it: PARITY UNPINNED). allowed[i] != 0 <=> the row
 * of block i passes the filter. allowed == NULL is the unfiltered search. */
int lmdo_search_filtered(const uint8_t *blocks, uint64_t n, uint64_t block_bytes, const int64_t *row_ids,
                         int64_t entry_row_id, int metric, int vec_type, uint64_t dims, uint32_t r,
                         uint32_t cfg_l_search, const float *queries, uint64_t nq, uint64_t k, uint64_t l_param,
                         int route, const uint8_t *allowed, int64_t *out_ids, float *out_dist, uint32_t *out_count,
                         uint64_t *out_stats, int nthreads) {
	return search_impl(blocks, n, block_bytes, row_ids, entry_row_id, metric, vec_type, dims, r, cfg_l_search, queries,
	                   nq, k, l_param, route, allowed, out_ids, out_dist, out_count, out_stats, nthreads);
}

/* ================================================================================================
 * Graph write side: GraphManager::RobustPrune and GraphManager::AddNode.
 * ================================================================================================ */
typedef struct {
	float d;
	int64_t id;
	uint64_t slot;
} pcand_t;

static int pcand_cmp(const void *a, const void *b) { /* std::sort of pair<float,row_t>, GraphManager.cpp:300 */
	const pcand_t *x = (const pcand_t *)a, *y = (const pcand_t *)b;
	if (x->d < y->d) {
		return -1;
	}
	if (y->d < x->d) {
		return 1;
	}
	return x->id < y->id ? -1 : (x->id > y->id ? 1 : 0);
}

/* GraphManager.cpp:259-457. node_vec = the float vector handed to RobustPrune (AddNode passes the caller's vector
 * for the new node, :140, and the block's converted vector for a neighbour, :185-197). Writes count + ids
 * (INT64_MAX padded) into the node's block (:388-452) and returns the selected ids in cand[0..count). */
static uint32_t robust_prune(index_t *ix, uint8_t *blocks, uint64_t node_slot, const float *node_vec, int64_t *cand,
                             uint64_t ncand, float alpha, pcand_t *work, float *sel_vecs, float *tmp) {
	if (ncand == 0) {
		return 0; /* :262-264: returns before touching the block */
	}
	const uint64_t dims = ix->dims;
	uint64_t m = 0;
	for (uint64_t i = 0; i < ncand; i++) { /* :272-297 */
		uint64_t slot;
		if (!try_get_slot(ix, cand[i], &slot)) {
			continue;
		}
		if (slot == node_slot) {
			continue; /* :280-282 self */
		}
		lmdo_convert_raw(blocks + slot * ix->block_bytes + ix->lay.vector_off, tmp, dims, ix->vec_type);
		work[m].d = lmdo_distance(node_vec, tmp, dims, ix->metric, NULL);
		work[m].id = cand[i];
		work[m].slot = slot;
		m++;
	}
	qsort(work, m, sizeof(pcand_t), pcand_cmp); /* :300 */
	uint64_t u = 0;
	for (uint64_t i = 0; i < m; i++) { /* :302-305 std::unique on row id */
		if (u == 0 || work[u - 1].id != work[i].id) {
			work[u++] = work[i];
		}
	}
	m = u;
	uint32_t nsel = 0;
	float *pvec = tmp;
	for (uint64_t i = 0; i < m && nsel < ix->r; i++) { /* :320-379 */
		lmdo_convert_raw(blocks + work[i].slot * ix->block_bytes + ix->lay.vector_off, pvec, dims, ix->vec_type);
		int keep = 1;
		for (uint32_t j = 0; j < nsel; j++) {
			float dpr = lmdo_distance(pvec, sel_vecs + (size_t)j * dims, dims, ix->metric, NULL);
			if (alpha * dpr < work[i].d) { /* :359 */
				keep = 0;
				break;
			}
		}
		if (keep) {
			memcpy(sel_vecs + (size_t)nsel * dims, pvec, dims * sizeof(float));
			cand[nsel++] = work[i].id;
		}
	}
	uint8_t *blk = blocks + node_slot * ix->block_bytes; /* :388-452 */
	uint16_t c16 = (uint16_t)nsel;
	memcpy(blk + ix->lay.count_off, &c16, 2);
	for (uint32_t j = 0; j < ix->r; j++) {
		int64_t v = j < nsel ? cand[j] : LMDO_EMPTY_ID;
		memcpy(blk + ix->lay.ids_off + (size_t)j * 8, &v, 8);
	}
	return nsel;
}

typedef struct {
	index_t ix;
	i64map_t rowmap;
	scratch_t s;
	pcand_t *work;
	float *sel_vecs, *tmp2, *nvec;
	int64_t *cand;
	uint64_t cand_cap;
} builder_t;

static int builder_init(builder_t *b, uint8_t *blocks, uint64_t block_bytes, int metric, int vec_type, uint64_t dims,
                        uint32_t r, uint32_t l_insert, int use_map) {
	memset(b, 0, sizeof(*b));
	if (lmdo_layout_calc(dims, vec_type, r, &b->ix.lay)) {
		return 2;
	}
	b->ix.blocks = blocks;
	b->ix.block_bytes = block_bytes;
	b->ix.metric = metric;
	b->ix.vec_type = vec_type;
	b->ix.route = LMDO_ROUTE_EXACT;
	b->ix.dims = dims;
	b->ix.r = r;
	b->ix.cfg_l_search = l_insert;
	b->ix.entry_row = LMDO_EMPTY_ID;
	if (use_map) {
		map_init(&b->rowmap, 1024);
		b->ix.rowmap = &b->rowmap;
	}
	map_init(&b->s.visited, 4096);
	b->s.tmp = (float *)malloc(dims * sizeof(float));
	b->s.qplanes = (uint64_t *)malloc(lmdo_words_per_plane(dims) * 16);
	b->s.qcodes = (int8_t *)malloc(dims + 16);
	b->cand_cap = (uint64_t)l_insert + r + 8;
	b->work = (pcand_t *)malloc(b->cand_cap * sizeof(pcand_t));
	b->cand = (int64_t *)malloc(b->cand_cap * sizeof(int64_t));
	b->sel_vecs = (float *)malloc((size_t)r * dims * sizeof(float));
	b->tmp2 = (float *)malloc(dims * sizeof(float));
	b->nvec = (float *)malloc(dims * sizeof(float));
	return 0;
}

static void builder_free(builder_t *b) {
	if (b->ix.rowmap) {
		map_free(&b->rowmap);
	}
	map_free(&b->s.visited);
	free(b->s.tmp);
	free(b->s.qplanes);
	free(b->s.qcodes);
	free(b->s.top.a);
	free(b->s.beam.a);
	free(b->s.fin.a);
	free(b->work);
	free(b->cand);
	free(b->sel_vecs);
	free(b->tmp2);
	free(b->nvec);
}

/* GraphManager.cpp:65-121: allocate + zero the block, store the vector, register the row id, first node = entry */
static void builder_add_vector(builder_t *b, uint8_t *blocks, uint64_t slot, int64_t row, const float *vec) {
	index_t *ix = &b->ix;
	uint8_t *blk = blocks + slot * ix->block_bytes;
	memset(blk, 0, ix->block_bytes); /* GraphNode.cpp:12-24 */
	if (ix->vec_type == LMDO_VEC_FLOAT32) {
		memcpy(blk + ix->lay.vector_off, vec, ix->dims * 4); /* :97-98 */
	} else {
		for (uint64_t i = 0; i < ix->dims; i++) {
			((int8_t *)(blk + ix->lay.vector_off))[i] = (int8_t)vec[i]; /* :104-107 plain cast */
		}
	}
	if (ix->rowmap) {
		/* rowid_to_node_ptr_map_[row_id] = ptr (:116): the latest mapping wins */
		uint64_t old;
		if (map_find(&b->rowmap, row, &old)) {
			size_t p = mix64((uint64_t)row) & (b->rowmap.cap - 1);
			while (b->rowmap.keys[p] != row || !b->rowmap.used[p]) {
				p = (p + 1) & (b->rowmap.cap - 1);
			}
			b->rowmap.vals[p] = slot;
		} else {
			map_put(&b->rowmap, row, slot);
		}
	}
	ix->n = slot + 1;
	if (slot == 0) { /* :119-121 */
		ix->entry_row = row;
		ix->entry_slot = 0;
		ix->entry_ok = 1;
	}
}

int lmdo_build_serial(const float *vectors, uint64_t n, uint64_t dims, int metric, int vec_type, uint32_t r,
                      uint32_t l_insert, float alpha, const int64_t *row_ids, uint64_t block_bytes, uint8_t *out_blocks,
                      int64_t *entry_row_out) {
	builder_t b;
	if (builder_init(&b, out_blocks, block_bytes, metric, vec_type, dims, r, l_insert, row_ids != NULL)) {
		return 2;
	}
	index_t *ix = &b.ix;
	for (uint64_t i = 0; i < n; i++) {
		const int64_t row = row_ids ? row_ids[i] : (int64_t)i;
		const float *vec = vectors + i * dims;
		builder_add_vector(&b, out_blocks, i, row, vec);
		/* :124-136 candidates: SearchForInitialCandidates(L = max(l_insert, l_insert)), Searcher.cpp:281-293 */
		uint64_t nc = 0;
		if (ix->entry_ok) {
			nc = search_core(ix, &b.s, vec, l_insert, l_insert, b.cand, 0);
			if (nc > l_insert) {
				nc = l_insert;
			}
		}
		/* :140 RobustPrune with the caller's float vector; :143-157 write the list (RobustPrune already did) */
		uint32_t deg = robust_prune(ix, out_blocks, i, vec, b.cand, nc, alpha, b.work, b.sel_vecs, b.tmp2);
		if (nc == 0) {
			/* RobustPrune returned early; AddNode still writes count 0 and INT64_MAX ids (:143-157) */
			uint8_t *blk = out_blocks + i * block_bytes;
			for (uint32_t j = 0; j < r; j++) {
				int64_t v = LMDO_EMPTY_ID;
				memcpy(blk + ix->lay.ids_off + (size_t)j * 8, &v, 8);
			}
		}
		/* :171-200 interconnect: every selected neighbour is re-pruned over its own list + the new node */
		int64_t *mine = (int64_t *)malloc((deg ? deg : 1) * sizeof(int64_t));
		memcpy(mine, b.cand, deg * sizeof(int64_t));
		for (uint32_t t = 0; t < deg; t++) {
			uint64_t vslot;
			if (!try_get_slot(ix, mine[t], &vslot)) {
				continue;
			}
			uint8_t *vblk = out_blocks + vslot * block_bytes;
			lmdo_convert_raw(vblk + ix->lay.vector_off, b.nvec, dims, vec_type); /* GetNodeVector */
			uint16_t cnt;
			memcpy(&cnt, vblk + ix->lay.count_off, 2);
			uint64_t m = 0;
			for (uint16_t j = 0; j < cnt; j++) { /* GetNeighbors, GraphManager.cpp:230-257 */
				int64_t id;
				memcpy(&id, vblk + ix->lay.ids_off + (size_t)j * 8, 8);
				if (id != LMDO_EMPTY_ID) {
					b.cand[m++] = id;
				}
			}
			b.cand[m++] = row;
			robust_prune(ix, out_blocks, vslot, b.nvec, b.cand, m, alpha, b.work, b.sel_vecs, b.tmp2);
		}
		free(mine);
	}
	if (entry_row_out) {
		*entry_row_out = ix->entry_ok ? ix->entry_row : LMDO_EMPTY_ID;
	}
	builder_free(&b);
	return 0;
}

int lmdo_robust_prune(const float *vectors, uint64_t n, uint64_t dims, int metric, int vec_type, uint32_t r, float alpha,
                      const int64_t *row_ids, uint64_t m, const uint64_t *node_idx, const int64_t *cands, uint64_t c,
                      int64_t *out_ids, uint32_t *out_count) {
	lmdo_layout lay;
	if (lmdo_layout_calc(dims, vec_type, r, &lay)) {
		return 2;
	}
	const uint64_t bb = lmdo_block_bytes(lay.total_node_size);
	uint8_t *blocks = (uint8_t *)malloc(n * bb);
	builder_t b;
	if (!blocks || builder_init(&b, blocks, bb, metric, vec_type, dims, r, (uint32_t)c, row_ids != NULL)) {
		free(blocks);
		return 2;
	}
	for (uint64_t i = 0; i < n; i++) {
		builder_add_vector(&b, blocks, i, row_ids ? row_ids[i] : (int64_t)i, vectors + i * dims);
	}
	for (uint64_t j = 0; j < m; j++) {
		uint64_t nc = 0;
		for (uint64_t t = 0; t < c; t++) {
			if (cands[j * c + t] != LMDO_EMPTY_ID) {
				b.cand[nc++] = cands[j * c + t];
			}
		}
		lmdo_convert_raw(blocks + node_idx[j] * bb + lay.vector_off, b.nvec, dims, vec_type);
		uint32_t deg = robust_prune(&b.ix, blocks, node_idx[j], b.nvec, b.cand, nc, alpha, b.work, b.sel_vecs, b.tmp2);
		out_count[j] = deg;
		for (uint32_t t = 0; t < r; t++) {
			out_ids[j * r + t] = t < deg ? b.cand[t] : LMDO_EMPTY_ID;
		}
	}
	builder_free(&b);
	free(blocks);
	return 0;
}
