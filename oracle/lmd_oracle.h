/* TEST INFRASTRUCTURE ONLY (oracle/). Never linked into, imported by or executed from the product
 * library; only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs
 * may load it, and only as the checker or the timed CPU baseline.
 *
 * Plain-C restatement of the reference's LM_DISKANN query hot path, written
 * from a reading of the reference sources; every function cites the file:line it follows
 * (paths relative to /root/reference/src/lm_diskann/).
 *
 * PARITY PINNING. The reference ships no golden vectors, known-answer tests or fixtures for this
 * path, so this port is pinned against the reference's OWN sources compiled
 * here (oracle/_ref/liblmd_ref_{literal,fixed}.so, built by oracle/Makefile from the files where they
 * lie under /root/reference):
 *   - layout / distance / convert / ternary encode+dot / approx distance: against the untouched
 *     reference functions, bit-exact (tests/test_oracle.py);
 *   - search, routing mode LMDO_ROUTE_EXACT: against Searcher.cpp + patches/searcher_fixed.patch
 *     (H1-H4, the "ORACLE_FIXED" of SURVEY.md section 8c), ids/distances/counters bit-exact;
 *     frozen outputs of that build are committed under tests/golden/.
 *   - search, routing mode LMDO_ROUTE_TERNARY: the reference never executes this mode
 *     (Searcher.cpp:168-171 is a TODO; nothing writes planes, GraphManager.cpp:402-444 throws), so
 *     its CONTROL FLOW IS "parity unpinned"; only its arithmetic pieces are pinned (above).
 */
#ifndef LMD_ORACLE_H
#define LMD_ORACLE_H
#include <stddef.h>
#include <stdint.h>
#ifdef __cplusplus
extern "C" {
#endif

/* common/ann.hpp:16-32 */
enum { LMDO_METRIC_L2 = 1, LMDO_METRIC_COSINE = 2, LMDO_METRIC_IP = 3 };
enum { LMDO_VEC_FLOAT32 = 1, LMDO_VEC_INT8 = 2 };
enum { LMDO_ROUTE_EXACT = 0, LMDO_ROUTE_TERNARY = 1, LMDO_ROUTE_INT8 = 2 };
/* int8 edge codec grid for LMDO_ROUTE_INT8 (row f4, parity unpinned: no reference code executes it) */
void lmdo_set_codec(const float *off, float scale);

#define LMDO_EMPTY_ID INT64_MAX /* GraphManager.cpp:154-156 */

/* core/index_config.hpp:78-87 */
typedef struct {
	uint64_t count_off, vector_off, ids_off, pos_off, neg_off, total_node_size, ternary_edge_bytes;
} lmdo_layout;

int lmdo_layout_calc(uint64_t dims, int vec_type, uint32_t r, lmdo_layout *out);
uint64_t lmdo_block_bytes(uint64_t total_node_size);
float lmdo_distance(const float *a, const float *b, uint64_t dims, int metric, int *err);
int lmdo_convert_raw(const uint8_t *raw, float *out, uint64_t dims, int vec_type);
uint64_t lmdo_words_per_plane(uint64_t dims);
void lmdo_encode_ternary(const float *v, uint64_t dims, uint64_t *pos, uint64_t *neg);
int64_t lmdo_ternary_dot(const uint64_t *qp, const uint64_t *qn, const uint64_t *vp, const uint64_t *vn,
                         uint64_t words);
float lmdo_approx_distance(const float *q, const uint8_t *compressed, uint64_t dims, int metric, int *err);

/* Same contract as lmdref_search in ref_harness.cpp, plus `route`.
 * out_ids [nq*k] (-1 padded), out_dist [nq*k] exact distance of each returned id, out_count [nq],
 * out_stats [nq*2] = {expansions, vector fetches}. nthreads <= 0: all host threads. */
int lmdo_search(const uint8_t *blocks, uint64_t n, uint64_t block_bytes, const int64_t *row_ids, int64_t entry_row_id,
                int metric, int vec_type, uint64_t dims, uint32_t r, uint32_t cfg_l_search, const float *queries,
                uint64_t nq, uint64_t k, uint64_t l_param, int route, int64_t *out_ids, float *out_dist,
                uint32_t *out_count, uint64_t *out_stats, int nthreads);

/* Filtered search (the dual-heap design of the reference's "Consolidated Diskann Proposal
 * v1", which no reference code implements: PARITY UNPINNED). allowed [n] bytes, one per block position (non-zero =
 * the row passes the filter), NULL = lmdo_search. The traversal is the unfiltered one; results come from a second
 * bounded heap that only admits allowed rows. */
int lmdo_search_filtered(const uint8_t *blocks, uint64_t n, uint64_t block_bytes, const int64_t *row_ids,
                         int64_t entry_row_id, int metric, int vec_type, uint64_t dims, uint32_t r,
                         uint32_t cfg_l_search, const float *queries, uint64_t nq, uint64_t k, uint64_t l_param,
                         int route, const uint8_t *allowed, int64_t *out_ids, float *out_dist, uint32_t *out_count,
                         uint64_t *out_stats, int nthreads);

/* Graph write side, same contracts as lmdref_build_serial / lmdref_robust_prune in
 * ref_build_harness.cpp. core/GraphManager.cpp:65-203 (AddNode) and :259-457 (RobustPrune). */
int lmdo_build_serial(const float *vectors, uint64_t n, uint64_t dims, int metric, int vec_type, uint32_t r,
                      uint32_t l_insert, float alpha, const int64_t *row_ids, uint64_t block_bytes, uint8_t *out_blocks,
                      int64_t *entry_row_out);
int lmdo_robust_prune(const float *vectors, uint64_t n, uint64_t dims, int metric, int vec_type, uint32_t r, float alpha,
                      const int64_t *row_ids, uint64_t m, const uint64_t *node_idx, const int64_t *cands, uint64_t c,
                      int64_t *out_ids, uint32_t *out_count);

#ifdef __cplusplus
}
#endif
#endif
