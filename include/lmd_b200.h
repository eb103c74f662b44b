/* lmd_b200.h - C ABI of the B200-native LM_DISKANN query path.
 *
 * This is the whole drop-in boundary: plain C types, caller-owned host buffers, library-owned device
 * memory, integer status codes + lmd_last_error(); no exception crosses it and no torch / DuckDB type
 * appears in it. The reference has no FFI of its own (it is one C++ process); each
 * entry point below names the reference interface it replaces (paths relative to
 * /root/reference/src/lm_diskann/). INTEGRATION.md shows the reference-side adapter
 * (GpuSearcher : core::ISearcher, DeviceNodeStore fed from core::IStorageManager) that binds them.
 *
 * There is NO CPU fallback: every entry point that computes needs a CUDA device (sm_100a build) and
 * returns LMD_ERR_CUDA otherwise.
 */
#ifndef LMD_B200_H
#define LMD_B200_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define LMD_ABI_VERSION 3

typedef int32_t lmd_status;
enum {
	LMD_OK = 0,
	LMD_ERR_INVALID_ARG = 1, /* reference: InvalidInputException / BinderException */
	LMD_ERR_CUDA = 2,        /* reference: IOException (Scan rewraps, db/LmDiskannIndex.cpp:718-721) */
	LMD_ERR_OOM = 3,
	LMD_ERR_STATE = 4,       /* reference: std::runtime_error "index not loaded" (core/Coordinator.cpp:66-74) */
	LMD_ERR_UNSUPPORTED = 5  /* reference: NotImplementedException / the L2 ternary throw (core/distance.hpp:236-239) */
};

/* common/ann.hpp:16-32 - same numeric values as the reference enums */
enum { LMD_METRIC_L2 = 1, LMD_METRIC_COSINE = 2, LMD_METRIC_IP = 3 };
enum { LMD_VEC_FLOAT32 = 1, LMD_VEC_INT8 = 2 };

/* Routing distance used while traversing.
 * EXACT   = what core/Searcher.cpp:145-173 executes today: exact fp32 distance to each neighbour's own vector.
 * TERNARY = the designed LM-DiskANN mode (Searcher.cpp:168-171 TODO): ternary planes of the neighbours held in
 *           the EXPANDED block (core/distance.hpp:212-243), exact rerank afterwards. COSINE / IP only.
 * INT8    = the L2-capable edge codec the libSQL original declares (libsql/lm_diskann.hpp:55-62 LmDiskannEdgeType INT8;
 *           routing loop libsql/lm_diskann.cpp:1596-1625): every neighbour's vector as int8 codes on one global grid
 *           (per-dimension offset, one scale), held in a device sidecar shaped like the block's plane region; routing
 *           distance = || q^ - x^ || by integer dot products, exact rerank afterwards. L2 only. No reference code
 *           executes it: PARITY UNPINNED (checked against oracle/lmd_oracle.c's restatement only). */
enum { LMD_ROUTE_EXACT = 0, LMD_ROUTE_TERNARY = 1, LMD_ROUTE_INT8 = 2 };
enum { LMD_CODEC_NONE = 0, LMD_CODEC_INT8 = 1 };

/* lmd_search_batch flags */
enum {
	LMD_SEARCH_SEPARATE_RERANK = 1u << 0, /* run the rerank kernel even when traversal distances are already exact */
	LMD_SEARCH_PER_QUERY_STATS = 1u << 1, /* fill out_qstats */
	LMD_SEARCH_ASYNC = 1u << 2,           /* device-resident or pinned buffers only: enqueue on the store's stream and
	                                         return without synchronising; `stats` must be NULL */
	LMD_SEARCH_STRICT_FP = 1u << 3,       /* parity tool, never the timed path: every exact distance is accumulated in the
	                                         reference's own order - one lane per row, element by element, multiply then
	                                         add, no FMA (core/distance.hpp:56-99) - so that the traversal takes the
	                                         reference's decisions bit for bit (ids, distances, expansion counts). Also
	                                         lmd_set_tuning(s, "strict_fp", 1) for lmd_robust_prune_batch / lmd_build_graph */
	LMD_SEARCH_NO_PREFILTER = 1u << 4,    /* exact routing on a store with the int8 codec enabled: fetch every new
	                                         neighbour's row (the plain traversal) instead of letting the int8 lower
	                                         bound skip rows that cannot enter the list. Same results either way. */
	LMD_SEARCH_TENSOR_RERANK = 1u << 5    /* run the separate rerank as the tcgen05 (3xTF32) contraction staged through
	                                         bulk async copies instead of the FMA gather kernel (implies SEPARATE_RERANK) */
};

#define LMD_EMPTY_ROW_ID INT64_MAX /* empty neighbour slot, core/GraphManager.cpp:154-156,451-453 */

/* core/index_config.hpp:78-87 (NodeLayoutOffsets) */
typedef struct {
	uint64_t count_off;          /* uint16 neighbour count */
	uint64_t vector_off;         /* node vector, dims * sizeof(T) */
	uint64_t ids_off;            /* int64 row ids [r] */
	uint64_t pos_off;            /* positive planes [r][words] u64 */
	uint64_t neg_off;            /* negative planes [r][words] u64 */
	uint64_t total_node_size;    /* live bytes, before sector alignment */
	uint64_t ternary_edge_bytes; /* 2 * words * 8 */
} lmd_layout;

/* Subset of core/index_config.hpp:49-69 (LmDiskannConfig) that the query path reads. */
typedef struct {
	uint32_t dims;        /* config.dimensions */
	uint32_t r;           /* config.r */
	uint32_t metric;      /* LMD_METRIC_* */
	uint32_t vec_type;    /* LMD_VEC_* (config.node_vector_type) */
	uint32_t l_search;    /* config.l_search, used when a search passes L = 0 */
	uint32_t reserved;    /* must be 0 */
	uint64_t block_bytes; /* 0 = AlignValue<4096>(total_node_size) as db/LmDiskannIndex.cpp:161-162 */
} lmd_config;

typedef struct {
	uint64_t queries;
	uint64_t expansions;      /* E: node blocks expanded (GetNeighbors calls) */
	uint64_t distance_evals;  /* V: routing distances evaluated, entry included */
	uint64_t reranked;        /* C: exact distances computed by the rerank kernel */
	uint64_t filter_evictions;/* visited-filter entries overwritten (costs recomputation, never changes results) */
	uint64_t tie_overflows;   /* tied frontier entries dropped because the tie region was full (0 = oracle-exact ties) */
	uint64_t merge_fallbacks; /* candidate batches replayed one by one because of a distance tie (statistics) */
	uint64_t algorithmic_bytes; /* SURVEY.md section 8d formula for this call */
	uint64_t exact_fetches;   /* node vectors actually gathered by the traversal (== distance_evals unless the int8
	                             pre-filter skipped rows; ternary / int8 routing: the entry point only) */
	float traverse_ms;        /* CUDA-event time of the traversal kernel */
	float rerank_ms;          /* CUDA-event time of the rerank kernel (0 when fused away) */
	float h2d_ms, d2h_ms;     /* copies done by the call (0 for device-resident buffers) */
	float total_ms;           /* CUDA-event time of everything the call enqueued */
	uint32_t launches;        /* kernels launched by the call */
	uint32_t l_used;          /* L after Searcher.cpp:261-265 resolution */
} lmd_stats;

typedef struct lmd_store lmd_store;
typedef struct lmd_ctx lmd_ctx;     /* a per-caller search context, see lmd_ctx_create */
typedef struct lmd_group lmd_group; /* several stores on several devices searched as one, see lmd_group_create */

/* --- layout (core/index_config.cpp:104-148 CalculateLayoutInternal; host code, no device needed) --- */
lmd_status lmd_layout_calc(uint64_t dims, uint32_t vec_type, uint32_t r, lmd_layout *out);
uint64_t lmd_block_bytes(uint64_t total_node_size);

/* --- device node store: replaces core/IStorageManager.hpp:156 GetNodeBlockData + the row_id -> IndexPointer
 *     map of core/GraphManager.hpp:122 / GraphManager.cpp:523-531 on the query path --- */
lmd_status lmd_store_create(const lmd_config *cfg, int device, uint64_t capacity_nodes, lmd_store **out);
lmd_status lmd_store_destroy(lmd_store *s);
/* Appends n node blocks (reference layout, block_bytes apart). row_ids == NULL: row id = slot number.
 * `blocks` / `row_ids` may be pageable or pinned host memory, or device memory on the store's device. */
lmd_status lmd_store_upload(lmd_store *s, const int64_t *row_ids, const void *blocks, uint64_t n);
/* Re-uploads one block that the host marked dirty (core/IStorageManager.hpp MarkBlockDirty). */
lmd_status lmd_store_update(lmd_store *s, int64_t row_id, const void *block);
/* Live mirror. Batch form of what the reference's graph code does to storage between two
 * searches - AllocateNodeBlock for new rows, GetMutableNodeBlockData + MarkBlockDirty for changed ones
 * (core/IStorageManager.hpp:142-177; callers core/GraphManager.cpp:81-160,430-444): the host passes every block it
 * created or dirtied. Known row ids overwrite their mirrored block, unknown ones are appended in batch order
 * (*out_appended, may be NULL); a row id given twice keeps its last block. row_ids: host memory; blocks: host
 * (pageable / pinned) or device memory. Only the views of the touched blocks are rebuilt unless a row id was new. */
lmd_status lmd_store_apply(lmd_store *s, const int64_t *row_ids, const void *blocks, uint64_t n,
                           uint64_t *out_appended);
/* core/GraphManager.cpp:476-499 HandleNodeDeletion + FreeNode: the row ids leave the row-id map. Neighbour entries
 * naming them are skipped from the next search on (core/Searcher.cpp:120-122,146-156); removing the entry point
 * leaves the store without one (empty results, Searcher.cpp:74-77) until lmd_store_set_entry. Unknown row ids are
 * ignored; *out_removed (may be NULL) = how many were present. The slots are not reused. row_ids: host memory. */
lmd_status lmd_store_remove(lmd_store *s, const int64_t *row_ids, uint64_t n, uint64_t *out_removed);
/* core/GraphManager.cpp:468-474 GetEntryPointRowId/Pointer. Unknown row id => searches return empty
 * results (Searcher.cpp:74-77). */
lmd_status lmd_store_set_entry(lmd_store *s, int64_t row_id);
/* Builds the derived search views (packed vectors, uint32 neighbour-slot sidecar). Called implicitly by the
 * first search after an upload/update. */
lmd_status lmd_store_finalize(lmd_store *s);
/* Reads one mirrored block back (verification / tests). */
lmd_status lmd_store_download_block(lmd_store *s, int64_t row_id, void *out_block);
/* Runs all later work of this store on the caller's CUDA stream (a cudaStream_t passed as void*; NULL = back to
 * the store's own stream). Lets a host with its own streams / events order and time the searches. */
lmd_status lmd_store_set_stream(lmd_store *s, void *cuda_stream);
/* int8 edge codec (row f4). Builds (and from then on keeps up to date, lazily, before the next search that needs it)
 * the code sidecar: 12 R + R * dims bytes per node. With it, LMD_ROUTE_INT8 searches work (L2 stores), and
 * LMD_ROUTE_EXACT searches (any metric) use the codes as a LOWER BOUND that decides which neighbours' vectors
 * are fetched at all - results, distances and expansion counts stay bit-identical to the plain traversal (see
 * LMD_SEARCH_NO_PREFILTER). L2: codes of x on a per-dimension-centred grid, bound from the grid distance; COSINE: codes
 * of x / || x || on a zero-offset grid, IP: codes of x on a zero-offset grid with the query scaled to the rows' RMS
 * length - bound from the grid dot product. */
lmd_status lmd_store_enable_codec(lmd_store *s, uint32_t codec);
/* The codec's grid: offsets [dims] (host buffer) and the scale; x ~ offsets[d] + scale * code[d]. LMD_ERR_STATE while no
 * codec is built. */
lmd_status lmd_store_get_codec(lmd_store *s, float *out_offsets, float *out_scale);
/* Frees the raw block mirror of a store that will only be searched (the search kernels read the derived views and
 * sidecars, never the blocks). Builds every view first. Afterwards upload / update / apply / download / build return
 * LMD_ERR_STATE; lmd_store_device_bytes reflects the release. */
lmd_status lmd_store_release_blocks(lmd_store *s);
uint64_t lmd_store_node_count(const lmd_store *s);
uint64_t lmd_store_device_bytes(const lmd_store *s); /* core/ISearcher.hpp:43 GetInMemorySize */

/* --- search: replaces core/ISearcher.hpp:30-34 Search -> core/Searcher.cpp:256-272 -> PerformSearch :45-246,
 *     batched (ISearcher.hpp:36-37 anticipates it) ---
 * queries  [nq][dims] fp32: device memory, pinned host memory (read in place over PCIe - no staging copy) or pageable
 *          host memory (staged through the store's device scratch).
 * k, L     as Searcher::Search: L == 0 -> cfg.l_search, then (k > 20 ? k + 20 : 40) if still 0, then max(L, k).
 * out_ids  [nq][k] row ids ascending by (distance, id), padded with -1; out_dist [nq][k] exact distances (may be
 *          NULL); out_count [nq] (may be NULL). Device, pinned host (written in place) or pageable host memory; the
 *          outputs must all be pageable or all be device / pinned.
 * out_qstats [nq][2] = {expansions, distance evals} when LMD_SEARCH_PER_QUERY_STATS (may be NULL).
 * Thread-safety: lmd_search_batch calls on one store are serialised on the store's own context; use one lmd_ctx per
 * host thread for searches that run side by side. Searches hold the store shared, uploads / apply / remove / build
 * exclusively, so either may be called at any time.
 * Stream ordering: a call reads `queries` on the store's (or context's) stream. Device-resident queries produced on
 * ANOTHER stream must be complete (or that stream bound with lmd_store_set_stream / lmd_ctx_set_stream) before the call.
 * Tie order: equal distances are ordered by slot (= arrival order of the row in the store; = row id while row ids are
 * slot numbers), which is the (distance, id) order of oracle hunk H4 for identity row ids only. */
lmd_status lmd_search_batch(lmd_store *s, const float *queries, uint64_t nq, uint32_t k, uint32_t L, uint32_t route,
                            uint32_t flags, int64_t *out_ids, float *out_dist, uint32_t *out_count,
                            uint32_t *out_qstats, lmd_stats *stats);

/* --- per-caller search contexts: the reference's Search takes no lock (db/LmDiskannIndex.cpp:677-724 ->
 *     core/Coordinator.cpp:63-102) and DuckDB runs scans from many threads. A context owns a stream, statistics
 *     counters, staging scratch and a small pinned block on the store's device; searches on different contexts of one
 *     store run concurrently on the GPU. `filter` may be NULL. Same arguments and semantics as lmd_search_batch. --- */
lmd_status lmd_ctx_create(lmd_store *s, lmd_ctx **out);
lmd_status lmd_ctx_destroy(lmd_ctx *c);
lmd_status lmd_ctx_set_stream(lmd_ctx *c, void *cuda_stream);
struct lmd_filter;
lmd_status lmd_ctx_search_batch(lmd_ctx *c, const struct lmd_filter *filter, const float *queries, uint64_t nq,
                                uint32_t k, uint32_t L, uint32_t route, uint32_t flags, int64_t *out_ids,
                                float *out_dist, uint32_t *out_count, uint32_t *out_qstats, lmd_stats *stats);

/* --- filtered search ---
 * The reference's plan ("Consolidated Diskann Proposal v1", section "Filtered Searches"; stubs in
 * libsql/lm_diskann.hpp) hands the index an `allowed_ids` set and searches with two heaps: exploration over ALL rows,
 * results over allowed rows only. No reference code implements it (parity is against oracle/lmd_oracle.c's
 * restatement of that design only). lmd_filter_create resolves the host's allowed row ids (host or device memory;
 * unknown ids ignored, duplicates harmless) to a bitmap over the store's slots; the filter stays valid until rows are
 * added to or removed from the store (LMD_ERR_STATE afterwards). lmd_search_batch_filtered is lmd_search_batch with
 * that filter: same traversal, results (and rerank candidates) restricted to allowed rows, out_count[q] < k when
 * fewer than k allowed rows were met - the host then warns or widens L, as the plan says. */
typedef struct lmd_filter lmd_filter;
lmd_status lmd_filter_create(lmd_store *s, const int64_t *allowed_row_ids, uint64_t n, lmd_filter **out);
lmd_status lmd_filter_destroy(lmd_filter *f);
uint64_t lmd_filter_count(const lmd_filter *f); /* distinct allowed row ids present in the store */
lmd_status lmd_search_batch_filtered(lmd_store *s, const lmd_filter *filter, const float *queries, uint64_t nq,
                                     uint32_t k, uint32_t L, uint32_t route, uint32_t flags, int64_t *out_ids,
                                     float *out_dist, uint32_t *out_count, uint32_t *out_qstats, lmd_stats *stats);

/* --- dataset sharding: k-way merge of G gathered per-shard results.
 * ids [G][nq][k], dist [G][nq][k] (device or host, as produced by lmd_search_batch on each shard and
 * all-gathered by the caller) -> out [nq][k] smallest by (distance, id); -1 ids are padding. */
lmd_status lmd_merge_topk(int device, const int64_t *ids, const float *dist, uint32_t shards, uint64_t nq, uint32_t k,
                          int64_t *out_ids, float *out_dist);

/* The same merge enqueued on the caller's stream (a cudaStream_t passed as void*, NULL = the default stream) without
 * synchronising: device buffers only. This is the form the sharded searcher uses - search (LMD_SEARCH_ASYNC on a store
 * bound to that stream with lmd_store_set_stream), all-gather and merge then run back to back on ONE stream with no host
 * round trip in between. */
lmd_status lmd_merge_topk_async(int device, const int64_t *ids, const float *dist, uint32_t shards, uint64_t nq,
                                uint32_t k, int64_t *out_ids, float *out_dist, void *cuda_stream);

/* --- device groups: ONE process, several devices. The reference's host is a single multi-threaded
 *     process and core/ISearcher.hpp:30-37 is the seam the searcher sits behind, so both multi-GPU modes are reachable
 *     through the C ABI without one process per GPU:
 *       LMD_GROUP_QUERY_SHARDED    every store holds the whole index (replicas on different devices); member g searches
 *                                  queries [lo_g, hi_g) (contiguous, balanced) into that slice of the caller's buffers.
 *       LMD_GROUP_DATASET_SHARDED  store s holds shard s (shard-local graph and entry point, global row ids); every
 *                                  member searches the full batch, then one kernel on the first member's device reads
 *                                  the G per-shard [nq][k] lists in place (NVLink peer loads) and keeps the k smallest by
 *                                  (distance, row id): the all-gather and the merge of the multi-process path
 *                                  (duckdb-lm-diskann_b200/sharded.py, NCCL) fused into one kernel.
 *     The stores stay the caller's (create / upload / finalize them first; destroy them after the group). queries,
 *     out_ids, out_dist (may be NULL), out_count (may be NULL): host (pinned or pageable) or device memory. The call
 *     returns when the results are in the caller's buffers. --- */
enum { LMD_GROUP_QUERY_SHARDED = 0, LMD_GROUP_DATASET_SHARDED = 1 };
typedef struct {
	uint32_t members, mode;
	uint32_t peer_access;     /* bit i: the merge read member i's lists in place (same device or peer loads) */
	uint32_t reserved;
	uint64_t queries[16];     /* queries searched by member i */
	float search_ms[16];      /* CUDA-event time of queries-in + search on member i's stream */
	float search_ms_max;
	float merge_ms;           /* dataset sharding: the gather + merge kernel */
	float wall_ms;            /* host clock around the whole call */
	uint32_t reserved2;
	uint64_t merge_input_bytes; /* dataset sharding: G * nq * k * 12 */
} lmd_group_stats;
lmd_status lmd_group_create(lmd_store *const *stores, uint32_t n, uint32_t mode, lmd_group **out);
lmd_status lmd_group_destroy(lmd_group *g);
uint32_t lmd_group_size(const lmd_group *g);
lmd_status lmd_group_search_batch(lmd_group *g, const float *queries, uint64_t nq, uint32_t k, uint32_t L, uint32_t route,
                                  uint32_t flags, int64_t *out_ids, float *out_dist, uint32_t *out_count,
                                  lmd_group_stats *stats);

/* --- graph write side: replaces core/GraphManager.cpp:65-203 (AddNode, one row at a
 *     time) and :259-457 (RobustPrune) by a batch-parallel build on the device. Writes the same node blocks, including
 *     the ternary planes of every neighbour, which nothing in the reference writes (GraphManager.cpp:402-444). --- */
typedef struct {
	uint32_t l_insert;  /* config.l_insert (core/index_config.hpp:33-38 default 128): candidate list per inserted node */
	float alpha;        /* config.alpha (default 1.2) */
	uint32_t max_batch; /* nodes inserted per round, 0 = automatic; a round sees the graph of all earlier rounds */
	uint32_t reserved;  /* must be 0 */
} lmd_build_params;

typedef struct {
	uint64_t nodes, rounds;
	uint64_t search_expansions, search_evals; /* candidate searches (traversal kernel) */
	uint64_t prune_evals;                     /* distances computed by the prune kernel */
	uint64_t backlink_prunes;                 /* neighbour lists re-pruned because a new node chose them */
	uint64_t inbox_overflows;                 /* back-link candidates dropped (more than 2R new nodes chose one node in a round) */
	float search_ms, prune_ms, backlink_ms, finalize_ms, total_ms; /* CUDA-event times */
	uint32_t launches;
	uint32_t reserved;
} lmd_build_stats;

/* Appends n nodes given only their vectors (fp32 [n][dims]; stored as the store's vec_type exactly as AddNode stores
 * them, GraphManager.cpp:96-108: memcpy or a plain cast to int8). Blocks get neighbour count 0. */
lmd_status lmd_store_upload_vectors(lmd_store *s, const int64_t *row_ids, const float *vectors, uint64_t n);
/* (Re)builds the graph over every node in the store; the first node is the entry point (GraphManager.cpp:119-121).
 * Rewrites count / ids / planes of every mirrored block and the search views. stats may be NULL.
 * LMD_ERR_UNSUPPORTED on a store that had rows removed (their slots still hold blocks). */
lmd_status lmd_build_graph(lmd_store *s, const lmd_build_params *params, lmd_build_stats *stats);
/* RobustPrune of m nodes in one launch. node_row_ids [m]; cand_row_ids [m][c] (INT64_MAX = no candidate; unknown row
 * ids, duplicates and the node itself are skipped as the reference skips them); out_ids [m][r] INT64_MAX padded;
 * out_count [m]. Host buffers. Does not modify the store. */
lmd_status lmd_robust_prune_batch(lmd_store *s, const int64_t *node_row_ids, const int64_t *cand_row_ids, uint64_t m,
                                  uint32_t c, float alpha, int64_t *out_ids, uint32_t *out_count);
/* Reads n mirrored blocks starting at slot `first_slot` back (host or device buffer): how a built graph reaches the
 * host's storage (IStorageManager::AllocateNodeBlock + memcpy). */
lmd_status lmd_store_download_blocks(lmd_store *s, uint64_t first_slot, uint64_t n, void *out_blocks);
/* The row ids of the same slots (host or device buffer): a block does not carry its own id, so whoever hands downloaded
 * blocks back to the host's storage pairs them through this (core/GraphManager.hpp:122 is the reference's row id map). */
lmd_status lmd_store_download_row_ids(lmd_store *s, uint64_t first_slot, uint64_t n, int64_t *out_row_ids);

/* --- searching without allocating. NO COUNTERPART IN THE REFERENCE: added for
 *     callers that cannot call cudaMalloc (code under CUDA graph capture, tight serving loops, harnesses that forbid it).
 * The search of lmd_search_batch (same kernels, same results) with the kernels' scratch - the work counter / statistics
 * block and, for the ternary and int8 routes, the nq * L candidate keys handed to the rerank kernel - in ONE file-scope
 * __device__ array of LMD_STATIC_SCRATCH_BYTES that every device loading the library holds until the process exits. The
 * call allocates nothing: no cudaMalloc / cudaFree / cudaHostAlloc, no allocating library routine.
 *   - queries [nq][dims], out_ids / out_dist [nq][k], out_count [nq]: device memory or pinned (mapped) host memory, all
 *     four required; pageable memory is refused (staging it would allocate);
 *   - limit: 256 + nq * L * 8 <= LMD_STATIC_SCRATCH_BYTES for the ternary / int8 routes (16 383 queries at L = 64); the
 *     exact route uses the 256-byte counter block only; beyond the limit -> LMD_ERR_INVALID_ARG with the sizes in
 *     lmd_last_error();
 *   - the store must already be prepared for the route - one earlier lmd_search_batch on that route (any batch size)
 *     after the last change to the store does it: building a view or sidecar allocates, so an unprepared route ->
 *     LMD_ERR_STATE;
 *   - cuda_stream: the cudaStream_t to enqueue on (NULL = the default stream); flags: 0, or LMD_SEARCH_ASYNC to return
 *     without synchronising;
 *   - ALL calls on a device share the array and it keeps the last call's contents (the counter block is cleared on the
 *     stream before each launch; the candidate region is rewritten before it is read): two calls that overlap on the GPU
 *     RACE - serialise them (one stream, or events between streams). */
#define LMD_STATIC_SCRATCH_BYTES (8u << 20)
uint64_t lmd_static_scratch_bytes(void);
lmd_status lmd_search_batch_static(lmd_store *s, const float *queries, uint64_t nq, uint32_t k, uint32_t L, uint32_t route,
                                   uint32_t flags, int64_t *out_ids, float *out_dist, uint32_t *out_count,
                                   void *cuda_stream);

/* --- misc --- */
const char *lmd_last_error(void);   /* thread-local message of the last non-OK status */
uint32_t lmd_abi_version(void);
/* Tuning knobs for experiments (not needed for correct use): "filter_log2", "filter_mode" (0 u32 open addressing, 1 two-
 * choice 16-bit buckets, 2 direct-mapped 16-bit, 3 direct-mapped by slot modulo "filter_entries"), "warps_per_cta",
 * "ctas_per_sm", "tie_slots", "merge_min", "prefetch_planes", "latency_mode" (-1 auto by batch size / 0 the throughput
 * kernels / 1 the small-batch kernels that gather all neighbours of an expansion in one pass / 2 the mid-size-batch
 * kernels with 4 rows in flight per team; 128-d and 96-d rows, exact routing). */
lmd_status lmd_set_tuning(lmd_store *s, const char *key, int64_t value);

#ifdef __cplusplus
}
#endif
#endif /* LMD_B200_H */
