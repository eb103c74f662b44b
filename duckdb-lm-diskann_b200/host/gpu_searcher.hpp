// Host-side adapter of the B200 LM_DISKANN query path, written against the reference's OWN interface headers
// (paths relative to /root/reference/src/lm_diskann/):
//
//   GpuSearcher      : core::ISearcher              core/ISearcher.hpp:14-74   replaces core::Searcher (core/Searcher.cpp)
//   DeviceNodeStore                                  the HBM mirror behind core/IStorageManager.hpp:156 GetNodeBlockData
//                                                    and core/IGraphManager.hpp:86 TryGetNodePointer, fed THROUGH those
//                                                    two interfaces (so any storage backend the reference grows works)
//   MirroredStorage  : core::IStorageManager       decorator over the host's storage manager that keeps the HBM mirror
//                                                    live: AllocateNodeBlock / MarkBlockDirty
//                                                    (core/IStorageManager.hpp:142-177) and EnqueueDeletion (:115) are
//                                                    forwarded AND recorded; Flush() replays them on the device
//   GpuIndexScan / GpuScanState                      db/LmDiskannIndex.cpp:639-724 InitializeScan / Scan and
//                                                    db/LmDiskannScanState.hpp:34-60, minus the DuckDB Vector types
//   GpuCoordinatorSearch                             core/Coordinator.cpp:63-102 Search (guards + L clamp + rethrow)
//
// Everything below the adapter is the C ABI of include/lmd_b200.h; this file contains no CUDA and no arithmetic.
// Error behaviour follows the reference: std::runtime_error from the coordinator guards, BinderException for a bad
// query (dimension / k == 0), IOException wrapping anything the scan throws; a missing entry point gives an empty
// result (core/Searcher.cpp:74-83), not an error.
#pragma once

#include "core/IGraphManager.hpp"
#include "core/ISearcher.hpp"
#include "core/IStorageManager.hpp"
#include "core/index_config.hpp"

#include "lmd_b200.h"

#include <atomic>
#include <cstdint>
#include <condition_variable>
#include <memory>
#include <mutex>
#include <string>
#include <thread>
#include <unordered_map>
#include <unordered_set>
#include <vector>

namespace diskann {
namespace b200 {

// Thrown for any non-OK status of the C ABI that has no closer reference exception.
std::string LastError();

// ---------------------------------------------------------------------------------------------------------------
// DeviceNodeStore: RAII owner of one lmd_store.
// ---------------------------------------------------------------------------------------------------------------
class DeviceNodeStore {
public:
	// block_bytes = 0: AlignValue<SECTOR_SIZE>(total_node_size), db/LmDiskannIndex.cpp:161-162
	DeviceNodeStore(const core::LmDiskannConfig &config, common::idx_t capacity_nodes, int device = 0,
	                common::idx_t block_bytes = 0);
	~DeviceNodeStore();
	DeviceNodeStore(const DeviceNodeStore &) = delete;
	DeviceNodeStore &operator=(const DeviceNodeStore &) = delete;

	// Streams the blocks of `row_ids` to HBM: TryGetNodePointer -> GetNodeBlockData -> pinned-staged upload, in
	// batches of `batch_nodes` blocks. Row ids that do not resolve are skipped (the reference skips them at search
	// time, core/Searcher.cpp:120-122). Returns the number of blocks mirrored.
	common::idx_t MirrorFrom(const core::IStorageManager &storage, const core::IGraphManager &graph,
	                         const std::vector<common::row_t> &row_ids, common::idx_t batch_nodes = 4096);
	// Direct bulk load of n contiguous blocks (row_ids == nullptr: row id = position).
	void Upload(const common::row_t *row_ids, common::const_data_ptr_t blocks, common::idx_t n);
	// Batch upsert (lmd_store_apply): blocks [n][BlockBytes()] of row_ids [n]; returns how many were new.
	common::idx_t Apply(const common::row_t *row_ids, common::const_data_ptr_t blocks, common::idx_t n);
	// core/GraphManager.cpp:476-499 HandleNodeDeletion / FreeNode for a batch; returns how many were present.
	common::idx_t Remove(const common::row_t *row_ids, common::idx_t n);
	// core/IStorageManager.hpp:177 MarkBlockDirty: re-mirror one block after the host changed it.
	void RefreshBlock(const core::IStorageManager &storage, const core::IGraphManager &graph, common::row_t row_id);
	// core/GraphManager.cpp:468-474: entry point = whatever the graph manager reports now.
	void SyncEntryPoint(const core::IGraphManager &graph);
	void SetEntryPoint(common::row_t row_id);

	// ---- graph write side ----
	// Appends nodes from raw fp32 vectors (what GraphManager::AddNode stores, core/GraphManager.cpp:81-121).
	void UploadVectors(const common::row_t *row_ids, const float *vectors, common::idx_t n);
	// Batch-parallel replacement of the AddNode loop (core/GraphManager.cpp:65-203 + RobustPrune :259-457) over every
	// node in the store; config.l_insert / config.alpha / config.r as the reference reads them.
	lmd_build_stats BuildGraph(common::idx_t max_batch = 0);
	// Copies the built blocks (neighbour ids, counts, ternary planes) into the host's blocks:
	// TryGetNodePointer -> GetMutableNodeBlockData -> memcpy -> MarkBlockDirty, and moves the graph's entry point to
	// the first node as AddNode does (:119-121). `row_ids` in upload order. Returns the number of blocks written.
	common::idx_t WriteBack(core::IStorageManager &storage, core::IGraphManager &graph,
	                        const std::vector<common::row_t> &row_ids, common::idx_t batch_nodes = 4096);

	common::idx_t NodeCount() const;
	common::idx_t DeviceBytes() const;
	common::idx_t BlockBytes() const {
		return block_bytes_;
	}
	// int8 code sidecar (lmd_store_enable_codec): exact searches then fetch only the rows an int8 lower bound cannot
	// exclude - same ids and distances, a quarter to a third of the vector bytes. Costs R * dims bytes per node of device
	// memory.
	void SetEdgeCodec(bool on);
	lmd_store *Handle() const {
		return store_;
	}
	const core::LmDiskannConfig &Config() const {
		return config_;
	}

private:
	core::LmDiskannConfig config_;
	common::idx_t block_bytes_ = 0;
	lmd_store *store_ = nullptr;
};

// ---------------------------------------------------------------------------------------------------------------
// MirroredStorage: the live feeder. Wraps the storage manager the host already has (for DuckDB: the one whose
// GetNodeBlockData pins a FixedSizeAllocator segment through the BufferManager, libsql/lm_diskann.cpp:720-734) and is
// handed to GraphManager / Coordinator IN ITS PLACE. Every call is forwarded unchanged; the three calls that change what
// a search may see are also recorded:
//   AllocateNodeBlock(row_id, ptr, data)  -> (row_id, ptr) becomes known and dirty
//   MarkBlockDirty(ptr)                   -> ptr dirty
//   EnqueueDeletion(row_id, head)         -> row_id removed (Coordinator::Delete, core/Coordinator.cpp:186-212)
// Flush() reads the dirty blocks through the inner GetNodeBlockData and replays them with one lmd_store_apply +
// one lmd_store_remove. GpuSearcher calls it before a search when attached (AttachLiveMirror), so the reference's own
// insert loop - AddNode -> SearchForInitialCandidates -> RobustPrune -> MarkBlockDirty - runs against an up-to-date
// mirror without knowing about it.
// ---------------------------------------------------------------------------------------------------------------
class MirroredStorage : public core::IStorageManager {
public:
	MirroredStorage(core::IStorageManager *inner, std::shared_ptr<DeviceNodeStore> store);

	// Initial bulk mirror of `row_ids` (DeviceNodeStore::MirrorFrom) + remembers pointer -> row id for later dirty marks.
	common::idx_t MirrorAll(const core::IGraphManager &graph, const std::vector<common::row_t> &row_ids,
	                        common::idx_t batch_nodes = 4096);
	bool HasPending() const {
		return !dirty_.empty() || !deleted_.empty();
	}
	// Replays the recorded changes on the device. Returns the number of blocks sent.
	common::idx_t Flush();
	common::idx_t FlushCount() const {
		return flushes_;
	}
	common::idx_t BlocksSent() const {
		return blocks_sent_;
	}
	DeviceNodeStore &Store() {
		return *store_;
	}

	// ---- core::IStorageManager: forwarded ----
	void LoadIndexContents(const std::string &index_path, core::LmDiskannConfig &config_out,
	                       core::IGraphManager *graph_manager_out, common::IndexPointer &entry_point_ptr_out,
	                       common::row_t &entry_point_rowid_out, common::IndexPointer &delete_queue_head_out) override;
	void InitializeNewStorage(const std::string &index_path, const core::LmDiskannConfig &config) override;
	void SaveIndexContents(const std::string &index_path, const core::LmDiskannConfig &config,
	                       const core::IGraphManager *graph_manager, common::IndexPointer entry_point_ptr,
	                       common::row_t entry_point_rowid, common::IndexPointer delete_queue_head) override;
	common::idx_t GetInMemorySize() const override;
	void ProcessDeletionQueue(common::IndexPointer &delete_queue_head_ptr) override;
	common::const_data_ptr_t GetNodeBlockData(common::IndexPointer node_ptr) const override;
	common::data_ptr_t GetMutableNodeBlockData(common::IndexPointer node_ptr) override;
	// ---- forwarded and recorded ----
	void EnqueueDeletion(common::row_t row_id, common::IndexPointer &delete_queue_head_ptr) override;
	bool AllocateNodeBlock(common::row_t row_id, common::IndexPointer &node_ptr_out,
	                       common::data_ptr_t &node_data_out) override;
	void MarkBlockDirty(common::IndexPointer node_ptr) override;

private:
	core::IStorageManager *inner_;
	std::shared_ptr<DeviceNodeStore> store_;
	std::unordered_map<uint64_t, common::row_t> ptr_to_row_; // IndexPointer::Get() -> row id
	std::unordered_map<common::row_t, uint64_t> row_to_ptr_;
	std::vector<uint64_t> dirty_order_;                      // first-dirtied order (appends keep insertion order)
	std::unordered_set<uint64_t> dirty_;
	std::vector<common::row_t> deleted_;
	std::vector<uint8_t> stage_;
	std::vector<common::row_t> stage_ids_;
	common::idx_t flushes_ = 0, blocks_sent_ = 0;
};

// ---------------------------------------------------------------------------------------------------------------
// AllowedIds: the `allowed_ids` set of a filtered scan (the reference's plan passes it as
// `struct ScanOptions { SelectionVector *allowed_ids; idx_t count; }`), resolved once to the device store's slots.
// Valid until rows are added to / removed from the store.
// ---------------------------------------------------------------------------------------------------------------
class AllowedIds {
public:
	AllowedIds(DeviceNodeStore &store, const common::row_t *row_ids, common::idx_t count);
	~AllowedIds();
	AllowedIds(const AllowedIds &) = delete;
	AllowedIds &operator=(const AllowedIds &) = delete;
	// distinct allowed row ids that exist in the index
	common::idx_t Count() const;
	const lmd_filter *Handle() const {
		return filter_;
	}

private:
	lmd_filter *filter_ = nullptr;
};

// ---------------------------------------------------------------------------------------------------------------
// GpuSearcher: drop-in for core::Searcher behind core::ISearcher.
// ---------------------------------------------------------------------------------------------------------------
class GpuSearcher : public core::ISearcher {
public:
	enum class Route : uint32_t { EXACT = LMD_ROUTE_EXACT, TERNARY = LMD_ROUTE_TERNARY };

	explicit GpuSearcher(std::shared_ptr<DeviceNodeStore> store, Route route = Route::EXACT);
	~GpuSearcher() override;

	// core/ISearcher.hpp:30-34. `graph_manager` is used only to keep the device entry point equal to the graph's
	// (the traversal itself reads the mirror). One query = a batch of one.
	void Search(const float *query_vector, const core::LmDiskannConfig &config, core::IGraphManager *graph_manager,
	            common::idx_t k_neighbors, std::vector<common::row_t> &result_row_ids,
	            common::idx_t search_list_size = 0) override;

	// core/ISearcher.hpp:43
	common::idx_t GetInMemorySize() const override;

	// core/ISearcher.hpp:68-73. The reference's implementation (core/Searcher.cpp:275-294) calls PerformSearch with
	// final_pass = false and therefore returns nothing; this one returns the
	// num_candidates_to_find closest nodes, which is what GraphManager::AddNode expects to receive.
	void SearchForInitialCandidates(const float *new_node_vector, common::idx_t dimensions,
	                                const core::LmDiskannConfig &config, core::IGraphManager *graph_manager,
	                                common::idx_t num_candidates_to_find, common::IndexPointer current_entry_point,
	                                std::vector<common::row_t> &candidate_row_ids_out) override;

	// The batched entry core/ISearcher.hpp:36-37 anticipates. queries [nq][dims]; out_ids [nq][k] (-1 padded),
	// out_dist [nq][k] (may be null), out_count [nq] (may be null). Host or device pointers (all on one side).
	void SearchBatch(const float *queries, common::idx_t nq, common::idx_t k_neighbors, common::idx_t search_list_size,
	                 common::row_t *out_ids, float *out_dist, uint32_t *out_count, lmd_stats *stats = nullptr);

	// Filtered form (dual heap: exploration over all rows, results over `allowed` only). When `widen` is set the list
	// starts at FilteredListSize(L, allowed rows, all rows) and queries that still met fewer than k allowed rows are
	// searched again with L doubled, up to the build's limit of 512 - the plan's "L_search ... automatically increased
	// for filtered queries". Returns the number of queries that still
	// have fewer than k results (the host warns the user, as the plan says).
	common::idx_t SearchBatchFiltered(const AllowedIds &allowed, const float *queries, common::idx_t nq,
	                                  common::idx_t k_neighbors, common::idx_t search_list_size, common::row_t *out_ids,
	                                  float *out_dist, uint32_t *out_count, bool widen = true);

	// L widened by 1 / sqrt(allowed fraction), at most 8 x and at most 512.
	static uint32_t FilteredListSize(uint32_t L, common::idx_t allowed_rows, common::idx_t total_rows);

	DeviceNodeStore &Store() {
		return *store_;
	}
	// Pending changes of `mirror` are flushed to the device before every search (nullptr detaches).
	void AttachLiveMirror(MirroredStorage *mirror) {
		mirror_ = mirror;
	}
	// Concurrent Search() calls (DuckDB scans from many threads): up to MAX_LEADERS of them are on the device side by
	// side, each through its thread's own lmd_ctx; calls beyond that wait, and the next free leader carries all of them as
	// ONE batch. A lone caller pays nothing. false = every call goes to the device on its own context, however many.
	void SetCombining(bool on) {
		combine_ = on;
	}
	// Batches led so far and queries they carried (statistics: queries / batches = the combining factor)
	void CombiningStats(uint64_t &batches, uint64_t &queries) const {
		batches = comb_batches_.load();
		queries = comb_queries_.load();
	}

private:
	// The reference's Search takes no lock and DuckDB scans from many threads (db/LmDiskannIndex.cpp:677-724): every
	// host thread gets its own lmd_ctx (stream, counters, scratch), so concurrent Search calls run side by side on the
	// device instead of queueing on the store's default context.
	lmd_ctx *ThreadContext();

	std::shared_ptr<DeviceNodeStore> store_;
	MirroredStorage *mirror_ = nullptr;
	Route route_;
	std::mutex mu_; // entry-point sync, mirror flush, the context table
	std::unordered_map<std::thread::id, lmd_ctx *> contexts_;
	common::row_t synced_entry_ = 0;
	bool entry_synced_ = false;

	// flat combining of single-query searches
	struct PendingSearch {
		const float *query;
		common::idx_t k, L;
		std::vector<common::row_t> *out;
		bool taken = false, done = false; // taken: part of some leader's batch
		std::exception_ptr error;
	};
	static constexpr int MAX_LEADERS = 16;
	void LeadBatch(std::vector<PendingSearch *> &batch);
	bool combine_ = true;
	std::mutex comb_mu_;
	std::condition_variable comb_cv_;
	std::vector<PendingSearch *> comb_queue_;
	int comb_leaders_ = 0;
	std::atomic<uint64_t> comb_batches_{0}, comb_queries_{0};
};

// ---------------------------------------------------------------------------------------------------------------
// GpuSearcherGroup: the same seam over SEVERAL devices of this one process (lmd_group_*).
//   QUERY    every store is a replica of the whole index; a batch is split into contiguous slices, one per device.
//            A single Search() is one query and goes to ONE replica, chosen round-robin, through a per-thread context.
//   DATASET  store s is shard s (shard-local graph, its own entry point, global row ids); every device searches the
//            whole batch and one kernel merges the per-shard lists over NVLink peer loads. Search() is a batch of one.
// ---------------------------------------------------------------------------------------------------------------
class GpuSearcherGroup : public core::ISearcher {
public:
	enum class ShardMode : uint32_t { QUERY = LMD_GROUP_QUERY_SHARDED, DATASET = LMD_GROUP_DATASET_SHARDED };

	GpuSearcherGroup(std::vector<std::shared_ptr<DeviceNodeStore>> stores, ShardMode mode,
	                 GpuSearcher::Route route = GpuSearcher::Route::EXACT);
	~GpuSearcherGroup() override;
	// Replicas of one host index on `devices`: each mirrored through the reference's storage / graph interfaces.
	static std::vector<std::shared_ptr<DeviceNodeStore>> MirrorReplicas(const core::LmDiskannConfig &config,
	                                                                    const core::IStorageManager &storage,
	                                                                    const core::IGraphManager &graph,
	                                                                    const std::vector<common::row_t> &row_ids,
	                                                                    const std::vector<int> &devices,
	                                                                    common::idx_t block_bytes = 0);

	void Search(const float *query_vector, const core::LmDiskannConfig &config, core::IGraphManager *graph_manager,
	            common::idx_t k_neighbors, std::vector<common::row_t> &result_row_ids,
	            common::idx_t search_list_size = 0) override;
	common::idx_t GetInMemorySize() const override; // all devices together
	void SearchForInitialCandidates(const float *new_node_vector, common::idx_t dimensions,
	                                const core::LmDiskannConfig &config, core::IGraphManager *graph_manager,
	                                common::idx_t num_candidates_to_find, common::IndexPointer current_entry_point,
	                                std::vector<common::row_t> &candidate_row_ids_out) override;
	// queries [nq][dims] -> out_ids [nq][k] (-1 padded), out_dist / out_count may be null; host or device pointers
	void SearchBatch(const float *queries, common::idx_t nq, common::idx_t k_neighbors, common::idx_t search_list_size,
	                 common::row_t *out_ids, float *out_dist, uint32_t *out_count, lmd_group_stats *stats = nullptr);
	common::idx_t Members() const {
		return stores_.size();
	}
	ShardMode Mode() const {
		return mode_;
	}

private:
	std::vector<std::shared_ptr<DeviceNodeStore>> stores_;
	std::vector<std::unique_ptr<GpuSearcher>> singles_; // QUERY mode: one searcher per replica for point queries
	ShardMode mode_;
	GpuSearcher::Route route_;
	lmd_group *group_ = nullptr;
	std::atomic<uint64_t> next_{0};
};

// ---------------------------------------------------------------------------------------------------------------
// Callers, DuckDB types replaced by plain pointers.
// ---------------------------------------------------------------------------------------------------------------
// core/Coordinator.cpp:63-102
void GpuCoordinatorSearch(core::ISearcher *searcher, core::IGraphManager *graph_manager,
                          const core::LmDiskannConfig &config, bool index_loaded, const float *query_vector,
                          common::idx_t k_neighbors, std::vector<common::row_t> &result_row_ids,
                          common::idx_t search_list_size);

// db/LmDiskannScanState.hpp:34-60
struct GpuScanState {
	std::vector<float> query_vector_storage;
	const float *query_vector_ptr = nullptr;
	common::idx_t k = 0;
	common::idx_t l_search = 0;
	std::vector<common::row_t> result_row_ids;
	bool searched = false; // the reference re-runs the search on every Scan call (db/LmDiskannIndex.cpp:695); one
	                       // search per scan state returns the same rows for k <= STANDARD_VECTOR_SIZE and pages
	                       // correctly beyond it
};

class GpuIndexScan {
public:
	// DuckDB's STANDARD_VECTOR_SIZE (a macro in vector_size.hpp, 2048 in v1.2.x): used when its headers are present
#ifdef STANDARD_VECTOR_SIZE
	static constexpr common::idx_t VECTOR_SIZE = STANDARD_VECTOR_SIZE;
#else
	static constexpr common::idx_t VECTOR_SIZE = 2048;
#endif

	GpuIndexScan(core::ISearcher *searcher, core::IGraphManager *graph_manager, const core::LmDiskannConfig &config)
	    : searcher_(searcher), graph_(graph_manager), config_(config) {
	}
	// db/LmDiskannIndex.cpp:639-667: BinderException on dimension mismatch or k == 0
	std::unique_ptr<GpuScanState> InitializeScan(const float *query, common::idx_t query_dims, common::idx_t k) const;
	// db/LmDiskannIndex.cpp:677-724: up to STANDARD_VECTOR_SIZE row ids per call, 0 when drained; errors rewrapped
	// as IOException("Error during LM_DISKANN scan: ...")
	common::idx_t Scan(GpuScanState &state, common::row_t *result, common::idx_t result_capacity = VECTOR_SIZE);
	// Batch form of the template's index join (hnsw_optimize_join.cpp:113-170): nq scans with the same k in one
	// device launch; out [nq][k], -1 padded.
	void MultiScan(const float *queries, common::idx_t nq, common::idx_t query_dims, common::idx_t k,
	               common::row_t *out_ids, uint32_t *out_count);
	// MultiScan restricted to `allowed` (a WHERE clause on the same table, resolved by the host to row ids). Returns
	// the number of scans that produced fewer than k rows after widening.
	common::idx_t MultiScanFiltered(const AllowedIds &allowed, const float *queries, common::idx_t nq,
	                                common::idx_t query_dims, common::idx_t k, common::row_t *out_ids,
	                                uint32_t *out_count);

private:
	core::ISearcher *searcher_;
	core::IGraphManager *graph_;
	core::LmDiskannConfig config_;
};

} // namespace b200
} // namespace diskann
