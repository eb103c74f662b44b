// TEST CODE. C entry points (for ctypes) that drive the product's host adapter
// (duckdb-lm-diskann_b200/host/gpu_searcher.{hpp,cpp}) the way the reference drives core::Searcher: through
// core::ISearcher / core::IStorageManager / core::IGraphManager objects. The reference ships no working
// implementation of the latter two (StorageManager.cpp:294-348 throws, GraphManager needs AddNode), so this file
// provides small in-memory ones over a caller-supplied array of node blocks - the "mock for testing" that
// store/IShadowStorageService.hpp:13-16 anticipates. Built by duckdb-lm-diskann_b200/host/Makefile.
#include "gpu_searcher.hpp"

#include "core/GraphManager.hpp"
#include "core/GraphNode.hpp"

#include <algorithm>
#include <atomic>
#include <chrono>
#include <cstring>
#include <memory>
#include <mutex>
#include <string>
#include <thread>
#include <unordered_map>

using namespace diskann;

namespace {

thread_local std::string g_err;
thread_local int g_err_type = 0; // 0 none, 1 std::runtime_error, 2 Binder, 3 IO, 4 InvalidInput, 5 NotImplemented, 6 other

class MemStorage : public core::IStorageManager {
public:
	MemStorage(const uint8_t *blocks, uint64_t n, uint64_t bb) : data_(blocks, blocks + n * bb), n_(n), bb_(bb) {
	}
	common::const_data_ptr_t GetNodeBlockData(common::IndexPointer p) const override {
		uint64_t v = p.Get();
		return (v == 0 || v > n_) ? nullptr : data_.data() + (v - 1) * bb_;
	}
	common::data_ptr_t GetMutableNodeBlockData(common::IndexPointer p) override {
		uint64_t v = p.Get();
		return (v == 0 || v > n_) ? nullptr : data_.data() + (v - 1) * bb_;
	}
	void LoadIndexContents(const std::string &, core::LmDiskannConfig &, core::IGraphManager *, common::IndexPointer &,
	                       common::row_t &, common::IndexPointer &) override {
	}
	void InitializeNewStorage(const std::string &, const core::LmDiskannConfig &) override {
	}
	void SaveIndexContents(const std::string &, const core::LmDiskannConfig &, const core::IGraphManager *,
	                       common::IndexPointer, common::row_t, common::IndexPointer) override {
	}
	common::idx_t GetInMemorySize() const override {
		return data_.size();
	}
	void EnqueueDeletion(common::row_t, common::IndexPointer &) override {
	}
	void ProcessDeletionQueue(common::IndexPointer &) override {
	}
	bool AllocateNodeBlock(common::row_t, common::IndexPointer &, common::data_ptr_t &) override {
		return false;
	}
	void MarkBlockDirty(common::IndexPointer) override {
	}

private:
	std::vector<uint8_t> data_;
	uint64_t n_, bb_;
};

class MemGraph : public core::IGraphManager {
public:
	MemGraph(const MemStorage *st, const core::NodeLayoutOffsets &lay, const int64_t *row_ids, uint64_t n, int64_t entry)
	    : st_(st), lay_(lay), n_(n) {
		for (uint64_t i = 0; i < n; i++) {
			map_[row_ids ? row_ids[i] : (int64_t)i] = i + 1;
		}
		SetEntryByRow(entry);
	}
	void SetEntryByRow(int64_t entry) {
		entry_row_ = entry;
		if (!TryGetNodePointer(entry, entry_ptr_)) {
			entry_row_ = common::NumericLimits<common::row_t>::Maximum();
			entry_ptr_ = common::IndexPointer();
		}
	}
	bool TryGetNodePointer(common::row_t r, common::IndexPointer &out) const override {
		auto it = map_.find(r);
		if (it == map_.end()) {
			out = common::IndexPointer();
			return false;
		}
		out = common::IndexPointer(it->second);
		return true;
	}
	bool GetNeighbors(common::IndexPointer p, std::vector<common::row_t> &out) const override {
		out.clear();
		common::const_data_ptr_t blk = st_->GetNodeBlockData(p);
		if (!blk) {
			return false;
		}
		uint16_t count = core::GraphNode::GetNeighborCount(blk, lay_);
		const common::row_t *ids = core::GraphNode::GetNeighborIDsPtr(blk, lay_);
		for (uint16_t i = 0; i < count; i++) {
			if (ids[i] != common::NumericLimits<common::row_t>::Maximum()) {
				out.push_back(ids[i]);
			}
		}
		return true;
	}
	common::IndexPointer GetEntryPointPointer() const override {
		return entry_ptr_;
	}
	common::row_t GetEntryPointRowId() const override {
		return entry_row_;
	}
	void InitializeEmptyGraph(const core::LmDiskannConfig &, common::IndexPointer &, common::row_t &) override {
	}
	bool AddNode(common::row_t, const float *, common::idx_t, common::IndexPointer &) override {
		return false;
	}
	bool GetNodeVector(common::IndexPointer, float *, common::idx_t) const override {
		return false;
	}
	void RobustPrune(common::IndexPointer, const float *, std::vector<common::row_t> &,
	                 const core::LmDiskannConfig &) override {
	}
	void SetEntryPoint(common::IndexPointer p, common::row_t r) override {
		entry_ptr_ = p;
		entry_row_ = r;
	}
	common::row_t SelectEntryPointForSearch(common::RandomEngine &) override {
		return entry_row_;
	}
	void HandleNodeDeletion(common::row_t) override {
	}
	void FreeNode(common::row_t) override {
	}
	common::idx_t GetNodeCount() const override {
		return n_;
	}
	common::idx_t GetInMemorySize() const override {
		return 0;
	}
	void Reset() override {
	}

private:
	const MemStorage *st_;
	core::NodeLayoutOffsets lay_;
	uint64_t n_;
	int64_t entry_row_ = 0;
	common::IndexPointer entry_ptr_;
	std::unordered_map<int64_t, uint64_t> map_;
};

struct Host {
	core::LmDiskannConfig cfg;
	std::unique_ptr<MemStorage> storage;
	std::unique_ptr<MemGraph> graph;
	std::shared_ptr<b200::DeviceNodeStore> store;
	std::unique_ptr<core::ISearcher> searcher; // held through the reference's interface, as Coordinator.hpp:115-118 does
	std::unique_ptr<b200::GpuIndexScan> scan;
	bool loaded = true;
};

// Several devices behind ONE core::ISearcher (b200::GpuSearcherGroup): member i has its own storage / graph fakes (the
// same index in QUERY mode, shard i in DATASET mode) and a DeviceNodeStore mirrored through them.
struct GroupHost {
	core::LmDiskannConfig cfg;
	std::vector<std::unique_ptr<MemStorage>> storages;
	std::vector<std::unique_ptr<MemGraph>> graphs;
	std::unique_ptr<b200::GpuSearcherGroup> group;
	core::ISearcher *searcher = nullptr; // the group through the reference's interface
};

// Growable storage for the live-index test: blocks handed out of one buffer, IndexPointer value = slot + 1.
class GrowStorage : public core::IStorageManager {
public:
	GrowStorage(uint64_t capacity, uint64_t bb) : data_(capacity * bb), cap_(capacity), bb_(bb) {
	}
	bool AllocateNodeBlock(common::row_t, common::IndexPointer &ptr_out, common::data_ptr_t &data_out) override {
		if (used_ >= cap_) {
			return false;
		}
		ptr_out = common::IndexPointer(used_ + 1);
		data_out = data_.data() + used_ * bb_;
		used_++;
		return true;
	}
	common::const_data_ptr_t GetNodeBlockData(common::IndexPointer p) const override {
		uint64_t v = p.Get();
		return (v == 0 || v > used_) ? nullptr : data_.data() + (v - 1) * bb_;
	}
	common::data_ptr_t GetMutableNodeBlockData(common::IndexPointer p) override {
		uint64_t v = p.Get();
		return (v == 0 || v > used_) ? nullptr : data_.data() + (v - 1) * bb_;
	}
	void MarkBlockDirty(common::IndexPointer) override {
		marks_++;
	}
	void LoadIndexContents(const std::string &, core::LmDiskannConfig &, core::IGraphManager *, common::IndexPointer &,
	                       common::row_t &, common::IndexPointer &) override {
	}
	void InitializeNewStorage(const std::string &, const core::LmDiskannConfig &) override {
	}
	void SaveIndexContents(const std::string &, const core::LmDiskannConfig &, const core::IGraphManager *,
	                       common::IndexPointer, common::row_t, common::IndexPointer) override {
	}
	common::idx_t GetInMemorySize() const override {
		return used_ * bb_;
	}
	void EnqueueDeletion(common::row_t, common::IndexPointer &) override {
		deletions_++;
	}
	void ProcessDeletionQueue(common::IndexPointer &) override {
	}
	uint64_t marks_ = 0, deletions_ = 0;

private:
	std::vector<uint8_t> data_;
	uint64_t cap_, bb_, used_ = 0;
};

// The reference's OWN core::GraphManager (AddNode, RobustPrune, HandleNodeDeletion ...) running on top of
// MirroredStorage(GrowStorage) with GpuSearcher as its ISearcher: the index a host keeps while rows come and go.
struct LiveHost {
	core::LmDiskannConfig cfg;
	core::NodeLayoutOffsets lay;
	uint64_t bb = 0;
	std::unique_ptr<GrowStorage> inner;
	std::shared_ptr<b200::DeviceNodeStore> store;
	std::unique_ptr<b200::MirroredStorage> storage;
	std::unique_ptr<b200::GpuSearcher> searcher;
	std::unique_ptr<core::GraphManager> graph;
	std::vector<common::row_t> inserted; // insertion order
	common::IndexPointer delete_queue_head;
};

template <typename F>
int Guard(F &&f) {
	g_err.clear();
	g_err_type = 0;
	try {
		f();
		return 0;
	} catch (const ::duckdb::BinderException &e) {
		g_err_type = 2;
		g_err = e.what();
	} catch (const ::duckdb::IOException &e) {
		g_err_type = 3;
		g_err = e.what();
	} catch (const ::duckdb::InvalidInputException &e) {
		g_err_type = 4;
		g_err = e.what();
	} catch (const ::duckdb::NotImplementedException &e) {
		g_err_type = 5;
		g_err = e.what();
	} catch (const ::duckdb::Exception &e) {
		g_err_type = 6;
		g_err = e.what();
	} catch (const std::runtime_error &e) {
		g_err_type = 1;
		g_err = e.what();
	} catch (const std::exception &e) {
		g_err_type = 6;
		g_err = e.what();
	}
	return g_err_type;
}

} // namespace

extern "C" {

const char *lmdhost_last_error() {
	return g_err.c_str();
}

// Builds: in-memory IStorageManager + IGraphManager over `blocks`, a DeviceNodeStore mirrored THROUGH those two
// interfaces (batches of `mirror_batch` blocks), a GpuSearcher behind core::ISearcher, and the scan mirror.
void *lmdhost_create(const uint8_t *blocks, uint64_t n, uint64_t block_bytes, const int64_t *row_ids, int64_t entry_row,
                     int metric, int vec_type, uint64_t dims, uint32_t r, uint32_t l_search, int route,
                     uint64_t mirror_batch, int device) {
	Host *h = nullptr;
	int rc = Guard([&]() {
		auto host = std::make_unique<Host>();
		host->cfg.metric_type = static_cast<common::LmDiskannMetricType>(metric);
		host->cfg.node_vector_type = static_cast<common::LmDiskannVectorType>(vec_type);
		host->cfg.dimensions = dims;
		host->cfg.r = r;
		host->cfg.l_search = l_search;
		core::NodeLayoutOffsets lay = core::CalculateLayoutInternal(host->cfg);
		host->storage = std::make_unique<MemStorage>(blocks, n, block_bytes);
		host->graph = std::make_unique<MemGraph>(host->storage.get(), lay, row_ids, n, entry_row);
		host->store = std::make_shared<b200::DeviceNodeStore>(host->cfg, n ? n : 1, device, block_bytes);
		std::vector<common::row_t> all(n);
		for (uint64_t i = 0; i < n; i++) {
			all[i] = row_ids ? row_ids[i] : (int64_t)i;
		}
		host->store->MirrorFrom(*host->storage, *host->graph, all, mirror_batch ? mirror_batch : 4096);
		host->store->SyncEntryPoint(*host->graph);
		host->searcher = std::make_unique<b200::GpuSearcher>(
		    host->store, route ? b200::GpuSearcher::Route::TERNARY : b200::GpuSearcher::Route::EXACT);
		host->scan = std::make_unique<b200::GpuIndexScan>(host->searcher.get(), host->graph.get(), host->cfg);
		h = host.release();
	});
	return rc == 0 ? h : nullptr;
}

void lmdhost_destroy(void *p) {
	delete static_cast<Host *>(p);
}

uint64_t lmdhost_in_memory_size(void *p) {
	return static_cast<Host *>(p)->searcher->GetInMemorySize();
}

void lmdhost_set_loaded(void *p, int loaded) {
	static_cast<Host *>(p)->loaded = loaded != 0;
}

void lmdhost_set_entry(void *p, int64_t entry_row) {
	static_cast<Host *>(p)->graph->SetEntryByRow(entry_row);
}

// core::ISearcher::Search through the virtual interface, one query. Returns the error class (0 = ok).
int lmdhost_isearcher_search(void *p, const float *query, uint64_t k, uint64_t L, int64_t *out_ids, uint64_t *out_n) {
	Host *h = static_cast<Host *>(p);
	return Guard([&]() {
		std::vector<common::row_t> res;
		h->searcher->Search(query, h->cfg, h->graph.get(), k, res, L);
		*out_n = res.size();
		std::copy(res.begin(), res.end(), out_ids);
	});
}

// Coordinator::Search mirror (guards + clamp), one query.
int lmdhost_coordinator_search(void *p, const float *query, uint64_t k, uint64_t L, int64_t *out_ids, uint64_t *out_n) {
	Host *h = static_cast<Host *>(p);
	return Guard([&]() {
		std::vector<common::row_t> res;
		b200::GpuCoordinatorSearch(h->searcher.get(), h->graph.get(), h->cfg, h->loaded, query, k, res, L);
		*out_n = res.size();
		std::copy(res.begin(), res.end(), out_ids);
	});
}

// InitializeScan + Scan until drained; `calls` receives the number of Scan calls that returned rows.
int lmdhost_scan_all(void *p, const float *query, uint64_t query_dims, uint64_t k, int64_t *out_ids, uint64_t *out_n,
                     uint64_t *calls) {
	Host *h = static_cast<Host *>(p);
	return Guard([&]() {
		auto st = h->scan->InitializeScan(query, query_dims, k);
		uint64_t total = 0, ncalls = 0;
		for (;;) {
			common::idx_t got = h->scan->Scan(*st, out_ids + total);
			if (got == 0) {
				break;
			}
			total += got;
			ncalls++;
		}
		*out_n = total;
		*calls = ncalls;
	});
}

int lmdhost_multi_scan(void *p, const float *queries, uint64_t nq, uint64_t query_dims, uint64_t k, int64_t *out_ids,
                       uint32_t *out_count) {
	Host *h = static_cast<Host *>(p);
	return Guard([&]() { h->scan->MultiScan(queries, nq, query_dims, k, out_ids, out_count); });
}

// Filtered index join: the host resolved a WHERE clause to `allowed` row ids; *under = scans still short of k.
int lmdhost_multi_scan_filtered(void *p, const int64_t *allowed, uint64_t n_allowed, const float *queries, uint64_t nq,
                                uint64_t query_dims, uint64_t k, int64_t *out_ids, uint32_t *out_count,
                                uint64_t *allowed_present, uint64_t *under) {
	Host *h = static_cast<Host *>(p);
	return Guard([&]() {
		b200::AllowedIds set(*h->store, allowed, n_allowed);
		*allowed_present = set.Count();
		*under = h->scan->MultiScanFiltered(set, queries, nq, query_dims, k, out_ids, out_count);
	});
}

int lmdhost_initial_candidates(void *p, const float *vec, uint64_t dims, uint64_t want, int64_t *out_ids,
                               uint64_t *out_n) {
	Host *h = static_cast<Host *>(p);
	return Guard([&]() {
		std::vector<common::row_t> res;
		h->searcher->SearchForInitialCandidates(vec, dims, h->cfg, h->graph.get(), want,
		                                        h->graph->GetEntryPointPointer(), res);
		*out_n = res.size();
		std::copy(res.begin(), res.end(), out_ids);
	});
}

// Graph write side through the adapter: the host's blocks hold only vectors; BuildGraph on the device, WriteBack
// through IStorageManager::GetMutableNodeBlockData / MarkBlockDirty; out_blocks = the HOST's blocks afterwards.
int lmdhost_build_and_write_back(void *p, uint32_t l_insert, float alpha, uint64_t max_batch, uint8_t *out_blocks,
                                 uint64_t n, uint64_t block_bytes, const int64_t *row_ids, int64_t *entry_out) {
	Host *h = static_cast<Host *>(p);
	return Guard([&]() {
		h->cfg.l_insert = l_insert;
		h->cfg.alpha = alpha;
		auto store = std::make_shared<b200::DeviceNodeStore>(h->cfg, n, 0, block_bytes);
		std::vector<common::row_t> all(n);
		for (uint64_t i = 0; i < n; i++) {
			all[i] = row_ids ? row_ids[i] : (int64_t)i;
		}
		store->MirrorFrom(*h->storage, *h->graph, all, 1000); // vectors only: the blocks have no neighbours yet
		store->BuildGraph(max_batch);
		store->WriteBack(*h->storage, *h->graph, all, 777);
		for (uint64_t i = 0; i < n; i++) {
			common::IndexPointer ptr;
			h->graph->TryGetNodePointer(all[i], ptr);
			std::memcpy(out_blocks + i * block_bytes, h->storage->GetNodeBlockData(ptr), block_bytes);
		}
		*entry_out = h->graph->GetEntryPointRowId();
		h->store = store; // searches now run on the freshly built store
		h->searcher = std::make_unique<b200::GpuSearcher>(h->store);
		h->scan = std::make_unique<b200::GpuIndexScan>(h->searcher.get(), h->graph.get(), h->cfg);
	});
}

// MarkBlockDirty path: overwrite one block on the host side, refresh its mirror.
int lmdhost_rewrite_block(void *p, int64_t row_id, const uint8_t *block, uint64_t block_bytes) {
	Host *h = static_cast<Host *>(p);
	return Guard([&]() {
		common::IndexPointer ptr;
		if (!h->graph->TryGetNodePointer(row_id, ptr)) {
			throw ::duckdb::InvalidInputException("unknown row id");
		}
		std::memcpy(h->storage->GetMutableNodeBlockData(ptr), block, block_bytes);
		h->storage->MarkBlockDirty(ptr);
		h->store->RefreshBlock(*h->storage, *h->graph, row_id);
	});
}

// ---- live index ---------------------------------------------------------------------
void *lmdlive_create(int metric, int vec_type, uint64_t dims, uint32_t r, uint32_t l_insert, float alpha,
                     uint32_t l_search, uint64_t capacity, int device) {
	LiveHost *out = nullptr;
	int rc = Guard([&]() {
		auto h = std::make_unique<LiveHost>();
		h->cfg.metric_type = static_cast<common::LmDiskannMetricType>(metric);
		h->cfg.node_vector_type = static_cast<common::LmDiskannVectorType>(vec_type);
		h->cfg.dimensions = dims;
		h->cfg.r = r;
		h->cfg.l_insert = l_insert;
		h->cfg.alpha = alpha;
		h->cfg.l_search = l_search;
		h->lay = core::CalculateLayoutInternal(h->cfg);
		h->bb = (h->lay.total_node_size + 4095) / 4096 * 4096; // db/LmDiskannIndex.cpp:161-162
		h->inner = std::make_unique<GrowStorage>(capacity, h->bb);
		h->store = std::make_shared<b200::DeviceNodeStore>(h->cfg, capacity, device, h->bb);
		h->storage = std::make_unique<b200::MirroredStorage>(h->inner.get(), h->store);
		h->searcher = std::make_unique<b200::GpuSearcher>(h->store);
		h->searcher->AttachLiveMirror(h->storage.get());
		// GraphManager sees only the two interfaces, exactly as Coordinator wires them (core/Coordinator.hpp:115-118)
		h->graph = std::make_unique<core::GraphManager>(h->cfg, h->lay, h->bb, h->storage.get(), h->searcher.get());
		out = h.release();
	});
	return rc == 0 ? out : nullptr;
}

void lmdlive_destroy(void *p) {
	delete static_cast<LiveHost *>(p);
}

// core::GraphManager::AddNode per row (core/Coordinator.cpp Insert -> GraphManager.cpp:65-203)
int lmdlive_insert(void *p, const int64_t *row_ids, const float *vectors, uint64_t n) {
	LiveHost *h = static_cast<LiveHost *>(p);
	return Guard([&]() {
		for (uint64_t i = 0; i < n; i++) {
			common::IndexPointer ptr;
			if (!h->graph->AddNode(row_ids[i], vectors + i * h->cfg.dimensions, h->cfg.dimensions, ptr)) {
				throw std::runtime_error("AddNode failed");
			}
			// a row id that was deleted and comes back counts once, at its new position
			h->inserted.erase(std::remove(h->inserted.begin(), h->inserted.end(), row_ids[i]), h->inserted.end());
			h->inserted.push_back(row_ids[i]);
		}
	});
}

// core/Coordinator.cpp:186-212 Delete, the three steps that concern storage and graph
int lmdlive_delete(void *p, const int64_t *row_ids, uint64_t n) {
	LiveHost *h = static_cast<LiveHost *>(p);
	return Guard([&]() {
		for (uint64_t i = 0; i < n; i++) {
			h->storage->EnqueueDeletion(row_ids[i], h->delete_queue_head);
			h->graph->HandleNodeDeletion(row_ids[i]);
			h->graph->FreeNode(row_ids[i]);
		}
	});
}

int lmdlive_set_entry(void *p, int64_t row_id) {
	LiveHost *h = static_cast<LiveHost *>(p);
	return Guard([&]() {
		common::IndexPointer ptr;
		if (!h->graph->TryGetNodePointer(row_id, ptr)) {
			throw ::duckdb::InvalidInputException("unknown row id");
		}
		h->graph->SetEntryPoint(ptr, row_id);
	});
}

// Batched search on the live mirror; the entry point is whatever the reference's graph manager reports now.
int lmdlive_search(void *p, const float *queries, uint64_t nq, uint64_t k, uint64_t L, int64_t *out_ids,
                   uint32_t *out_count) {
	LiveHost *h = static_cast<LiveHost *>(p);
	return Guard([&]() {
		h->storage->Flush();
		h->store->SyncEntryPoint(*h->graph);
		h->searcher->SearchBatch(queries, nq, k, L, out_ids, nullptr, out_count, nullptr);
	});
}

// One query through core::ISearcher::Search (flushes pending changes by itself).
int lmdlive_search_one(void *p, const float *query, uint64_t k, uint64_t L, int64_t *out_ids, uint64_t *out_n) {
	LiveHost *h = static_cast<LiveHost *>(p);
	return Guard([&]() {
		std::vector<common::row_t> res;
		core::ISearcher *s = h->searcher.get();
		s->Search(query, h->cfg, h->graph.get(), k, res, L);
		*out_n = res.size();
		std::copy(res.begin(), res.end(), out_ids);
	});
}

// The HOST's view: blocks of every live row id in insertion order + the graph manager's entry point.
int lmdlive_export(void *p, uint8_t *out_blocks, int64_t *out_row_ids, uint64_t *out_n, int64_t *entry_out,
                   uint64_t *block_bytes_out) {
	LiveHost *h = static_cast<LiveHost *>(p);
	return Guard([&]() {
		uint64_t m = 0;
		for (common::row_t rid : h->inserted) {
			common::IndexPointer ptr;
			if (!h->graph->TryGetNodePointer(rid, ptr)) {
				continue; // deleted
			}
			if (out_blocks) {
				std::memcpy(out_blocks + m * h->bb, h->inner->GetNodeBlockData(ptr), h->bb);
				out_row_ids[m] = rid;
			}
			m++;
		}
		*out_n = m;
		*entry_out = h->graph->GetEntryPointPointer() == common::IndexPointer()
		                 ? common::NumericLimits<common::row_t>::Maximum()
		                 : h->graph->GetEntryPointRowId();
		*block_bytes_out = h->bb;
	});
}

// ---- one searcher, many host threads (DuckDB scans from many threads; the reference's Search takes no lock) ----------
// `threads` threads each run `per_thread` single-query core::ISearcher::Search calls; returns wall seconds, and the number
// of answers that differ from `expect` [nq][k] (-1 padded) in *mismatch.
int lmdhost_concurrent_searches(void *p, const float *queries, uint64_t nq, uint64_t dims, uint64_t k, uint64_t L,
                                uint32_t threads, uint64_t per_thread, const int64_t *expect, double *seconds,
                                uint64_t *mismatch) {
	Host *h = static_cast<Host *>(p);
	return Guard([&]() {
		std::atomic<uint64_t> bad{0};
		std::atomic<int> failed{0};
		std::string first_error;
		std::mutex err_mu;
		{ // warm every thread's context outside the timed region is not possible (threads are new): time includes it
		}
		const auto t0 = std::chrono::steady_clock::now();
		std::vector<std::thread> pool;
		for (uint32_t t = 0; t < threads; t++) {
			pool.emplace_back([&, t]() {
				try {
					std::vector<common::row_t> res;
					for (uint64_t i = 0; i < per_thread; i++) {
						const uint64_t q = (t * 7919ull + i) % nq;
						h->searcher->Search(queries + q * dims, h->cfg, h->graph.get(), k, res, L);
						if (expect) {
							for (uint64_t j = 0; j < k; j++) {
								const int64_t want = expect[q * k + j];
								const int64_t got = j < res.size() ? res[j] : -1;
								if (want != got) {
									bad++;
									break;
								}
							}
						}
					}
				} catch (const std::exception &e) {
					std::lock_guard<std::mutex> g(err_mu);
					if (!failed.exchange(1)) {
						first_error = e.what();
					}
				}
			});
		}
		for (auto &th : pool) {
			th.join();
		}
		*seconds = std::chrono::duration<double>(std::chrono::steady_clock::now() - t0).count();
		*mismatch = bad.load();
		if (failed.load()) {
			throw std::runtime_error(first_error);
		}
	});
}

// 0 = ok, 1 = threw (message through lmdhost_last_error)
int lmdhost_set_edge_codec(void *p, int on) {
	return Guard([&]() {
		if (auto *g = dynamic_cast<b200::GpuSearcher *>(static_cast<Host *>(p)->searcher.get())) {
			g->Store().SetEdgeCodec(on != 0);
		}
	});
}

void lmdhost_set_combining(void *p, int on) {
	if (auto *g = dynamic_cast<b200::GpuSearcher *>(static_cast<Host *>(p)->searcher.get())) {
		g->SetCombining(on != 0);
	}
}

void lmdhost_combining_stats(void *p, uint64_t *batches, uint64_t *queries) {
	*batches = *queries = 0;
	if (auto *g = dynamic_cast<b200::GpuSearcher *>(static_cast<Host *>(p)->searcher.get())) {
		g->CombiningStats(*batches, *queries);
	}
}

// ---- device groups ------------------------------------------------------------------------------------------------
// members: blocks[i] (n[i] blocks), row_ids[i] (may be null), entry_rows[i], devices[i]. mode 0 = QUERY, 1 = DATASET.
void *lmdgroup_create(const uint8_t *const *blocks, const uint64_t *n, uint64_t block_bytes, const int64_t *const *row_ids,
                      const int64_t *entry_rows, uint32_t members, const int *devices, int mode, int metric, int vec_type,
                      uint64_t dims, uint32_t r, uint32_t l_search, int route) {
	GroupHost *out = nullptr;
	int rc = Guard([&]() {
		auto g = std::make_unique<GroupHost>();
		g->cfg.metric_type = static_cast<common::LmDiskannMetricType>(metric);
		g->cfg.node_vector_type = static_cast<common::LmDiskannVectorType>(vec_type);
		g->cfg.dimensions = dims;
		g->cfg.r = r;
		g->cfg.l_search = l_search;
		core::NodeLayoutOffsets lay = core::CalculateLayoutInternal(g->cfg);
		std::vector<std::shared_ptr<b200::DeviceNodeStore>> stores;
		for (uint32_t i = 0; i < members; i++) {
			g->storages.push_back(std::make_unique<MemStorage>(blocks[i], n[i], block_bytes));
			g->graphs.push_back(std::make_unique<MemGraph>(g->storages.back().get(), lay, row_ids ? row_ids[i] : nullptr,
			                                               n[i], entry_rows[i]));
			std::vector<common::row_t> all(n[i]);
			for (uint64_t j = 0; j < n[i]; j++) {
				all[j] = (row_ids && row_ids[i]) ? row_ids[i][j] : (int64_t)j;
			}
			auto one = b200::GpuSearcherGroup::MirrorReplicas(g->cfg, *g->storages.back(), *g->graphs.back(), all,
			                                                  {devices[i]}, block_bytes);
			stores.push_back(one[0]);
		}
		g->group = std::make_unique<b200::GpuSearcherGroup>(
		    std::move(stores), mode ? b200::GpuSearcherGroup::ShardMode::DATASET : b200::GpuSearcherGroup::ShardMode::QUERY,
		    route ? b200::GpuSearcher::Route::TERNARY : b200::GpuSearcher::Route::EXACT);
		g->searcher = g->group.get();
		out = g.release();
	});
	return rc == 0 ? out : nullptr;
}

void lmdgroup_destroy(void *p) {
	delete static_cast<GroupHost *>(p);
}

uint64_t lmdgroup_in_memory_size(void *p) {
	return static_cast<GroupHost *>(p)->searcher->GetInMemorySize();
}

int lmdgroup_search_batch(void *p, const float *queries, uint64_t nq, uint64_t k, uint64_t L, int64_t *out_ids,
                          float *out_dist, uint32_t *out_count, float *ms3) {
	GroupHost *g = static_cast<GroupHost *>(p);
	return Guard([&]() {
		lmd_group_stats st;
		g->group->SearchBatch(queries, nq, k, L, out_ids, out_dist, out_count, &st);
		if (ms3) {
			ms3[0] = st.search_ms_max;
			ms3[1] = st.merge_ms;
			ms3[2] = st.wall_ms;
		}
	});
}

// core::ISearcher::Search through the virtual interface (one query)
int lmdgroup_isearcher_search(void *p, const float *query, uint64_t k, uint64_t L, int64_t *out_ids, uint64_t *out_n) {
	GroupHost *g = static_cast<GroupHost *>(p);
	return Guard([&]() {
		std::vector<common::row_t> res;
		g->searcher->Search(query, g->cfg, g->graphs[0].get(), k, res, L);
		*out_n = res.size();
		std::copy(res.begin(), res.end(), out_ids);
	});
}

void lmdlive_counters(void *p, uint64_t *flushes, uint64_t *blocks_sent, uint64_t *marks, uint64_t *device_nodes) {
	LiveHost *h = static_cast<LiveHost *>(p);
	*flushes = h->storage->FlushCount();
	*blocks_sent = h->storage->BlocksSent();
	*marks = h->inner->marks_;
	*device_nodes = h->store->NodeCount();
}

} // extern "C"
