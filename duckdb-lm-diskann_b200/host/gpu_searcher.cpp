// See gpu_searcher.hpp. Host orchestration only: every distance, comparison and graph step happens in liblmd_b200.so.
#include "gpu_searcher.hpp"

#include <algorithm>
#include <cmath>
#include <cstdio>
#include <cstring>
#include <iostream>
#include <stdexcept>
#include <unordered_set>

namespace diskann {
namespace b200 {

std::string LastError() {
	const char *e = lmd_last_error();
	return e ? std::string(e) : std::string();
}

namespace {

// Status -> the reference's exception for the same condition (include/lmd_b200.h documents the mapping).
[[noreturn]] void ThrowStatus(lmd_status st, const char *where) {
	const std::string msg = std::string(where) + ": " + LastError();
	switch (st) {
	case LMD_ERR_INVALID_ARG:
		throw ::duckdb::InvalidInputException(msg);
	case LMD_ERR_UNSUPPORTED:
		throw ::duckdb::NotImplementedException(msg);
	case LMD_ERR_STATE:
		throw std::runtime_error(msg);
	default: // LMD_ERR_CUDA, LMD_ERR_OOM
		throw ::duckdb::IOException(msg);
	}
}

void Check(lmd_status st, const char *where) {
	if (st != LMD_OK) {
		ThrowStatus(st, where);
	}
}

lmd_config ToAbiConfig(const core::LmDiskannConfig &c, common::idx_t block_bytes) {
	lmd_config out;
	std::memset(&out, 0, sizeof(out));
	out.dims = static_cast<uint32_t>(c.dimensions);
	out.r = c.r;
	out.metric = static_cast<uint32_t>(c.metric_type);        // same numeric values, common/ann.hpp:16-24
	out.vec_type = static_cast<uint32_t>(c.node_vector_type); // common/ann.hpp:26-32
	out.l_search = c.l_search;
	out.block_bytes = block_bytes;
	return out;
}

} // namespace

// ---------------------------------------------------------------------------------------------------------------
DeviceNodeStore::DeviceNodeStore(const core::LmDiskannConfig &config, common::idx_t capacity_nodes, int device,
                                 common::idx_t block_bytes)
    : config_(config) {
	lmd_config cfg = ToAbiConfig(config, block_bytes);
	Check(lmd_store_create(&cfg, device, capacity_nodes, &store_), "DeviceNodeStore");
	lmd_layout lay;
	Check(lmd_layout_calc(cfg.dims, cfg.vec_type, cfg.r, &lay), "DeviceNodeStore");
	block_bytes_ = block_bytes ? block_bytes : lmd_block_bytes(lay.total_node_size);
}

DeviceNodeStore::~DeviceNodeStore() {
	if (store_) {
		lmd_store_destroy(store_);
	}
}

common::idx_t DeviceNodeStore::MirrorFrom(const core::IStorageManager &storage, const core::IGraphManager &graph,
                                          const std::vector<common::row_t> &row_ids, common::idx_t batch_nodes) {
	if (batch_nodes == 0) {
		batch_nodes = 1;
	}
	std::vector<uint8_t> stage(static_cast<size_t>(batch_nodes) * block_bytes_);
	std::vector<common::row_t> stage_ids(batch_nodes);
	common::idx_t filled = 0, total = 0;
	auto flush = [&]() {
		if (filled) {
			Check(lmd_store_upload(store_, stage_ids.data(), stage.data(), filled), "DeviceNodeStore::MirrorFrom");
			total += filled;
			filled = 0;
		}
	};
	for (common::row_t rid : row_ids) {
		common::IndexPointer ptr;
		if (!graph.TryGetNodePointer(rid, ptr)) {
			continue;
		}
		common::const_data_ptr_t blk = storage.GetNodeBlockData(ptr);
		if (!blk) {
			continue;
		}
		std::memcpy(stage.data() + static_cast<size_t>(filled) * block_bytes_, blk, block_bytes_);
		stage_ids[filled++] = rid;
		if (filled == batch_nodes) {
			flush();
		}
	}
	flush();
	Check(lmd_store_finalize(store_), "DeviceNodeStore::MirrorFrom");
	return total;
}

void DeviceNodeStore::Upload(const common::row_t *row_ids, common::const_data_ptr_t blocks, common::idx_t n) {
	Check(lmd_store_upload(store_, row_ids, blocks, n), "DeviceNodeStore::Upload");
	Check(lmd_store_finalize(store_), "DeviceNodeStore::Upload");
}

common::idx_t DeviceNodeStore::Apply(const common::row_t *row_ids, common::const_data_ptr_t blocks, common::idx_t n) {
	uint64_t appended = 0;
	Check(lmd_store_apply(store_, row_ids, blocks, n, &appended), "DeviceNodeStore::Apply");
	return appended;
}

common::idx_t DeviceNodeStore::Remove(const common::row_t *row_ids, common::idx_t n) {
	uint64_t removed = 0;
	Check(lmd_store_remove(store_, row_ids, n, &removed), "DeviceNodeStore::Remove");
	return removed;
}

void DeviceNodeStore::RefreshBlock(const core::IStorageManager &storage, const core::IGraphManager &graph,
                                   common::row_t row_id) {
	common::IndexPointer ptr;
	if (!graph.TryGetNodePointer(row_id, ptr)) {
		throw ::duckdb::InvalidInputException("DeviceNodeStore::RefreshBlock: unknown row id");
	}
	common::const_data_ptr_t blk = storage.GetNodeBlockData(ptr);
	if (!blk) {
		throw ::duckdb::IOException("DeviceNodeStore::RefreshBlock: GetNodeBlockData returned null");
	}
	Check(lmd_store_update(store_, row_id, blk), "DeviceNodeStore::RefreshBlock");
}

void DeviceNodeStore::SyncEntryPoint(const core::IGraphManager &graph) {
	// core/Searcher.cpp:70-77: an invalid pointer or the INT64_MAX row id means "no entry point" -> empty results
	common::IndexPointer ptr = graph.GetEntryPointPointer();
	common::row_t rid = graph.GetEntryPointRowId();
	if (ptr == common::IndexPointer()) {
		rid = common::NumericLimits<common::row_t>::Maximum();
	}
	SetEntryPoint(rid);
}

void DeviceNodeStore::SetEntryPoint(common::row_t row_id) {
	Check(lmd_store_set_entry(store_, row_id), "DeviceNodeStore::SetEntryPoint");
}

void DeviceNodeStore::SetEdgeCodec(bool on) {
	Check(lmd_store_enable_codec(store_, on ? LMD_CODEC_INT8 : LMD_CODEC_NONE), "DeviceNodeStore::SetEdgeCodec");
}

void DeviceNodeStore::UploadVectors(const common::row_t *row_ids, const float *vectors, common::idx_t n) {
	Check(lmd_store_upload_vectors(store_, row_ids, vectors, n), "DeviceNodeStore::UploadVectors");
}

lmd_build_stats DeviceNodeStore::BuildGraph(common::idx_t max_batch) {
	lmd_build_params bp;
	std::memset(&bp, 0, sizeof(bp));
	bp.l_insert = config_.l_insert;
	bp.alpha = config_.alpha;
	bp.max_batch = static_cast<uint32_t>(max_batch);
	lmd_build_stats st;
	Check(lmd_build_graph(store_, &bp, &st), "DeviceNodeStore::BuildGraph");
	return st;
}

common::idx_t DeviceNodeStore::WriteBack(core::IStorageManager &storage, core::IGraphManager &graph,
                                         const std::vector<common::row_t> &row_ids, common::idx_t batch_nodes) {
	if (batch_nodes == 0) {
		batch_nodes = 1;
	}
	// Blocks carry no id of their own: every downloaded block is paired with the row id the STORE holds for its slot
	// (stores filled by MirrorFrom / Apply may have skipped or reordered rows), and only rows the caller asked for are
	// written.
	const std::unordered_set<common::row_t> wanted(row_ids.begin(), row_ids.end());
	const common::idx_t total = NodeCount();
	std::vector<uint8_t> stage(static_cast<size_t>(batch_nodes) * block_bytes_);
	std::vector<common::row_t> slot_ids(batch_nodes);
	common::idx_t written = 0;
	for (common::idx_t first = 0; first < total; first += batch_nodes) {
		const common::idx_t now = std::min<common::idx_t>(batch_nodes, total - first);
		Check(lmd_store_download_blocks(store_, first, now, stage.data()), "DeviceNodeStore::WriteBack");
		Check(lmd_store_download_row_ids(store_, first, now, slot_ids.data()), "DeviceNodeStore::WriteBack");
		for (common::idx_t i = 0; i < now; i++) {
			common::IndexPointer ptr;
			if (!wanted.count(slot_ids[i]) || !graph.TryGetNodePointer(slot_ids[i], ptr)) {
				continue;
			}
			common::data_ptr_t dst = storage.GetMutableNodeBlockData(ptr);
			if (!dst) {
				continue;
			}
			std::memcpy(dst, stage.data() + static_cast<size_t>(i) * block_bytes_, block_bytes_);
			storage.MarkBlockDirty(ptr);
			written++;
		}
	}
	if (total > 0) {
		// AddNode makes the first inserted node the entry point (core/GraphManager.cpp:119-121): slot 0's row
		common::row_t first_row = 0;
		Check(lmd_store_download_row_ids(store_, 0, 1, &first_row), "DeviceNodeStore::WriteBack");
		common::IndexPointer first_ptr;
		if (graph.TryGetNodePointer(first_row, first_ptr)) {
			graph.SetEntryPoint(first_ptr, first_row);
		}
	}
	return written;
}

common::idx_t DeviceNodeStore::NodeCount() const {
	return lmd_store_node_count(store_);
}

common::idx_t DeviceNodeStore::DeviceBytes() const {
	return lmd_store_device_bytes(store_);
}

// ---------------------------------------------------------------------------------------------------------------
MirroredStorage::MirroredStorage(core::IStorageManager *inner, std::shared_ptr<DeviceNodeStore> store)
    : inner_(inner), store_(std::move(store)) {
	if (!inner_ || !store_) {
		throw ::duckdb::InternalException("MirroredStorage requires a storage manager and a DeviceNodeStore");
	}
}

common::idx_t MirroredStorage::MirrorAll(const core::IGraphManager &graph, const std::vector<common::row_t> &row_ids,
                                         common::idx_t batch_nodes) {
	for (common::row_t rid : row_ids) {
		common::IndexPointer ptr;
		if (graph.TryGetNodePointer(rid, ptr)) {
			ptr_to_row_[ptr.Get()] = rid;
			row_to_ptr_[rid] = ptr.Get();
		}
	}
	return store_->MirrorFrom(*inner_, graph, row_ids, batch_nodes);
}

common::idx_t MirroredStorage::Flush() {
	common::idx_t sent = 0;
	if (!deleted_.empty()) {
		// deletions first: a row id deleted and re-inserted since the last flush must come back as a NEW node
		store_->Remove(deleted_.data(), deleted_.size());
		deleted_.clear();
	}
	if (!dirty_order_.empty()) {
		const common::idx_t bb = store_->BlockBytes();
		stage_.resize(static_cast<size_t>(dirty_order_.size()) * bb);
		stage_ids_.clear();
		for (uint64_t p : dirty_order_) {
			auto it = ptr_to_row_.find(p);
			if (it == ptr_to_row_.end()) {
				continue; // a pointer this mirror never saw allocated or mirrored: not a node block of this index
			}
			common::const_data_ptr_t blk = inner_->GetNodeBlockData(common::IndexPointer(p));
			if (!blk) {
				continue;
			}
			std::memcpy(stage_.data() + stage_ids_.size() * bb, blk, bb);
			stage_ids_.push_back(it->second);
		}
		dirty_order_.clear();
		dirty_.clear();
		if (!stage_ids_.empty()) {
			store_->Apply(stage_ids_.data(), stage_.data(), stage_ids_.size());
			sent = stage_ids_.size();
		}
	}
	flushes_++;
	blocks_sent_ += sent;
	return sent;
}

void MirroredStorage::LoadIndexContents(const std::string &index_path, core::LmDiskannConfig &config_out,
                                        core::IGraphManager *graph_manager_out, common::IndexPointer &entry_point_ptr_out,
                                        common::row_t &entry_point_rowid_out,
                                        common::IndexPointer &delete_queue_head_out) {
	inner_->LoadIndexContents(index_path, config_out, graph_manager_out, entry_point_ptr_out, entry_point_rowid_out,
	                          delete_queue_head_out);
}

void MirroredStorage::InitializeNewStorage(const std::string &index_path, const core::LmDiskannConfig &config) {
	inner_->InitializeNewStorage(index_path, config);
}

void MirroredStorage::SaveIndexContents(const std::string &index_path, const core::LmDiskannConfig &config,
                                        const core::IGraphManager *graph_manager, common::IndexPointer entry_point_ptr,
                                        common::row_t entry_point_rowid, common::IndexPointer delete_queue_head) {
	inner_->SaveIndexContents(index_path, config, graph_manager, entry_point_ptr, entry_point_rowid, delete_queue_head);
}

common::idx_t MirroredStorage::GetInMemorySize() const {
	return inner_->GetInMemorySize();
}

void MirroredStorage::ProcessDeletionQueue(common::IndexPointer &delete_queue_head_ptr) {
	inner_->ProcessDeletionQueue(delete_queue_head_ptr);
}

common::const_data_ptr_t MirroredStorage::GetNodeBlockData(common::IndexPointer node_ptr) const {
	return inner_->GetNodeBlockData(node_ptr);
}

common::data_ptr_t MirroredStorage::GetMutableNodeBlockData(common::IndexPointer node_ptr) {
	return inner_->GetMutableNodeBlockData(node_ptr);
}

void MirroredStorage::EnqueueDeletion(common::row_t row_id, common::IndexPointer &delete_queue_head_ptr) {
	inner_->EnqueueDeletion(row_id, delete_queue_head_ptr);
	deleted_.push_back(row_id);
	auto it = row_to_ptr_.find(row_id);
	if (it != row_to_ptr_.end()) {
		// a block dirtied and then deleted before the next search is not sent at all
		if (dirty_.erase(it->second)) {
			dirty_order_.erase(std::find(dirty_order_.begin(), dirty_order_.end(), it->second));
		}
		ptr_to_row_.erase(it->second);
		row_to_ptr_.erase(it);
	}
}

bool MirroredStorage::AllocateNodeBlock(common::row_t row_id, common::IndexPointer &node_ptr_out,
                                        common::data_ptr_t &node_data_out) {
	if (!inner_->AllocateNodeBlock(row_id, node_ptr_out, node_data_out)) {
		return false;
	}
	ptr_to_row_[node_ptr_out.Get()] = row_id;
	row_to_ptr_[row_id] = node_ptr_out.Get();
	MarkBlockDirty(node_ptr_out); // GraphManager fills the block right after this call; it is read at Flush() time
	return true;
}

void MirroredStorage::MarkBlockDirty(common::IndexPointer node_ptr) {
	inner_->MarkBlockDirty(node_ptr);
	if (dirty_.insert(node_ptr.Get()).second) {
		dirty_order_.push_back(node_ptr.Get());
	}
}

// ---------------------------------------------------------------------------------------------------------------
GpuSearcher::GpuSearcher(std::shared_ptr<DeviceNodeStore> store, Route route) : store_(std::move(store)), route_(route) {
	if (!store_) {
		throw ::duckdb::InternalException("GpuSearcher requires a DeviceNodeStore");
	}
}

GpuSearcher::~GpuSearcher() {
	for (auto &kv : contexts_) {
		lmd_ctx_destroy(kv.second);
	}
}

lmd_ctx *GpuSearcher::ThreadContext() {
	std::lock_guard<std::mutex> g(mu_);
	auto it = contexts_.find(std::this_thread::get_id());
	if (it != contexts_.end()) {
		return it->second;
	}
	lmd_ctx *c = nullptr;
	Check(lmd_ctx_create(store_->Handle(), &c), "GpuSearcher: search context");
	contexts_[std::this_thread::get_id()] = c;
	return c;
}

void GpuSearcher::Search(const float *query_vector, const core::LmDiskannConfig &config,
                         core::IGraphManager *graph_manager, common::idx_t k_neighbors,
                         std::vector<common::row_t> &result_row_ids, common::idx_t search_list_size) {
	result_row_ids.clear();
	if (mirror_ && mirror_->HasPending()) {
		std::lock_guard<std::mutex> g(mu_);
		if (mirror_->HasPending()) {
			mirror_->Flush(); // blocks the host allocated / dirtied / deleted since the last search
		}
	}
	if (graph_manager) {
		std::lock_guard<std::mutex> g(mu_);
		// the entry point can move between searches (first insert / deletion, core/GraphManager.cpp:119-121)
		common::row_t rid = graph_manager->GetEntryPointRowId();
		if (graph_manager->GetEntryPointPointer() == common::IndexPointer()) {
			rid = common::NumericLimits<common::row_t>::Maximum();
		}
		if (!entry_synced_ || rid != synced_entry_) {
			store_->SetEntryPoint(rid);
			synced_entry_ = rid;
			entry_synced_ = true;
		}
	}
	// core/Searcher.cpp:261-265 is applied inside lmd_search_batch with the STORE's l_search; honour a differing
	// per-call config here so that both resolve identically
	common::idx_t L = search_list_size > 0 ? search_list_size : config.l_search;
	if (L == 0) {
		L = k_neighbors > 20 ? k_neighbors + 20 : 40;
	}
	if (k_neighbors == 0) {
		return; // PerformSearch with k = 0 keeps nothing (core/Searcher.cpp:224-227)
	}
	if (!combine_) {
		std::vector<common::row_t> ids(k_neighbors);
		uint32_t count = 0;
		SearchBatch(query_vector, 1, k_neighbors, L, ids.data(), nullptr, &count, nullptr);
		result_row_ids.assign(ids.begin(), ids.begin() + count);
		return;
	}
	// Combining: up to MAX_LEADERS searches are on the device at once, each on its thread's own context (that alone
	// scales to ~16-32 host threads, profiles/r2_06_concurrent_point_queries.json); a call that finds every leader slot
	// taken queues, and the next free leader carries everything queued so far as ONE batch.
	PendingSearch mine {query_vector, k_neighbors, L, &result_row_ids};
	std::unique_lock<std::mutex> lk(comb_mu_);
	comb_queue_.push_back(&mine);
	while (!mine.done) {
		if (!mine.taken && comb_leaders_ < MAX_LEADERS) {
			comb_leaders_++;
			std::vector<PendingSearch *> batch;
			batch.swap(comb_queue_);
			for (auto *p : batch) {
				p->taken = true;
			}
			lk.unlock();
			LeadBatch(batch); // never throws: errors are stored per request
			lk.lock();
			for (auto *p : batch) {
				p->done = true;
			}
			comb_leaders_--;
			comb_cv_.notify_all();
		} else {
			comb_cv_.wait(lk);
		}
	}
	lk.unlock();
	if (mine.error) {
		std::rethrow_exception(mine.error);
	}
}

// One device launch per distinct (k, L) among the queued requests (normally one: scans of a session share both).
void GpuSearcher::LeadBatch(std::vector<PendingSearch *> &batch) {
	const common::idx_t dims = store_->Config().dimensions;
	std::vector<bool> taken(batch.size(), false);
	std::vector<float> queries;
	std::vector<common::row_t> ids;
	std::vector<uint32_t> counts;
	std::vector<size_t> members;
	for (size_t first = 0; first < batch.size(); first++) {
		if (taken[first]) {
			continue;
		}
		members.clear();
		for (size_t i = first; i < batch.size(); i++) {
			if (!taken[i] && batch[i]->k == batch[first]->k && batch[i]->L == batch[first]->L) {
				taken[i] = true;
				members.push_back(i);
			}
		}
		const common::idx_t k = batch[first]->k, n = members.size();
		try {
			const float *q = batch[first]->query;
			if (n > 1) {
				queries.resize(n * dims);
				for (size_t m = 0; m < n; m++) {
					std::memcpy(queries.data() + m * dims, batch[members[m]]->query, dims * sizeof(float));
				}
				q = queries.data();
			}
			ids.assign(n * k, -1);
			counts.assign(n, 0);
			SearchBatch(q, n, k, batch[first]->L, ids.data(), nullptr, counts.data(), nullptr);
			for (size_t m = 0; m < n; m++) {
				auto &out = *batch[members[m]]->out;
				out.assign(ids.begin() + m * k, ids.begin() + m * k + counts[m]);
			}
			comb_batches_.fetch_add(1, std::memory_order_relaxed);
			comb_queries_.fetch_add(n, std::memory_order_relaxed);
		} catch (...) {
			for (size_t m : members) {
				batch[m]->error = std::current_exception();
			}
		}
	}
}

common::idx_t GpuSearcher::GetInMemorySize() const {
	return store_->DeviceBytes();
}

void GpuSearcher::SearchForInitialCandidates(const float *new_node_vector, common::idx_t dimensions,
                                             const core::LmDiskannConfig &config, core::IGraphManager *graph_manager,
                                             common::idx_t num_candidates_to_find,
                                             common::IndexPointer current_entry_point,
                                             std::vector<common::row_t> &candidate_row_ids_out) {
	candidate_row_ids_out.clear();
	if (dimensions != config.dimensions || num_candidates_to_find == 0 || current_entry_point == common::IndexPointer()) {
		return; // core/Searcher.cpp:74-77: nothing to start from
	}
	// core/Searcher.cpp:281-282: L = max(config.l_insert, num_candidates_to_find)
	Search(new_node_vector, config, graph_manager, num_candidates_to_find, candidate_row_ids_out,
	       std::max<common::idx_t>(num_candidates_to_find, config.l_insert));
}

void GpuSearcher::SearchBatch(const float *queries, common::idx_t nq, common::idx_t k_neighbors,
                              common::idx_t search_list_size, common::row_t *out_ids, float *out_dist,
                              uint32_t *out_count, lmd_stats *stats) {
	if (mirror_ && mirror_->HasPending()) {
		std::lock_guard<std::mutex> g(mu_);
		if (mirror_->HasPending()) {
			mirror_->Flush();
		}
	}
	Check(lmd_ctx_search_batch(ThreadContext(), nullptr, queries, nq, static_cast<uint32_t>(k_neighbors),
	                           static_cast<uint32_t>(search_list_size), static_cast<uint32_t>(route_), 0, out_ids,
	                           out_dist, out_count, nullptr, stats),
	      "GpuSearcher::Search");
}

// ---------------------------------------------------------------------------------------------------------------
GpuSearcherGroup::GpuSearcherGroup(std::vector<std::shared_ptr<DeviceNodeStore>> stores, ShardMode mode,
                                   GpuSearcher::Route route)
    : stores_(std::move(stores)), mode_(mode), route_(route) {
	if (stores_.empty()) {
		throw ::duckdb::InternalException("GpuSearcherGroup requires at least one DeviceNodeStore");
	}
	std::vector<lmd_store *> handles;
	for (auto &s : stores_) {
		if (!s) {
			throw ::duckdb::InternalException("GpuSearcherGroup: null DeviceNodeStore");
		}
		handles.push_back(s->Handle());
	}
	Check(lmd_group_create(handles.data(), static_cast<uint32_t>(handles.size()), static_cast<uint32_t>(mode), &group_),
	      "GpuSearcherGroup");
	if (mode_ == ShardMode::QUERY) {
		for (auto &s : stores_) {
			singles_.push_back(std::make_unique<GpuSearcher>(s, route_));
		}
	}
}

GpuSearcherGroup::~GpuSearcherGroup() {
	singles_.clear();
	if (group_) {
		lmd_group_destroy(group_);
	}
}

std::vector<std::shared_ptr<DeviceNodeStore>> GpuSearcherGroup::MirrorReplicas(
    const core::LmDiskannConfig &config, const core::IStorageManager &storage, const core::IGraphManager &graph,
    const std::vector<common::row_t> &row_ids, const std::vector<int> &devices, common::idx_t block_bytes) {
	std::vector<std::shared_ptr<DeviceNodeStore>> out;
	for (int dev : devices) {
		auto s = std::make_shared<DeviceNodeStore>(config, row_ids.empty() ? 1 : row_ids.size(), dev, block_bytes);
		s->MirrorFrom(storage, graph, row_ids);
		s->SyncEntryPoint(graph);
		out.push_back(std::move(s));
	}
	return out;
}

void GpuSearcherGroup::Search(const float *query_vector, const core::LmDiskannConfig &config,
                              core::IGraphManager *graph_manager, common::idx_t k_neighbors,
                              std::vector<common::row_t> &result_row_ids, common::idx_t search_list_size) {
	result_row_ids.clear();
	if (mode_ == ShardMode::QUERY) {
		// one query cannot be split: it goes to one replica (round-robin spreads concurrent scans over the devices)
		const uint64_t pick = next_.fetch_add(1, std::memory_order_relaxed) % singles_.size();
		singles_[pick]->Search(query_vector, config, graph_manager, k_neighbors, result_row_ids, search_list_size);
		return;
	}
	// shard-local entry points are part of each shard's store; the host graph manager has no single one to sync
	common::idx_t L = search_list_size > 0 ? search_list_size : config.l_search;
	if (L == 0) {
		L = k_neighbors > 20 ? k_neighbors + 20 : 40; // core/Searcher.cpp:261-265
	}
	if (k_neighbors == 0) {
		return;
	}
	std::vector<common::row_t> ids(k_neighbors);
	uint32_t count = 0;
	SearchBatch(query_vector, 1, k_neighbors, L, ids.data(), nullptr, &count, nullptr);
	result_row_ids.assign(ids.begin(), ids.begin() + count);
}

common::idx_t GpuSearcherGroup::GetInMemorySize() const {
	common::idx_t total = 0;
	for (auto &s : stores_) {
		total += s->DeviceBytes();
	}
	return total;
}

void GpuSearcherGroup::SearchForInitialCandidates(const float *new_node_vector, common::idx_t dimensions,
                                                  const core::LmDiskannConfig &config,
                                                  core::IGraphManager *graph_manager,
                                                  common::idx_t num_candidates_to_find,
                                                  common::IndexPointer current_entry_point,
                                                  std::vector<common::row_t> &candidate_row_ids_out) {
	candidate_row_ids_out.clear();
	if (dimensions != config.dimensions || num_candidates_to_find == 0 || current_entry_point == common::IndexPointer()) {
		return; // core/Searcher.cpp:74-77
	}
	Search(new_node_vector, config, graph_manager, num_candidates_to_find, candidate_row_ids_out,
	       std::max<common::idx_t>(num_candidates_to_find, config.l_insert)); // core/Searcher.cpp:281-282
}

void GpuSearcherGroup::SearchBatch(const float *queries, common::idx_t nq, common::idx_t k_neighbors,
                                   common::idx_t search_list_size, common::row_t *out_ids, float *out_dist,
                                   uint32_t *out_count, lmd_group_stats *stats) {
	Check(lmd_group_search_batch(group_, queries, nq, static_cast<uint32_t>(k_neighbors),
	                             static_cast<uint32_t>(search_list_size), static_cast<uint32_t>(route_), 0, out_ids,
	                             out_dist, out_count, stats),
	      "GpuSearcherGroup::SearchBatch");
}

// ---------------------------------------------------------------------------------------------------------------
AllowedIds::AllowedIds(DeviceNodeStore &store, const common::row_t *row_ids, common::idx_t count) {
	Check(lmd_filter_create(store.Handle(), row_ids, count, &filter_), "AllowedIds");
}

AllowedIds::~AllowedIds() {
	if (filter_) {
		lmd_filter_destroy(filter_);
	}
}

common::idx_t AllowedIds::Count() const {
	return lmd_filter_count(filter_);
}

// The plan: "L_search ... potentially automatically increased for filtered queries". The exploration list holds the L
// closest rows of ANY kind, of which only a fraction f passes the filter, so the list is widened by 1 / sqrt(f), at most
// eightfold and never beyond the build's limit of 512. Measured on 1 M x 128 (profiles/r27_filtered_1m_d128.json):
// recall over the allowed rows at f = 0.1 is 0.85 at L = 64 and 0.97 at L = 205; at f = 0.01, 0.53 and 0.93 (L = 512).
uint32_t GpuSearcher::FilteredListSize(uint32_t L, common::idx_t allowed_rows, common::idx_t total_rows) {
	if (allowed_rows == 0 || total_rows == 0 || allowed_rows >= total_rows) {
		return L;
	}
	double scale = std::sqrt(static_cast<double>(total_rows) / static_cast<double>(allowed_rows));
	scale = std::min(8.0, scale);
	const double wide = std::ceil(static_cast<double>(L) * scale);
	return static_cast<uint32_t>(std::min(512.0, std::max(static_cast<double>(L), wide)));
}

common::idx_t GpuSearcher::SearchBatchFiltered(const AllowedIds &allowed, const float *queries, common::idx_t nq,
                                               common::idx_t k_neighbors, common::idx_t search_list_size,
                                               common::row_t *out_ids, float *out_dist, uint32_t *out_count,
                                               bool widen) {
	if (mirror_ && mirror_->HasPending()) {
		mirror_->Flush(); // note: new or deleted rows make `allowed` stale; the call below then reports it
	}
	const common::idx_t dims = store_->Config().dimensions;
	std::vector<uint32_t> count_storage;
	if (!out_count) {
		count_storage.resize(nq);
		out_count = count_storage.data();
	}
	uint32_t L = static_cast<uint32_t>(search_list_size ? search_list_size : store_->Config().l_search);
	if (L == 0) {
		L = k_neighbors > 20 ? static_cast<uint32_t>(k_neighbors) + 20 : 40; // core/Searcher.cpp:261-265
	}
	L = std::max<uint32_t>(L, static_cast<uint32_t>(k_neighbors));
	if (widen) {
		L = FilteredListSize(L, allowed.Count(), store_->NodeCount());
	}
	Check(lmd_ctx_search_batch(ThreadContext(), allowed.Handle(), queries, nq, static_cast<uint32_t>(k_neighbors),
	                                L, static_cast<uint32_t>(route_), 0, out_ids, out_dist, out_count, nullptr, nullptr),
	      "GpuSearcher::SearchBatchFiltered");
	const common::idx_t reachable = std::min<common::idx_t>(allowed.Count(), k_neighbors);
	std::vector<common::idx_t> shortq;
	for (common::idx_t q = 0; q < nq; q++) {
		if (out_count[q] < reachable) {
			shortq.push_back(q);
		}
	}
	// widen: only the queries that came back short, with twice the list each round
	std::vector<float> sub_q;
	std::vector<common::row_t> sub_ids;
	std::vector<float> sub_dist;
	std::vector<uint32_t> sub_cnt;
	while (widen && !shortq.empty() && L < 512) {
		L = std::min<uint32_t>(512, L * 2);
		const common::idx_t m = shortq.size();
		sub_q.resize(m * dims);
		sub_ids.resize(m * k_neighbors);
		sub_dist.resize(m * k_neighbors);
		sub_cnt.resize(m);
		for (common::idx_t i = 0; i < m; i++) {
			std::memcpy(sub_q.data() + i * dims, queries + shortq[i] * dims, dims * sizeof(float));
		}
		Check(lmd_ctx_search_batch(ThreadContext(), allowed.Handle(), sub_q.data(), m,
		                                static_cast<uint32_t>(k_neighbors), L, static_cast<uint32_t>(route_), 0,
		                                sub_ids.data(), sub_dist.data(), sub_cnt.data(), nullptr, nullptr),
		      "GpuSearcher::SearchBatchFiltered");
		std::vector<common::idx_t> still;
		for (common::idx_t i = 0; i < m; i++) {
			const common::idx_t q = shortq[i];
			std::memcpy(out_ids + q * k_neighbors, sub_ids.data() + i * k_neighbors, k_neighbors * sizeof(common::row_t));
			if (out_dist) {
				std::memcpy(out_dist + q * k_neighbors, sub_dist.data() + i * k_neighbors, k_neighbors * sizeof(float));
			}
			out_count[q] = sub_cnt[i];
			if (sub_cnt[i] < reachable) {
				still.push_back(q);
			}
		}
		shortq.swap(still);
	}
	common::idx_t under = 0;
	for (common::idx_t q = 0; q < nq; q++) {
		under += out_count[q] < k_neighbors ? 1 : 0;
	}
	return under;
}

// ---------------------------------------------------------------------------------------------------------------
void GpuCoordinatorSearch(core::ISearcher *searcher, core::IGraphManager *graph_manager,
                          const core::LmDiskannConfig &config, bool index_loaded, const float *query_vector,
                          common::idx_t k_neighbors, std::vector<common::row_t> &result_row_ids,
                          common::idx_t search_list_size) {
	// core/Coordinator.cpp:66-74, same order and messages
	if (!index_loaded) {
		throw std::runtime_error("Coordinator: Error - Index not loaded for search.");
	}
	if (!searcher || !graph_manager) {
		throw std::runtime_error("Coordinator: Searcher or GraphManager not initialized.");
	}
	if (k_neighbors == 0) {
		throw std::runtime_error("Coordinator: k_neighbors cannot be 0 for search.");
	}
	result_row_ids.clear();                                                            // :77
	common::idx_t L = search_list_size > 0 ? search_list_size : config.l_search;       // :80
	if (L < k_neighbors) {                                                             // :81-85
		L = k_neighbors;
	}
	try {
		searcher->Search(query_vector, config, graph_manager, k_neighbors, result_row_ids, L); // :91
	} catch (const std::exception &e) {                                                        // :97-101
		std::cerr << "Coordinator: Error during Search: " << e.what() << std::endl;
		result_row_ids.clear();
		throw;
	}
}

std::unique_ptr<GpuScanState> GpuIndexScan::InitializeScan(const float *query, common::idx_t query_dims,
                                                           common::idx_t k) const {
	if (!query) {
		throw ::duckdb::BinderException("LM_DISKANN query vector must be ARRAY<FLOAT>.");
	}
	if (query_dims != config_.dimensions) {
		throw ::duckdb::BinderException(::duckdb::StringUtil::Format(
		    "Query vector dimension (%d) does not match index dimension (%d).", (int)query_dims, (int)config_.dimensions));
	}
	if (k == 0) {
		throw ::duckdb::BinderException("Cannot perform index scan with k=0");
	}
	auto st = std::make_unique<GpuScanState>();
	st->query_vector_storage.assign(query, query + query_dims); // db/LmDiskannScanState.cpp:9-32 copies the query
	st->query_vector_ptr = st->query_vector_storage.data();
	st->k = k;
	st->l_search = config_.l_search;
	return st;
}

common::idx_t GpuIndexScan::Scan(GpuScanState &state, common::row_t *result, common::idx_t result_capacity) {
	if (!searcher_) {
		throw ::duckdb::InternalException("Coordinator not initialized in Scan");
	}
	if (!state.query_vector_ptr) {
		throw ::duckdb::InternalException("Query vector pointer not set in scan state");
	}
	common::idx_t output_count = 0;
	try {
		if (!state.searched) {
			GpuCoordinatorSearch(searcher_, graph_, config_, true, state.query_vector_ptr, state.k,
			                     state.result_row_ids, state.l_search);
			state.searched = true;
		}
		const common::idx_t cap = std::min<common::idx_t>(result_capacity, VECTOR_SIZE);
		output_count = std::min<common::idx_t>(cap, state.result_row_ids.size());
		std::copy(state.result_row_ids.begin(), state.result_row_ids.begin() + output_count, result);
		state.result_row_ids.erase(state.result_row_ids.begin(), state.result_row_ids.begin() + output_count);
	} catch (const std::exception &e) {
		throw ::duckdb::IOException(::duckdb::StringUtil::Format("Error during LM_DISKANN scan: %s", e.what()));
	}
	return output_count;
}

common::idx_t GpuIndexScan::MultiScanFiltered(const AllowedIds &allowed, const float *queries, common::idx_t nq,
                                              common::idx_t query_dims, common::idx_t k, common::row_t *out_ids,
                                              uint32_t *out_count) {
	if (query_dims != config_.dimensions) {
		throw ::duckdb::BinderException(::duckdb::StringUtil::Format(
		    "Query vector dimension (%d) does not match index dimension (%d).", (int)query_dims, (int)config_.dimensions));
	}
	if (k == 0) {
		throw ::duckdb::BinderException("Cannot perform index scan with k=0");
	}
	auto *gpu = dynamic_cast<GpuSearcher *>(searcher_);
	if (!gpu) {
		// the reference's ISearcher has no filtered entry (the plan's dual-heap search was never written)
		throw ::duckdb::NotImplementedException("filtered LM_DISKANN scan needs the GpuSearcher");
	}
	try {
		if (graph_) {
			gpu->Store().SyncEntryPoint(*graph_);
		}
		const common::idx_t L = std::max<common::idx_t>(config_.l_search, k);
		const common::idx_t under = gpu->SearchBatchFiltered(allowed, queries, nq, k, L, out_ids, nullptr, out_count);
		if (under) {
			// the plan: "issue a warning ... suggesting to increase L_search"
			// (plain stdio: formatting an integer through std::cerr crashed inside a Python host whose other extension
			// modules carry their own libstdc++ locale state)
			std::fprintf(stderr,
			             "LM_DISKANN: %llu filtered scan(s) found fewer than k=%llu rows that satisfy the filter; "
			             "consider increasing L_SEARCH\n",
			             (unsigned long long)under, (unsigned long long)k);
		}
		return under;
	} catch (const std::exception &e) {
		throw ::duckdb::IOException(::duckdb::StringUtil::Format("Error during LM_DISKANN scan: %s", e.what()));
	}
}

void GpuIndexScan::MultiScan(const float *queries, common::idx_t nq, common::idx_t query_dims, common::idx_t k,
                             common::row_t *out_ids, uint32_t *out_count) {
	if (query_dims != config_.dimensions) {
		throw ::duckdb::BinderException(::duckdb::StringUtil::Format(
		    "Query vector dimension (%d) does not match index dimension (%d).", (int)query_dims, (int)config_.dimensions));
	}
	if (k == 0) {
		throw ::duckdb::BinderException("Cannot perform index scan with k=0");
	}
	auto *gpu = dynamic_cast<GpuSearcher *>(searcher_);
	try {
		if (gpu) {
			if (graph_) {
				gpu->Store().SyncEntryPoint(*graph_);
			}
			common::idx_t L = std::max<common::idx_t>(config_.l_search, k); // core/Coordinator.cpp:80-85
			gpu->SearchBatch(queries, nq, k, L, out_ids, nullptr, out_count, nullptr);
			return;
		}
		// any other ISearcher: the template's loop, one scan per row (hnsw_optimize_join.cpp:113-170)
		std::vector<common::row_t> ids;
		for (common::idx_t q = 0; q < nq; q++) {
			GpuCoordinatorSearch(searcher_, graph_, config_, true, queries + q * query_dims, k, ids, config_.l_search);
			for (common::idx_t j = 0; j < k; j++) {
				out_ids[q * k + j] = j < ids.size() ? ids[j] : -1;
			}
			if (out_count) {
				out_count[q] = static_cast<uint32_t>(std::min<common::idx_t>(ids.size(), k));
			}
		}
	} catch (const std::exception &e) {
		throw ::duckdb::IOException(::duckdb::StringUtil::Format("Error during LM_DISKANN scan: %s", e.what()));
	}
}

} // namespace b200
} // namespace diskann
