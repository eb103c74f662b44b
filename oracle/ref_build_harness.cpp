// TEST INFRASTRUCTURE ONLY (oracle/). Not part of the product path.
//
// C-ABI harness around the reference's OWN graph WRITE side, compiled where the files lie
// under /root/reference by oracle/Makefile into oracle/_ref/liblmd_ref_{literal,fixed}.so:
//   core/GraphManager.cpp  AddNode (:65-203), RobustPrune (:259-457), GetNeighbors, TryGetNodePointer, ...
//   core/Searcher.cpp      SearchForInitialCandidates (:275-294) -> PerformSearch(final_pass = false)
// Only IStorageManager is faked (the reference's implementation throws, StorageManager.cpp:294-348): blocks are
// handed out of one caller-owned buffer, IndexPointer value = slot + 1.
//
// Note on the two builds: with the untouched Searcher.cpp, PerformSearch(final_pass = false) returns no ids
// (Searcher.cpp:189), so the literal library builds an EDGELESS graph; the fixed library
// (oracle/patches/searcher_fixed.patch, hunk H5) hands the top list to AddNode as the code around it intends.
#include "core/GraphManager.hpp"
#include "core/GraphNode.hpp"
#include "core/IStorageManager.hpp"
#include "core/Searcher.hpp"
#include "core/index_config.hpp"

#include <cstdint>
#include <cstring>
#include <vector>

using namespace diskann;
using core::IGraphManager;
using core::IStorageManager;
using core::LmDiskannConfig;
using core::NodeLayoutOffsets;

namespace {

class BuildStorage : public IStorageManager {
public:
	BuildStorage(uint8_t *base, uint64_t capacity, uint64_t block_bytes) : base_(base), cap_(capacity), bb_(block_bytes) {
	}
	bool AllocateNodeBlock(common::row_t, common::IndexPointer &node_ptr_out, common::data_ptr_t &node_data_out) override {
		if (used_ >= cap_) {
			return false;
		}
		node_ptr_out = common::IndexPointer(used_ + 1);
		node_data_out = base_ + used_ * bb_;
		used_++;
		return true;
	}
	common::const_data_ptr_t GetNodeBlockData(common::IndexPointer p) const override {
		uint64_t v = p.Get();
		return (v == 0 || v > used_) ? nullptr : base_ + (v - 1) * bb_;
	}
	common::data_ptr_t GetMutableNodeBlockData(common::IndexPointer p) override {
		uint64_t v = p.Get();
		return (v == 0 || v > used_) ? nullptr : base_ + (v - 1) * bb_;
	}
	void MarkBlockDirty(common::IndexPointer) override {
	}
	void LoadIndexContents(const std::string &, LmDiskannConfig &, IGraphManager *, common::IndexPointer &,
	                       common::row_t &, common::IndexPointer &) override {
	}
	void InitializeNewStorage(const std::string &, const LmDiskannConfig &) override {
	}
	void SaveIndexContents(const std::string &, const LmDiskannConfig &, const IGraphManager *, common::IndexPointer,
	                       common::row_t, common::IndexPointer) override {
	}
	common::idx_t GetInMemorySize() const override {
		return used_ * bb_;
	}
	void EnqueueDeletion(common::row_t, common::IndexPointer &) override {
	}
	void ProcessDeletionQueue(common::IndexPointer &) override {
	}

private:
	uint8_t *base_;
	uint64_t cap_, bb_, used_ = 0;
};

// Finds nothing: AddNode then only allocates the block, copies the vector and registers the row id.
class NullSearcher : public core::ISearcher {
public:
	void Search(const float *, const LmDiskannConfig &, IGraphManager *, common::idx_t, std::vector<common::row_t> &r,
	            common::idx_t) override {
		r.clear();
	}
	common::idx_t GetInMemorySize() const override {
		return 0;
	}
	void SearchForInitialCandidates(const float *, common::idx_t, const LmDiskannConfig &, IGraphManager *, common::idx_t,
	                                common::IndexPointer, std::vector<common::row_t> &r) override {
		r.clear();
	}
};

LmDiskannConfig BuildConfig(int metric, int vec_type, uint64_t dims, uint32_t r, uint32_t l_insert, float alpha) {
	LmDiskannConfig c;
	c.metric_type = static_cast<common::LmDiskannMetricType>(metric);
	c.node_vector_type = static_cast<common::LmDiskannVectorType>(vec_type);
	c.dimensions = dims;
	c.r = r;
	c.l_insert = l_insert;
	c.alpha = alpha;
	c.l_search = l_insert;
	return c;
}

} // namespace

extern "C" {

// Serial build: GraphManager::AddNode for rows 0..n-1 in order (row id = row_ids[i], or i). out_blocks [n][block_bytes]
// receives the node blocks exactly as the reference leaves them (ternary planes stay zero: nothing in the reference
// writes them, GraphManager.cpp:402-444). Returns 0 on success.
int lmdref_build_serial(const float *vectors, uint64_t n, uint64_t dims, int metric, int vec_type, uint32_t r,
                        uint32_t l_insert, float alpha, const int64_t *row_ids, uint64_t block_bytes, uint8_t *out_blocks,
                        int64_t *entry_row_out) {
	try {
		LmDiskannConfig cfg = BuildConfig(metric, vec_type, dims, r, l_insert, alpha);
		NodeLayoutOffsets layout = core::CalculateLayoutInternal(cfg);
		BuildStorage storage(out_blocks, n, block_bytes);
		core::Searcher searcher(&storage);
		core::GraphManager gm(cfg, layout, block_bytes, &storage, &searcher);
		for (uint64_t i = 0; i < n; i++) {
			common::IndexPointer p;
			if (!gm.AddNode(row_ids ? row_ids[i] : (int64_t)i, vectors + i * dims, dims, p)) {
				return 3;
			}
		}
		if (entry_row_out) {
			*entry_row_out = gm.GetEntryPointRowId();
		}
		return 0;
	} catch (...) {
		return 2;
	}
}

// GraphManager::RobustPrune (GraphManager.cpp:259-457) for m nodes over a fixed set of n vectors.
// node_idx [m] = position of the node to connect; cands [m][c] candidate ROW IDS (INT64_MAX = no candidate);
// out_ids [m][r] (INT64_MAX padded, as the reference writes the block), out_count [m].
int lmdref_robust_prune(const float *vectors, uint64_t n, uint64_t dims, int metric, int vec_type, uint32_t r, float alpha,
                        const int64_t *row_ids, uint64_t m, const uint64_t *node_idx, const int64_t *cands, uint64_t c,
                        int64_t *out_ids, uint32_t *out_count) {
	try {
		LmDiskannConfig cfg = BuildConfig(metric, vec_type, dims, r, 1, alpha);
		NodeLayoutOffsets layout = core::CalculateLayoutInternal(cfg);
		const uint64_t bb = ::duckdb::AlignValue<uint64_t, ::duckdb::Storage::SECTOR_SIZE>(layout.total_node_size);
		std::vector<uint8_t> blocks(n * bb);
		BuildStorage storage(blocks.data(), n, bb);
		NullSearcher searcher;
		core::GraphManager gm(cfg, layout, bb, &storage, &searcher);
		std::vector<common::IndexPointer> ptrs(n);
		for (uint64_t i = 0; i < n; i++) {
			if (!gm.AddNode(row_ids ? row_ids[i] : (int64_t)i, vectors + i * dims, dims, ptrs[i])) {
				return 3;
			}
		}
		std::vector<float> self(dims);
		for (uint64_t j = 0; j < m; j++) {
			const uint64_t i = node_idx[j];
			std::vector<common::row_t> cand;
			for (uint64_t t = 0; t < c; t++) {
				if (cands[j * c + t] != common::NumericLimits<common::row_t>::Maximum()) {
					cand.push_back(cands[j * c + t]);
				}
			}
			// the node's own vector as the reference sees it (through the block: INT8 nodes are cast on write)
			if (!gm.GetNodeVector(ptrs[i], self.data(), dims)) {
				return 4;
			}
			gm.RobustPrune(ptrs[i], self.data(), cand, cfg);
			out_count[j] = (uint32_t)cand.size();
			for (uint32_t t = 0; t < r; t++) {
				out_ids[j * r + t] = t < cand.size() ? cand[t] : common::NumericLimits<common::row_t>::Maximum();
			}
		}
		return 0;
	} catch (...) {
		return 2;
	}
}

} // extern "C"
