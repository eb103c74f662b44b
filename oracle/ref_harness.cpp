// TEST INFRASTRUCTURE ONLY (oracle/). Not part of the product path.
//
// C-ABI harness around the reference's OWN hot-path sources, compiled where they lie under
// /root/reference by oracle/Makefile into oracle/_ref/liblmd_ref_{literal,fixed}.so:
//   core/Searcher.cpp      (literal, or patched on the fly with oracle/patches/searcher_fixed.patch)
//   core/index_config.cpp  (CalculateLayoutInternal)
//   core/GraphNode.cpp     (block accessors)
//   core/distance.hpp, core/ternary_quantization.hpp, common/ann.hpp (header-only arithmetic)
// The only things faked are the two seams the reference leaves unimplemented or DuckDB-bound:
//   IStorageManager::GetNodeBlockData  (reference impl throws, StorageManager.cpp:311-327)
//   IGraphManager read API             (reference impl needs AddNode to fill its std::map)
// Both fakes follow the reference's read-side code (GraphManager.cpp:230-257, :523-531).
#include "core/GraphNode.hpp"
#include "core/IGraphManager.hpp"
#include "core/IStorageManager.hpp"
#include "core/Searcher.hpp"
#include "core/distance.hpp"
#include "core/index_config.hpp"
#include "core/ternary_quantization.hpp"

#include <cstdint>
#include <cstring>
#include <unordered_map>
#include <vector>
#include <atomic>
#include <thread>

using namespace diskann;
using core::IGraphManager;
using core::IStorageManager;
using core::LmDiskannConfig;
using core::NodeLayoutOffsets;

namespace {

struct Counters {
	uint64_t block_fetches = 0;
	uint64_t expansions = 0;
};
thread_local Counters g_cnt;

// One contiguous host buffer of n blocks; IndexPointer value = slot + 1 (0 is "invalid").
class FakeStorage : public IStorageManager {
public:
	FakeStorage(const uint8_t *base, uint64_t n, uint64_t block_bytes) : base_(base), n_(n), bb_(block_bytes) {
	}
	common::const_data_ptr_t GetNodeBlockData(common::IndexPointer p) const override {
		uint64_t v = p.Get();
		if (v == 0 || v > n_) {
			return nullptr;
		}
		g_cnt.block_fetches++;
		return base_ + (v - 1) * bb_;
	}
	// Unused on the query path.
	void LoadIndexContents(const std::string &, LmDiskannConfig &, IGraphManager *, common::IndexPointer &,
	                       common::row_t &, common::IndexPointer &) override {
	}
	void InitializeNewStorage(const std::string &, const LmDiskannConfig &) override {
	}
	void SaveIndexContents(const std::string &, const LmDiskannConfig &, const IGraphManager *, common::IndexPointer,
	                       common::row_t, common::IndexPointer) override {
	}
	common::idx_t GetInMemorySize() const override {
		return 0;
	}
	void EnqueueDeletion(common::row_t, common::IndexPointer &) override {
	}
	void ProcessDeletionQueue(common::IndexPointer &) override {
	}
	bool AllocateNodeBlock(common::row_t, common::IndexPointer &, common::data_ptr_t &) override {
		return false;
	}
	common::data_ptr_t GetMutableNodeBlockData(common::IndexPointer) override {
		return nullptr;
	}
	void MarkBlockDirty(common::IndexPointer) override {
	}

private:
	const uint8_t *base_;
	uint64_t n_, bb_;
};

class FakeGraph : public IGraphManager {
public:
	FakeGraph(const FakeStorage *st, const NodeLayoutOffsets &layout, const int64_t *row_ids, uint64_t n,
	          int64_t entry_row_id)
	    : st_(st), layout_(layout), n_(n), entry_row_(entry_row_id) {
		if (row_ids) {
			map_.reserve(n * 2);
			for (uint64_t i = 0; i < n; i++) {
				map_[row_ids[i]] = i;
			}
			identity_ = false;
		}
		if (!TryGetNodePointer(entry_row_, entry_ptr_)) {
			entry_row_ = common::NumericLimits<common::row_t>::Maximum();
		}
	}
	// GraphManager.cpp:523-531 (std::map::find) — identity or hash map here.
	bool TryGetNodePointer(common::row_t r, common::IndexPointer &out) const override {
		if (identity_) {
			if (r < 0 || (uint64_t)r >= n_) {
				out = common::IndexPointer();
				return false;
			}
			out = common::IndexPointer((uint64_t)r + 1);
			return true;
		}
		auto it = map_.find(r);
		if (it == map_.end()) {
			out = common::IndexPointer();
			return false;
		}
		out = common::IndexPointer(it->second + 1);
		return true;
	}
	// GraphManager.cpp:230-257: read count, walk the first `count` ids, drop INT64_MAX sentinels.
	bool GetNeighbors(common::IndexPointer p, std::vector<common::row_t> &out) const override {
		out.clear();
		if (p == common::IndexPointer()) {
			return false;
		}
		common::const_data_ptr_t blk = st_->GetNodeBlockData(p);
		if (!blk) {
			return false;
		}
		g_cnt.block_fetches--; // the id read is accounted as an expansion, not a vector fetch
		g_cnt.expansions++;
		uint16_t count = core::GraphNode::GetNeighborCount(blk, layout_);
		const common::row_t *ids = core::GraphNode::GetNeighborIDsPtr(blk, layout_);
		out.reserve(count);
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
	// Unused on the query path.
	void InitializeEmptyGraph(const LmDiskannConfig &, common::IndexPointer &, common::row_t &) override {
	}
	bool AddNode(common::row_t, const float *, common::idx_t, common::IndexPointer &) override {
		return false;
	}
	bool GetNodeVector(common::IndexPointer, float *, common::idx_t) const override {
		return false;
	}
	void RobustPrune(common::IndexPointer, const float *, std::vector<common::row_t> &,
	                 const LmDiskannConfig &) override {
	}
	void SetEntryPoint(common::IndexPointer, common::row_t) override {
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
	const FakeStorage *st_;
	NodeLayoutOffsets layout_;
	uint64_t n_;
	int64_t entry_row_;
	common::IndexPointer entry_ptr_;
	bool identity_ = true;
	std::unordered_map<int64_t, uint64_t> map_;
};

LmDiskannConfig MakeConfig(int metric, int vec_type, uint64_t dims, uint32_t r, uint32_t l_search) {
	LmDiskannConfig c;
	c.metric_type = static_cast<common::LmDiskannMetricType>(metric);
	c.node_vector_type = static_cast<common::LmDiskannVectorType>(vec_type);
	c.dimensions = dims;
	c.r = r;
	c.l_search = l_search;
	return c;
}

} // namespace

extern "C" {

#ifdef LMD_REF_FIXED
int lmdref_is_fixed() {
	return 1;
}
#else
int lmdref_is_fixed() {
	return 0;
}
#endif

// out[7] = count@, vector@, ids@, pos@, neg@, total_node_size, ternary_edge_size_bytes
int lmdref_layout(uint64_t dims, int vec_type, uint32_t r, uint64_t *out) {
	try {
		LmDiskannConfig c = MakeConfig(1, vec_type, dims, r, 0);
		NodeLayoutOffsets l = core::CalculateLayoutInternal(c);
		out[0] = l.neighbor_count_offset;
		out[1] = l.node_vector_offset;
		out[2] = l.neighbor_ids_offset;
		out[3] = l.neighbor_pos_planes_offset;
		out[4] = l.neighbor_neg_planes_offset;
		out[5] = l.total_node_size;
		out[6] = l.ternary_edge_size_bytes;
		return 0;
	} catch (...) {
		return 1;
	}
}

// LmDiskannIndex.cpp:161-162: block = AlignValue<SECTOR_SIZE>(total_node_size)
uint64_t lmdref_block_bytes(uint64_t total_node_size) {
	return ::duckdb::AlignValue<uint64_t, ::duckdb::Storage::SECTOR_SIZE>(total_node_size);
}

float lmdref_distance(const float *a, const float *b, uint64_t dims, int metric, int *err) {
	try {
		if (err) {
			*err = 0;
		}
		return core::ComputeExactDistanceFloat(a, b, dims, static_cast<common::LmDiskannMetricType>(metric));
	} catch (...) {
		if (err) {
			*err = 1;
		}
		return 0.0f;
	}
}

int lmdref_convert_raw(const uint8_t *raw, float *out, uint64_t dims, int vec_type) {
	return common::ConvertRawVectorToFloat(raw, out, dims, static_cast<common::LmDiskannVectorType>(vec_type)) ? 0 : 1;
}

uint64_t lmdref_words_per_plane(uint64_t dims) {
	return core::WordsPerPlane(dims);
}

void lmdref_encode_ternary(const float *v, uint64_t dims, uint64_t *pos, uint64_t *neg) {
	core::EncodeTernary<float>(v, pos, neg, dims);
}

int64_t lmdref_ternary_dot(const uint64_t *qp, const uint64_t *qn, const uint64_t *vp, const uint64_t *vn,
                           uint64_t words) {
	return core::GetDotKernel()(qp, qn, vp, vn, words);
}

int64_t lmdref_ternary_dot_scalar(const uint64_t *qp, const uint64_t *qn, const uint64_t *vp, const uint64_t *vn,
                                  uint64_t words) {
	return core::Dot_scalar(qp, qn, vp, vn, words);
}

// distance.hpp:253-270 — [pos‖neg] contiguous per neighbour
int lmdref_compress_edge(const float *v, uint8_t *out, uint64_t dims) {
	LmDiskannConfig c = MakeConfig(2, 1, dims, 1, 0);
	return core::CompressVectorForEdge(v, out, c) ? 0 : 1;
}

// distance.hpp:212-243
float lmdref_approx_distance(const float *q, const uint8_t *compressed, uint64_t dims, int metric, int *err) {
	try {
		if (err) {
			*err = 0;
		}
		LmDiskannConfig c = MakeConfig(metric, 1, dims, 1, 0);
		return core::CalculateApproxDistance(q, compressed, c);
	} catch (...) {
		if (err) {
			*err = 1;
		}
		return 0.0f;
	}
}

// Runs Searcher::Search (Searcher.cpp:256-272 → PerformSearch :45-246) per query.
// out_ids [nq*k] padded with -1; out_dist [nq*k] = ComputeExactDistanceFloat of each returned id
// (the reference returns ids only); out_count [nq]; out_stats [nq*2] = {expansions, vector fetches}.
int lmdref_search(const uint8_t *blocks, uint64_t n, uint64_t block_bytes, const int64_t *row_ids,
                  int64_t entry_row_id, int metric, int vec_type, uint64_t dims, uint32_t r, uint32_t cfg_l_search,
                  const float *queries, uint64_t nq, uint64_t k, uint64_t l_param, int64_t *out_ids, float *out_dist,
                  uint32_t *out_count, uint64_t *out_stats, int nthreads) {
	try {
		LmDiskannConfig cfg = MakeConfig(metric, vec_type, dims, r, cfg_l_search);
		NodeLayoutOffsets layout = core::CalculateLayoutInternal(cfg);
		FakeStorage storage(blocks, n, block_bytes);
		FakeGraph graph(&storage, layout, row_ids, n, entry_row_id);
		std::atomic<int> failed {0};
		std::atomic<int64_t> next {0};
		if (nthreads <= 0) {
			nthreads = (int)std::thread::hardware_concurrency();
			if (nthreads <= 0) {
				nthreads = 1;
			}
		}
		// Queries are independent (Searcher.cpp:63-67: all state is per call); dynamic chunks of 8.
		auto worker = [&]() {
			for (;;) {
				int64_t q0 = next.fetch_add(8);
				if (q0 >= (int64_t)nq) {
					break;
				}
				int64_t q1 = q0 + 8 < (int64_t)nq ? q0 + 8 : (int64_t)nq;
				for (int64_t qi = q0; qi < q1; qi++) {
					try {
						core::Searcher searcher(&storage);
						std::vector<common::row_t> res;
						g_cnt = Counters();
						const float *q = queries + qi * dims;
						searcher.Search(q, cfg, &graph, k, res, l_param);
						uint64_t cnt = res.size() < k ? res.size() : k;
						if (out_count) {
							out_count[qi] = (uint32_t)cnt;
						}
						if (out_stats) {
							out_stats[qi * 2 + 0] = g_cnt.expansions;
							out_stats[qi * 2 + 1] = g_cnt.block_fetches;
						}
						std::vector<float> tmp(dims);
						for (uint64_t j = 0; j < k; j++) {
							int64_t id = j < cnt ? res[j] : -1;
							out_ids[qi * k + j] = id;
							if (out_dist) {
								float d = 0.0f;
								common::IndexPointer p;
								if (id >= 0 && graph.TryGetNodePointer(id, p)) {
									common::const_data_ptr_t blk = storage.GetNodeBlockData(p);
									common::ConvertRawVectorToFloat(blk + layout.node_vector_offset, tmp.data(), dims,
									                                cfg.node_vector_type);
									d = core::ComputeExactDistanceFloat(q, tmp.data(), dims, cfg.metric_type);
								}
								out_dist[qi * k + j] = d;
							}
						}
					} catch (...) {
						failed = 1;
					}
				}
			}
		};
		if (nthreads == 1) {
			worker();
		} else {
			std::vector<std::thread> pool;
			for (int t = 0; t < nthreads; t++) {
				pool.emplace_back(worker);
			}
			for (auto &t : pool) {
				t.join();
			}
		}
		return failed.load();
	} catch (...) {
		return 2;
	}
}

} // extern "C"
