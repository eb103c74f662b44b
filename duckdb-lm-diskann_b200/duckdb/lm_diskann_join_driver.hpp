// The chunk loop of the LM_DISKANN index join, free of DuckDB types so that it can be driven by a mock host
// (tests/test_sql_surface.py -> tests/sql_surface_harness.cpp) as well as by PhysicalLmDiskannIndexJoin::Execute.
//
// The template it replaces (hnsw_optimize_join.cpp:113-170) calls ExecuteMultiScan once per OUTER ROW - a batch of
// one. Here the outer rows of a chunk that fit one output vector (STANDARD_VECTOR_SIZE / limit of them) go to the index
// as ONE batch: one GpuIndexScan::MultiScan = one lmd_search_batch = one traversal launch with up to 2 048 queries.
#pragma once

#include <algorithm>
#include <cstdint>
#include <vector>

namespace diskann {
namespace b200 {

struct JoinChunkResult {
	uint64_t outer_rows = 0;   // outer rows consumed by this call
	uint64_t output_rows = 0;  // matches produced (<= vector_size)
	bool need_more_input = false; // the whole input chunk has been consumed
};

// State kept between Execute calls on the same input chunk.
struct JoinChunkState {
	uint64_t input_idx = 0;
	std::vector<int64_t> row_ids;   // [batch][limit], -1 padded (MultiScan output)
	std::vector<uint32_t> counts;   // [batch]
	std::vector<int64_t> match_rows; // compacted row ids of this call, output order
	uint64_t multi_scans = 0;       // MultiScan calls so far (statistics / tests)
};

// One Execute step. `vectors` = the outer chunk's query vectors [input_rows][dims] (flattened); `multi_scan(queries, nq,
// out_ids, out_count)` = GpuIndexScan::MultiScan bound to (dims, limit). For every match i of outer row r, in order,
// `emit(output_index, r, rank)` is called (rank is 1-based, the window's row_number()); state.match_rows holds the
// matching inner row ids for the one Fetch that follows.
template <class MultiScan, class Emit>
JoinChunkResult JoinChunkStep(JoinChunkState &state, const float *vectors, uint64_t input_rows, uint64_t dims,
                              uint64_t limit, uint64_t vector_size, MultiScan &&multi_scan, Emit &&emit) {
	JoinChunkResult res;
	state.match_rows.clear();
	if (limit == 0 || limit > vector_size) {
		res.need_more_input = true;
		state.input_idx = 0;
		return res;
	}
	const uint64_t left = input_rows - state.input_idx;
	const uint64_t batch = std::min<uint64_t>(left, vector_size / limit);
	if (batch > 0) {
		state.row_ids.assign(batch * limit, -1);
		state.counts.assign(batch, 0);
		multi_scan(vectors + state.input_idx * dims, batch, state.row_ids.data(), state.counts.data());
		state.multi_scans++;
		for (uint64_t b = 0; b < batch; b++) {
			const uint64_t outer = state.input_idx + b;
			const uint64_t m = std::min<uint64_t>(state.counts[b], limit);
			for (uint64_t i = 0; i < m; i++) {
				emit(res.output_rows, outer, i + 1);
				state.match_rows.push_back(state.row_ids[b * limit + i]);
				res.output_rows++;
			}
		}
		state.input_idx += batch;
		res.outer_rows = batch;
	}
	if (state.input_idx == input_rows) {
		state.input_idx = 0;
		res.need_more_input = true;
	}
	return res;
}

} // namespace b200
} // namespace diskann
