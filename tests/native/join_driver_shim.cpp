// TEST CODE: C entry point over duckdb-lm-diskann_b200/duckdb/lm_diskann_join_driver.hpp with a mock MultiScan (tests/
// test_sql_surface.py). The mock "index" answers query row q (first float = an integer tag t) with min(t % (limit + 2),
// limit) matches whose row ids are 1000 * t + rank, and records every call's batch size.
#include "lm_diskann_join_driver.hpp"

#include <cstring>

extern "C" {

// Runs Execute steps over one input chunk until it is consumed. out_* arrays are sized input_rows * limit.
// Returns the number of MultiScan calls; *steps = Execute calls; batch_sizes[i] = queries of call i (up to 64 recorded).
uint64_t join_driver_run(const float *vectors, uint64_t input_rows, uint64_t dims, uint64_t limit, uint64_t vector_size,
                         uint64_t *out_outer, uint64_t *out_rank, int64_t *out_row_id, uint64_t *out_total,
                         uint64_t *steps, uint64_t *batch_sizes, uint64_t *max_step_rows) {
	diskann::b200::JoinChunkState state;
	uint64_t total = 0, nsteps = 0, calls = 0, max_rows = 0;
	for (;;) {
		std::vector<uint64_t> outer_of, rank_of;
		auto res = diskann::b200::JoinChunkStep(
		    state, vectors, input_rows, dims, limit, vector_size,
		    [&](const float *q, uint64_t nq, int64_t *ids, uint32_t *counts) {
			    if (calls < 64) {
				    batch_sizes[calls] = nq;
			    }
			    calls++;
			    for (uint64_t b = 0; b < nq; b++) {
				    const uint64_t tag = (uint64_t)q[b * dims];
				    const uint64_t m = std::min<uint64_t>(tag % (limit + 2), limit);
				    counts[b] = (uint32_t)m;
				    for (uint64_t i = 0; i < m; i++) {
					    ids[b * limit + i] = (int64_t)(1000 * tag + i + 1);
				    }
			    }
		    },
		    [&](uint64_t out, uint64_t outer_row, uint64_t r) {
			    if (out != outer_of.size()) {
				    outer_of.push_back(~0ull); // out of order: the test will see the hole
			    }
			    outer_of.push_back(outer_row);
			    rank_of.push_back(r);
		    });
		nsteps++;
		max_rows = std::max(max_rows, res.output_rows);
		for (uint64_t i = 0; i < res.output_rows; i++) {
			out_outer[total + i] = outer_of[i];
			out_rank[total + i] = rank_of[i];
			out_row_id[total + i] = state.match_rows[i];
		}
		total += res.output_rows;
		if (res.need_more_input || nsteps > 100000) {
			break;
		}
	}
	*out_total = total;
	*steps = nsteps;
	*max_step_rows = max_rows;
	return state.multi_scans;
}
}
