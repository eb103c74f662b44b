// SQL surface of the B200 LM_DISKANN query path: the DuckDB-typed sources a maintainer adds
// to the extension next to src/lm_diskann/. The reference ships this surface only for its HNSW template
// (/root/reference/src/hnsw/); these files give LM_DISKANN the same four hooks, written for the batched GPU searcher:
//
//   lm_diskann_index_type.cpp     CREATE INDEX ... USING LM_DISKANN: IndexType{name, create_instance, create_plan}
//                                 registration + the session settings         (template: hnsw_index.cpp:655-679, :286-299)
//   lm_diskann_index_scan.cpp     table function lm_diskann_index_scan: InitGlobal -> InitializeScan, Execute -> Scan +
//                                 Fetch                                         (template: hnsw_index_scan.cpp:39-153)
//   lm_diskann_optimize_scan.cpp  ORDER BY array_distance(col, const) LIMIT k   ->  index scan
//                                                                               (template: hnsw_optimize_scan.cpp:29-258)
//   lm_diskann_optimize_join.cpp  LATERAL top-k join -> LM_DISKANN_INDEX_JOIN, ONE GpuIndexScan::MultiScan per chunk of
//                                 up to STANDARD_VECTOR_SIZE / k outer rows     (template: hnsw_optimize_join.cpp:113-170)
//
// They compile against DuckDB v1.2.x headers (DUCKDB_INC=<duckdb>/src/include) plus the reference's db/LmDiskannIndex.hpp
// with patches/lm_diskann_index_gpu.patch applied (the accessor the batch path needs); this image has no DuckDB, so here
// they are syntax-checked against include/duckdb_shim (`make -C duckdb-lm-diskann_b200/duckdb check`). DuckDB CLI:
// unavailable on this box - nothing in this directory is executed here except lm_diskann_join_driver.hpp (plain C++,
// driven by tests/test_sql_surface.py).
#pragma once

#include "duckdb/catalog/catalog_entry/duck_table_entry.hpp"
#include "duckdb/common/types/data_chunk.hpp"
#include "duckdb/common/types/value.hpp"
#include "duckdb/function/table_function.hpp"
#include "duckdb/main/client_context.hpp"
#include "duckdb/main/database.hpp"
#include "duckdb/planner/expression.hpp"
#include "duckdb/planner/operator/logical_get.hpp"

#include "db/LmDiskannIndex.hpp" // the reference's index class (+ GetCoordinator(), patches/lm_diskann_index_gpu.patch)
#include "gpu_searcher.hpp"      // duckdb-lm-diskann_b200/host/

namespace duckdb {

using LmDiskannIndex = ::diskann::db::LmDiskannIndex;

// Session settings (SET lm_diskann_l_search = 128; ...)
struct LmDiskannSettings {
	static constexpr const char *L_SEARCH = "lm_diskann_l_search";   // search list size override, 0 / unset = index default
	static constexpr const char *GPU_DEVICES = "lm_diskann_gpu_devices"; // how many devices a new index mirrors itself to
	// int8 code sidecar on the device mirror (DeviceNodeStore::SetEdgeCodec): same results, fewer vector bytes per search
	static constexpr const char *EDGE_CODEC = "lm_diskann_edge_codec";
	static bool EdgeCodec(ClientContext &context); // the setting; true when unset
	// L for a scan of `index` in `context`: the setting when it is a positive BIGINT, else config.l_search
	// (core/Searcher.cpp:261-265 resolves 0 further)
	static idx_t ListSize(ClientContext &context, const ::diskann::core::LmDiskannConfig &config);
};

// What the planner needs to know about an index to match a query against it.
struct LmDiskannMatch {
	// Names of the distance functions / operators whose ORDER BY an index with this metric answers
	// (common/ann.hpp:16-24; array_distance is true L2 like core/distance.hpp:56-67, so the order is the index's own).
	static const char *const *DistanceFunctions(::diskann::common::LmDiskannMetricType metric, idx_t &count);
	// `expr` is metric_fn(a, b) over two FLOAT[dims] arguments -> bindings = {expr, a, b}
	static bool MatchDistanceFunction(const LmDiskannIndex &index, Expression &expr, vector<reference<Expression>> &bindings);
	// The index's own (unbound) key expression re-bound to the columns of `get`; false when `get` does not produce it
	static bool BindIndexExpression(const LmDiskannIndex &index, LogicalGet &get, unique_ptr<Expression> &result);
	// Of bindings[1] / bindings[2]: the one that is the index expression goes to [2]; false when neither is
	static bool OrientBindings(const Expression &index_expr, vector<reference<Expression>> &bindings);
};

// Bind data of the index-scan table function: one query vector, one limit.
struct LmDiskannIndexScanBindData final : public TableFunctionData {
	LmDiskannIndexScanBindData(DuckTableEntry &table_p, LmDiskannIndex &index_p, idx_t limit_p,
	                           unsafe_unique_array<float> query_p)
	    : table(table_p), index(index_p), limit(limit_p), query(std::move(query_p)) {
	}
	DuckTableEntry &table;
	LmDiskannIndex &index;
	idx_t limit;
	unsafe_unique_array<float> query;
};

struct LmDiskannIndexScanFunction {
	static constexpr const char *NAME = "lm_diskann_index_scan";
	static TableFunction GetFunction();
};

// The batched searcher behind an index, when its coordinator was built with one; nullptr = the index
// still runs the reference's CPU searcher and only the single-query path applies.
::diskann::b200::GpuSearcher *LmDiskannGpuSearcherOf(LmDiskannIndex &index);

struct LmDiskannModule {
	static void RegisterIndexType(DatabaseInstance &db);
	static void RegisterIndexScan(DatabaseInstance &db);
	static void RegisterScanOptimizer(DatabaseInstance &db);
	static void RegisterJoinOptimizer(DatabaseInstance &db);
	// everything above, in the order the extension's Load() calls it (src/lm_diskann_extension.cpp)
	static void Register(DatabaseInstance &db) {
		RegisterIndexType(db);
		RegisterIndexScan(db);
		RegisterScanOptimizer(db);
		RegisterJoinOptimizer(db);
	}
};

} // namespace duckdb
