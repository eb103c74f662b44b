// CREATE INDEX ... USING LM_DISKANN: index-type registration, the session settings, and the planner-side matching
// helpers shared by the two optimizer files. Template: /root/reference/src/hnsw/hnsw_index.cpp:655-679 (RegisterIndex),
// :286-299 (per-session list size), :566-648 (TryBindIndexExpression / TryMatchDistanceFunction).
// The index class itself is the reference's own db/LmDiskannIndex (options METRIC, R, L_INSERT, ALPHA, L_SEARCH parsed
// at db/LmDiskannIndex.cpp:72-110); with patches/lm_diskann_index_gpu.patch its coordinator owns a GpuSearcher.
#include "lm_diskann_sql.hpp"

#include "duckdb/execution/index/index_type.hpp"
#include "duckdb/main/config.hpp"
#include "duckdb/planner/expression/bound_columnref_expression.hpp"
#include "duckdb/planner/expression/bound_function_expression.hpp"
#include "duckdb/planner/expression_iterator.hpp"

#include "core/Coordinator.hpp"

namespace duckdb {

// ---------------------------------------------------------------------------------------------------------------
// Settings
// ---------------------------------------------------------------------------------------------------------------
idx_t LmDiskannSettings::ListSize(ClientContext &context, const ::diskann::core::LmDiskannConfig &config) {
	Value opt;
	if (context.TryGetCurrentSetting(L_SEARCH, opt) && !opt.IsNull() && opt.type() == LogicalType::BIGINT) {
		const auto v = opt.GetValue<int64_t>();
		if (v > 0) {
			return static_cast<idx_t>(v);
		}
	}
	return config.l_search;
}

bool LmDiskannSettings::EdgeCodec(ClientContext &context) {
	Value opt;
	if (context.TryGetCurrentSetting(EDGE_CODEC, opt) && !opt.IsNull() && opt.type() == LogicalType::BOOLEAN) {
		return opt.GetValue<bool>();
	}
	return true;
}

::diskann::b200::GpuSearcher *LmDiskannGpuSearcherOf(LmDiskannIndex &index) {
	auto *coordinator = index.GetCoordinator();
	if (!coordinator) {
		return nullptr;
	}
	return dynamic_cast<::diskann::b200::GpuSearcher *>(coordinator->GetSearcher());
}

// ---------------------------------------------------------------------------------------------------------------
// Matching a query against an index
// ---------------------------------------------------------------------------------------------------------------
const char *const *LmDiskannMatch::DistanceFunctions(::diskann::common::LmDiskannMetricType metric, idx_t &count) {
	static const char *const l2[] = {"array_distance", "<->"};
	static const char *const cosine[] = {"array_cosine_distance", "<=>"};
	static const char *const ip[] = {"array_negative_inner_product", "<#>"};
	count = 2;
	switch (metric) {
	case ::diskann::common::LmDiskannMetricType::L2:
		return l2;
	case ::diskann::common::LmDiskannMetricType::COSINE:
		return cosine;
	case ::diskann::common::LmDiskannMetricType::IP:
		return ip;
	default:
		count = 0;
		return nullptr;
	}
}

static bool IsFloatArrayOf(const LogicalType &type, idx_t dims) {
	return type.id() == LogicalTypeId::ARRAY && ArrayType::GetChildType(type).id() == LogicalTypeId::FLOAT &&
	       ArrayType::GetSize(type) == dims;
}

bool LmDiskannMatch::MatchDistanceFunction(const LmDiskannIndex &index, Expression &expr,
                                           vector<reference<Expression>> &bindings) {
	bindings.clear();
	auto *coordinator = index.GetCoordinator();
	if (!coordinator || expr.type != ExpressionType::BOUND_FUNCTION) {
		return false;
	}
	auto &fn = expr.Cast<BoundFunctionExpression>();
	if (fn.children.size() != 2) {
		return false;
	}
	idx_t n_names = 0;
	const auto *names = DistanceFunctions(coordinator->GetConfig().metric_type, n_names);
	bool known = false;
	for (idx_t i = 0; i < n_names; i++) {
		known |= fn.function.name == names[i];
	}
	const auto dims = coordinator->GetConfig().dimensions;
	if (!known || !IsFloatArrayOf(fn.children[0]->return_type, dims) || !IsFloatArrayOf(fn.children[1]->return_type, dims)) {
		return false;
	}
	bindings.emplace_back(expr);
	bindings.emplace_back(*fn.children[0]);
	bindings.emplace_back(*fn.children[1]);
	return true;
}

// The index stores its key as an unbound expression over the table's columns; re-point every column reference in a
// copy of it at the position that column has in `get`'s output.
static void RebindColumns(Expression &expr, idx_t table_index, const vector<column_t> &index_columns,
                          const vector<ColumnIndex> &get_columns, bool &ok, bool &saw_column) {
	if (expr.type == ExpressionType::BOUND_COLUMN_REF) {
		saw_column = true;
		auto &ref = expr.Cast<BoundColumnRefExpression>();
		const auto wanted = index_columns[ref.binding.column_index];
		ref.binding.table_index = table_index;
		for (idx_t i = 0; i < get_columns.size(); i++) {
			if (get_columns[i].GetPrimaryIndex() == wanted) {
				ref.binding.column_index = i;
				return;
			}
		}
		ok = false; // the scan does not produce the indexed column
		return;
	}
	ExpressionIterator::EnumerateChildren(
	    expr, [&](Expression &child) { RebindColumns(child, table_index, index_columns, get_columns, ok, saw_column); });
}

bool LmDiskannMatch::BindIndexExpression(const LmDiskannIndex &index, LogicalGet &get, unique_ptr<Expression> &result) {
	if (index.unbound_expressions.empty()) {
		return false;
	}
	auto copy = index.unbound_expressions.back()->Copy();
	bool ok = true, saw_column = false;
	RebindColumns(*copy, get.table_index, index.GetColumnIds(), get.GetColumnIds(), ok, saw_column);
	if (!ok || !saw_column) {
		return false;
	}
	result = std::move(copy);
	return true;
}

bool LmDiskannMatch::OrientBindings(const Expression &index_expr, vector<reference<Expression>> &bindings) {
	if (bindings.size() != 3) {
		return false;
	}
	if (index_expr.Equals(bindings[2].get())) {
		return !index_expr.Equals(bindings[1].get()); // distance of the column to itself is not a search
	}
	if (index_expr.Equals(bindings[1].get())) {
		std::swap(bindings[1], bindings[2]);
		return true;
	}
	return false;
}

// ---------------------------------------------------------------------------------------------------------------
// Index type
// ---------------------------------------------------------------------------------------------------------------
static unique_ptr<BoundIndex> CreateLmDiskannIndex(CreateIndexInput &input) {
	// estimated cardinality is unknown at this point; the device store grows with the first bulk Append
	return make_uniq<LmDiskannIndex>(input.name, input.constraint_type, input.column_ids, input.table_io_manager,
	                                 input.unbound_expressions, input.db, input.options, input.storage_info, idx_t(0));
}

void LmDiskannModule::RegisterIndexType(DatabaseInstance &db) {
	IndexType type;
	type.name = LmDiskannIndex::TYPE_NAME; // "LM_DISKANN", db/LmDiskannIndex.hpp:82
	type.create_instance = CreateLmDiskannIndex;
	// create_plan stays null: DuckDB then builds the index with its generic sink, which feeds LmDiskannIndex::Append
	// (db/LmDiskannIndex.cpp:350-376); the batch-parallel build (lmd_build_graph) is called from there.
	type.create_plan = nullptr;

	db.config.AddExtensionOption(LmDiskannSettings::L_SEARCH,
	                             "LM_DISKANN: search list size (L) used by index scans of this session; 0 = the index's own",
	                             LogicalType::BIGINT, Value::BIGINT(0));
	db.config.AddExtensionOption(LmDiskannSettings::GPU_DEVICES,
	                             "LM_DISKANN: number of GPUs a newly opened index is mirrored to (query-sharded replicas)",
	                             LogicalType::BIGINT, Value::BIGINT(1));
	db.config.AddExtensionOption(LmDiskannSettings::EDGE_CODEC,
	                             "LM_DISKANN: keep the int8 code sidecar on the device mirror (exact searches fetch only the "
	                             "rows its lower bound cannot exclude; same results; R * dims bytes per node)",
	                             LogicalType::BOOLEAN, Value::BOOLEAN(true));
	db.config.GetIndexTypes().RegisterIndexType(type);
}

} // namespace duckdb
