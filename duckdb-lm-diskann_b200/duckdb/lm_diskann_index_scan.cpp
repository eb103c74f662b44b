// Table function lm_diskann_index_scan: what a rewritten `ORDER BY array_distance(col, q) LIMIT k` executes.
// Template: /root/reference/src/hnsw/hnsw_index_scan.cpp:39-153. InitGlobal -> LmDiskannIndex::InitializeScan
// (db/LmDiskannIndex.cpp:639-667), Execute -> LmDiskannIndex::Scan (:677-724) + one Fetch of the row ids it returned.
#include "lm_diskann_sql.hpp"

#include "duckdb/common/insertion_order_preserving_map.hpp"
#include "duckdb/main/extension_util.hpp"
#include "duckdb/storage/data_table.hpp"
#include "duckdb/storage/table/scan_state.hpp"
#include "duckdb/transaction/duck_transaction.hpp"
#include "duckdb/transaction/local_storage.hpp"

#include "core/Coordinator.hpp"

namespace duckdb {

namespace {

struct ScanGlobalState final : public GlobalTableFunctionState {
	ColumnFetchState fetch_state;
	TableScanState local_storage_state;
	vector<StorageIndex> storage_columns;
	unique_ptr<IndexScanState> index_state;
	Vector row_ids = Vector(LogicalType::ROW_TYPE);
	Vector query = Vector(LogicalType::ARRAY(LogicalType::FLOAT, 1)); // re-typed in Init
};

unique_ptr<GlobalTableFunctionState> InitGlobal(ClientContext &context, TableFunctionInitInput &input) {
	auto &bind = input.bind_data->Cast<LmDiskannIndexScanBindData>();
	auto state = make_uniq<ScanGlobalState>();

	// logical column ids of the projection -> storage column ids
	state->storage_columns.reserve(input.column_ids.size());
	for (auto id : input.column_ids) {
		storage_t storage_id = id;
		if (id != DConstants::INVALID_INDEX) {
			storage_id = bind.table.GetColumn(LogicalIndex(id)).StorageOid();
		}
		state->storage_columns.emplace_back(storage_id);
	}
	state->local_storage_state.Initialize(state->storage_columns, input.filters.get());
	LocalStorage::Get(context, bind.table.catalog)
	    .InitializeScan(bind.table.GetStorage(), state->local_storage_state.local_state, input.filters);

	// the query vector as the FLOAT[dims] Vector LmDiskannIndex::InitializeScan validates
	const auto dims = bind.index.GetDimensions();
	state->query = Vector(LogicalType::ARRAY(LogicalType::FLOAT, dims));
	auto *dst = FlatVector::GetData<float>(ArrayVector::GetEntry(state->query));
	for (idx_t i = 0; i < dims; i++) {
		dst[i] = bind.query[i];
	}
	state->index_state = bind.index.InitializeScan(context, state->query, bind.limit);
	// per-session list size (SET lm_diskann_l_search = ...), applied on top of config.l_search
	if (auto *coordinator = bind.index.GetCoordinator()) {
		auto &scan = state->index_state->Cast<::diskann::db::LmDiskannScanState>();
		scan.l_search = LmDiskannSettings::ListSize(context, coordinator->GetConfig());
	}
	return std::move(state);
}

void Execute(ClientContext &context, TableFunctionInput &data, DataChunk &output) {
	auto &bind = data.bind_data->Cast<LmDiskannIndexScanBindData>();
	auto &state = data.global_state->Cast<ScanGlobalState>();
	const auto rows = bind.index.Scan(*state.index_state, state.row_ids);
	if (rows == 0) {
		output.SetCardinality(0);
		return;
	}
	auto &transaction = DuckTransaction::Get(context, bind.table.catalog);
	bind.table.GetStorage().Fetch(transaction, output, state.storage_columns, state.row_ids, rows, state.fetch_state);
}

unique_ptr<BaseStatistics> Statistics(ClientContext &context, const FunctionData *bind_data, column_t column_id) {
	auto &bind = bind_data->Cast<LmDiskannIndexScanBindData>();
	if (LocalStorage::Get(context, bind.table.catalog).Find(bind.table.GetStorage())) {
		return nullptr; // transaction-local rows are not covered by the table statistics
	}
	return bind.table.GetStatistics(context, column_id);
}

void Dependency(LogicalDependencyList &entries, const FunctionData *bind_data) {
	entries.AddDependency(bind_data->Cast<LmDiskannIndexScanBindData>().table);
}

unique_ptr<NodeStatistics> Cardinality(ClientContext &, const FunctionData *bind_data) {
	const auto k = bind_data->Cast<LmDiskannIndexScanBindData>().limit;
	return make_uniq<NodeStatistics>(k, k);
}

InsertionOrderPreservingMap<string> ToString(TableFunctionToStringInput &input) {
	InsertionOrderPreservingMap<string> out;
	auto &bind = input.bind_data->Cast<LmDiskannIndexScanBindData>();
	out["Table"] = bind.table.name;
	out["LM_DISKANN Index"] = bind.index.GetIndexName();
	out["k"] = to_string(bind.limit);
	return out;
}

BindInfo GetBindInfo(const optional_ptr<FunctionData> bind_data) {
	return BindInfo(bind_data->Cast<LmDiskannIndexScanBindData>().table);
}

} // namespace

TableFunction LmDiskannIndexScanFunction::GetFunction() {
	TableFunction fn(NAME, {}, Execute);
	fn.init_global = InitGlobal;
	fn.init_local = nullptr;
	fn.statistics = Statistics;
	fn.dependency = Dependency;
	fn.cardinality = Cardinality;
	fn.to_string = ToString;
	fn.get_bind_info = GetBindInfo;
	fn.projection_pushdown = true;
	fn.filter_pushdown = false; // table filters are pulled above the scan by the optimizer (lm_diskann_optimize_scan.cpp)
	return fn;
}

void LmDiskannModule::RegisterIndexScan(DatabaseInstance &db) {
	ExtensionUtil::RegisterFunction(db, LmDiskannIndexScanFunction::GetFunction());
}

} // namespace duckdb
