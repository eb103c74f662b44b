// Index join: `SELECT ... FROM outer, LATERAL (SELECT ... FROM inner ORDER BY metric_fn(inner.vec, outer.vec) LIMIT k)`
// becomes one LM_DISKANN_INDEX_JOIN operator that answers a whole chunk of outer rows with ONE batched search.
// Template: /root/reference/src/hnsw/hnsw_optimize_join.cpp (physical operator :32-186, logical operator :192-326, plan
// match :352-693). The difference that matters: the template's Execute calls ExecuteMultiScan once per outer row
// (:142-155); here the chunk loop is lm_diskann_join_driver.hpp and every step is one GpuIndexScan::MultiScan over up to
// STANDARD_VECTOR_SIZE / k query vectors - the batch the traversal kernel wants (SURVEY.md section 8b "who calls it").
#include "lm_diskann_sql.hpp"
#include "lm_diskann_join_driver.hpp"

#include "duckdb/common/insertion_order_preserving_map.hpp"
#include "duckdb/execution/physical_operator.hpp"
#include "duckdb/execution/physical_plan_generator.hpp"
#include "duckdb/main/config.hpp"
#include "duckdb/optimizer/column_binding_replacer.hpp"
#include "duckdb/optimizer/optimizer.hpp"
#include "duckdb/optimizer/optimizer_extension.hpp"
#include "duckdb/planner/binder.hpp"
#include "duckdb/planner/expression/bound_columnref_expression.hpp"
#include "duckdb/planner/expression/bound_comparison_expression.hpp"
#include "duckdb/planner/expression/bound_constant_expression.hpp"
#include "duckdb/planner/expression/bound_window_expression.hpp"
#include "duckdb/planner/expression_iterator.hpp"
#include "duckdb/planner/operator/logical_cross_product.hpp"
#include "duckdb/planner/operator/logical_delim_get.hpp"
#include "duckdb/planner/operator/logical_extension_operator.hpp"
#include "duckdb/planner/operator/logical_filter.hpp"
#include "duckdb/planner/operator/logical_join.hpp"
#include "duckdb/planner/operator/logical_projection.hpp"
#include "duckdb/planner/operator/logical_window.hpp"
#include "duckdb/storage/data_table.hpp"
#include "duckdb/storage/table/scan_state.hpp"
#include "duckdb/transaction/duck_transaction.hpp"
#include "duckdb/transaction/local_storage.hpp"

#include "core/Coordinator.hpp"

namespace duckdb {

namespace {

// What both operators carry.
struct JoinSpec {
	DuckTableEntry *table = nullptr;
	LmDiskannIndex *index = nullptr;
	idx_t limit = 0;
	vector<column_t> inner_column_ids;   // logical ids of the inner table's columns the join returns
	vector<idx_t> inner_projection_ids;
	vector<LogicalType> inner_returned_types;
	idx_t outer_vector_column = 0;       // position of the query vector in the outer chunk
};

// ---------------------------------------------------------------------------------------------------------------
// Physical operator
// ---------------------------------------------------------------------------------------------------------------
class JoinOperatorState final : public OperatorState {
public:
	::diskann::b200::JoinChunkState chunk;
	ColumnFetchState fetch_state;
	TableScanState local_storage_state;
	vector<StorageIndex> storage_columns;
	SelectionVector outer_sel;
	Vector row_ids = Vector(LogicalType::ROW_TYPE);
	unique_ptr<::diskann::b200::GpuIndexScan> scan; // guards + batching over the index's searcher
	idx_t list_size = 0;
};

class PhysicalLmDiskannIndexJoin final : public PhysicalOperator {
public:
	PhysicalLmDiskannIndexJoin(const vector<LogicalType> &types, idx_t estimated_cardinality, JoinSpec spec_p)
	    : PhysicalOperator(PhysicalOperatorType::EXTENSION, types, estimated_cardinality), spec(std::move(spec_p)) {
	}
	JoinSpec spec;

	string GetName() const override {
		return "LM_DISKANN_INDEX_JOIN";
	}
	bool ParallelOperator() const override {
		return false; // one operator state at a time; the parallelism is inside the batched search
	}

	unique_ptr<OperatorState> GetOperatorState(ExecutionContext &context) const override {
		auto state = make_uniq<JoinOperatorState>();
		state->storage_columns.reserve(spec.inner_column_ids.size());
		for (auto id : spec.inner_column_ids) {
			storage_t storage_id = id;
			if (id != DConstants::INVALID_INDEX) {
				storage_id = spec.table->GetColumn(LogicalIndex(id)).StorageOid();
			}
			state->storage_columns.emplace_back(storage_id);
		}
		state->outer_sel.Initialize();
		state->local_storage_state.Initialize(state->storage_columns, nullptr);
		LocalStorage::Get(context.client, spec.table->catalog)
		    .InitializeScan(spec.table->GetStorage(), state->local_storage_state.local_state, nullptr);
		auto *coordinator = spec.index->GetCoordinator();
		if (!coordinator) {
			throw InternalException("LM_DISKANN index join: the index has no coordinator");
		}
		auto config = coordinator->GetConfig();
		config.l_search = static_cast<uint32_t>(LmDiskannSettings::ListSize(context.client, config));
		state->scan = make_uniq<::diskann::b200::GpuIndexScan>(coordinator->GetSearcher(), coordinator->GetGraphManager(),
		                                                       config);
		return std::move(state);
	}

	OperatorResultType Execute(ExecutionContext &context, DataChunk &input, DataChunk &chunk, GlobalOperatorState &,
	                           OperatorState &ostate) const override {
		auto &state = ostate.Cast<JoinOperatorState>();
		input.Flatten();
		// output layout: [inner columns][row_number][outer columns]
		const idx_t rank_column = spec.inner_column_ids.size();
		const idx_t outer_offset = rank_column + 1;
		auto &vectors = input.data[spec.outer_vector_column];
		const idx_t dims = ArrayType::GetSize(vectors.GetType());
		const float *queries = FlatVector::GetData<float>(ArrayVector::GetEntry(vectors));
		auto *rank = FlatVector::GetData<idx_t>(chunk.data[rank_column]);

		auto &scan = *state.scan;
		const auto step = ::diskann::b200::JoinChunkStep(
		    state.chunk, queries, input.size(), dims, spec.limit, STANDARD_VECTOR_SIZE,
		    [&](const float *q, uint64_t nq, int64_t *ids, uint32_t *counts) {
			    scan.MultiScan(q, nq, dims, spec.limit, ids, counts); // ONE device launch for the whole batch
		    },
		    [&](uint64_t out, uint64_t outer_row, uint64_t r) {
			    state.outer_sel.set_index(out, outer_row);
			    rank[out] = r;
		    });

		// one Fetch of the matched inner rows, then the outer columns sliced next to them
		auto *ids = FlatVector::GetData<row_t>(state.row_ids);
		for (idx_t i = 0; i < step.output_rows; i++) {
			ids[i] = state.chunk.match_rows[i];
		}
		auto &transaction = DuckTransaction::Get(context.client, spec.table->catalog);
		spec.table->GetStorage().Fetch(transaction, chunk, state.storage_columns, state.row_ids, step.output_rows,
		                               state.fetch_state);
		chunk.Slice(input, state.outer_sel, step.output_rows, outer_offset);
		chunk.SetCardinality(step.output_rows);
		return step.need_more_input ? OperatorResultType::NEED_MORE_INPUT : OperatorResultType::HAVE_MORE_OUTPUT;
	}

	InsertionOrderPreservingMap<string> ParamsToString() const override {
		InsertionOrderPreservingMap<string> out;
		out.insert("table", spec.table->name);
		out.insert("index", spec.index->name);
		out.insert("limit", to_string(spec.limit));
		out.insert("batch", to_string(STANDARD_VECTOR_SIZE / (spec.limit ? spec.limit : 1)) + " scans per launch");
		SetEstimatedCardinality(out, estimated_cardinality);
		return out;
	}
};

// ---------------------------------------------------------------------------------------------------------------
// Logical operator
// ---------------------------------------------------------------------------------------------------------------
class LogicalLmDiskannIndexJoin final : public LogicalExtensionOperator {
public:
	LogicalLmDiskannIndexJoin(idx_t table_index_p, JoinSpec spec_p) : table_index(table_index_p), spec(std::move(spec_p)) {
	}
	idx_t table_index;
	JoinSpec spec;

	string GetName() const override {
		return "LM_DISKANN_INDEX_JOIN";
	}

	void ResolveTypes() override {
		if (spec.inner_column_ids.empty()) {
			spec.inner_column_ids.push_back(COLUMN_IDENTIFIER_ROW_ID);
		}
		types.clear();
		auto type_of = [&](column_t id) {
			return id == COLUMN_IDENTIFIER_ROW_ID ? LogicalType(LogicalType::ROW_TYPE) : spec.inner_returned_types[id];
		};
		if (spec.inner_projection_ids.empty()) {
			for (auto id : spec.inner_column_ids) {
				types.push_back(type_of(id));
			}
		} else {
			for (auto p : spec.inner_projection_ids) {
				types.push_back(type_of(spec.inner_column_ids[p]));
			}
		}
		types.emplace_back(LogicalType::BIGINT); // row_number() of the match within its outer row
		for (auto &t : children[0]->types) {
			types.push_back(t);
		}
	}

	vector<ColumnBinding> GetColumnBindings() override {
		vector<ColumnBinding> out;
		if (spec.inner_projection_ids.empty()) {
			for (idx_t i = 0; i < spec.inner_column_ids.size(); i++) {
				out.emplace_back(table_index, i);
			}
		} else {
			for (auto p : spec.inner_projection_ids) {
				out.emplace_back(table_index, p);
			}
		}
		out.emplace_back(table_index, spec.inner_column_ids.size());
		for (auto &b : children[0]->GetColumnBindings()) {
			out.push_back(b);
		}
		return out;
	}

	unique_ptr<PhysicalOperator> CreatePlan(ClientContext &, PhysicalPlanGenerator &generator) override {
		auto op = make_uniq<PhysicalLmDiskannIndexJoin>(types, estimated_cardinality, spec);
		op->children.push_back(generator.CreatePlan(std::move(children[0])));
		return std::move(op);
	}

	idx_t EstimateCardinality(ClientContext &context) override {
		if (!has_estimated_cardinality) {
			estimated_cardinality = children[0]->EstimateCardinality(context) * spec.limit;
			has_estimated_cardinality = true;
		}
		return estimated_cardinality;
	}
};

// ---------------------------------------------------------------------------------------------------------------
// Plan match. The binder turns the LATERAL top-k subquery into
//   DELIM_JOIN( [PROJECTION]( FILTER(row_number <= k)( WINDOW(row_number() ORDER BY dist)( PROJECTION(dist, ...)(
//               CROSS_PRODUCT(DELIM_GET, GET inner) ) ) ) ),  GET outer )
// ---------------------------------------------------------------------------------------------------------------
struct LateralTopK {
	LogicalJoin *delim_join = nullptr;
	unique_ptr<LogicalOperator> *outer_get_slot = nullptr;
	LogicalGet *outer_get = nullptr;
	LogicalProjection *outer_projection = nullptr; // optional
	LogicalFilter *filter = nullptr;
	LogicalWindow *window = nullptr;
	LogicalProjection *inner_projection = nullptr;
	LogicalDelimGet *delim_get = nullptr;
	LogicalGet *inner_get = nullptr;
	idx_t k = 0;
	Expression *distance = nullptr;
};

bool IsOp(const unique_ptr<LogicalOperator> &op, LogicalOperatorType type, idx_t n_children) {
	return op && op->type == type && op->children.size() == n_children;
}

bool MatchLateralTopK(unique_ptr<LogicalOperator> &plan, LateralTopK &m) {
	if (!IsOp(plan, LogicalOperatorType::LOGICAL_DELIM_JOIN, 2) || !IsOp(plan->children[1], LogicalOperatorType::LOGICAL_GET, 0)) {
		return false;
	}
	m.delim_join = &plan->Cast<LogicalJoin>();
	m.outer_get_slot = &plan->children[1];
	m.outer_get = &plan->children[1]->Cast<LogicalGet>();

	auto *below = &plan->children[0];
	if ((*below)->type == LogicalOperatorType::LOGICAL_PROJECTION && (*below)->children.size() == 1) {
		m.outer_projection = &(*below)->Cast<LogicalProjection>(); // the distance is projected away above the filter
		below = &(*below)->children[0];
	}
	if (!IsOp(*below, LogicalOperatorType::LOGICAL_FILTER, 1) ||
	    !IsOp((*below)->children[0], LogicalOperatorType::LOGICAL_WINDOW, 1)) {
		return false;
	}
	m.filter = &(*below)->Cast<LogicalFilter>();
	m.window = &m.filter->children[0]->Cast<LogicalWindow>();
	if (!IsOp(m.window->children[0], LogicalOperatorType::LOGICAL_PROJECTION, 1) ||
	    !IsOp(m.window->children[0]->children[0], LogicalOperatorType::LOGICAL_CROSS_PRODUCT, 2)) {
		return false;
	}
	m.inner_projection = &m.window->children[0]->Cast<LogicalProjection>();
	auto &cross = *m.inner_projection->children[0];
	for (int side = 0; side < 2; side++) {
		auto &a = cross.children[side], &b = cross.children[1 - side];
		if (a->type == LogicalOperatorType::LOGICAL_DELIM_GET && b->type == LogicalOperatorType::LOGICAL_GET) {
			m.delim_get = &a->Cast<LogicalDelimGet>();
			m.inner_get = &b->Cast<LogicalGet>();
		}
	}
	if (!m.inner_get || m.inner_get->function.name != "seq_scan") {
		return false;
	}

	// FILTER: row_number <= constant k, 0 <= k < STANDARD_VECTOR_SIZE
	if (m.filter->expressions.size() != 1 || m.filter->expressions[0]->type != ExpressionType::COMPARE_LESSTHANOREQUALTO) {
		return false;
	}
	const auto &cmp = m.filter->expressions[0]->Cast<BoundComparisonExpression>();
	if (cmp.right->type != ExpressionType::VALUE_CONSTANT || cmp.left->type != ExpressionType::BOUND_COLUMN_REF ||
	    cmp.right->return_type != LogicalType::BIGINT) {
		return false;
	}
	const auto k = cmp.right->Cast<BoundConstantExpression>().value.GetValue<int64_t>();
	const auto &rank_ref = cmp.left->Cast<BoundColumnRefExpression>().binding;
	if (k <= 0 || k >= STANDARD_VECTOR_SIZE || rank_ref.table_index != m.window->window_index || rank_ref.column_index != 0) {
		return false;
	}
	m.k = static_cast<idx_t>(k);

	// WINDOW: row_number() OVER (ORDER BY <projected distance> ASC)
	if (m.window->expressions.size() != 1 || m.window->expressions[0]->type != ExpressionType::WINDOW_ROW_NUMBER) {
		return false;
	}
	const auto &win = m.window->expressions[0]->Cast<BoundWindowExpression>();
	if (win.orders.size() != 1 || win.orders[0].type != OrderType::ASCENDING ||
	    win.orders[0].expression->type != ExpressionType::BOUND_COLUMN_REF) {
		return false;
	}
	const auto &dist_ref = win.orders[0].expression->Cast<BoundColumnRefExpression>().binding;
	if (dist_ref.table_index != m.inner_projection->table_index ||
	    dist_ref.column_index >= m.inner_projection->expressions.size()) {
		return false;
	}
	m.distance = m.inner_projection->expressions[dist_ref.column_index].get();
	return true;
}

class ResetCardinalities final : public LogicalOperatorVisitor {
public:
	explicit ResetCardinalities(ClientContext &context_p) : context(context_p) {
	}
	ClientContext &context;
	void VisitOperator(LogicalOperator &op) override {
		op.has_estimated_cardinality = false;
		VisitOperatorChildren(op);
		op.EstimateCardinality(context);
	}
};

bool RewriteJoin(Binder &binder, ClientContext &context, unique_ptr<LogicalOperator> &root,
                 unique_ptr<LogicalOperator> &plan) {
	LateralTopK m;
	if (!MatchLateralTopK(plan, m)) {
		return false;
	}
	auto table = m.inner_get->GetTable();
	if (!table || !table->IsDuckTable()) {
		return false;
	}
	auto &info = *table->GetStorage().GetDataTableInfo();

	// an LM_DISKANN index on the inner table whose key is one side of the distance, with a GPU searcher behind it
	LmDiskannIndex *index = nullptr;
	vector<reference<Expression>> bindings;
	info.GetIndexes().BindAndScan<LmDiskannIndex>(context, info, [&](LmDiskannIndex &candidate) {
		unique_ptr<Expression> key;
		if (!LmDiskannGpuSearcherOf(candidate) || !LmDiskannMatch::MatchDistanceFunction(candidate, *m.distance, bindings) ||
		    !LmDiskannMatch::BindIndexExpression(candidate, *m.inner_get, key)) {
			return false;
		}
		// inside the subquery the outer row arrives through the DELIM_GET
		ExpressionIterator::EnumerateExpression(key, [&](Expression &e) {
			if (e.type == ExpressionType::BOUND_COLUMN_REF) {
				auto &b = e.Cast<BoundColumnRefExpression>().binding;
				if (b.table_index == m.outer_get->table_index) {
					b.table_index = m.delim_get->table_index;
				}
			}
		});
		if (!LmDiskannMatch::OrientBindings(*key, bindings)) {
			return false;
		}
		index = &candidate;
		return true;
	});
	// plain column references on both sides only: [1] the outer vector, [2] the indexed column
	if (!index || bindings[1].get().type != ExpressionType::BOUND_COLUMN_REF ||
	    bindings[2].get().type != ExpressionType::BOUND_COLUMN_REF) {
		return false;
	}
	const auto &outer_ref = bindings[1].get().Cast<BoundColumnRefExpression>().binding;
	const auto &inner_ref = bindings[2].get().Cast<BoundColumnRefExpression>().binding;
	if (inner_ref.table_index != m.inner_get->table_index) {
		return false;
	}

	JoinSpec spec;
	spec.table = &table->Cast<DuckTableEntry>();
	spec.index = index;
	spec.limit = m.k;
	for (auto &id : m.inner_get->GetColumnIds()) {
		spec.inner_column_ids.emplace_back(id.GetPrimaryIndex());
	}
	spec.inner_projection_ids = m.inner_get->projection_ids;
	spec.inner_returned_types = m.inner_get->returned_types;
	spec.outer_vector_column = outer_ref.column_index;
	auto join = make_uniq<LogicalLmDiskannIndexJoin>(binder.GenerateTableIndex(), std::move(spec));
	const idx_t rank_position = join->spec.inner_column_ids.size();

	// A projection on top re-exports what the DELIM_JOIN exported (minus the window column), so nothing above changes
	// shape; references into the subquery are then inlined down to the new operator.
	ColumnBindingReplacer replacer;
	vector<unique_ptr<Expression>> exported;
	const idx_t top_index = binder.GenerateTableIndex();
	const auto old_bindings = m.delim_join->GetColumnBindings();
	idx_t next = 0;
	for (idx_t i = 0; i < old_bindings.size(); i++) {
		if (old_bindings[i].table_index == m.window->window_index) {
			continue; // row_number is consumed by the filter
		}
		exported.push_back(make_uniq<BoundColumnRefExpression>(m.delim_join->types[i], old_bindings[i]));
		replacer.replacement_bindings.emplace_back(old_bindings[i], ColumnBinding(top_index, next++));
	}
	const ColumnBinding window_binding(m.window->window_index, 0);
	exported.push_back(make_uniq<BoundColumnRefExpression>(LogicalType::BIGINT, window_binding));
	replacer.replacement_bindings.emplace_back(window_binding, ColumnBinding(top_index, next++));
	auto top = make_uniq<LogicalProjection>(top_index, std::move(exported));
	replacer.VisitOperator(*root);
	replacer.replacement_bindings.clear();

	for (auto &expr : top->expressions) {
		auto &ref = expr->Cast<BoundColumnRefExpression>();
		if (ref.binding.table_index == m.inner_get->table_index) {
			ref.binding.table_index = join->table_index;
		}
		if (m.outer_projection && ref.binding.table_index == m.outer_projection->table_index) {
			ref.binding = m.outer_projection->expressions[ref.binding.column_index]->Cast<BoundColumnRefExpression>().binding;
		}
		if (ref.binding.table_index == m.inner_projection->table_index) {
			expr = m.inner_projection->expressions[ref.binding.column_index]->Copy(); // may still name the DELIM_GET
		} else if (ref.binding.table_index == m.window->window_index) {
			expr = make_uniq<BoundColumnRefExpression>(LogicalType::BIGINT, ColumnBinding(join->table_index, rank_position));
		}
	}
	for (auto &b : m.delim_get->GetColumnBindings()) {
		replacer.replacement_bindings.emplace_back(b, ColumnBinding(m.outer_get->table_index, b.column_index));
	}
	for (auto &b : m.inner_get->GetColumnBindings()) {
		replacer.replacement_bindings.emplace_back(b, ColumnBinding(join->table_index, b.column_index));
	}
	replacer.VisitOperator(*top);

	join->children.emplace_back(std::move(*m.outer_get_slot));
	top->children.emplace_back(std::move(join));
	top->EstimateCardinality(context);
	plan = std::move(top);
	ResetCardinalities reset(context);
	reset.VisitOperator(*root);
	return true;
}

void RewriteJoins(OptimizerExtensionInput &input, unique_ptr<LogicalOperator> &root, unique_ptr<LogicalOperator> &plan) {
	if (RewriteJoin(input.optimizer.binder, input.context, root, plan)) {
		return;
	}
	for (auto &child : plan->children) {
		RewriteJoins(input, root, child);
	}
}

void OptimizeJoin(OptimizerExtensionInput &input, unique_ptr<LogicalOperator> &plan) {
	RewriteJoins(input, plan, plan);
}

} // namespace

void LmDiskannModule::RegisterJoinOptimizer(DatabaseInstance &db) {
	OptimizerExtension ext;
	ext.optimize_function = OptimizeJoin;
	db.config.optimizer_extensions.push_back(ext);
}

} // namespace duckdb
