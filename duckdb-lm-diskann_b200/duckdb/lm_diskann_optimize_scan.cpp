// Plan rewrite: TopN(k) over a projection of metric_fn(indexed_column, constant) over a sequential scan  ->
// lm_diskann_index_scan. Template: /root/reference/src/hnsw/hnsw_optimize_scan.cpp:29-258; conditions as SURVEY.md
// section 8b lists them (single ASC order key, the key is a projection column, the projection sits on a seq_scan without
// dynamic filters, one argument is a constant FLOAT[dims] and the other the indexed expression).
#include "lm_diskann_sql.hpp"

#include "duckdb/main/config.hpp"
#include "duckdb/optimizer/optimizer_extension.hpp"
#include "duckdb/planner/expression/bound_columnref_expression.hpp"
#include "duckdb/planner/expression/bound_constant_expression.hpp"
#include "duckdb/planner/expression_iterator.hpp"
#include "duckdb/planner/operator/logical_filter.hpp"
#include "duckdb/planner/operator/logical_projection.hpp"
#include "duckdb/planner/operator/logical_top_n.hpp"
#include "duckdb/storage/data_table.hpp"

namespace duckdb {

namespace {

// The shape the rewrite looks for, resolved to the operators it needs.
struct TopNOverScan {
	LogicalTopN *top_n = nullptr;
	LogicalProjection *projection = nullptr;
	unique_ptr<LogicalOperator> *get_slot = nullptr; // the projection's child pointer (replaced when filters are pulled up)
	LogicalGet *get = nullptr;
	Expression *distance = nullptr;                  // the projected expression the order key refers to
};

bool MatchShape(LogicalOperator &op, TopNOverScan &m) {
	if (op.type != LogicalOperatorType::LOGICAL_TOP_N || op.children.size() != 1) {
		return false;
	}
	m.top_n = &op.Cast<LogicalTopN>();
	if (m.top_n->orders.size() != 1 || m.top_n->orders[0].type != OrderType::ASCENDING ||
	    m.top_n->orders[0].expression->type != ExpressionType::BOUND_COLUMN_REF) {
		return false;
	}
	auto &child = *op.children[0];
	if (child.type != LogicalOperatorType::LOGICAL_PROJECTION || child.children.size() != 1) {
		return false;
	}
	m.projection = &child.Cast<LogicalProjection>();
	const auto &key = m.top_n->orders[0].expression->Cast<BoundColumnRefExpression>();
	if (key.binding.table_index != m.projection->table_index || key.binding.column_index >= m.projection->expressions.size()) {
		return false;
	}
	m.distance = m.projection->expressions[key.binding.column_index].get();
	m.get_slot = &child.children[0];
	if ((*m.get_slot)->type != LogicalOperatorType::LOGICAL_GET) {
		return false;
	}
	m.get = &(*m.get_slot)->Cast<LogicalGet>();
	if (m.get->function.name != "seq_scan") {
		return false;
	}
	if (m.get->dynamic_filters && m.get->dynamic_filters->HasFilters()) {
		return false; // a join pushed a filter into this scan at run time: the index cannot honour it
	}
	return true;
}

// The first LM_DISKANN index of the table that answers `m.distance`, as bind data of the index scan.
unique_ptr<LmDiskannIndexScanBindData> BindScan(ClientContext &context, TopNOverScan &m) {
	auto table = m.get->GetTable();
	if (!table || !table->IsDuckTable()) {
		return nullptr;
	}
	auto &duck_table = table->Cast<DuckTableEntry>();
	auto &info = *table->GetStorage().GetDataTableInfo();
	unique_ptr<LmDiskannIndexScanBindData> bind;
	vector<reference<Expression>> bindings;
	info.GetIndexes().BindAndScan<LmDiskannIndex>(context, info, [&](LmDiskannIndex &index) {
		unique_ptr<Expression> index_expr;
		if (!LmDiskannMatch::MatchDistanceFunction(index, *m.distance, bindings) ||
		    !LmDiskannMatch::BindIndexExpression(index, *m.get, index_expr) ||
		    !LmDiskannMatch::OrientBindings(*index_expr, bindings)) {
			return false;
		}
		auto &constant = bindings[1].get();
		if (constant.type != ExpressionType::VALUE_CONSTANT) {
			return false;
		}
		const auto dims = index.GetDimensions();
		const auto &elements = ArrayValue::GetChildren(constant.Cast<BoundConstantExpression>().value);
		if (elements.size() != dims) {
			return false;
		}
		auto query = make_unsafe_uniq_array<float>(dims);
		for (idx_t i = 0; i < dims; i++) {
			query[i] = elements[i].GetValue<float>();
		}
		bind = make_uniq<LmDiskannIndexScanBindData>(duck_table, index, m.top_n->limit, std::move(query));
		return true; // stop at the first index that matches
	});
	return bind;
}

// Table filters that were pushed into the sequential scan become a LogicalFilter above the index scan (which has no
// filter pushdown): the k rows the index returns are filtered afterwards, as the template does.
void PullUpTableFilters(TopNOverScan &m) {
	auto &get = *m.get;
	get.projection_ids.clear();
	get.types.clear();
	auto filter = make_uniq<LogicalFilter>();
	const auto &column_ids = get.GetColumnIds();
	for (auto &entry : get.table_filters.filters) {
		idx_t position = column_ids.size();
		for (idx_t i = 0; i < column_ids.size(); i++) {
			if (column_ids[i].GetPrimaryIndex() == entry.first) {
				position = i;
				break;
			}
		}
		if (position == column_ids.size()) {
			throw InternalException("LM_DISKANN scan rewrite: a table filter refers to a column the scan does not produce");
		}
		BoundColumnRefExpression column(get.returned_types[entry.first], ColumnBinding(get.table_index, position));
		filter->expressions.push_back(entry.second->ToExpression(column));
	}
	filter->children.push_back(std::move(*m.get_slot));
	filter->ResolveOperatorTypes();
	*m.get_slot = std::move(filter);
}

bool RewriteOne(ClientContext &context, unique_ptr<LogicalOperator> &plan) {
	TopNOverScan m;
	if (!MatchShape(*plan, m)) {
		return false;
	}
	auto bind = BindScan(context, m);
	if (!bind) {
		return false;
	}
	auto &get = *m.get;
	get.function = LmDiskannIndexScanFunction::GetFunction();
	const auto cardinality = get.function.cardinality(context, bind.get());
	get.has_estimated_cardinality = cardinality->has_estimated_cardinality;
	get.estimated_cardinality = cardinality->estimated_cardinality;
	get.bind_data = std::move(bind);
	if (!get.table_filters.filters.empty()) {
		PullUpTableFilters(m);
	}
	// the index scan already returns the k closest rows in order: the TopN goes away
	auto below = std::move(plan->children[0]);
	plan = std::move(below);
	return true;
}

bool RewriteTree(ClientContext &context, unique_ptr<LogicalOperator> &plan) {
	bool any = RewriteOne(context, plan);
	for (auto &child : plan->children) {
		any |= RewriteTree(context, child);
	}
	return any;
}

// After the rewrite a projection over the distance projection may keep expressions nobody reads; the distance itself
// would otherwise be evaluated for every fetched row. Unreferenced slots are replaced by a constant (positions stay).
void DropUnusedDistance(unique_ptr<LogicalOperator> &plan) {
	auto &op = *plan;
	if (op.type == LogicalOperatorType::LOGICAL_PROJECTION && !op.children.empty() &&
	    op.children[0]->type == LogicalOperatorType::LOGICAL_PROJECTION) {
		auto &inner = op.children[0]->Cast<LogicalProjection>();
		if (!inner.children.empty() && inner.children[0]->type == LogicalOperatorType::LOGICAL_GET &&
		    inner.children[0]->Cast<LogicalGet>().function.name == LmDiskannIndexScanFunction::NAME) {
			column_binding_set_t used;
			for (auto &expr : op.expressions) {
				ExpressionIterator::EnumerateExpression(expr, [&](Expression &e) {
					if (e.type == ExpressionType::BOUND_COLUMN_REF) {
						used.insert(e.Cast<BoundColumnRefExpression>().binding);
					}
				});
			}
			const auto produced = inner.GetColumnBindings();
			for (idx_t i = 0; i < inner.expressions.size(); i++) {
				if (used.find(produced[i]) == used.end()) {
					inner.expressions[i] = make_uniq_base<Expression, BoundConstantExpression>(Value(LogicalType::TINYINT));
				}
			}
			return;
		}
	}
	for (auto &child : op.children) {
		DropUnusedDistance(child);
	}
}

void OptimizeScan(OptimizerExtensionInput &input, unique_ptr<LogicalOperator> &plan) {
	if (RewriteTree(input.context, plan)) {
		DropUnusedDistance(plan);
	}
}

} // namespace

void LmDiskannModule::RegisterScanOptimizer(DatabaseInstance &db) {
	OptimizerExtension ext;
	ext.optimize_function = OptimizeScan;
	db.config.optimizer_extensions.push_back(ext);
}

} // namespace duckdb
