// BUILD-TIME STAND-IN, not product code (see lmd_duckdb_shim.hpp). Second part: the planner / table-function /
// operator declarations of DuckDB v1.2.x that the SQL-surface sources under duckdb-lm-diskann_b200/duckdb/ (SURVEY.md
// section 8 row f3) and the reference's db/LmDiskannIndex.hpp name. DuckDB is absent from this image, so those sources
// are SYNTAX-CHECKED against these declarations (`make -C duckdb-lm-diskann_b200/duckdb check`); with a real checkout
// pass DUCKDB_INC=<duckdb>/src/include and this directory is left out. Declarations only - nothing here runs; each item
// is the smallest declaration with the same name, members and call shape as the DuckDB one. Nothing is copied from DuckDB.
#pragma once

#include "duckdb/lmd_duckdb_shim.hpp"

#include <functional>
#include <map>
#include <set>
#include <unordered_set>

#define D_ASSERT(x) ((void)0)
#ifndef STANDARD_VECTOR_SIZE
#define STANDARD_VECTOR_SIZE 2048
#endif

namespace duckdb {

// ---- common ---------------------------------------------------------------------------------------------------------
using column_t = idx_t;
using storage_t = idx_t;
using std::unordered_set;
static constexpr const column_t COLUMN_IDENTIFIER_ROW_ID = column_t(-1);
template <class T>
using reference = std::reference_wrapper<T>;
template <class T>
using unsafe_unique_array = std::unique_ptr<T[]>;
template <class T>
using unique_array = std::unique_ptr<T[]>;
template <class T, class... A>
unique_ptr<T> make_uniq(A &&...a) {
	return unique_ptr<T>(new T(std::forward<A>(a)...));
}
template <class B, class T, class... A>
unique_ptr<B> make_uniq_base(A &&...a) {
	return unique_ptr<B>(new T(std::forward<A>(a)...));
}
template <class T>
unsafe_unique_array<T> make_unsafe_uniq_array(idx_t n) {
	return unsafe_unique_array<T>(new T[n]);
}
template <class T>
T MinValue(T a, T b) {
	return a < b ? a : b;
}
using std::to_string;
template <class T>
class optional_ptr {
public:
	optional_ptr() = default;
	optional_ptr(T *p) : ptr(p) {
	}
	T *operator->() const {
		return ptr;
	}
	T &operator*() const {
		return *ptr;
	}
	explicit operator bool() const {
		return ptr != nullptr;
	}
	T *get() const {
		return ptr;
	}

private:
	T *ptr = nullptr;
};
template <class V>
using case_insensitive_map_t = std::map<string, V>;
template <class V>
class InsertionOrderPreservingMap {
public:
	V &operator[](const string &k);
	void insert(const string &k, V v);
};
class ErrorData {
public:
	ErrorData() = default;
	explicit ErrorData(const string &msg);
};

// ---- types / values / vectors ---------------------------------------------------------------------------------------
enum class LogicalTypeId : uint8_t { INVALID, BOOLEAN, TINYINT, BIGINT, FLOAT, ARRAY, LIST };
class LogicalType {
public:
	LogicalType() = default;
	LogicalType(LogicalTypeId id_p) : id_(id_p) {
	}
	LogicalTypeId id() const {
		return id_;
	}
	bool operator==(const LogicalType &o) const;
	bool operator!=(const LogicalType &o) const;
	static LogicalType ARRAY(const LogicalType &child, idx_t size);
	static const LogicalType ROW_TYPE, BIGINT, FLOAT, BOOLEAN, TINYINT;

private:
	LogicalTypeId id_ = LogicalTypeId::INVALID;
};
struct ArrayType {
	static idx_t GetSize(const LogicalType &t);
	static const LogicalType &GetChildType(const LogicalType &t);
};
class Value {
public:
	Value() = default;
	explicit Value(LogicalType t);
	bool IsNull() const;
	const LogicalType &type() const;
	template <class T>
	T GetValue() const;
	static Value BOOLEAN(bool v);
	static Value BIGINT(int64_t v);
};
struct ArrayValue {
	static const vector<Value> &GetChildren(const Value &v);
};
class Vector {
public:
	explicit Vector(LogicalType t);
	const LogicalType &GetType() const;
};
struct FlatVector {
	template <class T>
	static T *GetData(Vector &v);
	template <class T>
	static const T *GetData(const Vector &v);
};
struct ArrayVector {
	static Vector &GetEntry(Vector &v);
};
class SelectionVector {
public:
	void Initialize(idx_t count = STANDARD_VECTOR_SIZE);
	void set_index(idx_t idx, idx_t loc);
};
class DataChunk {
public:
	vector<Vector> data;
	idx_t size() const;
	void SetCardinality(idx_t n);
	void Flatten();
	void Slice(DataChunk &other, const SelectionVector &sel, idx_t count, idx_t col_offset = 0);
};

// ---- expressions ----------------------------------------------------------------------------------------------------
enum class ExpressionType : uint8_t {
	INVALID,
	BOUND_COLUMN_REF,
	VALUE_CONSTANT,
	BOUND_FUNCTION,
	COMPARE_LESSTHANOREQUALTO,
	WINDOW_ROW_NUMBER
};
enum class OrderType : uint8_t { INVALID, ASCENDING, DESCENDING };
struct ColumnBinding {
	idx_t table_index = 0, column_index = 0;
	ColumnBinding() = default;
	ColumnBinding(idx_t t, idx_t c) : table_index(t), column_index(c) {
	}
	bool operator==(const ColumnBinding &o) const {
		return table_index == o.table_index && column_index == o.column_index;
	}
	bool operator<(const ColumnBinding &o) const {
		return table_index < o.table_index || (table_index == o.table_index && column_index < o.column_index);
	}
};
using column_binding_set_t = std::set<ColumnBinding>;
class Expression {
public:
	virtual ~Expression() = default;
	ExpressionType type = ExpressionType::INVALID;
	LogicalType return_type;
	virtual unique_ptr<Expression> Copy() const;
	virtual bool Equals(const Expression &other) const;
	template <class T>
	T &Cast() {
		return static_cast<T &>(*this);
	}
	template <class T>
	const T &Cast() const {
		return static_cast<const T &>(*this);
	}
};
class BoundColumnRefExpression : public Expression {
public:
	BoundColumnRefExpression(LogicalType type, ColumnBinding binding_p);
	ColumnBinding binding;
};
class BoundConstantExpression : public Expression {
public:
	explicit BoundConstantExpression(Value v);
	Value value;
};
struct ScalarFunction {
	string name;
};
class BoundFunctionExpression : public Expression {
public:
	ScalarFunction function;
	vector<unique_ptr<Expression>> children;
};
class BoundComparisonExpression : public Expression {
public:
	unique_ptr<Expression> left, right;
};
struct BoundOrderByNode {
	OrderType type = OrderType::INVALID;
	unique_ptr<Expression> expression;
};
class BoundWindowExpression : public Expression {
public:
	vector<BoundOrderByNode> orders;
};
struct ExpressionIterator {
	static void EnumerateChildren(Expression &e, const std::function<void(Expression &)> &f);
	static void EnumerateExpression(unique_ptr<Expression> &e, const std::function<void(Expression &)> &f);
};

// ---- catalog / storage ----------------------------------------------------------------------------------------------
class ClientContext;
class DatabaseInstance;
class AttachedDatabase;
class TableIOManager;
class FixedSizeAllocator;
class Catalog;
class BaseStatistics {};
class DuckTransaction {
public:
	static DuckTransaction &Get(ClientContext &c, Catalog &catalog);
};
struct ColumnIndex {
	idx_t GetPrimaryIndex() const;
};
struct StorageIndex {
	StorageIndex(idx_t i = 0);
};
struct LogicalIndex {
	explicit LogicalIndex(idx_t i);
};
struct ColumnDefinition {
	storage_t StorageOid() const;
};
class TableFilter {
public:
	unique_ptr<Expression> ToExpression(const Expression &column) const;
};
struct TableFilterSet {
	std::map<idx_t, unique_ptr<TableFilter>> filters;
};
struct DynamicTableFilterSet {
	bool HasFilters() const;
};
struct ColumnFetchState {};
struct LocalScanState {};
class TableScanState {
public:
	void Initialize(vector<StorageIndex> column_ids, optional_ptr<TableFilterSet> filters = nullptr);
	LocalScanState local_state;
};
class IndexLock {};
class IndexScanState {
public:
	virtual ~IndexScanState() = default;
	template <class T>
	T &Cast() {
		return static_cast<T &>(*this);
	}
};
enum class IndexConstraintType : uint8_t { NONE, UNIQUE, PRIMARY, FOREIGN };
enum class VerifyExistenceType : uint8_t { APPEND, APPEND_FK, DELETE_FK };
struct IndexStorageInfo {
	string name;
	idx_t root = 0;
	bool IsValid() const;
};
class BoundIndex {
public:
	BoundIndex(const string &name, const string &index_type, IndexConstraintType constraint_type,
	           const vector<column_t> &column_ids, TableIOManager &table_io_manager,
	           const vector<unique_ptr<Expression>> &unbound_expressions, AttachedDatabase &db);
	virtual ~BoundIndex() = default;
	string name;
	vector<unique_ptr<Expression>> unbound_expressions;
	const string &GetIndexName() const;
	const vector<column_t> &GetColumnIds() const;
	template <class T>
	T &Cast() {
		return static_cast<T &>(*this);
	}
	virtual ErrorData Append(IndexLock &lock, DataChunk &entries, Vector &row_identifiers) = 0;
	virtual void CommitDrop(IndexLock &index_lock) = 0;
	virtual void Delete(IndexLock &lock, DataChunk &entries, Vector &row_identifiers) = 0;
	virtual ErrorData Insert(IndexLock &lock, DataChunk &data, Vector &row_ids) = 0;
	virtual bool MergeIndexes(IndexLock &state, BoundIndex &other_index) = 0;
	virtual void Vacuum(IndexLock &state) = 0;
	virtual string VerifyAndToString(IndexLock &state, const bool only_verify) = 0;
	virtual void VerifyAllocations(IndexLock &state) = 0;
	virtual string GetConstraintViolationMessage(VerifyExistenceType verify_type, idx_t failed_index,
	                                             DataChunk &input) = 0;
};
class DataTableInfo;
class TableIndexList {
public:
	template <class T, class F>
	void BindAndScan(ClientContext &context, DataTableInfo &info, F &&callback) {
		(void)sizeof(decltype(callback(std::declval<T &>()))); // the callback is bool(T &)
	}
};
class DataTableInfo {
public:
	TableIndexList &GetIndexes();
};
class DataTable {
public:
	shared_ptr<DataTableInfo> &GetDataTableInfo();
	void Fetch(DuckTransaction &transaction, DataChunk &result, const vector<StorageIndex> &column_ids,
	           const Vector &row_ids, idx_t fetch_count, ColumnFetchState &state);
};
class LocalStorage {
public:
	static LocalStorage &Get(ClientContext &c, Catalog &catalog);
	void InitializeScan(DataTable &table, LocalScanState &state, optional_ptr<TableFilterSet> filters);
	bool Find(DataTable &table);
};
class TableCatalogEntry {
public:
	virtual ~TableCatalogEntry() = default;
	string name;
	TableCatalogEntry(Catalog &c);
	Catalog &catalog;
	bool IsDuckTable() const;
	DataTable &GetStorage();
	const ColumnDefinition &GetColumn(LogicalIndex idx);
	unique_ptr<BaseStatistics> GetStatistics(ClientContext &context, column_t column_id);
	template <class T>
	T &Cast() {
		return static_cast<T &>(*this);
	}
};
class DuckTableEntry : public TableCatalogEntry {};
struct LogicalDependencyList {
	void AddDependency(TableCatalogEntry &entry);
};

// ---- functions ------------------------------------------------------------------------------------------------------
struct FunctionData {
	virtual ~FunctionData() = default;
	template <class T>
	T &Cast() {
		return static_cast<T &>(*this);
	}
	template <class T>
	const T &Cast() const {
		return static_cast<const T &>(*this);
	}
};
struct TableFunctionData : public FunctionData {};
struct GlobalTableFunctionState {
	virtual ~GlobalTableFunctionState() = default;
	template <class T>
	T &Cast() {
		return static_cast<T &>(*this);
	}
};
struct LocalTableFunctionState {};
struct NodeStatistics {
	NodeStatistics(idx_t estimated, idx_t max);
	bool has_estimated_cardinality = false;
	idx_t estimated_cardinality = 0;
};
struct BindInfo {
	explicit BindInfo(TableCatalogEntry &table);
};
struct TableFunctionInitInput {
	optional_ptr<const FunctionData> bind_data;
	vector<column_t> column_ids;
	optional_ptr<TableFilterSet> filters;
};
struct TableFunctionInput {
	optional_ptr<const FunctionData> bind_data;
	optional_ptr<GlobalTableFunctionState> global_state;
};
struct TableFunctionToStringInput {
	optional_ptr<const FunctionData> bind_data;
};
class TableFunction {
public:
	using function_t = void (*)(ClientContext &, TableFunctionInput &, DataChunk &);
	TableFunction() = default;
	TableFunction(string name_p, vector<LogicalType> arguments, function_t function);
	string name;
	unique_ptr<GlobalTableFunctionState> (*init_global)(ClientContext &, TableFunctionInitInput &) = nullptr;
	unique_ptr<LocalTableFunctionState> (*init_local)(ClientContext &, TableFunctionInitInput &) = nullptr;
	unique_ptr<BaseStatistics> (*statistics)(ClientContext &, const FunctionData *, column_t) = nullptr;
	void (*dependency)(LogicalDependencyList &, const FunctionData *) = nullptr;
	unique_ptr<NodeStatistics> (*cardinality)(ClientContext &, const FunctionData *) = nullptr;
	void *pushdown_complex_filter = nullptr;
	InsertionOrderPreservingMap<string> (*to_string)(TableFunctionToStringInput &) = nullptr;
	void *table_scan_progress = nullptr;
	BindInfo (*get_bind_info)(const optional_ptr<FunctionData>) = nullptr;
	bool projection_pushdown = false, filter_pushdown = false;
};
struct ExtensionUtil {
	static void RegisterFunction(DatabaseInstance &db, TableFunction f);
};

// ---- logical plan ---------------------------------------------------------------------------------------------------
enum class LogicalOperatorType : uint8_t {
	LOGICAL_INVALID,
	LOGICAL_PROJECTION,
	LOGICAL_FILTER,
	LOGICAL_WINDOW,
	LOGICAL_TOP_N,
	LOGICAL_GET,
	LOGICAL_DELIM_GET,
	LOGICAL_DELIM_JOIN,
	LOGICAL_CROSS_PRODUCT,
	LOGICAL_EXTENSION_OPERATOR
};
class PhysicalOperator;
class PhysicalPlanGenerator;
class LogicalOperator {
public:
	virtual ~LogicalOperator() = default;
	LogicalOperatorType type = LogicalOperatorType::LOGICAL_INVALID;
	vector<unique_ptr<LogicalOperator>> children;
	vector<unique_ptr<Expression>> expressions;
	vector<LogicalType> types;
	idx_t estimated_cardinality = 0;
	bool has_estimated_cardinality = false;
	virtual vector<ColumnBinding> GetColumnBindings();
	virtual idx_t EstimateCardinality(ClientContext &context);
	void ResolveOperatorTypes();
	template <class T>
	T &Cast() {
		return static_cast<T &>(*this);
	}
};
class LogicalTopN : public LogicalOperator {
public:
	vector<BoundOrderByNode> orders;
	idx_t limit = 0, offset = 0;
};
class LogicalProjection : public LogicalOperator {
public:
	LogicalProjection(idx_t table_index_p, vector<unique_ptr<Expression>> select_list);
	idx_t table_index = 0;
};
class LogicalFilter : public LogicalOperator {};
class LogicalWindow : public LogicalOperator {
public:
	idx_t window_index = 0;
};
class LogicalCrossProduct : public LogicalOperator {};
class LogicalJoin : public LogicalOperator {};
class LogicalDelimGet : public LogicalOperator {
public:
	idx_t table_index = 0;
};
class LogicalGet : public LogicalOperator {
public:
	idx_t table_index = 0;
	TableFunction function;
	unique_ptr<FunctionData> bind_data;
	vector<LogicalType> returned_types;
	vector<idx_t> projection_ids;
	TableFilterSet table_filters;
	shared_ptr<DynamicTableFilterSet> dynamic_filters;
	optional_ptr<TableCatalogEntry> GetTable() const;
	const vector<ColumnIndex> &GetColumnIds() const;
};
class LogicalExtensionOperator : public LogicalOperator {
public:
	virtual string GetName() const;
	virtual void ResolveTypes() = 0;
	virtual unique_ptr<PhysicalOperator> CreatePlan(ClientContext &context, PhysicalPlanGenerator &generator) = 0;
};
class LogicalOperatorVisitor {
public:
	virtual ~LogicalOperatorVisitor() = default;
	virtual void VisitOperator(LogicalOperator &op);
	void VisitOperatorChildren(LogicalOperator &op);
};
struct ReplacementBinding {
	ReplacementBinding(ColumnBinding old_binding, ColumnBinding new_binding);
};
class ColumnBindingReplacer : public LogicalOperatorVisitor {
public:
	vector<ReplacementBinding> replacement_bindings;
};

// ---- optimizer / main -----------------------------------------------------------------------------------------------
class Binder {
public:
	idx_t GenerateTableIndex();
};
class Optimizer {
public:
	Binder &binder;
};
struct OptimizerExtensionInput {
	ClientContext &context;
	Optimizer &optimizer;
};
class OptimizerExtension {
public:
	void (*optimize_function)(OptimizerExtensionInput &input, unique_ptr<LogicalOperator> &plan) = nullptr;
};
struct CreateIndexInput {
	TableIOManager &table_io_manager;
	AttachedDatabase &db;
	IndexConstraintType constraint_type;
	const string &name;
	const vector<column_t> &column_ids;
	const vector<unique_ptr<Expression>> &unbound_expressions;
	const IndexStorageInfo &storage_info;
	const case_insensitive_map_t<Value> &options;
};
struct PlanIndexInput;
class IndexType {
public:
	string name;
	unique_ptr<BoundIndex> (*create_instance)(CreateIndexInput &input) = nullptr;
	unique_ptr<PhysicalOperator> (*create_plan)(PlanIndexInput &input) = nullptr;
};
class IndexTypeSet {
public:
	void RegisterIndexType(const IndexType &t);
};
class DBConfig {
public:
	void AddExtensionOption(const string &name, string description, LogicalType type, const Value &default_value = Value());
	IndexTypeSet &GetIndexTypes();
	vector<OptimizerExtension> optimizer_extensions;
};
class DatabaseInstance {
public:
	DBConfig config;
};
class ClientContext {
public:
	bool TryGetCurrentSetting(const string &key, Value &result) const;
};

// ---- physical plan --------------------------------------------------------------------------------------------------
enum class PhysicalOperatorType : uint8_t { INVALID, EXTENSION };
enum class OperatorResultType : uint8_t { NEED_MORE_INPUT, HAVE_MORE_OUTPUT, FINISHED };
struct ExecutionContext {
	ClientContext &client;
};
class OperatorState {
public:
	virtual ~OperatorState() = default;
	template <class T>
	T &Cast() {
		return static_cast<T &>(*this);
	}
};
class GlobalOperatorState {};
class PhysicalOperator {
public:
	PhysicalOperator(PhysicalOperatorType type, vector<LogicalType> types, idx_t estimated_cardinality);
	virtual ~PhysicalOperator() = default;
	vector<unique_ptr<PhysicalOperator>> children;
	vector<LogicalType> types;
	idx_t estimated_cardinality = 0;
	virtual string GetName() const;
	virtual bool ParallelOperator() const;
	virtual unique_ptr<OperatorState> GetOperatorState(ExecutionContext &context) const;
	virtual OperatorResultType Execute(ExecutionContext &context, DataChunk &input, DataChunk &chunk,
	                                   GlobalOperatorState &gstate, OperatorState &state) const;
	virtual InsertionOrderPreservingMap<string> ParamsToString() const;
	static void SetEstimatedCardinality(InsertionOrderPreservingMap<string> &result, idx_t estimated_cardinality);
};
class PhysicalPlanGenerator {
public:
	unique_ptr<PhysicalOperator> CreatePlan(unique_ptr<LogicalOperator> op);
};

} // namespace duckdb
