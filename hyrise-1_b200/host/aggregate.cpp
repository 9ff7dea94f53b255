// aggregate.cpp — host half of the device Aggregate: output schema, names and NULL rules of the reference's Aggregate
// (src/lib/operators/aggregate.cpp:30-75 surface, :760-815 output columns) stay here; the pass over each aggregate column
// is one hj_aggregate_relation() call on the GPU (all of COUNT / SUM / MIN / MAX / AVG of a column come out of one pass,
// so several aggregates on the same column cost one call).
#include "aggregate.hpp"

#include <map>
#include <sstream>

#include "hj.h"

namespace opossum {

namespace hjhost {  // join_hash.cpp
hj_ctx* context();
std::shared_ptr<void> flatten_column(const Table& table, ColumnID column_id, const hj_relation** relation);
}  // namespace hjhost

namespace {

const char* function_name(AggregateFunction f) {  // constant_mappings.cpp:61-69
  switch (f) {
    case AggregateFunction::Min: return "MIN";
    case AggregateFunction::Max: return "MAX";
    case AggregateFunction::Sum: return "SUM";
    case AggregateFunction::Avg: return "AVG";
    case AggregateFunction::Count: return "COUNT";
    default: return "COUNT DISTINCT";
  }
}

bool numeric(DataType t) { return t == DataType::Int || t == DataType::Long || t == DataType::Float || t == DataType::Double; }

// aggregate_traits.hpp:16-66
DataType aggregate_data_type(AggregateFunction f, DataType column_type) {
  switch (f) {
    case AggregateFunction::Count:
    case AggregateFunction::CountDistinct: return DataType::Long;
    case AggregateFunction::Avg: return DataType::Double;
    case AggregateFunction::Sum: return (column_type == DataType::Int || column_type == DataType::Long) ? DataType::Long : DataType::Double;
    default: return column_type;  // MIN / MAX
  }
}

}  // namespace

Aggregate::Aggregate(const std::shared_ptr<AbstractOperator>& in, const std::vector<AggregateColumnDefinition>& aggregates,
                     const std::vector<ColumnID>& groupby_column_ids)
    : AbstractReadOnlyOperator(OperatorType::Aggregate, in), _aggregates(aggregates), _groupby_column_ids(groupby_column_ids) {
  Assert(!(aggregates.empty() && groupby_column_ids.empty()), "Neither aggregate nor groupby columns have been specified");
}

const std::vector<AggregateColumnDefinition>& Aggregate::aggregates() const { return _aggregates; }
const std::vector<ColumnID>& Aggregate::groupby_column_ids() const { return _groupby_column_ids; }
const std::string Aggregate::name() const { return "Aggregate"; }

const std::string Aggregate::description(DescriptionMode) const {  // aggregate.cpp:47-68
  std::stringstream desc;
  desc << "[Aggregate] GroupBy ColumnIDs: ";
  for (size_t i = 0; i < _groupby_column_ids.size(); ++i) desc << _groupby_column_ids[i] << (i + 1 < _groupby_column_ids.size() ? ", " : "");
  desc << " Aggregates: ";
  for (size_t i = 0; i < _aggregates.size(); ++i) {
    desc << function_name(_aggregates[i].function);
    if (_aggregates[i].column) desc << "(Column #" << *_aggregates[i].column << ")"; else desc << "(*)";
    if (i + 1 < _aggregates.size()) desc << ", ";
  }
  return desc.str();
}

bool Aggregate::supported(const Table& table, const std::vector<AggregateColumnDefinition>& aggregates,
                          const std::vector<ColumnID>& groupby_column_ids) {
  if (!groupby_column_ids.empty() || aggregates.empty()) return false;
  for (const auto& a : aggregates) {
    if (a.function == AggregateFunction::CountDistinct) return false;
    if (!a.column) {
      if (a.function != AggregateFunction::Count) return false;  // only COUNT(*) has no column (aggregate.cpp:424)
      continue;
    }
    if (!numeric(table.column_data_type(*a.column))) return false;
  }
  return true;
}

std::shared_ptr<const Table> Aggregate::_on_execute() {
  const auto in_table = input_table_left();
  Assert(in_table != nullptr, "Aggregate: input has not been executed");
  if (!supported(*in_table, _aggregates, _groupby_column_ids))
    Fail("Aggregate on the device covers MIN / MAX / SUM / AVG / COUNT of numeric columns and COUNT(*) without GROUP BY; this "
         "aggregate is not supported and there is no CPU fallback");

  hj_ctx* ctx = hjhost::context();
  std::map<ColumnID, hj_aggregate> per_column;  // one pass per distinct aggregate column
  for (const auto& a : _aggregates) {
    if (!a.column || per_column.count(*a.column)) continue;
    hj_aggregate result{};
    result.is_floating = in_table->column_data_type(*a.column) == DataType::Float || in_table->column_data_type(*a.column) == DataType::Double;
    if (in_table->chunk_count() > 0) {
      const hj_relation* rel = nullptr;
      const auto keep_alive = hjhost::flatten_column(*in_table, *a.column, &rel);
      if (hj_aggregate_relation(ctx, rel, nullptr, &result) != HJ_OK) Fail(std::string("Aggregate: ") + hj_last_error(ctx));
    }
    per_column.emplace(*a.column, result);
  }

  // one output row; names and nullability as in write_aggregate_output (aggregate.cpp:775-790)
  TableColumnDefinitions definitions;
  std::vector<AllTypeVariant> row;
  for (const auto& a : _aggregates) {
    std::stringstream column_name;
    column_name << function_name(a.function) << "(" << (a.column ? in_table->column_name(*a.column) : std::string("*")) << ")";
    if (!a.column) {  // COUNT(*): the row count (aggregate.cpp:424-431, 579-586)
      definitions.emplace_back(column_name.str(), DataType::Long, false);
      row.emplace_back(static_cast<int64_t>(in_table->row_count()));
      continue;
    }
    const DataType column_type = in_table->column_data_type(*a.column);
    const DataType out_type = aggregate_data_type(a.function, column_type);
    const hj_aggregate& r = per_column.at(*a.column);
    const bool needs_null = a.function != AggregateFunction::Count;
    definitions.emplace_back(column_name.str(), out_type, needs_null);
    if (a.function == AggregateFunction::Count) {
      row.emplace_back(static_cast<int64_t>(r.n_values));
    } else if (r.n_values == 0) {
      row.emplace_back(NULL_VALUE);  // no row contributed: NULL (aggregate.cpp:635-642, 805-811)
    } else if (a.function == AggregateFunction::Avg) {
      row.emplace_back(r.sum_double / static_cast<double>(r.n_values));
    } else if (a.function == AggregateFunction::Sum) {
      if (r.is_floating) row.emplace_back(r.sum_double); else row.emplace_back(static_cast<int64_t>(r.sum_int));
    } else {
      const bool is_min = a.function == AggregateFunction::Min;
      switch (column_type) {
        case DataType::Int: row.emplace_back(static_cast<int32_t>(is_min ? r.min_int : r.max_int)); break;
        case DataType::Long: row.emplace_back(static_cast<int64_t>(is_min ? r.min_int : r.max_int)); break;
        case DataType::Float: row.emplace_back(static_cast<float>(is_min ? r.min_double : r.max_double)); break;
        default: row.emplace_back(is_min ? r.min_double : r.max_double);
      }
    }
  }
  auto output = std::make_shared<Table>(definitions, TableType::Data);
  output->append(row);
  return output;
}

std::shared_ptr<AbstractOperator> Aggregate::_on_deep_copy(const std::shared_ptr<AbstractOperator>& copied_input_left,
                                                           const std::shared_ptr<AbstractOperator>&) const {
  return std::make_shared<Aggregate>(copied_input_left, _aggregates, _groupby_column_ids);
}

}  // namespace opossum
