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

// the value of one aggregate from the library's per-column result, typed like write_aggregate_values (aggregate.cpp:622-712)
AllTypeVariant aggregate_value(AggregateFunction f, DataType column_type, const hj_aggregate& r) {
  if (f == AggregateFunction::Count) return static_cast<int64_t>(r.n_values);
  if (r.n_values == 0) return NULL_VALUE;  // no row contributed: NULL (aggregate.cpp:635-642, 805-811)
  if (f == AggregateFunction::Avg) return r.sum_double / static_cast<double>(r.n_values);
  if (f == AggregateFunction::Sum) {
    if (r.is_floating) return r.sum_double;
    return static_cast<int64_t>(r.sum_int);
  }
  const bool is_min = f == AggregateFunction::Min;
  switch (column_type) {
    case DataType::Int: return static_cast<int32_t>(is_min ? r.min_int : r.max_int);
    case DataType::Long: return static_cast<int64_t>(is_min ? r.min_int : r.max_int);
    case DataType::Float: return static_cast<float>(is_min ? r.min_double : r.max_double);
    default: return is_min ? r.min_double : r.max_double;
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
  if (groupby_column_ids.size() > 1) return false;
  if (groupby_column_ids.empty() && aggregates.empty()) return false;
  if (groupby_column_ids.size() == 1) {
    const DataType t = table.column_data_type(groupby_column_ids[0]);
    if (t != DataType::Int && t != DataType::Long) return false;
  }
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
    Fail("Aggregate on the device covers MIN / MAX / SUM / AVG / COUNT of numeric columns and COUNT(*), without GROUP BY or "
         "grouped by one int / long column; this aggregate is not supported and there is no CPU fallback");
  if (!_groupby_column_ids.empty()) return _execute_grouped(*in_table);

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
    definitions.emplace_back(column_name.str(), aggregate_data_type(a.function, column_type), a.function != AggregateFunction::Count);
    row.emplace_back(aggregate_value(a.function, column_type, per_column.at(*a.column)));
  }
  auto output = std::make_shared<Table>(definitions, TableType::Data);
  output->append(row);
  return output;
}

// GROUP BY one int / long column: one hj_aggregate_grouped_relation call per distinct aggregate column (or one for the
// groups alone); group-by column first, then the aggregates (aggregate.cpp:540-560), one row per group.
std::shared_ptr<const Table> Aggregate::_execute_grouped(const Table& in_table) {
  const ColumnID group_id = _groupby_column_ids[0];
  const DataType group_type = in_table.column_data_type(group_id);
  TableColumnDefinitions definitions;
  definitions.emplace_back(in_table.column_name(group_id), group_type, in_table.column_is_nullable(group_id));
  for (const auto& a : _aggregates) {
    std::stringstream column_name;
    column_name << function_name(a.function) << "(" << (a.column ? in_table.column_name(*a.column) : std::string("*")) << ")";
    if (!a.column) definitions.emplace_back(column_name.str(), DataType::Long, false);
    else definitions.emplace_back(column_name.str(), aggregate_data_type(a.function, in_table.column_data_type(*a.column)),
                                  a.function != AggregateFunction::Count);
  }
  auto output = std::make_shared<Table>(definitions, TableType::Data);
  if (in_table.chunk_count() == 0) return output;  // no rows, no groups

  using GroupKey = std::pair<bool, int64_t>;  // (is NULL, value)
  hj_ctx* ctx = hjhost::context();
  const hj_relation* group_rel = nullptr;
  const auto keep_group = hjhost::flatten_column(in_table, group_id, &group_rel);
  auto run = [&](const hj_relation* value_rel) {
    hj_group_info info;
    if (hj_aggregate_grouped_relation(ctx, group_rel, value_rel, 0, nullptr, &info) != HJ_OK) Fail(std::string("Aggregate: ") + hj_last_error(ctx));
    std::vector<hj_group> groups(std::max<uint64_t>(1, info.n_groups));
    if (hj_aggregate_grouped_fetch(ctx, groups.data(), groups.size()) != HJ_OK) Fail(std::string("Aggregate: ") + hj_last_error(ctx));
    groups.resize(info.n_groups);
    return groups;
  };
  std::vector<GroupKey> order;                                   // the groups, in the order of the first call
  std::map<GroupKey, uint64_t> rows_per_group;                   // COUNT(*)
  std::map<ColumnID, std::map<GroupKey, hj_aggregate>> per_column;
  auto remember = [&](const std::vector<hj_group>& groups) {
    if (!order.empty() || groups.empty()) return;
    for (const auto& g : groups) {
      order.emplace_back(g.key_is_null != 0, g.key);
      rows_per_group[order.back()] = g.agg.n_rows;
    }
  };
  for (const auto& a : _aggregates) {
    if (!a.column || per_column.count(*a.column)) continue;
    const hj_relation* value_rel = nullptr;
    const auto keep_value = hjhost::flatten_column(in_table, *a.column, &value_rel);
    const auto groups = run(value_rel);
    remember(groups);
    auto& m = per_column[*a.column];
    for (const auto& g : groups) m[GroupKey(g.key_is_null != 0, g.key)] = g.agg;
  }
  if (order.empty()) remember(run(nullptr));  // no aggregate column: the groups (and COUNT(*)) alone

  for (const auto& key : order) {
    std::vector<AllTypeVariant> row;
    if (key.first) row.emplace_back(NULL_VALUE);
    else if (group_type == DataType::Int) row.emplace_back(static_cast<int32_t>(key.second));
    else row.emplace_back(static_cast<int64_t>(key.second));
    for (const auto& a : _aggregates) {
      if (!a.column) row.emplace_back(static_cast<int64_t>(rows_per_group.at(key)));
      else row.emplace_back(aggregate_value(a.function, in_table.column_data_type(*a.column), per_column.at(*a.column).at(key)));
    }
    output->append(row);
  }
  return output;
}

std::shared_ptr<AbstractOperator> Aggregate::_on_deep_copy(const std::shared_ptr<AbstractOperator>& copied_input_left,
                                                           const std::shared_ptr<AbstractOperator>&) const {
  return std::make_shared<Aggregate>(copied_input_left, _aggregates, _groupby_column_ids);
}

}  // namespace opossum
