// aggregate.cpp — host half of the device Aggregate: output schema, names and NULL rules of the reference's Aggregate
// (src/lib/operators/aggregate.cpp:30-75 surface, :760-815 output columns) stay here; the pass over each aggregate column
// is one hj_aggregate_relation() call on the GPU (all of COUNT / SUM / MIN / MAX / AVG of a column come out of one pass,
// so several aggregates on the same column cost one call).
#include "aggregate.hpp"

#include <algorithm>
#include <cstdio>
#include <type_traits>
#include <variant>
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
  if (groupby_column_ids.size() > 4) return false;
  if (groupby_column_ids.empty() && aggregates.empty()) return false;
  for (const ColumnID g : groupby_column_ids)
    if (!numeric(table.column_data_type(g))) return false;
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
         "grouped by up to four numeric columns; this aggregate is not supported and there is no CPU fallback");
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

// GROUP BY: one hj_aggregate_grouped_rows_relation call per distinct aggregate column (or one for the groups alone). A group
// comes back as a representative row of the input table; the group-by columns of the output are that row's values
// (_write_groupby_output, aggregate.cpp:722-735); group-by columns first, then the aggregates (:540-560).
std::shared_ptr<const Table> Aggregate::_execute_grouped(const Table& in_table) {
  TableColumnDefinitions definitions;
  for (const ColumnID g : _groupby_column_ids)
    definitions.emplace_back(in_table.column_name(g), in_table.column_data_type(g), in_table.column_is_nullable(g));
  for (const auto& a : _aggregates) {
    std::stringstream column_name;
    column_name << function_name(a.function) << "(" << (a.column ? in_table.column_name(*a.column) : std::string("*")) << ")";
    if (!a.column) definitions.emplace_back(column_name.str(), DataType::Long, false);
    else definitions.emplace_back(column_name.str(), aggregate_data_type(a.function, in_table.column_data_type(*a.column)),
                                  a.function != AggregateFunction::Count);
  }
  auto output = std::make_shared<Table>(definitions, TableType::Data);
  if (in_table.chunk_count() == 0) return output;  // no rows, no groups

  hj_ctx* ctx = hjhost::context();
  std::vector<std::shared_ptr<void>> keep_alive;
  std::vector<const hj_relation*> group_rels;
  for (const ColumnID g : _groupby_column_ids) {
    const hj_relation* rel = nullptr;
    keep_alive.push_back(hjhost::flatten_column(in_table, g, &rel));
    group_rels.push_back(rel);
  }
  auto run = [&](const hj_relation* value_rel) {
    hj_group_info info;
    if (hj_aggregate_grouped_rows_relation(ctx, group_rels.data(), static_cast<uint32_t>(group_rels.size()), value_rel, 0, nullptr, &info) != HJ_OK)
      Fail(std::string("Aggregate: ") + hj_last_error(ctx));
    std::vector<hj_group> groups(std::max<uint64_t>(1, info.n_groups));
    if (hj_aggregate_grouped_fetch(ctx, groups.data(), groups.size()) != HJ_OK) Fail(std::string("Aggregate: ") + hj_last_error(ctx));
    groups.resize(info.n_groups);
    return groups;
  };
  // row number in the table -> (chunk, offset)
  std::vector<uint64_t> chunk_begin(in_table.chunk_count() + 1, 0);
  for (ChunkID c = 0; c < in_table.chunk_count(); ++c) chunk_begin[c + 1] = chunk_begin[c] + in_table.get_chunk(c)->size();
  auto group_values = [&](uint64_t row) {
    const auto it = std::upper_bound(chunk_begin.begin(), chunk_begin.end(), row);
    const ChunkID c = static_cast<ChunkID>(it - chunk_begin.begin() - 1);
    const ChunkOffset offset = static_cast<ChunkOffset>(row - chunk_begin[c]);
    std::vector<AllTypeVariant> values;
    for (const ColumnID g : _groupby_column_ids) values.push_back((*in_table.get_chunk(c)->get_column(g))[offset]);
    return values;
  };
  // the representative row differs from call to call: groups are matched across calls by their group-by values (as text)
  auto key_of = [](const std::vector<AllTypeVariant>& values) {
    std::string k;
    for (const auto& v : values) {
      if (variant_is_null(v)) k += "~|";
      else std::visit([&](const auto& x) {
        using X = std::decay_t<decltype(x)>;
        if constexpr (std::is_same_v<X, std::string>) k += x + "|";
        else if constexpr (std::is_same_v<X, NullValue>) k += "~|";
        else if constexpr (std::is_floating_point_v<X>) { char buf[40]; std::snprintf(buf, sizeof buf, "%a|", static_cast<double>(x)); k += buf; }
        else k += std::to_string(x) + "|";
      }, v);
    }
    return k;
  };
  std::vector<std::string> order;
  std::map<std::string, std::vector<AllTypeVariant>> values_of_group;
  std::map<std::string, uint64_t> rows_per_group;  // COUNT(*)
  std::map<ColumnID, std::map<std::string, hj_aggregate>> per_column;
  auto absorb = [&](const std::vector<hj_group>& groups, std::map<std::string, hj_aggregate>* into) {
    for (const auto& g : groups) {
      auto values = group_values(static_cast<uint64_t>(g.key));
      const std::string key = key_of(values);
      if (!values_of_group.count(key)) {
        order.push_back(key);
        values_of_group.emplace(key, std::move(values));
        rows_per_group[key] = g.agg.n_rows;
      }
      if (into) (*into)[key] = g.agg;
    }
  };
  for (const auto& a : _aggregates) {
    if (!a.column || per_column.count(*a.column)) continue;
    const hj_relation* value_rel = nullptr;
    const auto keep_value = hjhost::flatten_column(in_table, *a.column, &value_rel);
    absorb(run(value_rel), &per_column[*a.column]);
  }
  if (per_column.empty()) absorb(run(nullptr), nullptr);  // no aggregate column: the groups (and COUNT(*)) alone

  for (const auto& key : order) {
    std::vector<AllTypeVariant> row = values_of_group.at(key);
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
