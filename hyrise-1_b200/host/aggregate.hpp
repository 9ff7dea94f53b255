// aggregate.hpp — Hyrise's Aggregate operator with the reference's class surface (src/lib/operators/aggregate.hpp:85-99),
// whose _on_execute() runs aggregates over numeric columns — without group-by columns (hj_aggregate_relation) or grouped
// by up to four numeric columns (hj_aggregate_grouped_rows_relation) — on a B200 through include/hj.h (SURVEY.md §8f N4:
// the consumer after the join). COUNT(DISTINCT) and string columns are refused with a message —
// there is no CPU fallback in this operator.
#pragma once
#include <memory>
#include <optional>
#include <string>
#include <vector>

#include "opossum_min.hpp"

namespace opossum {

enum class AggregateFunction { Min, Max, Sum, Avg, Count, CountDistinct };  // expression/aggregate_expression.hpp:7

// operators/aggregate.hpp:36-42: COUNT(*) is the one aggregate without a column
struct AggregateColumnDefinition final {
  AggregateColumnDefinition(const std::optional<ColumnID>& column_, const AggregateFunction function_)
      : column(column_), function(function_) {}
  std::optional<ColumnID> column;
  AggregateFunction function;
};

class Aggregate : public AbstractReadOnlyOperator {
 public:
  Aggregate(const std::shared_ptr<AbstractOperator>& in, const std::vector<AggregateColumnDefinition>& aggregates,
            const std::vector<ColumnID>& groupby_column_ids);

  const std::vector<AggregateColumnDefinition>& aggregates() const;
  const std::vector<ColumnID>& groupby_column_ids() const;

  const std::string name() const override;
  const std::string description(DescriptionMode description_mode) const override;

  // Not in the reference: can this aggregate run on the device? (else _on_execute fails; callers keep their CPU operator)
  static bool supported(const Table& table, const std::vector<AggregateColumnDefinition>& aggregates,
                        const std::vector<ColumnID>& groupby_column_ids);

 protected:
  std::shared_ptr<const Table> _on_execute() override;
  std::shared_ptr<const Table> _execute_grouped(const Table& in_table);
  std::shared_ptr<AbstractOperator> _on_deep_copy(const std::shared_ptr<AbstractOperator>& copied_input_left,
                                                  const std::shared_ptr<AbstractOperator>& copied_input_right) const override;

  const std::vector<AggregateColumnDefinition> _aggregates;
  const std::vector<ColumnID> _groupby_column_ids;
};

}  // namespace opossum
