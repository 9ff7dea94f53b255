// table_scan.hpp — Hyrise's TableScan operator with the reference's class surface (src/lib/operators/table_scan.hpp:24-60),
// whose _on_execute() runs the single-column scans the device library covers (integer column vs
// literal, IS [NOT] NULL) on a B200 through hj_scan() of include/hj.h. Everything else the reference's TableScan can do
// (column-vs-column, LIKE, float/string columns) is refused with a message — there is no CPU fallback in this operator.
#pragma once
#include <memory>
#include <string>
#include <vector>

#include "opossum_min.hpp"

namespace opossum {

// all_parameter_variant.hpp: the right side of a scan is a literal or another column of the same table.
struct AllParameterVariant {
  AllParameterVariant(AllTypeVariant v) : value(std::move(v)) {}                     // NOLINT: implicit like the variant
  AllParameterVariant(int v) : value(static_cast<int32_t>(v)) {}                     // NOLINT
  AllParameterVariant(long v) : value(static_cast<int64_t>(v)) {}                    // NOLINT
  AllParameterVariant(long long v) : value(static_cast<int64_t>(v)) {}               // NOLINT
  AllParameterVariant(double v) : value(v) {}                                        // NOLINT
  static AllParameterVariant column(ColumnID c) {
    AllParameterVariant p{AllTypeVariant{NullValue{}}};
    p.is_column = true;
    p.column_id = c;
    return p;
  }
  AllTypeVariant value;
  bool is_column{false};
  ColumnID column_id{0};
};

class TableScan : public AbstractReadOnlyOperator {
 public:
  TableScan(const std::shared_ptr<const AbstractOperator>& in, ColumnID left_column_id,
            const PredicateCondition predicate_condition, const AllParameterVariant& right_parameter);
  ~TableScan() override;

  // Chunks the scan skips (table_scan.cpp:42: used by the optimizer's chunk pruning)
  void set_excluded_chunk_ids(const std::vector<ChunkID>& chunk_ids);

  ColumnID left_column_id() const;
  PredicateCondition predicate_condition() const;
  const AllParameterVariant& right_parameter() const;

  const std::string name() const override;
  const std::string description(DescriptionMode description_mode) const override;

  // Not in the reference: can this scan run on the device? (else _on_execute fails; callers keep their CPU operator)
  static bool supported(const Table& table, ColumnID column_id, PredicateCondition predicate_condition,
                        const AllParameterVariant& right_parameter);

 protected:
  std::shared_ptr<const Table> _on_execute() override;
  std::shared_ptr<AbstractOperator> _on_deep_copy(const std::shared_ptr<AbstractOperator>& copied_input_left,
                                                  const std::shared_ptr<AbstractOperator>& copied_input_right) const override;
  void _on_cleanup() override;

  const ColumnID _left_column_id;
  const PredicateCondition _predicate_condition;
  AllParameterVariant _right_parameter;
  std::vector<ChunkID> _excluded_chunk_ids;
};

}  // namespace opossum
