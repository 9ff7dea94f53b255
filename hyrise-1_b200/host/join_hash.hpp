// join_hash.hpp — the drop-in: Hyrise's JoinHash operator with the reference's exact class surface
// (src/lib/operators/join_hash.hpp:24-45), whose _on_execute() runs on a B200 through the C ABI of include/hj.h.
#pragma once
#include <memory>
#include <string>

#include "opossum_min.hpp"

struct hj_ctx;

namespace opossum {

/**
 * This operator joins two tables using one column of each table.
 * The output is a new table with referenced columns for all columns of the two inputs and filtered pos_lists.
 * As with the reference, positions are not stable: the row order of the output is unspecified (join_hash.hpp:19-20).
 */
class JoinHash : public AbstractJoinOperator {
 public:
  // Same signature as the reference. `radix_bits` is accepted and — like the reference, which overwrites it with its own
  // heuristic (join_hash.cpp:640-668, quirk Q1) — not used to size the partitions; libhj picks the bits for a
  // shared-memory table (hj_radix_bits).
  JoinHash(const std::shared_ptr<const AbstractOperator>& left, const std::shared_ptr<const AbstractOperator>& right,
           const JoinMode mode, const ColumnIDPair& column_ids, const PredicateCondition predicate_condition,
           const size_t radix_bits = 9);
  ~JoinHash() override;

  const std::string name() const override;

  // Not in the reference: which CUDA device the operator uses and the quirk switches (include/hj.h HJ_FLAG_*).
  static void set_device(int device_id);
  static void set_library_flags(uint32_t flags);
  // Quirk Q3 (join_hash.cpp:61-63): the reference lets the size rule swap a Right join's inputs, which turns it into a
  // left-outer join. true (default) reproduces that literally; false always probes with the right input.
  static void set_literal_right_join_swap(bool literal);

 protected:
  std::shared_ptr<const Table> _on_execute() override;
  std::shared_ptr<AbstractOperator> _on_deep_copy(const std::shared_ptr<AbstractOperator>& copied_input_left,
                                                  const std::shared_ptr<AbstractOperator>& copied_input_right) const override;
  void _on_cleanup() override;

  hj_ctx* _ctx{nullptr};
  const size_t _radix_bits;
};

}  // namespace opossum
