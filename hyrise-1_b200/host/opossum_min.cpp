// opossum_min.cpp — out-of-line parts of the Hyrise stand-in types (see opossum_min.hpp).
#include "opossum_min.hpp"

#include <chrono>
#include <map>

namespace opossum {

const AllTypeVariant ReferenceColumn::operator[](const ChunkOffset chunk_offset) const {
  const auto row_id = (*_pos_list)[chunk_offset];  // reference_column.cpp:26-35
  if (row_id.is_null()) return NULL_VALUE;
  return (*_referenced_table->get_chunk(row_id.chunk_id)->get_column(_referenced_column_id))[row_id.chunk_offset];
}

namespace {
template <typename T>
std::shared_ptr<BaseColumn> make_value_column(bool nullable) {
  return std::make_shared<ValueColumn<T>>(nullable);
}
std::shared_ptr<BaseColumn> make_value_column(DataType t, bool nullable) {
  switch (t) {
    case DataType::Int: return make_value_column<int32_t>(nullable);
    case DataType::Long: return make_value_column<int64_t>(nullable);
    case DataType::Float: return make_value_column<float>(nullable);
    case DataType::Double: return make_value_column<double>(nullable);
    case DataType::String: return make_value_column<std::string>(nullable);
    default: Fail("cannot create a column of DataType::Null");
  }
}
template <typename T>
bool try_append(BaseColumn& c, const AllTypeVariant& v) {
  if (auto* vc = dynamic_cast<ValueColumn<T>*>(&c)) {
    vc->append(v);
    return true;
  }
  return false;
}
}  // namespace

void Table::append(const std::vector<AllTypeVariant>& values) {  // storage/table.cpp: append + create_new_chunk
  Assert(_type == TableType::Data, "append() works on Data tables only");
  Assert(values.size() == _column_definitions.size(), "append(): wrong number of values");
  if (_chunks.empty() || _chunks.back()->size() >= _max_chunk_size) {
    ChunkColumns columns;
    for (const auto& def : _column_definitions) columns.push_back(make_value_column(def.data_type, def.nullable));
    _chunks.push_back(std::make_shared<Chunk>(columns));
  }
  auto& chunk = *_chunks.back();
  for (size_t c = 0; c < values.size(); ++c) {
    auto& col = *chunk.get_column(static_cast<ColumnID>(c));
    const bool ok = try_append<int32_t>(col, values[c]) || try_append<int64_t>(col, values[c]) ||
                    try_append<float>(col, values[c]) || try_append<double>(col, values[c]) ||
                    try_append<std::string>(col, values[c]);
    Assert(ok, "append(): not a ValueColumn");
  }
}

void AbstractOperator::execute() {  // abstract_operator.cpp:25-54 (no transaction context: JoinHash is read-only)
  DebugAssert(!_input_left || _input_left->get_output(), "Left input has not yet been executed");
  DebugAssert(!_input_right || _input_right->get_output(), "Right input has not yet been executed");
  DebugAssert(!_output, "Operator has already been executed");
  const auto t0 = std::chrono::steady_clock::now();
  _output = _on_execute();
  _on_cleanup();  // release any temporary data if possible
  _performance_data.walltime_ns =
      static_cast<uint64_t>(std::chrono::duration_cast<std::chrono::nanoseconds>(std::chrono::steady_clock::now() - t0).count());
}

std::shared_ptr<AbstractOperator> AbstractOperator::deep_copy() const {
  const auto left = _input_left ? _input_left->deep_copy() : nullptr;
  const auto right = _input_right ? _input_right->deep_copy() : nullptr;
  return _on_deep_copy(left, right);
}

AbstractJoinOperator::AbstractJoinOperator(const OperatorType type, const std::shared_ptr<const AbstractOperator>& left,
                                           const std::shared_ptr<const AbstractOperator>& right, const JoinMode mode,
                                           const ColumnIDPair& column_ids, const PredicateCondition predicate_condition)
    : AbstractReadOnlyOperator(type, left, right), _mode(mode), _column_ids(column_ids), _predicate_condition(predicate_condition) {
  DebugAssert(mode != JoinMode::Cross, "Specified JoinMode not supported by an AbstractJoin, use Product etc. instead.");
}

const std::string& join_mode_to_string(JoinMode mode) {
  static const std::map<JoinMode, std::string> names = {
      {JoinMode::Cross, "Cross"}, {JoinMode::Inner, "Inner"}, {JoinMode::Left, "Left"}, {JoinMode::Outer, "Outer"},
      {JoinMode::Right, "Right"}, {JoinMode::Semi, "Semi"},   {JoinMode::Anti, "Anti"}};
  return names.at(mode);
}

const std::string AbstractJoinOperator::description(DescriptionMode description_mode) const {  // abstract_join_operator.cpp:29-41
  std::string column_name_left = std::string("Col #") + std::to_string(_column_ids.first);
  std::string column_name_right = std::string("Col #") + std::to_string(_column_ids.second);
  if (_input_left && input_table_left()) column_name_left = input_table_left()->column_name(_column_ids.first);
  if (_input_right && input_table_right()) column_name_right = input_table_right()->column_name(_column_ids.second);
  const auto separator = description_mode == DescriptionMode::MultiLine ? "\n" : " ";
  return name() + separator + "(" + join_mode_to_string(_mode) + " Join where " + column_name_left + " = " + column_name_right + ")";
}

}  // namespace opossum
