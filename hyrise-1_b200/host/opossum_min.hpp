// opossum_min.hpp — the slice of Hyrise's host-side types that JoinHash touches, with the reference's names, layouts and
// semantics, so that join_hash.{hpp,cpp} in this directory reads (and is called) like the reference operator.
//
// The reference cannot be compiled in this image (Boost/TBB), so these are stand-ins, not copies:
// std::vector instead of pmr/tbb vectors, std::variant instead of boost::variant. In a real Hyrise checkout this header is
// dropped and join_hash.cpp includes the genuine headers (INTEGRATION.md §2 lists the 1:1 mapping).
//
// Mirrors (paths relative to the reference's src/lib):
//   types.hpp:38,92-99,138,150,166-189   ChunkID/ColumnID/ChunkOffset, RowID, PosList, NULL_ROW_ID, JoinMode, PredicateCondition
//   all_type_variant.hpp:34-39            DataType, AllTypeVariant
//   utils/assert.hpp:49-70                Fail / Assert / DebugAssert (throw std::logic_error)
//   storage/base_column.hpp, value_column.hpp:15-73, dictionary_column.hpp:20-72, reference_column.hpp:19-50,
//   storage/vector_compression/fixed_size_byte_aligned/fixed_size_byte_aligned_vector.hpp:21-51,
//   storage/chunk.hpp, storage/table.hpp, storage/table_column_definition.hpp
//   operators/abstract_operator.hpp:56-172, abstract_read_only_operator.hpp, abstract_join_operator.hpp:25-53,
//   operators/table_wrapper.hpp
#pragma once
#include <cstdint>
#include <limits>
#include <memory>
#include <stdexcept>
#include <string>
#include <utility>
#include <variant>
#include <vector>

namespace opossum {

// ---- utils/assert.hpp ------------------------------------------------------------------------------------------------
[[noreturn]] inline void Fail(const std::string& msg) { throw std::logic_error(msg); }
inline void Assert(bool expr, const std::string& msg) {
  if (!expr) Fail(msg);
}
#ifndef NDEBUG
#define HJ_IS_DEBUG 1
inline void DebugAssert(bool expr, const std::string& msg) { Assert(expr, msg); }
#else
#define HJ_IS_DEBUG 0
inline void DebugAssert(bool, const std::string&) {}
#endif

// ---- types.hpp ---------------------------------------------------------------------------------------------------------
using ChunkID = uint32_t;     // STRONG_TYPEDEF(uint32_t, ChunkID), types.hpp:38
using ColumnID = uint16_t;    // STRONG_TYPEDEF(uint16_t, ColumnID)
using ValueID = uint32_t;
using ChunkOffset = uint32_t;
constexpr ChunkOffset INVALID_CHUNK_OFFSET{std::numeric_limits<ChunkOffset>::max()};
constexpr ChunkID INVALID_CHUNK_ID{std::numeric_limits<ChunkID>::max()};

struct RowID {
  ChunkID chunk_id{INVALID_CHUNK_ID};
  ChunkOffset chunk_offset{INVALID_CHUNK_OFFSET};
  RowID() = default;
  RowID(const ChunkID chunk_id_, const ChunkOffset chunk_offset_) : chunk_id(chunk_id_), chunk_offset(chunk_offset_) {}
  bool is_null() const { return chunk_offset == INVALID_CHUNK_OFFSET; }
  bool operator==(const RowID& o) const { return chunk_id == o.chunk_id && chunk_offset == o.chunk_offset; }
};
static_assert(sizeof(RowID) == 8, "RowID must stay {uint32, uint32}: libhj writes hj_row_id straight into PosLists");
const RowID NULL_ROW_ID = RowID{INVALID_CHUNK_ID, INVALID_CHUNK_OFFSET};
using PosList = std::vector<RowID>;
using ColumnIDPair = std::pair<ColumnID, ColumnID>;

enum class PredicateCondition { Equals, NotEquals, LessThan, LessThanEquals, GreaterThan, GreaterThanEquals, Between, In,
                                Like, NotLike, IsNull, IsNotNull };
enum class JoinMode { Inner, Left, Right, Outer, Cross, Semi, Anti };  // integer values are part of the C ABI (hj.h)
enum class TableType { References, Data };
enum class DescriptionMode { SingleLine, MultiLine };
enum class OperatorType { JoinHash, TableWrapper, TableScan, Aggregate, Mock };

// ---- all_type_variant.hpp ------------------------------------------------------------------------------------------------
enum class DataType : uint8_t { Null, Int, Long, Float, Double, String };
struct NullValue {
  bool operator==(const NullValue&) const { return true; }
};
using AllTypeVariant = std::variant<NullValue, int32_t, int64_t, float, double, std::string>;
inline bool variant_is_null(const AllTypeVariant& v) { return v.index() == 0; }
const AllTypeVariant NULL_VALUE = NullValue{};

template <typename T> constexpr DataType data_type_of();
template <> constexpr DataType data_type_of<int32_t>() { return DataType::Int; }
template <> constexpr DataType data_type_of<int64_t>() { return DataType::Long; }
template <> constexpr DataType data_type_of<float>() { return DataType::Float; }
template <> constexpr DataType data_type_of<double>() { return DataType::Double; }
template <> constexpr DataType data_type_of<std::string>() { return DataType::String; }

// ---- storage ---------------------------------------------------------------------------------------------------------------
class BaseColumn {
 public:
  virtual ~BaseColumn() = default;
  virtual const AllTypeVariant operator[](const ChunkOffset chunk_offset) const = 0;
  virtual size_t size() const = 0;
};

// storage/value_column.hpp:15-73 — values + optional null flags (1 byte each). A NULL row still stores a value (T{}).
template <typename T>
class ValueColumn : public BaseColumn {
 public:
  explicit ValueColumn(bool nullable = false) : _nullable(nullable) {}
  ValueColumn(std::vector<T> values, std::vector<uint8_t> null_values)
      : _values(std::move(values)), _null_values(std::move(null_values)), _nullable(true) {}
  explicit ValueColumn(std::vector<T> values) : _values(std::move(values)), _nullable(false) {}

  const AllTypeVariant operator[](const ChunkOffset i) const override {
    if (_nullable && _null_values[i]) return NULL_VALUE;
    return _values[i];
  }
  void append(const AllTypeVariant& val) {  // value_column.cpp:70-90
    const bool is_null = variant_is_null(val);
    if (_nullable) {
      _null_values.push_back(is_null);
      _values.push_back(is_null ? T{} : std::get<T>(val));
      return;
    }
    Assert(!is_null, "ValueColumns is not nullable but value passed is null.");
    _values.push_back(std::get<T>(val));
  }
  const std::vector<T>& values() const { return _values; }
  bool is_nullable() const { return _nullable; }
  const std::vector<uint8_t>& null_values() const { return _null_values; }
  size_t size() const override { return _values.size(); }

 private:
  std::vector<T> _values;
  std::vector<uint8_t> _null_values;
  bool _nullable;
};

// fixed_size_byte_aligned_vector.hpp:21-51 — the attribute vector, width chosen by fixed_size_byte_aligned_compressor.cpp:21-30
class BaseCompressedVector {
 public:
  virtual ~BaseCompressedVector() = default;
  virtual size_t size() const = 0;
  virtual uint32_t get(size_t i) const = 0;
  virtual const void* raw_data() const = 0;
  virtual uint8_t width_bytes() const = 0;
};
template <typename UnsignedIntType>
class FixedSizeByteAlignedVector : public BaseCompressedVector {
 public:
  explicit FixedSizeByteAlignedVector(std::vector<UnsignedIntType> data) : _data(std::move(data)) {}
  size_t size() const override { return _data.size(); }
  uint32_t get(size_t i) const override { return _data[i]; }
  const void* raw_data() const override { return _data.data(); }
  uint8_t width_bytes() const override { return sizeof(UnsignedIntType); }
  const std::vector<UnsignedIntType>& data() const { return _data; }

 private:
  std::vector<UnsignedIntType> _data;
};

// storage/dictionary_column.hpp:20-72
class BaseDictionaryColumn : public BaseColumn {
 public:
  virtual const BaseCompressedVector* attribute_vector_raw() const = 0;
  virtual ValueID null_value_id() const = 0;
  virtual size_t unique_values_count() const = 0;
};
template <typename T>
class DictionaryColumn : public BaseDictionaryColumn {
 public:
  DictionaryColumn(std::shared_ptr<const std::vector<T>> dictionary,
                   std::shared_ptr<const BaseCompressedVector> attribute_vector, ValueID null_value_id)
      : _dictionary(std::move(dictionary)), _attribute_vector(std::move(attribute_vector)), _null_value_id(null_value_id) {}
  const AllTypeVariant operator[](const ChunkOffset i) const override {
    const auto value_id = _attribute_vector->get(i);
    if (value_id == _null_value_id) return NULL_VALUE;
    return (*_dictionary)[value_id];
  }
  std::shared_ptr<const std::vector<T>> dictionary() const { return _dictionary; }
  std::shared_ptr<const BaseCompressedVector> attribute_vector() const { return _attribute_vector; }
  const BaseCompressedVector* attribute_vector_raw() const override { return _attribute_vector.get(); }
  ValueID null_value_id() const override { return _null_value_id; }
  size_t unique_values_count() const override { return _dictionary->size(); }
  size_t size() const override { return _attribute_vector->size(); }

 private:
  std::shared_ptr<const std::vector<T>> _dictionary;
  std::shared_ptr<const BaseCompressedVector> _attribute_vector;
  ValueID _null_value_id;
};

class Table;

// storage/reference_column.hpp:19-50
class ReferenceColumn : public BaseColumn {
 public:
  ReferenceColumn(std::shared_ptr<const Table> referenced_table, ColumnID referenced_column_id,
                  std::shared_ptr<const PosList> pos)
      : _referenced_table(std::move(referenced_table)), _referenced_column_id(referenced_column_id), _pos_list(std::move(pos)) {}
  const AllTypeVariant operator[](const ChunkOffset chunk_offset) const override;
  size_t size() const override { return _pos_list->size(); }
  const std::shared_ptr<const PosList> pos_list() const { return _pos_list; }
  const std::shared_ptr<const Table> referenced_table() const { return _referenced_table; }
  ColumnID referenced_column_id() const { return _referenced_column_id; }

 private:
  std::shared_ptr<const Table> _referenced_table;
  ColumnID _referenced_column_id;
  std::shared_ptr<const PosList> _pos_list;
};

using ChunkColumns = std::vector<std::shared_ptr<BaseColumn>>;

class Chunk {
 public:
  explicit Chunk(ChunkColumns columns) : _columns(std::move(columns)) {}
  std::shared_ptr<BaseColumn> get_column(ColumnID column_id) const { return _columns.at(column_id); }
  const ChunkColumns& columns() const { return _columns; }
  uint16_t column_count() const { return static_cast<uint16_t>(_columns.size()); }
  uint32_t size() const { return _columns.empty() ? 0u : static_cast<uint32_t>(_columns.front()->size()); }
  void replace_column(size_t column_id, std::shared_ptr<BaseColumn> column) { _columns.at(column_id) = std::move(column); }

 private:
  ChunkColumns _columns;
};

struct TableColumnDefinition {
  TableColumnDefinition() = default;
  TableColumnDefinition(std::string name_, DataType data_type_, bool nullable_ = false)
      : name(std::move(name_)), data_type(data_type_), nullable(nullable_) {}
  std::string name;
  DataType data_type{DataType::Int};
  bool nullable{false};
};
using TableColumnDefinitions = std::vector<TableColumnDefinition>;
inline TableColumnDefinitions concatenated(const TableColumnDefinitions& a, const TableColumnDefinitions& b) {
  auto r = a;
  r.insert(r.end(), b.begin(), b.end());
  return r;
}

// storage/table.hpp
class Table : public std::enable_shared_from_this<Table> {
 public:
  Table(TableColumnDefinitions column_definitions, TableType type, uint32_t max_chunk_size = std::numeric_limits<uint32_t>::max() - 1)
      : _column_definitions(std::move(column_definitions)), _type(type), _max_chunk_size(max_chunk_size) {}
  static std::shared_ptr<Table> create_dummy_table(const TableColumnDefinitions& defs) {
    return std::make_shared<Table>(defs, TableType::Data);
  }
  const TableColumnDefinitions& column_definitions() const { return _column_definitions; }
  TableType type() const { return _type; }
  uint16_t column_count() const { return static_cast<uint16_t>(_column_definitions.size()); }
  const std::string& column_name(ColumnID c) const { return _column_definitions.at(c).name; }
  DataType column_data_type(ColumnID c) const { return _column_definitions.at(c).data_type; }
  bool column_is_nullable(ColumnID c) const { return _column_definitions.at(c).nullable; }
  uint64_t row_count() const {
    uint64_t n = 0;
    for (const auto& c : _chunks) n += c->size();
    return n;
  }
  ChunkID chunk_count() const { return static_cast<ChunkID>(_chunks.size()); }
  const std::vector<std::shared_ptr<Chunk>>& chunks() const { return _chunks; }
  std::shared_ptr<Chunk> get_chunk(ChunkID chunk_id) const { return _chunks.at(chunk_id); }
  uint32_t max_chunk_size() const { return _max_chunk_size; }
  void append_chunk(const ChunkColumns& columns) { _chunks.push_back(std::make_shared<Chunk>(columns)); }
  void append(const std::vector<AllTypeVariant>& values);  // Data tables of ValueColumns only

 private:
  TableColumnDefinitions _column_definitions;
  TableType _type;
  uint32_t _max_chunk_size;
  std::vector<std::shared_ptr<Chunk>> _chunks;
};

// ---- operators -----------------------------------------------------------------------------------------------------------
struct OperatorPerformanceData {
  uint64_t walltime_ns{0};
};

// operators/abstract_operator.hpp:56-172 (lifecycle only: execute() -> _on_execute() -> _on_cleanup())
class AbstractOperator : public std::enable_shared_from_this<AbstractOperator> {
 public:
  AbstractOperator(OperatorType type, std::shared_ptr<const AbstractOperator> left = nullptr,
                   std::shared_ptr<const AbstractOperator> right = nullptr)
      : _type(type), _input_left(std::move(left)), _input_right(std::move(right)) {}
  virtual ~AbstractOperator() = default;
  AbstractOperator(const AbstractOperator&) = delete;
  AbstractOperator& operator=(const AbstractOperator&) = delete;

  OperatorType type() const { return _type; }
  void execute();  // abstract_operator.cpp:25-54
  std::shared_ptr<const Table> get_output() const { return _output; }
  void clear_output() { _output = nullptr; }
  virtual const std::string name() const = 0;
  virtual const std::string description(DescriptionMode = DescriptionMode::SingleLine) const { return name(); }
  std::shared_ptr<AbstractOperator> deep_copy() const;
  std::shared_ptr<const Table> input_table_left() const { return _input_left->get_output(); }
  std::shared_ptr<const Table> input_table_right() const { return _input_right->get_output(); }
  std::shared_ptr<const AbstractOperator> input_left() const { return _input_left; }
  std::shared_ptr<const AbstractOperator> input_right() const { return _input_right; }
  const OperatorPerformanceData& performance_data() const { return _performance_data; }

 protected:
  virtual std::shared_ptr<const Table> _on_execute() = 0;
  virtual void _on_cleanup() {}
  virtual std::shared_ptr<AbstractOperator> _on_deep_copy(const std::shared_ptr<AbstractOperator>& copied_input_left,
                                                          const std::shared_ptr<AbstractOperator>& copied_input_right) const = 0;
  const OperatorType _type;
  std::shared_ptr<const AbstractOperator> _input_left, _input_right;
  std::shared_ptr<const Table> _output;
  OperatorPerformanceData _performance_data;
};
using AbstractReadOnlyOperator = AbstractOperator;  // abstract_read_only_operator.hpp: ignores the transaction context

// operators/abstract_join_operator.hpp:25-53
class AbstractJoinOperator : public AbstractReadOnlyOperator {
 public:
  AbstractJoinOperator(const OperatorType type, const std::shared_ptr<const AbstractOperator>& left,
                       const std::shared_ptr<const AbstractOperator>& right, const JoinMode mode,
                       const ColumnIDPair& column_ids, const PredicateCondition predicate_condition);
  JoinMode mode() const { return _mode; }
  const ColumnIDPair& column_ids() const { return _column_ids; }
  PredicateCondition predicate_condition() const { return _predicate_condition; }
  const std::string description(DescriptionMode description_mode) const override;

 protected:
  const JoinMode _mode;
  const ColumnIDPair _column_ids;
  const PredicateCondition _predicate_condition;
};

// operators/table_wrapper.hpp — hands an existing table to an operator tree
class TableWrapper : public AbstractReadOnlyOperator {
 public:
  explicit TableWrapper(std::shared_ptr<const Table> table) : AbstractReadOnlyOperator(OperatorType::TableWrapper), _table(std::move(table)) {}
  const std::string name() const override { return "TableWrapper"; }

 protected:
  std::shared_ptr<const Table> _on_execute() override { return _table; }
  std::shared_ptr<AbstractOperator> _on_deep_copy(const std::shared_ptr<AbstractOperator>&,
                                                  const std::shared_ptr<AbstractOperator>&) const override {
    return std::make_shared<TableWrapper>(_table);
  }
  std::shared_ptr<const Table> _table;
};

const std::string& join_mode_to_string(JoinMode mode);

}  // namespace opossum
