// table_scan.cpp — host half of the device TableScan: what the reference's TableScan::_on_execute does around scan_chunk
// (src/lib/operators/table_scan.cpp:84-171: one output chunk per input chunk with matches, ReferenceColumns over the data
// table, PosLists of reference-table inputs filtered once per distinct list) stays here; scan_chunk for all chunks is one
// hj_scan() call on the GPU.
#include "table_scan.hpp"

#include <cstring>

#include <algorithm>
#include <map>
#include <unordered_set>

#include "hj.h"

namespace opossum {

static_assert(static_cast<int>(PredicateCondition::Equals) == HJ_PRED_EQUALS &&
                  static_cast<int>(PredicateCondition::GreaterThanEquals) == HJ_PRED_GREATER_THAN_EQUALS &&
                  static_cast<int>(PredicateCondition::Between) == HJ_PRED_BETWEEN &&
                  static_cast<int>(PredicateCondition::IsNull) == HJ_PRED_IS_NULL &&
                  static_cast<int>(PredicateCondition::IsNotNull) == HJ_PRED_IS_NOT_NULL,
              "PredicateCondition values are part of the C ABI (types.hpp:166-179)");

namespace hjhost {  // join_hash.cpp
hj_ctx* context();
std::shared_ptr<void> flatten_column(const Table& table, ColumnID column_id, const hj_relation** relation);
}  // namespace hjhost

namespace {

constexpr int kScanSlot = 0;

bool is_null_predicate(PredicateCondition p) { return p == PredicateCondition::IsNull || p == PredicateCondition::IsNotNull; }

// The literal for hj_scan: an integer, or (is_double) the bits of a double. The library converts it to the column's type
// the way type_cast<ColumnDataType>(_right_value) does (single_column_table_scan_impl.cpp:36, type_cast.hpp:39-58).
bool numeric_literal(const AllTypeVariant& v, int64_t* out, bool* is_double) {
  *is_double = false;
  if (const auto* p = std::get_if<int32_t>(&v)) { *out = *p; return true; }
  if (const auto* p = std::get_if<int64_t>(&v)) { *out = *p; return true; }
  double d;
  if (const auto* p = std::get_if<float>(&v)) d = static_cast<double>(*p);
  else if (const auto* p = std::get_if<double>(&v)) d = *p;
  else return false;
  std::memcpy(out, &d, sizeof(d));
  *is_double = true;
  return true;
}

// String columns: the literal as std::string (type_cast<std::string>: integers as decimal text; floating literals are not
// reproduced — lexical_cast's digits), and every chunk a DictionaryColumn<std::string> of a data table, so that the scan
// can run on value ids (hj_scan_value_ids) with the dictionary searched here.
bool string_literal(const AllTypeVariant& v, std::string* out) {
  if (const auto* p = std::get_if<std::string>(&v)) { *out = *p; return true; }
  if (const auto* p = std::get_if<int32_t>(&v)) { *out = std::to_string(*p); return true; }
  if (const auto* p = std::get_if<int64_t>(&v)) { *out = std::to_string(*p); return true; }
  return false;
}

bool all_string_dictionary_chunks(const Table& table, ColumnID column_id) {
  if (table.type() != TableType::Data) return false;
  for (ChunkID c = 0; c < table.chunk_count(); ++c) {
    const auto column = std::dynamic_pointer_cast<const DictionaryColumn<std::string>>(table.get_chunk(c)->get_column(column_id));
    if (!column) return false;
    const uint8_t width = column->attribute_vector_raw()->width_bytes();
    if (width != 1 && width != 2 && width != 4) return false;
  }
  return true;
}

}  // namespace

TableScan::TableScan(const std::shared_ptr<const AbstractOperator>& in, ColumnID left_column_id,
                     const PredicateCondition predicate_condition, const AllParameterVariant& right_parameter)
    : AbstractReadOnlyOperator{OperatorType::TableScan, in},
      _left_column_id{left_column_id},
      _predicate_condition{predicate_condition},
      _right_parameter{right_parameter} {}

TableScan::~TableScan() = default;

void TableScan::set_excluded_chunk_ids(const std::vector<ChunkID>& chunk_ids) { _excluded_chunk_ids = chunk_ids; }
ColumnID TableScan::left_column_id() const { return _left_column_id; }
PredicateCondition TableScan::predicate_condition() const { return _predicate_condition; }
const AllParameterVariant& TableScan::right_parameter() const { return _right_parameter; }
const std::string TableScan::name() const { return "TableScan"; }

const std::string TableScan::description(DescriptionMode description_mode) const {
  std::string column_name = std::string("Col #") + std::to_string(_left_column_id);
  if (_input_left && input_table_left()) column_name = input_table_left()->column_name(_left_column_id);
  const char* ops[] = {"=", "!=", "<", "<=", ">", ">=", "BETWEEN", "IN", "LIKE", "NOT LIKE", "IS NULL", "IS NOT NULL"};
  const auto separator = description_mode == DescriptionMode::MultiLine ? "\n" : " ";
  return name() + separator + "(" + column_name + " " + ops[static_cast<int>(_predicate_condition)] + ")";
}

bool TableScan::supported(const Table& table, ColumnID column_id, PredicateCondition predicate_condition,
                          const AllParameterVariant& right_parameter) {
  const DataType type = table.column_data_type(column_id);
  if (type == DataType::String) {
    if (!all_string_dictionary_chunks(table, column_id)) return false;
    if (is_null_predicate(predicate_condition)) return true;
    if (static_cast<int>(predicate_condition) > static_cast<int>(PredicateCondition::GreaterThanEquals)) return false;
    std::string text;
    return !right_parameter.is_column && (variant_is_null(right_parameter.value) || string_literal(right_parameter.value, &text));
  }
  if (type != DataType::Int && type != DataType::Long && type != DataType::Float && type != DataType::Double) return false;
  if (is_null_predicate(predicate_condition)) return true;
  if (static_cast<int>(predicate_condition) > static_cast<int>(PredicateCondition::GreaterThanEquals)) return false;
  if (right_parameter.is_column) return false;
  int64_t literal;
  bool is_double;
  return variant_is_null(right_parameter.value) || numeric_literal(right_parameter.value, &literal, &is_double);
}

std::shared_ptr<const Table> TableScan::_on_execute() {
  const auto in_table = input_table_left();
  Assert(in_table != nullptr, "TableScan: input has not been executed");
  if (!supported(*in_table, _left_column_id, _predicate_condition, _right_parameter))
    Fail("TableScan on the device covers numeric columns (and dictionary-encoded string columns of data tables) compared with "
         "a literal and IS [NOT] NULL; this scan is not supported and there is no CPU fallback");
  int64_t literal = 0;
  bool literal_is_double = false;
  const bool null_literal = !is_null_predicate(_predicate_condition) && variant_is_null(_right_parameter.value);
  if (!is_null_predicate(_predicate_condition) && !null_literal) numeric_literal(_right_parameter.value, &literal, &literal_is_double);

  auto output = std::make_shared<Table>(in_table->column_definitions(), TableType::References);
  if (in_table->chunk_count() == 0) return output;

  const bool string_column = in_table->column_data_type(_left_column_id) == DataType::String;
  const auto excluded = std::unordered_set<ChunkID>{_excluded_chunk_ids.cbegin(), _excluded_chunk_ids.cend()};
  std::vector<hj_segment> segs;
  std::vector<ChunkID> chunk_of_seg;
  hj_ctx* ctx = hjhost::context();
  hj_scan_info info;
  if (string_column) {
    // dictionary chunks of strings: the two binary searches per chunk run here (single_column_table_scan_impl.cpp:135-150),
    // the attribute vectors are scanned on the device
    std::string text;
    if (!is_null_predicate(_predicate_condition) && !null_literal) string_literal(_right_parameter.value, &text);
    std::vector<uint32_t> bounds;
    uint64_t n_rows = 0;
    for (ChunkID c = 0; c < in_table->chunk_count(); ++c) {
      if (excluded.count(c)) continue;
      const auto column = std::static_pointer_cast<const DictionaryColumn<std::string>>(in_table->get_chunk(c)->get_column(_left_column_id));
      const auto& dictionary = *column->dictionary();
      hj_segment sg{};
      sg.chunk_id = c;
      sg.n_rows = static_cast<uint32_t>(column->size());
      sg.kind = HJ_SEG_DICT;
      sg.key_type = HJ_KEY_INT32;
      sg.attr = column->attribute_vector_raw()->raw_data();
      sg.attr_width = column->attribute_vector_raw()->width_bytes();
      sg.null_value_id = column->null_value_id();
      segs.push_back(sg);
      chunk_of_seg.push_back(c);
      bounds.push_back(static_cast<uint32_t>(std::lower_bound(dictionary.cbegin(), dictionary.cend(), text) - dictionary.cbegin()));
      bounds.push_back(static_cast<uint32_t>(std::upper_bound(dictionary.cbegin(), dictionary.cend(), text) - dictionary.cbegin()));
      n_rows += sg.n_rows;
    }
    hj_relation rel{};
    rel.segs = segs.data();
    rel.n_segs = static_cast<uint32_t>(segs.size());
    rel.n_rows = n_rows;
    rel.location = HJ_LOC_HOST;
    if (null_literal) {
      // comparing with NULL matches nothing (single_column_table_scan_impl.cpp:25-34)
      return output;
    }
    if (hj_scan_value_ids(ctx, &rel, static_cast<int>(_predicate_condition), bounds.data(), kScanSlot, 0u, nullptr, &info) != HJ_OK)
      Fail(std::string("TableScan: ") + hj_last_error(ctx));
  } else {
    // the scanned column, one descriptor per chunk; excluded chunks (table_scan.cpp:93-96) are simply not handed over
    const hj_relation* all = nullptr;
    const auto keep_alive = hjhost::flatten_column(*in_table, _left_column_id, &all);
    for (uint32_t i = 0; i < all->n_segs; ++i) {
      if (excluded.count(static_cast<ChunkID>(i))) continue;
      segs.push_back(all->segs[i]);
      chunk_of_seg.push_back(static_cast<ChunkID>(i));
    }
    hj_relation rel = *all;
    rel.segs = segs.data();
    rel.n_segs = static_cast<uint32_t>(segs.size());
    rel.n_rows = 0;
    for (const auto& s : segs) rel.n_rows += s.n_rows;
    if (hj_scan(ctx, &rel, static_cast<int>(_predicate_condition), literal, null_literal ? 1 : 0, kScanSlot,
                literal_is_double ? HJ_SCAN_VALUE_IS_DOUBLE : 0u, nullptr, &info) != HJ_OK)
      Fail(std::string("TableScan: ") + hj_last_error(ctx));
  }
  std::vector<RowID> matches(std::max<uint64_t>(1, info.n_rows));
  std::vector<uint32_t> begin(segs.size() + 1, 0);
  static_assert(sizeof(RowID) == sizeof(hj_row_id), "PosList::data() is handed to hj_scan_fetch");
  if (hj_scan_fetch(ctx, kScanSlot, reinterpret_cast<hj_row_id*>(matches.data()), matches.size(), begin.data()) != HJ_OK)
    Fail(std::string("TableScan: ") + hj_last_error(ctx));

  for (size_t s = 0; s < segs.size(); ++s) {
    if (begin[s + 1] == begin[s]) continue;  // :97 chunks without matches produce no output chunk
    const ChunkID chunk_id = chunk_of_seg[s];
    const RowID* first = matches.data() + begin[s];
    const RowID* last = matches.data() + begin[s + 1];
    ChunkColumns out_columns;
    if (in_table->type() == TableType::References) {
      // :98-131: the matches are offsets into this reference chunk; every DISTINCT input PosList is filtered once
      const auto chunk_in = in_table->get_chunk(chunk_id);
      std::map<std::shared_ptr<const PosList>, std::shared_ptr<PosList>> filtered_pos_lists;
      for (ColumnID column_id = 0; column_id < in_table->column_count(); ++column_id) {
        const auto ref_column_in = std::dynamic_pointer_cast<const ReferenceColumn>(chunk_in->get_column(column_id));
        DebugAssert(ref_column_in != nullptr, "All columns should be of type ReferenceColumn.");
        const auto pos_list_in = ref_column_in->pos_list();
        auto& filtered = filtered_pos_lists[pos_list_in];
        if (!filtered) {
          filtered = std::make_shared<PosList>();
          filtered->reserve(static_cast<size_t>(last - first));
          for (const RowID* m = first; m != last; ++m) filtered->push_back((*pos_list_in)[m->chunk_offset]);
        }
        out_columns.push_back(std::make_shared<ReferenceColumn>(ref_column_in->referenced_table(),
                                                                ref_column_in->referenced_column_id(), filtered));
      }
    } else {
      // :132-137: the matches ARE the PosList; flatten numbers the chunks by their index, so chunk_id is already right
      auto matches_out = std::make_shared<PosList>(first, last);
      for (auto& row : *matches_out) row.chunk_id = chunk_id;
      for (ColumnID column_id = 0; column_id < in_table->column_count(); ++column_id)
        out_columns.push_back(std::make_shared<ReferenceColumn>(in_table, column_id, matches_out));
    }
    output->append_chunk(out_columns);
  }
  return output;
}

std::shared_ptr<AbstractOperator> TableScan::_on_deep_copy(const std::shared_ptr<AbstractOperator>& copied_input_left,
                                                           const std::shared_ptr<AbstractOperator>&) const {
  // like the reference (table_scan.cpp:72-76), the copy does not inherit the excluded chunk ids
  return std::make_shared<TableScan>(copied_input_left, _left_column_id, _predicate_condition, _right_parameter);
}

void TableScan::_on_cleanup() {
  // the result has been copied into the output's PosLists; the device copy in the slot is no longer needed
  hj_scan_release(hjhost::context(), kScanSlot);
}

}  // namespace opossum
