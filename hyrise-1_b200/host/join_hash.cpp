// join_hash.cpp — host half of the B200 JoinHash: everything the reference does on the host around its hot loops
// (side selection join_hash.cpp:49-87, output assembly :533-613 and :690-708, :829-855) stays here in C++; the hot loops
// themselves (materialize/partition/build/probe, :127-527) run on the GPU behind hj_join().
#include "join_hash.hpp"

#include <atomic>
#include <cstdio>
#include <cstring>
#include <map>
#include <memory>
#include <tuple>
#include <type_traits>

#include "hj.h"

namespace opossum {

static_assert(sizeof(RowID) == sizeof(hj_row_id), "PosList::data() is handed to hj_fetch");
static_assert(static_cast<int>(JoinMode::Inner) == HJ_MODE_INNER && static_cast<int>(JoinMode::Left) == HJ_MODE_LEFT &&
                  static_cast<int>(JoinMode::Right) == HJ_MODE_RIGHT && static_cast<int>(JoinMode::Outer) == HJ_MODE_OUTER &&
                  static_cast<int>(JoinMode::Cross) == HJ_MODE_CROSS && static_cast<int>(JoinMode::Semi) == HJ_MODE_SEMI &&
                  static_cast<int>(JoinMode::Anti) == HJ_MODE_ANTI,
              "JoinMode values are part of the C ABI (types.hpp:189)");

namespace {

std::atomic<int> g_device{0};
std::atomic<uint32_t> g_flags{0};
std::atomic<bool> g_literal_right_swap{true};

// One libhj context per calling thread: operators may run concurrently on scheduler workers
// (node_queue_scheduler.cpp:94-121); a context is not shared between threads (hj.h). Device scratch stays cached across
// operators of the thread so that consecutive joins do not re-allocate.
struct ThreadCtx {
  hj_ctx* ctx = nullptr;
  int device = -1;
  uint32_t flags = 0;
  ~ThreadCtx() {
    if (ctx) hj_destroy(ctx);
  }
};
hj_ctx* acquire_ctx() {
  thread_local ThreadCtx tc;
  const int device = g_device.load();
  const uint32_t flags = g_flags.load();
  if (tc.ctx && (tc.device != device || tc.flags != flags)) {
    hj_destroy(tc.ctx);
    tc.ctx = nullptr;
  }
  if (!tc.ctx) {
    hj_options opts;
    std::memset(&opts, 0, sizeof(opts));
    opts.abi_version = HJ_ABI_VERSION;
    opts.flags = flags;
    opts.radix_bits = -1;
    if (hj_create(device, &opts, &tc.ctx) != HJ_OK) Fail(std::string("JoinHash: ") + hj_last_error(nullptr));
    tc.device = device;
    tc.flags = flags;
  }
  return tc.ctx;
}

// The key column of `table`, one hj_segment per chunk. A ReferenceColumn is handed over as it is stored:
// its PosList plus the flattened key column of the referenced data table (HJ_SEG_REFERENCE); the library gathers
// (value, is_null) on the device (reference_column_iterable.hpp:59-90: NULL RowID / NULL value -> T{} + null flag).
struct FlatColumn {
  std::vector<hj_segment> segs;
  hj_relation rel{};
  // referenced data tables' key columns, flattened once per (table, column)
  std::vector<std::tuple<const Table*, ColumnID, std::unique_ptr<FlatColumn>>> referenced;
};

template <typename T>
void flatten_typed(const Table& table, ColumnID column_id, FlatColumn& out);

template <typename T>
void reference_segment(const ReferenceColumn& ref, hj_segment& seg, FlatColumn& out) {
  const Table* table = ref.referenced_table().get();
  const ColumnID ref_column = ref.referenced_column_id();
  FlatColumn* flat = nullptr;
  for (auto& [t, c, f] : out.referenced)
    if (t == table && c == ref_column) flat = f.get();
  if (!flat) {
    auto fresh = std::make_unique<FlatColumn>();
    flatten_typed<T>(*table, ref_column, *fresh);
    for (const hj_segment& q : fresh->segs)
      Assert(q.kind == HJ_SEG_VALUE || q.kind == HJ_SEG_DICT, "JoinHash: a ReferenceColumn must reference a data table");
    flat = fresh.get();
    out.referenced.emplace_back(table, ref_column, std::move(fresh));
  }
  static_assert(sizeof(RowID) == sizeof(hj_row_id), "PosList::data() is handed to the library as hj_row_id[]");
  seg.kind = HJ_SEG_REFERENCE;
  seg.values = &flat->rel;
  seg.attr = ref.pos_list()->data();
  seg.nulls = nullptr;
  seg.row_offsets = nullptr;  // RowIDs carry the index inside the ReferenceColumn (the intent of join_hash.cpp:257-263; quirk Q2)
}

template <typename T>
void flatten_typed(const Table& table, ColumnID column_id, FlatColumn& out) {
  const auto n_chunks = table.chunk_count();
  out.segs.resize(n_chunks);
  uint64_t rows = 0;
  for (ChunkID chunk_id = 0; chunk_id < n_chunks; ++chunk_id) {
    const auto column = table.get_chunk(chunk_id)->get_column(column_id);
    hj_segment& seg = out.segs[chunk_id];
    std::memset(&seg, 0, sizeof(seg));
    seg.chunk_id = chunk_id;
    seg.n_rows = static_cast<uint32_t>(column->size());
    seg.key_type = std::is_same_v<T, float> ? HJ_KEY_FLOAT : std::is_same_v<T, double> ? HJ_KEY_DOUBLE
                   : sizeof(T) == 8 ? HJ_KEY_INT64 : HJ_KEY_INT32;
    if (const auto* vc = dynamic_cast<const ValueColumn<T>*>(column.get())) {
      seg.kind = HJ_SEG_VALUE;
      seg.values = vc->values().data();
      seg.nulls = vc->is_nullable() ? vc->null_values().data() : nullptr;
    } else if (const auto* dc = dynamic_cast<const DictionaryColumn<T>*>(column.get())) {
      seg.kind = HJ_SEG_DICT;
      seg.values = dc->dictionary()->data();
      seg.dict_n = static_cast<uint32_t>(dc->dictionary()->size());
      seg.null_value_id = dc->null_value_id();
      seg.attr = dc->attribute_vector_raw()->raw_data();
      seg.attr_width = dc->attribute_vector_raw()->width_bytes();
    } else if (const auto* rc = dynamic_cast<const ReferenceColumn*>(column.get())) {
      reference_segment<T>(*rc, seg, out);
    } else {
      Fail("JoinHash: unsupported column encoding for the join key (Value, Dictionary/FixedSizeByteAligned and Reference "
           "columns are supported)");
    }
    rows += seg.n_rows;
  }
  out.rel.segs = out.segs.data();
  out.rel.n_segs = n_chunks;
  out.rel.location = HJ_LOC_HOST;
  out.rel.n_rows = rows;
}

// ---- string keys (hash_traits.hpp:37-41: a join with a string column hashes std::string; the other side is
// lexical_cast to text) -------------------------------------------------------------------------------------------------
// The library joins fixed-width keys, so the operator codes the texts: one dictionary over BOTH key columns maps each
// distinct text to an int32, the device joins the codes. Equal texts <=> equal codes, so the result is the reference's.
// This is the one place where key values are touched on the host (SURVEY §8f N3 asks for dictionary-code joins).
struct CodedColumn {
  std::vector<std::vector<int32_t>> codes;   // per chunk
  std::vector<std::vector<uint8_t>> nulls;   // per chunk, 1 = NULL
  std::vector<hj_segment> segs;
  hj_relation rel{};
};

std::string key_text(const AllTypeVariant& v) {
  // boost::lexical_cast<std::string> of the variant's value (type_cast.hpp:44-48): integers as decimal text, floating
  // point with lcast_precision digits (9 for float, 17 for double)
  if (const auto* s = std::get_if<std::string>(&v)) return *s;
  if (const auto* i = std::get_if<int32_t>(&v)) return std::to_string(*i);
  if (const auto* l = std::get_if<int64_t>(&v)) return std::to_string(*l);
  char buf[64];
  if (const auto* f = std::get_if<float>(&v)) std::snprintf(buf, sizeof(buf), "%.9g", static_cast<double>(*f));
  else std::snprintf(buf, sizeof(buf), "%.17g", std::get<double>(v));
  return buf;
}

void code_column(const Table& table, ColumnID column_id, std::map<std::string, int32_t>& dictionary, CodedColumn& out) {
  const auto n_chunks = table.chunk_count();
  out.codes.resize(n_chunks);
  out.nulls.resize(n_chunks);
  out.segs.resize(n_chunks);
  uint64_t rows = 0;
  for (ChunkID chunk_id = 0; chunk_id < n_chunks; ++chunk_id) {
    const auto column = table.get_chunk(chunk_id)->get_column(column_id);
    const auto n = static_cast<uint32_t>(column->size());
    auto& codes = out.codes[chunk_id];
    auto& nulls = out.nulls[chunk_id];
    codes.assign(n, 0);
    nulls.assign(n, 0);
    for (ChunkOffset i = 0; i < n; ++i) {
      const AllTypeVariant v = (*column)[i];
      if (std::holds_alternative<NullValue>(v)) {
        nulls[i] = 1;
        continue;
      }
      const auto it = dictionary.emplace(key_text(v), static_cast<int32_t>(dictionary.size())).first;
      codes[i] = it->second;
    }
    hj_segment& seg = out.segs[chunk_id];
    std::memset(&seg, 0, sizeof(seg));
    seg.chunk_id = chunk_id;
    seg.n_rows = n;
    seg.kind = dynamic_cast<const ReferenceColumn*>(column.get()) ? HJ_SEG_GATHERED : HJ_SEG_VALUE;
    seg.key_type = HJ_KEY_INT32;
    seg.values = codes.data();
    seg.nulls = nulls.data();
    rows += n;
  }
  out.rel.segs = out.segs.data();
  out.rel.n_segs = n_chunks;
  out.rel.location = HJ_LOC_HOST;
  out.rel.n_rows = rows;
}

void flatten(const Table& table, ColumnID column_id, FlatColumn& out) {
  switch (table.column_data_type(column_id)) {
    case DataType::Int: flatten_typed<int32_t>(table, column_id, out); break;
    case DataType::Long: flatten_typed<int64_t>(table, column_id, out); break;
    case DataType::Float: flatten_typed<float>(table, column_id, out); break;    // hashed as float/double, hash_traits.hpp:14-35
    case DataType::Double: flatten_typed<double>(table, column_id, out); break;
    default:
      Fail("JoinHash: string key columns are dictionary-coded by _on_execute, not flattened");
  }
}

// ---- output assembly, join_hash.cpp:529-613 ---------------------------------------------------------------------------
using PosLists = std::vector<std::shared_ptr<const PosList>>;
using PosListsByColumn = std::vector<std::shared_ptr<PosLists>>;

PosListsByColumn setup_pos_lists_by_column(const std::shared_ptr<const Table>& input_table) {
  DebugAssert(input_table->type() == TableType::References, "Function only works for reference tables");
  std::map<PosLists, std::shared_ptr<PosLists>> shared_pos_lists_by_pos_lists;
  PosListsByColumn pos_lists_by_column(input_table->column_count());
  for (ColumnID column_id{0}; column_id < input_table->column_count(); ++column_id) {
    auto pos_list_ptrs = std::make_shared<PosLists>(input_table->chunk_count());
    for (ChunkID chunk_id{0}; chunk_id < input_table->chunk_count(); chunk_id++) {
      const auto ref_column = std::static_pointer_cast<const ReferenceColumn>(input_table->get_chunk(chunk_id)->get_column(column_id));
      (*pos_list_ptrs)[chunk_id] = ref_column->pos_list();
    }
    pos_lists_by_column[column_id] = shared_pos_lists_by_pos_lists.emplace(*pos_list_ptrs, pos_list_ptrs).first->second;
  }
  return pos_lists_by_column;
}

void write_output_columns(hj_ctx* ctx, ChunkColumns& output_columns, const std::shared_ptr<const Table>& input_table,
                          const PosListsByColumn& input_pos_list_ptrs_sptrs_by_column, std::shared_ptr<PosList> pos_list) {
  std::map<std::shared_ptr<PosLists>, std::shared_ptr<PosList>> output_pos_list_cache;
  std::shared_ptr<Table> dummy_table;
  for (ColumnID column_id{0}; column_id < input_table->column_count(); ++column_id) {
    if (input_table->type() == TableType::References) {
      if (input_table->chunk_count() > 0) {
        const auto& input_table_pos_lists = input_pos_list_ptrs_sptrs_by_column[column_id];
        auto iter = output_pos_list_cache.find(input_table_pos_lists);
        if (iter == output_pos_list_cache.end()) {
          // dereference the emitted RowIDs through the input's PosLists on the device (join_hash.cpp:584-592)
          auto new_pos_list = std::make_shared<PosList>(*pos_list);
          std::vector<const hj_row_id*> lists(input_table_pos_lists->size());
          std::vector<uint32_t> sizes(input_table_pos_lists->size());
          for (size_t c = 0; c < lists.size(); ++c) {
            lists[c] = reinterpret_cast<const hj_row_id*>((*input_table_pos_lists)[c]->data());
            sizes[c] = static_cast<uint32_t>((*input_table_pos_lists)[c]->size());
          }
          if (hj_translate(ctx, reinterpret_cast<hj_row_id*>(new_pos_list->data()), new_pos_list->size(), lists.data(),
                           sizes.data(), static_cast<uint32_t>(lists.size())) != HJ_OK)
            Fail(std::string("JoinHash: ") + hj_last_error(ctx));
          iter = output_pos_list_cache.emplace(input_table_pos_lists, new_pos_list).first;
        }
        auto ref_col = std::static_pointer_cast<const ReferenceColumn>(input_table->get_chunk(ChunkID{0})->get_column(column_id));
        output_columns.push_back(std::make_shared<ReferenceColumn>(ref_col->referenced_table(), ref_col->referenced_column_id(), iter->second));
      } else {
        // no chunks: pos_list holds only NULL_ROW_IDs; reference a dummy table (join_hash.cpp:601-607)
        if (!dummy_table) dummy_table = Table::create_dummy_table(input_table->column_definitions());
        output_columns.push_back(std::make_shared<ReferenceColumn>(dummy_table, column_id, pos_list));
      }
    } else {
      output_columns.push_back(std::make_shared<ReferenceColumn>(input_table, column_id, pos_list));
    }
  }
}

}  // namespace

JoinHash::JoinHash(const std::shared_ptr<const AbstractOperator>& left, const std::shared_ptr<const AbstractOperator>& right,
                   const JoinMode mode, const ColumnIDPair& column_ids, const PredicateCondition predicate_condition,
                   const size_t radix_bits)
    : AbstractJoinOperator(OperatorType::JoinHash, left, right, mode, column_ids, predicate_condition), _radix_bits(radix_bits) {
  DebugAssert(predicate_condition == PredicateCondition::Equals, "Operator not supported by Hash Join.");
}

JoinHash::~JoinHash() = default;

const std::string JoinHash::name() const { return "JoinHash"; }

void JoinHash::set_device(int device_id) { g_device = device_id; }
void JoinHash::set_library_flags(uint32_t flags) { g_flags = flags; }
void JoinHash::set_literal_right_join_swap(bool literal) { g_literal_right_swap = literal; }

std::shared_ptr<AbstractOperator> JoinHash::_on_deep_copy(const std::shared_ptr<AbstractOperator>& copied_input_left,
                                                          const std::shared_ptr<AbstractOperator>& copied_input_right) const {
  return std::make_shared<JoinHash>(copied_input_left, copied_input_right, _mode, _column_ids, _predicate_condition);
}

void JoinHash::_on_cleanup() { _ctx = nullptr; }  // the context (and its cached device scratch) belongs to the thread

std::shared_ptr<const Table> JoinHash::_on_execute() {
  Assert(_mode != JoinMode::Outer && _mode != JoinMode::Cross, "JoinHash supports Inner, Left, Right, Semi and Anti joins");
  Assert(_predicate_condition == PredicateCondition::Equals, "Operator not supported by Hash Join.");

  // (1) outer relation becomes probe relation (swap only for Left); (2) Semi/Anti always swap; (3) else the smaller input builds
  bool inputs_swapped = (_mode == JoinMode::Left || _mode == JoinMode::Anti || _mode == JoinMode::Semi);
  const bool size_rule_applies = _mode == JoinMode::Inner || (_mode == JoinMode::Right && g_literal_right_swap.load());
  if (!inputs_swapped && size_rule_applies && input_table_left()->row_count() > input_table_right()->row_count()) {
    inputs_swapped = true;
  }
  const auto build_input = inputs_swapped ? input_table_right() : input_table_left();
  const auto probe_input = inputs_swapped ? input_table_left() : input_table_right();
  const ColumnID build_column_id = inputs_swapped ? _column_ids.second : _column_ids.first;
  const ColumnID probe_column_id = inputs_swapped ? _column_ids.first : _column_ids.second;

  // output column definitions, join_hash.cpp:690-708 (left_in_table == build, right_in_table == probe there)
  const bool semi_anti = _mode == JoinMode::Semi || _mode == JoinMode::Anti;
  TableColumnDefinitions output_column_definitions;
  if (inputs_swapped) {
    output_column_definitions = semi_anti ? probe_input->column_definitions()
                                          : concatenated(probe_input->column_definitions(), build_input->column_definitions());
  } else {
    output_column_definitions = concatenated(build_input->column_definitions(), probe_input->column_definitions());
  }
  auto output_table = std::make_shared<Table>(output_column_definitions, TableType::References);

  {
    // a Float x Double pair is widened by the reference through boost::lexical_cast's 9-digit decimal text
    // (type_cast.hpp:44-48), not by the IEEE cast the library uses: refuse it rather than differ in the last bits
    const auto bt = build_input->column_data_type(build_column_id), pt = probe_input->column_data_type(probe_column_id);
    Assert(!((bt == DataType::Float && pt == DataType::Double) || (bt == DataType::Double && pt == DataType::Float)),
           "JoinHash: float x double key pairs are not supported by the B200 path");
  }
  _ctx = acquire_ctx();
  FlatColumn build_flat, probe_flat;
  CodedColumn build_coded, probe_coded;
  const hj_relation* build_rel = &build_flat.rel;
  const hj_relation* probe_rel = &probe_flat.rel;
  if (build_input->column_data_type(build_column_id) == DataType::String ||
      probe_input->column_data_type(probe_column_id) == DataType::String) {
    std::map<std::string, int32_t> dictionary;  // shared by both sides
    code_column(*build_input, build_column_id, dictionary, build_coded);
    code_column(*probe_input, probe_column_id, dictionary, probe_coded);
    build_rel = &build_coded.rel;
    probe_rel = &probe_coded.rel;
  } else {
    flatten(*build_input, build_column_id, build_flat);
    flatten(*probe_input, probe_column_id, probe_flat);
  }

  hj_result_info info;
  if (hj_join(_ctx, build_rel, probe_rel, static_cast<int>(_mode), nullptr, &info) != HJ_OK)
    Fail(std::string("JoinHash: ") + hj_last_error(_ctx));
  if (info.n_rows == 0) return output_table;  // no empty chunks (abstract_operator.cpp:57-67)

  auto build_pos = std::make_shared<PosList>();
  auto probe_pos = std::make_shared<PosList>(info.n_rows);
  if (!semi_anti) build_pos->resize(info.n_rows);
  if (hj_fetch(_ctx, semi_anti ? nullptr : reinterpret_cast<hj_row_id*>(build_pos->data()),
               reinterpret_cast<hj_row_id*>(probe_pos->data()), info.n_rows) != HJ_OK)
    Fail(std::string("JoinHash: ") + hj_last_error(_ctx));

  PosListsByColumn build_pos_lists_by_column, probe_pos_lists_by_column;
  if (build_input->type() == TableType::References && !semi_anti) build_pos_lists_by_column = setup_pos_lists_by_column(build_input);
  if (probe_input->type() == TableType::References) probe_pos_lists_by_column = setup_pos_lists_by_column(probe_input);

  // one output chunk for the whole result (the reference appends one per partition; chunking is unspecified, join_hash.hpp:19-20)
  ChunkColumns output_columns;
  if (inputs_swapped) {
    write_output_columns(_ctx, output_columns, probe_input, probe_pos_lists_by_column, probe_pos);
    if (!semi_anti) write_output_columns(_ctx, output_columns, build_input, build_pos_lists_by_column, build_pos);
  } else {
    write_output_columns(_ctx, output_columns, build_input, build_pos_lists_by_column, build_pos);
    write_output_columns(_ctx, output_columns, probe_input, probe_pos_lists_by_column, probe_pos);
  }
  output_table->append_chunk(output_columns);
  return output_table;
}

// ---- shared with the other device operators of the shim (table_scan.cpp) ---------------------------------------------
namespace hjhost {
hj_ctx* context() { return acquire_ctx(); }
std::shared_ptr<void> flatten_column(const Table& table, ColumnID column_id, const hj_relation** relation) {
  auto flat = std::make_shared<FlatColumn>();
  flatten(table, column_id, *flat);
  *relation = &flat->rel;
  return flat;  // keeps the descriptors `relation` points into alive
}
}  // namespace hjhost

}  // namespace opossum
