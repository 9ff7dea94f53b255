// shim_test — replays the reference's own JoinHash test cases (tests/golden/join_fixtures.txt, generated from
// src/test/operators/join_{equi,null,semi_anti}_test.cpp and src/test/tables/**.tbl) through the C++ drop-in operator
// opossum::JoinHash of this directory, i.e. host shim -> C ABI -> B200 kernels, and compares the output tables with the
// reference's expectation tables the way EXPECT_TABLE_EQ_UNORDERED does (src/test/testing_assert.cpp:118-254: rows
// materialised through BaseColumn::operator[], sorted, floats compared with an epsilon).
//
// TEST INFRASTRUCTURE. usage: shim_test <join_fixtures.txt> [case-name-substring]; exit code = number of failed cases.
#include <algorithm>
#include <cmath>
#include <cstdio>
#include <cstdlib>
#include <fstream>
#include <iostream>
#include <map>
#include <sstream>

#include "join_hash.hpp"
#include "aggregate.hpp"
#include "table_scan.hpp"

using namespace opossum;

namespace {

struct TblText {
  std::vector<std::string> names, types;
  std::vector<std::vector<std::string>> rows;
};

std::vector<std::string> split(const std::string& s, char sep) {
  std::vector<std::string> out;
  std::string cur;
  for (char c : s) {
    if (c == sep) { out.push_back(cur); cur.clear(); } else { cur += c; }
  }
  out.push_back(cur);
  return out;
}

DataType parse_type(std::string t, bool* nullable) {
  *nullable = false;
  const std::string suffix = "_null";
  if (t.size() > suffix.size() && t.compare(t.size() - suffix.size(), suffix.size(), suffix) == 0) {
    *nullable = true;
    t.resize(t.size() - suffix.size());
  }
  if (t == "int") return DataType::Int;
  if (t == "long") return DataType::Long;
  if (t == "float") return DataType::Float;
  if (t == "double") return DataType::Double;
  if (t == "string") return DataType::String;
  Fail("unknown column type " + t);
}

AllTypeVariant parse_value(const std::string& s, DataType t) {
  if (s == "null") return NULL_VALUE;
  switch (t) {
    case DataType::Int: return static_cast<int32_t>(std::stol(s));
    case DataType::Long: return static_cast<int64_t>(std::stoll(s));
    case DataType::Float: return std::stof(s);
    case DataType::Double: return std::stod(s);
    default: return s;
  }
}

// utils/load_table.cpp:14-48
std::shared_ptr<Table> load_table(const TblText& txt, size_t chunk_size) {
  TableColumnDefinitions defs;
  std::vector<DataType> types;
  for (size_t c = 0; c < txt.names.size(); ++c) {
    bool nullable;
    const DataType t = parse_type(txt.types[c], &nullable);
    types.push_back(t);
    defs.emplace_back(txt.names[c], t, nullable);
  }
  auto table = std::make_shared<Table>(defs, TableType::Data, static_cast<uint32_t>(chunk_size));
  for (const auto& row : txt.rows) {
    std::vector<AllTypeVariant> values;
    for (size_t c = 0; c < row.size(); ++c) values.push_back(parse_value(row[c], types[c]));
    table->append(values);
  }
  return table;
}

// storage/dictionary_column/dictionary_encoder.hpp:58-130 + fixed_size_byte_aligned_compressor.cpp:21-30
template <typename T>
std::shared_ptr<BaseColumn> encode_dictionary(const ValueColumn<T>& vc) {
  const auto& values = vc.values();
  std::vector<T> dict;
  for (size_t i = 0; i < values.size(); ++i)
    if (!(vc.is_nullable() && vc.null_values()[i])) dict.push_back(values[i]);
  std::sort(dict.begin(), dict.end());
  dict.erase(std::unique(dict.begin(), dict.end()), dict.end());
  const ValueID null_value_id = static_cast<ValueID>(dict.size());
  std::vector<uint32_t> attr(values.size());
  for (size_t i = 0; i < values.size(); ++i) {
    if (vc.is_nullable() && vc.null_values()[i]) attr[i] = null_value_id;
    else attr[i] = static_cast<uint32_t>(std::lower_bound(dict.begin(), dict.end(), values[i]) - dict.begin());
  }
  const size_t max_value = dict.size() + 1;
  std::shared_ptr<const BaseCompressedVector> av;
  if (max_value <= 0xFF) av = std::make_shared<FixedSizeByteAlignedVector<uint8_t>>(std::vector<uint8_t>(attr.begin(), attr.end()));
  else if (max_value <= 0xFFFF) av = std::make_shared<FixedSizeByteAlignedVector<uint16_t>>(std::vector<uint16_t>(attr.begin(), attr.end()));
  else av = std::make_shared<FixedSizeByteAlignedVector<uint32_t>>(attr);
  return std::make_shared<DictionaryColumn<T>>(std::make_shared<const std::vector<T>>(std::move(dict)), av, null_value_id);
}

std::shared_ptr<BaseColumn> encode_column(const std::shared_ptr<BaseColumn>& col) {
  if (auto c = std::dynamic_pointer_cast<ValueColumn<int32_t>>(col)) return encode_dictionary(*c);
  if (auto c = std::dynamic_pointer_cast<ValueColumn<int64_t>>(col)) return encode_dictionary(*c);
  if (auto c = std::dynamic_pointer_cast<ValueColumn<float>>(col)) return encode_dictionary(*c);
  if (auto c = std::dynamic_pointer_cast<ValueColumn<double>>(col)) return encode_dictionary(*c);
  if (auto c = std::dynamic_pointer_cast<ValueColumn<std::string>>(col)) return encode_dictionary(*c);
  return col;
}

void encode_chunks(Table& table, const std::string& which) {
  std::vector<ChunkID> ids;
  if (which == "all") {
    for (ChunkID c = 0; c < table.chunk_count(); ++c) ids.push_back(c);
  } else {
    for (const auto& s : split(which, ',')) ids.push_back(static_cast<ChunkID>(std::stoul(s)));
  }
  for (ChunkID cid : ids) {
    if (cid >= table.chunk_count()) continue;
    auto chunk = table.get_chunk(cid);
    for (size_t c = 0; c < table.column_count(); ++c) chunk->replace_column(c, encode_column(chunk->get_column(static_cast<ColumnID>(c))));
  }
}

double as_double(const AllTypeVariant& v) {
  if (auto p = std::get_if<int32_t>(&v)) return *p;
  if (auto p = std::get_if<int64_t>(&v)) return static_cast<double>(*p);
  if (auto p = std::get_if<float>(&v)) return *p;
  if (auto p = std::get_if<double>(&v)) return *p;
  Fail("not numeric");
}
bool is_null(const AllTypeVariant& v) { return std::holds_alternative<NullValue>(v); }

// operators/table_scan.cpp:78-160: one output chunk per input chunk with matches; output ReferenceColumns point at the
// Data table (pos lists of reference inputs are resolved); NULL never matches.
std::shared_ptr<Table> table_scan(const std::shared_ptr<const Table>& in, ColumnID column_id, const std::string& op, double value) {
  auto out = std::make_shared<Table>(in->column_definitions(), TableType::References);
  for (ChunkID cid = 0; cid < in->chunk_count(); ++cid) {
    const auto chunk = in->get_chunk(cid);
    const auto col = chunk->get_column(column_id);
    std::vector<ChunkOffset> matches;
    for (ChunkOffset i = 0; i < col->size(); ++i) {
      const auto v = (*col)[i];
      if (is_null(v)) continue;
      const double d = as_double(v);
      const bool hit = op == ">=" ? d >= value : op == ">" ? d > value : op == "=" ? d == value : op == "<=" ? d <= value : d < value;
      if (hit) matches.push_back(i);
    }
    if (matches.empty()) continue;
    ChunkColumns cols;
    if (in->type() == TableType::References) {
      std::map<const PosList*, std::shared_ptr<PosList>> cache;
      for (size_t c = 0; c < in->column_count(); ++c) {
        const auto ref = std::dynamic_pointer_cast<const ReferenceColumn>(chunk->get_column(static_cast<ColumnID>(c)));
        auto& filtered = cache[ref->pos_list().get()];
        if (!filtered) {
          filtered = std::make_shared<PosList>();
          for (auto i : matches) filtered->push_back((*ref->pos_list())[i]);
        }
        cols.push_back(std::make_shared<ReferenceColumn>(ref->referenced_table(), ref->referenced_column_id(), filtered));
      }
    } else {
      auto pl = std::make_shared<PosList>();
      for (auto i : matches) pl->push_back(RowID{cid, i});
      for (size_t c = 0; c < in->column_count(); ++c) cols.push_back(std::make_shared<ReferenceColumn>(in, static_cast<ColumnID>(c), pl));
    }
    out->append_chunk(cols);
  }
  return out;
}

// Minimal JoinNestedLoop (Inner/Left) for the one reference case that feeds a nested-loop join's NULL-padded output into
// JoinHash (join_equi_test.cpp:356). Output: one chunk of ReferenceColumns over the Data tables, NULL_ROW_ID padding.
std::shared_ptr<Table> nested_loop_join(const std::shared_ptr<const Table>& left, const std::shared_ptr<const Table>& right,
                                        JoinMode mode, ColumnID lc, ColumnID rc) {
  Assert(left->type() == TableType::Data && right->type() == TableType::Data, "test nlj joins Data tables only");
  TableColumnDefinitions defs = left->column_definitions();
  for (auto d : right->column_definitions()) {
    if (mode == JoinMode::Left) d.nullable = true;
    defs.push_back(d);
  }
  auto lp = std::make_shared<PosList>(), rp = std::make_shared<PosList>();
  for (ChunkID lcid = 0; lcid < left->chunk_count(); ++lcid) {
    const auto lcol = left->get_chunk(lcid)->get_column(lc);
    for (ChunkOffset li = 0; li < lcol->size(); ++li) {
      const auto lv = (*lcol)[li];
      bool matched = false;
      for (ChunkID rcid = 0; rcid < right->chunk_count() && !is_null(lv); ++rcid) {
        const auto rcol = right->get_chunk(rcid)->get_column(rc);
        for (ChunkOffset ri = 0; ri < rcol->size(); ++ri) {
          const auto rv = (*rcol)[ri];
          if (!is_null(rv) && as_double(lv) == as_double(rv)) {
            lp->push_back(RowID{lcid, li});
            rp->push_back(RowID{rcid, ri});
            matched = true;
          }
        }
      }
      if (!matched && mode == JoinMode::Left) {
        lp->push_back(RowID{lcid, li});
        rp->push_back(NULL_ROW_ID);
      }
    }
  }
  auto out = std::make_shared<Table>(defs, TableType::References);
  if (!lp->empty()) {
    ChunkColumns cols;
    for (ColumnID c = 0; c < left->column_count(); ++c) cols.push_back(std::make_shared<ReferenceColumn>(left, c, lp));
    for (ColumnID c = 0; c < right->column_count(); ++c) cols.push_back(std::make_shared<ReferenceColumn>(right, c, rp));
    out->append_chunk(cols);
  }
  return out;
}

// testing_assert.cpp:118-254, unordered comparison
using Row = std::vector<AllTypeVariant>;
std::vector<Row> materialize(const Table& t) {
  std::vector<Row> rows;
  for (ChunkID cid = 0; cid < t.chunk_count(); ++cid) {
    const auto chunk = t.get_chunk(cid);
    for (ChunkOffset i = 0; i < chunk->size(); ++i) {
      Row r;
      for (ColumnID c = 0; c < t.column_count(); ++c) r.push_back((*chunk->get_column(c))[i]);
      rows.push_back(std::move(r));
    }
  }
  return rows;
}
std::string cell(const AllTypeVariant& v) {
  if (is_null(v)) return "~null";  // sorts after numbers
  if (auto p = std::get_if<std::string>(&v)) return *p;
  char buf[64];
  std::snprintf(buf, sizeof(buf), "%020.4f", as_double(v) + 1e12);  // order-preserving text for the sort
  return buf;
}
bool cells_equal(const AllTypeVariant& a, const AllTypeVariant& b) {
  if (is_null(a) || is_null(b)) return is_null(a) && is_null(b);
  if (std::holds_alternative<std::string>(a) || std::holds_alternative<std::string>(b))
    return std::holds_alternative<std::string>(a) && std::holds_alternative<std::string>(b) && std::get<std::string>(a) == std::get<std::string>(b);
  const double x = as_double(a), y = as_double(b);
  return std::fabs(x - y) <= 1e-4 + 1e-6 * std::fabs(y);
}
bool tables_equal_unordered(const Table& actual, const Table& expected, std::string* msg) {
  if (actual.column_count() != expected.column_count()) { *msg = "column count differs"; return false; }
  auto a = materialize(actual), e = materialize(expected);
  if (a.size() != e.size()) { *msg = "row count " + std::to_string(a.size()) + " != expected " + std::to_string(e.size()); return false; }
  auto key = [](const Row& r) { std::string k; for (const auto& v : r) k += cell(v) + "|"; return k; };
  auto by_key = [&](const Row& x, const Row& y) { return key(x) < key(y); };
  std::sort(a.begin(), a.end(), by_key);
  std::sort(e.begin(), e.end(), by_key);
  for (size_t i = 0; i < a.size(); ++i)
    for (size_t c = 0; c < a[i].size(); ++c)
      if (!cells_equal(a[i][c], e[i][c])) { *msg = "row " + std::to_string(i) + " differs: " + key(a[i]) + " vs " + key(e[i]); return false; }
  return true;
}

JoinMode parse_mode(const std::string& m) {
  if (m == "Inner") return JoinMode::Inner;
  if (m == "Left") return JoinMode::Left;
  if (m == "Right") return JoinMode::Right;
  if (m == "Semi") return JoinMode::Semi;
  if (m == "Anti") return JoinMode::Anti;
  Fail("mode " + m);
}

}  // namespace

int main(int argc, char** argv) {
  if (argc < 2) { std::fprintf(stderr, "usage: shim_test <join_fixtures.txt> [filter]\n"); return 2; }
  const std::string filter = argc > 2 ? argv[2] : "";
  std::ifstream in(argv[1]);
  if (!in) { std::fprintf(stderr, "cannot open %s\n", argv[1]); return 2; }
  std::map<std::string, TblText> tables;
  struct Case { std::string name, cite; std::vector<std::vector<std::string>> ops; };
  std::vector<Case> cases;
  std::string line;
  while (std::getline(in, line)) {
    if (line.empty() || line[0] == '#') continue;
    if (line.rfind("@table ", 0) == 0) {
      TblText t;
      const std::string name = line.substr(7);
      std::getline(in, line); t.names = split(line, '|');
      std::getline(in, line); t.types = split(line, '|');
      while (std::getline(in, line) && line != "@end") if (!line.empty()) t.rows.push_back(split(line, '|'));
      tables[name] = std::move(t);
    } else if (line.rfind("@case ", 0) == 0) {
      Case c;
      const auto parts = split(line, ' ');
      c.name = parts.at(1);
      c.cite = parts.size() > 2 ? parts[2] : "";
      while (std::getline(in, line) && line != "@end") if (!line.empty()) c.ops.push_back(split(line, ' '));
      cases.push_back(std::move(c));
    }
  }
  int failed = 0, ran = 0, skipped = 0, n_device_scans = 0;
  const char* ds = std::getenv("HJ_SHIM_DEVICE_SCAN");
  const bool device_scan = ds && ds[0] == '1';
  for (const auto& c : cases) {
    if (!filter.empty() && c.name.find(filter) == std::string::npos) continue;
    bool in_scope = true;
    if (!in_scope) { ++skipped; continue; }  // ("scope" lines name the key types beyond int/long; all of them run now)
    std::map<std::string, std::shared_ptr<const Table>> env;
    std::string msg;
    bool ok = false;
    try {
      for (const auto& op : c.ops) {
        if (op[0] == "base") {
          auto t = load_table(tables.at(op[2]), std::stoul(op[3]));
          if (op.size() > 5 && op[4] == "dict") encode_chunks(*t, op[5]);
          env[op[1]] = t;
        } else if (op[0] == "scan") {
          const auto in_table = env.at(op[2]);
          const auto column_id = static_cast<ColumnID>(std::stoul(op[3]));
          const std::map<std::string, PredicateCondition> conds = {
              {">=", PredicateCondition::GreaterThanEquals}, {">", PredicateCondition::GreaterThan}, {"=", PredicateCondition::Equals},
              {"<=", PredicateCondition::LessThanEquals}, {"<", PredicateCondition::LessThan}, {"!=", PredicateCondition::NotEquals}};
          if (in_table->column_data_type(column_id) == DataType::String) {
            // string columns exist on the device path only (dictionary chunks scanned on value ids, hj_scan_value_ids)
            if (!device_scan) Fail("a scan on a string column needs HJ_SHIM_DEVICE_SCAN=1");
            auto wrapper = std::make_shared<TableWrapper>(in_table);
            wrapper->execute();
            auto scan = std::make_shared<TableScan>(wrapper, column_id, conds.at(op[4]), AllParameterVariant{AllTypeVariant{op[5]}});
            scan->execute();
            env[op[1]] = scan->get_output();
            ++n_device_scans;
            continue;
          }
          const double literal = std::stod(op[5]);
          const bool integral = literal == static_cast<double>(static_cast<int64_t>(literal));
          // HJ_SHIM_DEVICE_SCAN=1: the scans the device operator covers run through opossum::TableScan -> hj_scan, so the
          // case is scan -> join on the GPU; the others keep the semantic helper. (A fractional literal on an INTEGER column
          // stays with the helper: the reference truncates it, type_cast.hpp:56, the helper compares as double.)
          const DataType column_type = in_table->column_data_type(column_id);
          const bool floating_column = column_type == DataType::Float || column_type == DataType::Double;
          if (device_scan && conds.count(op[4]) && (integral || floating_column) &&
              TableScan::supported(*in_table, column_id, conds.at(op[4]), AllParameterVariant{literal})) {
            auto wrapper = std::make_shared<TableWrapper>(in_table);
            wrapper->execute();
            const AllParameterVariant value = floating_column ? AllParameterVariant{literal}
                                              : column_type == DataType::Int ? AllParameterVariant{static_cast<int>(literal)}
                                                                             : AllParameterVariant{static_cast<int64_t>(literal)};
            auto scan = std::make_shared<TableScan>(wrapper, column_id, conds.at(op[4]), value);
            scan->execute();
            env[op[1]] = scan->get_output();
            ++n_device_scans;
          } else {
            env[op[1]] = table_scan(in_table, column_id, op[4], literal);
          }
        } else if (op[0] == "join") {
          auto l = std::make_shared<TableWrapper>(env.at(op[2])), r = std::make_shared<TableWrapper>(env.at(op[3]));
          l->execute();
          r->execute();
          auto join = std::make_shared<JoinHash>(l, r, parse_mode(op[4]),
                                                 ColumnIDPair{static_cast<ColumnID>(std::stoul(op[5])), static_cast<ColumnID>(std::stoul(op[6]))},
                                                 PredicateCondition::Equals);
          join->execute();
          env[op[1]] = join->get_output();
        } else if (op[0] == "aggregate") {
          // aggregate <out> <in> [groupby:<column id>] <FUNCTION>:<column id | *> ...   (opossum::Aggregate -> hj_aggregate_*relation)
          const std::map<std::string, AggregateFunction> functions = {
              {"MIN", AggregateFunction::Min}, {"MAX", AggregateFunction::Max}, {"SUM", AggregateFunction::Sum},
              {"AVG", AggregateFunction::Avg}, {"COUNT", AggregateFunction::Count}};
          std::vector<AggregateColumnDefinition> aggregates;
          std::vector<ColumnID> groupby;
          for (size_t k = 3; k < op.size(); ++k) {
            if (op[k].empty()) continue;  // trailing blank of a line without aggregates
            const auto parts = split(op[k], ':');
            if (parts.at(0) == "groupby") {  // groupby:<column id>
              groupby.push_back(static_cast<ColumnID>(std::stoul(parts.at(1))));
              continue;
            }
            std::optional<ColumnID> column;
            if (parts.at(1) != "*") column = static_cast<ColumnID>(std::stoul(parts[1]));
            aggregates.emplace_back(column, functions.at(parts[0]));
          }
          auto wrapper = std::make_shared<TableWrapper>(env.at(op[2]));
          wrapper->execute();
          auto aggregate = std::make_shared<Aggregate>(wrapper, aggregates, groupby);
          aggregate->execute();
          env[op[1]] = aggregate->get_output();
        } else if (op[0] == "nlj") {
          env[op[1]] = nested_loop_join(env.at(op[2]), env.at(op[3]), parse_mode(op[4]), static_cast<ColumnID>(std::stoul(op[5])),
                                        static_cast<ColumnID>(std::stoul(op[6])));
        } else if (op[0] == "expect") {
          const auto expected = load_table(tables.at(op[2]), 1u << 20);
          ok = tables_equal_unordered(*env.at(op[1]), *expected, &msg);
        }
      }
    } catch (const std::exception& e) {
      ok = false;
      msg = std::string("exception: ") + e.what();
    }
    ++ran;
    std::printf("%s %s (%s)%s%s\n", ok ? "PASS" : "FAIL", c.name.c_str(), c.cite.c_str(), ok ? "" : ": ", msg.c_str());
    failed += ok ? 0 : 1;
  }
  std::printf("shim_test: %d cases run, %d failed, %d out of scope\n", ran, failed, skipped);
  if (device_scan) std::printf("shim_test: %d scans ran on the device (opossum::TableScan -> hj_scan)\n", n_device_scans);
  return failed;
}
