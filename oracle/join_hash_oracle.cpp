// ORACLE — TEST INFRASTRUCTURE, NOT PRODUCT CODE.
//
// CPU restatement of the reference's JoinHash (src/lib/operators/join_hash.cpp, Hyrise snapshot 47ed4dc2).
// Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs may load this
// library; the product path (hyrise-1_b200/) never does and has no CPU fallback.
//
// Parity status: PINNED. The restatement is checked (tests/test_oracle_golden.py) against every JoinHash
// expectation table of the reference's own test-suite that lies in scope (join_equi_test.cpp,
// join_null_test.cpp, join_semi_anti_test.cpp; fixtures consolidated in tests/golden/join_fixtures.txt by
// tests/golden/make_golden.py), and its murmur2 against known answers produced by the reference's own
// src/lib/utils/murmur_hash.cpp compiled into oracle/_ref/ (tests/golden/murmur_kat.json).
//
// The reference itself cannot be compiled here (Boost + TBB are hard dependencies of types.hpp:3-6 and are not
// installed), so the data structures are restated with the standard library:
//   boost::variant<RowID, PosList> -> std::variant, tbb::concurrent_vector -> std::vector,
//   JobTask/CurrentScheduler -> a std::thread pool with the same task granularity
//   (one job per chunk for materialize/scatter, one per partition for build/probe).
// Inputs arrive as the same hj_relation segment descriptors libhj takes (include/hj.h), so both are driven by the
// same harness.

#include <algorithm>
#include <atomic>
#include <cmath>
#include <cstdint>
#include <cstdlib>
#include <cstring>
#include <functional>
#include <memory>
#include <optional>
#include <string>
#include <thread>
#include <type_traits>
#include <unordered_map>
#include <variant>
#include <vector>

#include "../include/hj.h"

namespace {

// ---- types.hpp:92-99,138,150 -------------------------------------------------------------------------------
using ChunkID = uint32_t;
using ChunkOffset = uint32_t;
constexpr ChunkOffset INVALID_CHUNK_OFFSET = 0xFFFFFFFFu;
constexpr ChunkID INVALID_CHUNK_ID = 0xFFFFFFFFu;
struct RowID {
  ChunkID chunk_id{INVALID_CHUNK_ID};
  ChunkOffset chunk_offset{INVALID_CHUNK_OFFSET};
};
static_assert(sizeof(RowID) == 8 && sizeof(RowID) == sizeof(hj_row_id), "RowID layout, types.hpp:97-99");
const RowID NULL_ROW_ID{};
using PosList = std::vector<RowID>;

// ---- utils/murmur_hash.cpp:21-73 (MurmurHash2, Austin Appleby), restated ------------------------------------
uint32_t murmur_hash2(const void* key, uint32_t len, uint32_t seed) {
  const uint32_t m = 0x5bd1e995u;
  const int r = 24;
  uint32_t h = seed ^ len;
  const auto* data = static_cast<const unsigned char*>(key);
  while (len >= 4) {
    uint32_t k;
    std::memcpy(&k, data, 4);
    k *= m;
    k ^= k >> r;
    k *= m;
    h *= m;
    h ^= k;
    data += 4;
    len -= 4;
  }
  switch (len) {
    case 3:
      h ^= static_cast<uint32_t>(data[2]) << 16u;
      [[fallthrough]];
    case 2:
      h ^= static_cast<uint32_t>(data[1]) << 8u;
      [[fallthrough]];
    case 1:
      h ^= data[0];
      h *= m;
  }
  h ^= h >> 13u;
  h *= m;
  h ^= h >> 15u;
  return h;
}

// utils/murmur_hash.hpp:12-15
template <typename T>
uint32_t murmur2(T key, uint32_t seed) {
  return murmur_hash2(&key, sizeof(T), seed);
}

using Hash = uint32_t;  // join_hash.cpp:92

// ---- join_hash.cpp:98-122 ----------------------------------------------------------------------------------
template <typename T>
struct PartitionedElement {
  PartitionedElement() : row_id(NULL_ROW_ID), partition_hash(0), value(T()) {}
  PartitionedElement(RowID row, Hash hash, T val) : row_id(row), partition_hash(hash), value(val) {}
  RowID row_id;
  Hash partition_hash{0};
  T value;
};
template <typename T>
using Partition = std::vector<PartitionedElement<T>>;
template <typename T>
using HashTable = std::unordered_map<T, std::variant<RowID, PosList>>;
template <typename T>
struct RadixContainer {
  std::shared_ptr<Partition<T>> elements;
  std::vector<size_t> partition_offsets;
};

// ---- JobTask / CurrentScheduler::wait_for_tasks stand-in ------------------------------------------------------
// Without a scheduler the reference runs each job inline (abstract_task.cpp:59-68) == n_threads 1.
void run_jobs(std::vector<std::function<void()>>& jobs, int n_threads) {
  if (n_threads <= 1 || jobs.size() <= 1) {
    for (auto& job : jobs) job();
    return;
  }
  std::atomic<size_t> next{0};
  const auto worker = [&]() {
    for (;;) {
      const size_t i = next.fetch_add(1);
      if (i >= jobs.size()) return;
      jobs[i]();
    }
  };
  std::vector<std::thread> threads;
  const size_t n = std::min<size_t>(static_cast<size_t>(n_threads), jobs.size());
  threads.reserve(n);
  for (size_t t = 0; t < n; ++t) threads.emplace_back(worker);
  for (auto& t : threads) t.join();
}

// ---- column iterables: what `iterable.for_each` yields per row -------------------------------------------------
// value_column_iterable.hpp:90-100   : {stored value, null flag, chunk_offset}
// dictionary_column_iterable.hpp:76-87: NULL (attr == null_value_id) -> {T{}, true, chunk_offset}
// reference_column_iterable.hpp:59-90 : NULL -> {T{}, true, ...}; the shim supplies the gathered values and the
//                                       chunk offsets to report (hj_segment.row_offsets)
template <typename T>
struct ColumnValue {
  T value;
  bool is_null;
  ChunkOffset chunk_offset;
};

template <typename T, typename F>
void for_each_in_segment(const hj_segment& seg, const F& f) {
  const auto n = seg.n_rows;
  if (seg.kind == HJ_SEG_VALUE || seg.kind == HJ_SEG_GATHERED) {
    const auto* values = static_cast<const T*>(seg.values);
    for (uint32_t i = 0; i < n; ++i) {
      const bool is_null = seg.nulls && seg.nulls[i] != 0;
      T value = values[i];
      if (seg.kind == HJ_SEG_GATHERED && is_null) value = T{};
      const ChunkOffset off = (seg.kind == HJ_SEG_GATHERED && seg.row_offsets) ? seg.row_offsets[i] : i;
      f(ColumnValue<T>{value, is_null, off});
    }
  } else {  // HJ_SEG_DICT
    const auto* dict = static_cast<const T*>(seg.values);
    for (uint32_t i = 0; i < n; ++i) {
      uint32_t value_id;
      switch (seg.attr_width) {
        case 1: value_id = static_cast<const uint8_t*>(seg.attr)[i]; break;
        case 2: value_id = static_cast<const uint16_t*>(seg.attr)[i]; break;
        default: value_id = static_cast<const uint32_t*>(seg.attr)[i]; break;
      }
      if (value_id == seg.null_value_id) {
        f(ColumnValue<T>{T{}, true, i});
      } else {
        f(ColumnValue<T>{dict[value_id], false, i});
      }
    }
  }
}

// ---- join_hash.cpp:203-285 materialize_input --------------------------------------------------------------------
// `sql_null_probe` (HJ_FLAG_SQL_NULL_PROBE) is NOT in the reference: kept NULLs get partition_hash/value as usual but
// are remembered so that probe() can skip the lookup (quirk Q4 switch).
template <typename T, typename HashedType>
std::shared_ptr<Partition<T>> materialize_input(const hj_relation& in_table,
                                                std::vector<std::shared_ptr<std::vector<size_t>>>& histograms,
                                                const size_t radix_bits, const unsigned int partitioning_seed,
                                                bool keep_nulls, int n_threads,
                                                std::vector<uint8_t>* null_flags /* per element, or nullptr */) {
  auto elements = std::make_shared<Partition<T>>();
  elements->resize(in_table.n_rows);
  if (null_flags) null_flags->assign(in_table.n_rows, 0);

  const size_t num_partitions = 1ull << radix_bits;
  size_t pass = 0;
  size_t mask = static_cast<uint32_t>(pow(2, radix_bits * (pass + 1)) - 1);

  auto chunk_offsets = std::vector<size_t>(in_table.n_segs);
  size_t output_offset = 0;
  for (ChunkID chunk_id{0}; chunk_id < in_table.n_segs; chunk_id++) {
    chunk_offsets[chunk_id] = output_offset;
    output_offset += in_table.segs[chunk_id].n_rows;
  }

  histograms = std::vector<std::shared_ptr<std::vector<size_t>>>();
  histograms.resize(chunk_offsets.size());

  std::vector<std::function<void()>> jobs;
  jobs.reserve(in_table.n_segs);
  for (ChunkID seg_idx{0}; seg_idx < in_table.n_segs; ++seg_idx) {
    jobs.emplace_back([&, seg_idx]() {
      auto elem_offset = chunk_offsets[seg_idx];
      auto output_iterator = elements->begin() + elem_offset;
      const hj_segment& seg = in_table.segs[seg_idx];
      const ChunkID chunk_id = seg.chunk_id;

      histograms[seg_idx] = std::make_shared<std::vector<size_t>>(num_partitions);
      auto& histogram = *histograms[seg_idx];

      for_each_in_segment<T>(seg, [&](const ColumnValue<T>& value) {
        if (!value.is_null || keep_nulls) {
          // hash_value<T, HashedType>, join_hash.cpp:191-201: cast to the hashed type first
          const Hash hashed_value = murmur2<HashedType>(static_cast<HashedType>(value.value), partitioning_seed);
          if (null_flags && value.is_null) (*null_flags)[output_iterator - elements->begin()] = 1;
          // join_hash.cpp:261 is a dead branch (quirk Q2): RowIDs always use value.chunk_offset()
          *(output_iterator++) = PartitionedElement<T>{RowID{chunk_id, value.chunk_offset}, hashed_value, value.value};
          const Hash radix = hashed_value & mask;
          histogram[radix]++;
        }
      });
    });
  }
  run_jobs(jobs, n_threads);
  return elements;
}

// ---- join_hash.cpp:287-355 partition_radix_parallel ---------------------------------------------------------------
template <typename T>
RadixContainer<T> partition_radix_parallel(const std::shared_ptr<Partition<T>>& materialized,
                                           const std::vector<size_t>& offsets,
                                           std::vector<std::shared_ptr<std::vector<size_t>>>& histograms,
                                           const size_t radix_bits, bool keep_nulls, int n_threads,
                                           const std::vector<uint8_t>* null_flags_in, std::vector<uint8_t>* null_flags_out) {
  const size_t num_partitions = 1ull << radix_bits;
  size_t pass = 0;
  size_t mask = static_cast<uint32_t>(pow(2, radix_bits * (pass + 1)) - 1);

  auto output = std::make_shared<Partition<T>>();
  output->resize(materialized->size());
  if (null_flags_out) null_flags_out->assign(materialized->size(), 0);

  RadixContainer<T> radix_output;
  radix_output.elements = output;
  radix_output.partition_offsets.resize(num_partitions + 1);

  // partition-major, chunk-minor exclusive prefix sum (join_hash.cpp:310-319)
  size_t offset = 0;
  std::vector<std::vector<size_t>> output_offsets_by_chunk(offsets.size(), std::vector<size_t>(num_partitions));
  for (size_t partition_id = 0; partition_id < num_partitions; ++partition_id) {
    radix_output.partition_offsets[partition_id] = offset;
    for (ChunkID chunk_id{0}; chunk_id < offsets.size(); ++chunk_id) {
      output_offsets_by_chunk[chunk_id][partition_id] = offset;
      offset += (*histograms[chunk_id])[partition_id];
    }
  }
  radix_output.partition_offsets[num_partitions] = offset;

  std::vector<std::function<void()>> jobs;
  jobs.reserve(offsets.size());
  for (ChunkID chunk_id{0}; chunk_id < offsets.size(); ++chunk_id) {
    jobs.emplace_back([&, chunk_id]() {
      size_t input_offset = offsets[chunk_id];
      auto& output_offsets = output_offsets_by_chunk[chunk_id];
      size_t input_size = 0;
      if (chunk_id < offsets.size() - 1) {
        input_size = offsets[chunk_id + 1] - input_offset;
      } else {
        input_size = materialized->size() - input_offset;
      }
      auto& out = *output;
      for (size_t column_offset = input_offset; column_offset < input_offset + input_size; ++column_offset) {
        auto& element = (*materialized)[column_offset];
        if (!keep_nulls && element.row_id.chunk_offset == INVALID_CHUNK_OFFSET) {
          continue;
        }
        const size_t radix = element.partition_hash & mask;
        if (null_flags_out) (*null_flags_out)[output_offsets[radix]] = (*null_flags_in)[column_offset];
        out[output_offsets[radix]++] = element;
      }
    });
  }
  run_jobs(jobs, n_threads);
  return radix_output;
}

// ---- join_hash.cpp:127-185 build ------------------------------------------------------------------------------------
template <typename LeftType, typename HashedType>
std::vector<std::optional<HashTable<HashedType>>> build(const RadixContainer<LeftType>& radix_container, int n_threads) {
  std::vector<std::optional<HashTable<HashedType>>> hashtables;
  hashtables.resize(radix_container.partition_offsets.size() - 1);

  std::vector<std::function<void()>> jobs;
  jobs.reserve(radix_container.partition_offsets.size() - 1);

  for (size_t current_partition_id = 0; current_partition_id < (radix_container.partition_offsets.size() - 1);
       ++current_partition_id) {
    const auto partition_left_begin = radix_container.partition_offsets[current_partition_id];
    const auto partition_left_end = radix_container.partition_offsets[current_partition_id + 1];
    const auto partition_size = partition_left_end - partition_left_begin;
    if (partition_size == 0) continue;  // empty partitions keep std::nullopt (join_hash.cpp:146-148)

    jobs.emplace_back([&, partition_left_begin, partition_left_end, current_partition_id, partition_size]() {
      auto& partition_left = *radix_container.elements;
      auto hashtable = HashTable<HashedType>(partition_size);
      for (size_t partition_offset = partition_left_begin; partition_offset < partition_left_end; ++partition_offset) {
        const auto& element = partition_left[partition_offset];
        const auto hash_key = static_cast<HashedType>(element.value);  // type_cast<HashedType>
        const auto it = hashtable.find(hash_key);
        if (it == hashtable.end()) {
          hashtable[hash_key] = element.row_id;
        } else {
          auto& map_entry = it->second;
          if (std::holds_alternative<RowID>(map_entry)) {
            hashtable[hash_key] = PosList{std::get<RowID>(map_entry), element.row_id};
          } else {
            std::get<PosList>(map_entry).push_back(element.row_id);
          }
        }
      }
      hashtables[current_partition_id] = std::move(hashtable);
    });
  }
  run_jobs(jobs, n_threads);
  return hashtables;
}

// ---- join_hash.cpp:362-466 probe --------------------------------------------------------------------------------------
template <typename RightType, typename HashedType>
void probe(const RadixContainer<RightType>& radix_container,
           const std::vector<std::optional<HashTable<HashedType>>>& hashtables, std::vector<PosList>& pos_lists_left,
           std::vector<PosList>& pos_lists_right, const int mode, int n_threads,
           const std::vector<uint8_t>* sql_null_flags) {
  std::vector<std::function<void()>> jobs;
  jobs.reserve(radix_container.partition_offsets.size() - 1);

  for (size_t current_partition_id = 0; current_partition_id < (radix_container.partition_offsets.size() - 1);
       ++current_partition_id) {
    const auto partition_begin = radix_container.partition_offsets[current_partition_id];
    const auto partition_end = radix_container.partition_offsets[current_partition_id + 1];
    if (partition_begin == partition_end) continue;

    jobs.emplace_back([&, partition_begin, partition_end, current_partition_id]() {
      auto& partition = *radix_container.elements;
      PosList pos_list_left_local;
      PosList pos_list_right_local;

      if (hashtables[current_partition_id].has_value()) {
        const auto& hashtable = hashtables.at(current_partition_id).value();
        const size_t expected_output_size =
            static_cast<size_t>(std::max(10.0, std::ceil((partition_end - partition_begin) / 2)));
        pos_list_left_local.reserve(expected_output_size);
        pos_list_right_local.reserve(expected_output_size);

        for (size_t partition_offset = partition_begin; partition_offset < partition_end; ++partition_offset) {
          auto& row = partition[partition_offset];
          if (mode == HJ_MODE_INNER && row.row_id.chunk_offset == INVALID_CHUNK_OFFSET) continue;

          auto rows_iter = hashtable.find(static_cast<HashedType>(row.value));
          // Q4 switch (not in the reference): SQL semantics, a NULL key matches nothing
          if (sql_null_flags && (*sql_null_flags)[partition_offset]) rows_iter = hashtable.end();

          if (rows_iter != hashtable.end()) {
            const auto& matching_rows_variant = rows_iter->second;
            if (std::holds_alternative<PosList>(matching_rows_variant)) {
              for (const auto& row_id : std::get<PosList>(matching_rows_variant)) {
                if (row_id.chunk_offset != INVALID_CHUNK_OFFSET) {
                  pos_list_left_local.emplace_back(row_id);
                  pos_list_right_local.emplace_back(row.row_id);
                }
              }
            } else {
              const auto row_id = std::get<RowID>(matching_rows_variant);
              if (row_id.chunk_offset != INVALID_CHUNK_OFFSET) {
                pos_list_left_local.emplace_back(row_id);
                pos_list_right_local.emplace_back(row.row_id);
              }
            }
          } else if (mode == HJ_MODE_LEFT || mode == HJ_MODE_RIGHT) {
            pos_list_left_local.emplace_back(NULL_ROW_ID);
            pos_list_right_local.emplace_back(row.row_id);
          }
        }
      } else if (mode == HJ_MODE_LEFT || mode == HJ_MODE_RIGHT) {
        pos_list_left_local.reserve(partition_end - partition_begin);
        pos_list_right_local.reserve(partition_end - partition_begin);
        for (size_t partition_offset = partition_begin; partition_offset < partition_end; ++partition_offset) {
          auto& row = partition[partition_offset];
          pos_list_left_local.emplace_back(NULL_ROW_ID);
          pos_list_right_local.emplace_back(row.row_id);
        }
      }

      if (!pos_list_left_local.empty()) {
        pos_lists_left[current_partition_id] = std::move(pos_list_left_local);
        pos_lists_right[current_partition_id] = std::move(pos_list_right_local);
      }
    });
  }
  run_jobs(jobs, n_threads);
}

// ---- join_hash.cpp:468-527 probe_semi_anti ---------------------------------------------------------------------------
template <typename RightType, typename HashedType>
void probe_semi_anti(const RadixContainer<RightType>& radix_container,
                     const std::vector<std::optional<HashTable<HashedType>>>& hashtables,
                     std::vector<PosList>& pos_lists, const int mode, int n_threads) {
  std::vector<std::function<void()>> jobs;
  jobs.reserve(radix_container.partition_offsets.size() - 1);

  for (size_t current_partition_id = 0; current_partition_id < (radix_container.partition_offsets.size() - 1);
       ++current_partition_id) {
    const auto partition_begin = radix_container.partition_offsets[current_partition_id];
    const auto partition_end = radix_container.partition_offsets[current_partition_id + 1];
    if (partition_begin == partition_end) continue;

    jobs.emplace_back([&, partition_begin, partition_end, current_partition_id]() {
      auto& partition = *radix_container.elements;
      PosList pos_list_local;
      if (hashtables[current_partition_id].has_value()) {
        for (size_t partition_offset = partition_begin; partition_offset < partition_end; ++partition_offset) {
          auto& row = partition[partition_offset];
          if (row.row_id.chunk_offset == INVALID_CHUNK_OFFSET) continue;
          const auto& hashtable = hashtables[current_partition_id].value();
          const auto it = hashtable.find(static_cast<HashedType>(row.value));
          if ((mode == HJ_MODE_SEMI && it != hashtable.end()) || (mode == HJ_MODE_ANTI && it == hashtable.end())) {
            pos_list_local.emplace_back(row.row_id);
          }
        }
      } else if (mode == HJ_MODE_ANTI) {
        pos_list_local.reserve(partition_end - partition_begin);
        for (size_t partition_offset = partition_begin; partition_offset < partition_end; ++partition_offset) {
          auto& row = partition[partition_offset];
          pos_list_local.emplace_back(row.row_id);
        }
      }
      if (!pos_list_local.empty()) pos_lists[current_partition_id] = std::move(pos_list_local);
    });
  }
  run_jobs(jobs, n_threads);
}

// ---- join_hash.cpp:640-668 radix-bit heuristic (sizeof(PosList) == 32 with boost's pmr vector) ----
template <typename LeftType>
size_t reference_radix_bits(size_t build_relation_size) {
  const auto l2_cache_size = 256'000;
  const size_t sizeof_pos_list = 32;
  const auto complete_hash_map_size = build_relation_size * (sizeof(LeftType) + sizeof(void*)) +
                                      (build_relation_size / 2) * (sizeof_pos_list + 2 * sizeof(RowID));
  const auto adaption_factor = 2.0f;
  const auto cluster_count = std::max(1.0f, (adaption_factor * complete_hash_map_size) / l2_cache_size);
  return static_cast<size_t>(std::ceil(std::log2(cluster_count)));
}

struct OracleResult {
  PosList left;   // build side
  PosList right;  // probe side
  size_t radix_bits = 0;
  uint64_t n_build_tuples = 0;
  uint64_t n_probe_tuples = 0;
};

// ---- join_hash.cpp:686-800 JoinHashImpl<L,R>::_on_execute, up to (not including) write_output_columns ----------------
template <typename LeftType, typename RightType>
OracleResult join_impl(const hj_relation& left_in_table, const hj_relation& right_in_table, int mode, int radix_bits_override,
                       uint32_t flags, int n_threads) {
  using HashedType = std::conditional_t<sizeof(LeftType) < sizeof(RightType), RightType, LeftType>;  // hash_traits.hpp:21-25
  const unsigned int partitioning_seed = 17;  // join_hash.cpp:680

  OracleResult result;
  size_t radix_bits = reference_radix_bits<LeftType>(left_in_table.n_rows);
  if (radix_bits_override >= 0) radix_bits = static_cast<size_t>(radix_bits_override);
  result.radix_bits = radix_bits;

  auto keep_nulls = (mode == HJ_MODE_LEFT || mode == HJ_MODE_RIGHT);  // join_hash.cpp:715
  const bool sql_null_probe = keep_nulls && (flags & HJ_FLAG_SQL_NULL_PROBE);

  std::vector<size_t> left_chunk_offsets(left_in_table.n_segs), right_chunk_offsets(right_in_table.n_segs);
  size_t offset_left = 0;
  for (uint32_t i = 0; i < left_in_table.n_segs; ++i) {
    left_chunk_offsets[i] = offset_left;
    offset_left += left_in_table.segs[i].n_rows;
  }
  size_t offset_right = 0;
  for (uint32_t i = 0; i < right_in_table.n_segs; ++i) {
    right_chunk_offsets[i] = offset_right;
    offset_right += right_in_table.segs[i].n_rows;
  }

  std::vector<std::shared_ptr<std::vector<size_t>>> histograms_left, histograms_right;
  std::vector<uint8_t> null_flags_mat, null_flags_part;
  auto materialized_left = materialize_input<LeftType, HashedType>(left_in_table, histograms_left, radix_bits,
                                                                   partitioning_seed, false, n_threads, nullptr);
  auto materialized_right =
      materialize_input<RightType, HashedType>(right_in_table, histograms_right, radix_bits, partitioning_seed, keep_nulls,
                                               n_threads, sql_null_probe ? &null_flags_mat : nullptr);

  auto radix_left = partition_radix_parallel<LeftType>(materialized_left, left_chunk_offsets, histograms_left, radix_bits,
                                                       false, n_threads, nullptr, nullptr);
  auto radix_right = partition_radix_parallel<RightType>(materialized_right, right_chunk_offsets, histograms_right,
                                                         radix_bits, keep_nulls, n_threads,
                                                         sql_null_probe ? &null_flags_mat : nullptr,
                                                         sql_null_probe ? &null_flags_part : nullptr);
  result.n_build_tuples = radix_left.partition_offsets.back();
  result.n_probe_tuples = radix_right.partition_offsets.back();

  auto hashtables = build<LeftType, HashedType>(radix_left, n_threads);

  const size_t partition_count = radix_right.partition_offsets.size() - 1;
  std::vector<PosList> left_pos_lists(partition_count), right_pos_lists(partition_count);
  if (mode == HJ_MODE_SEMI || mode == HJ_MODE_ANTI) {
    probe_semi_anti<RightType, HashedType>(radix_right, hashtables, right_pos_lists, mode, n_threads);
  } else {
    probe<RightType, HashedType>(radix_right, hashtables, left_pos_lists, right_pos_lists, mode, n_threads,
                                 sql_null_probe ? &null_flags_part : nullptr);
  }

  // join_hash.cpp:829-855 appends one output chunk per non-empty partition; the canonical comparison form
  //  is the concatenation.
  size_t total = 0;
  for (const auto& p : right_pos_lists) total += p.size();
  result.right.reserve(total);
  if (!(mode == HJ_MODE_SEMI || mode == HJ_MODE_ANTI)) result.left.reserve(total);
  for (size_t p = 0; p < partition_count; ++p) {
    result.right.insert(result.right.end(), right_pos_lists[p].begin(), right_pos_lists[p].end());
    result.left.insert(result.left.end(), left_pos_lists[p].begin(), left_pos_lists[p].end());
  }
  return result;
}

int relation_key_type(const hj_relation& r, int fallback) {
  for (uint32_t i = 0; i < r.n_segs; ++i) return r.segs[i].key_type;
  return fallback;
}

thread_local std::string g_error;

}  // namespace

extern "C" {

struct oracle_result {
  hj_row_id* build_side;  // malloc'ed, n_rows entries; NULL for Semi/Anti
  hj_row_id* probe_side;  // malloc'ed
  uint64_t n_rows;
  uint32_t radix_bits;
  uint32_t reserved;
  uint64_t n_build_tuples;
  uint64_t n_probe_tuples;
};

// MurmurHash2 as used by the reference for the radix bits (seed 17), exported for the known-answer test.
uint32_t oracle_murmur_hash2(const void* key, uint32_t len, uint32_t seed) { return murmur_hash2(key, len, seed); }

// join_hash.cpp:654-668 for int32 (key_type 0) / int64 (1) build keys.
int oracle_reference_radix_bits(uint64_t n_build_rows, int key_type) {
  return static_cast<int>(key_type == HJ_KEY_INT64 ? reference_radix_bits<int64_t>(n_build_rows)
                                                   : reference_radix_bits<int32_t>(n_build_rows));
}

const char* oracle_last_error() { return g_error.c_str(); }

// build/probe: relations after the operator's side selection (join_hash.cpp:58-76). radix_bits_override < 0 uses the
// reference heuristic. n_threads <= 1 is the reference without a scheduler.
int oracle_join(const hj_relation* build, const hj_relation* probe, int mode, int radix_bits_override, uint32_t flags,
                int n_threads, oracle_result* out) {
  if (!build || !probe || !out) {
    g_error = "null argument";
    return HJ_ERR_INVALID;
  }
  if (!(mode == HJ_MODE_INNER || mode == HJ_MODE_LEFT || mode == HJ_MODE_RIGHT || mode == HJ_MODE_SEMI ||
        mode == HJ_MODE_ANTI)) {
    g_error = "JoinHash supports Inner/Left/Right/Semi/Anti only";
    return HJ_ERR_INVALID;
  }
  const int bt = relation_key_type(*build, relation_key_type(*probe, HJ_KEY_INT32));
  const int pt = relation_key_type(*probe, bt);
  for (const auto* r : {build, probe}) {
    const int t = (r == build) ? bt : pt;
    for (uint32_t i = 0; i < r->n_segs; ++i) {
      if (r->segs[i].key_type != t) {
        g_error = "all chunks of a column share one data type";
        return HJ_ERR_INVALID;
      }
      // the oracle restates JoinHash over the values a column iterable yields: a ReferenceColumn / encoded column arrives
      // GATHERED (the caller dereferences it). Anything else would be read as a dictionary chunk with garbage pointers.
      const uint8_t kind = r->segs[i].kind;
      if (kind != HJ_SEG_VALUE && kind != HJ_SEG_DICT && kind != HJ_SEG_GATHERED) {
        g_error = "the oracle takes VALUE, DICT and GATHERED chunks only (gather REFERENCE / FOR / RLE chunks first)";
        return HJ_ERR_INVALID;
      }
    }
  }
  OracleResult r;
  if (bt == HJ_KEY_INT32 && pt == HJ_KEY_INT32) {
    r = join_impl<int32_t, int32_t>(*build, *probe, mode, radix_bits_override, flags, n_threads);
  } else if (bt == HJ_KEY_INT32 && pt == HJ_KEY_INT64) {
    r = join_impl<int32_t, int64_t>(*build, *probe, mode, radix_bits_override, flags, n_threads);
  } else if (bt == HJ_KEY_INT64 && pt == HJ_KEY_INT32) {
    r = join_impl<int64_t, int32_t>(*build, *probe, mode, radix_bits_override, flags, n_threads);
  } else {
    r = join_impl<int64_t, int64_t>(*build, *probe, mode, radix_bits_override, flags, n_threads);
  }
  const bool one_side = (mode == HJ_MODE_SEMI || mode == HJ_MODE_ANTI);
  out->n_rows = r.right.size();
  out->radix_bits = static_cast<uint32_t>(r.radix_bits);
  out->n_build_tuples = r.n_build_tuples;
  out->n_probe_tuples = r.n_probe_tuples;
  out->probe_side = static_cast<hj_row_id*>(std::malloc(std::max<size_t>(1, r.right.size()) * sizeof(hj_row_id)));
  std::memcpy(out->probe_side, r.right.data(), r.right.size() * sizeof(hj_row_id));
  out->build_side = nullptr;
  if (!one_side) {
    out->build_side = static_cast<hj_row_id*>(std::malloc(std::max<size_t>(1, r.left.size()) * sizeof(hj_row_id)));
    std::memcpy(out->build_side, r.left.data(), r.left.size() * sizeof(hj_row_id));
  }
  return HJ_OK;
}

void oracle_result_free(oracle_result* r) {
  if (!r) return;
  std::free(r->build_side);
  std::free(r->probe_side);
  r->build_side = r->probe_side = nullptr;
  r->n_rows = 0;
}

}  // extern "C"
