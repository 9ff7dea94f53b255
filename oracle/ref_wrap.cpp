// Thin C exports around reference code compiled IN PLACE from /root/reference (never copied into this repo).
// Built only by oracle/Makefile into oracle/_ref/ (git-ignored). Test infrastructure.
#include <cstdint>
#include <string>
#include <type_traits>

#include "operators/join_hash/hash_traits.hpp"  // -I/root/reference/src/lib
#include "utils/murmur_hash.hpp"

// join_hash_test.cpp:30-118 pins these; the oracle's and libhj's "wider integer" rule must agree with them.
static_assert(std::is_same_v<opossum::JoinHashTraits<int32_t, int32_t>::HashType, int32_t>);
static_assert(std::is_same_v<opossum::JoinHashTraits<int32_t, int64_t>::HashType, int64_t>);
static_assert(std::is_same_v<opossum::JoinHashTraits<int64_t, int32_t>::HashType, int64_t>);
static_assert(std::is_same_v<opossum::JoinHashTraits<int64_t, int64_t>::HashType, int64_t>);

extern "C" uint32_t ref_murmur_hash2(const void* key, uint32_t len, uint32_t seed) {
  return opossum::murmur_hash2(key, len, seed);
}
