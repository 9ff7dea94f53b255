// C exports around the reference's SimdBp128Packing, compiled IN PLACE from /root/reference (never copied). Pins the
// SimdBp128 layout that tests/test_gpu_encodings.py packs with numpy and libhj decodes on the device.
// Built only by oracle/Makefile into oracle/_ref/ (git-ignored). Test infrastructure.
#include <cstdint>

#include "storage/vector_compression/simd_bp128/simd_bp128_packing.hpp"  // -I/root/reference/src/lib

extern "C" void ref_bp128_pack_block(const uint32_t* in128, uint32_t* out_words, uint8_t bit_size) {
  // out_words: bit_size 16-byte words
  opossum::SimdBp128Packing::pack_block(in128, reinterpret_cast<opossum::uint128_t*>(out_words), bit_size);
}

extern "C" void ref_bp128_unpack_block(const uint32_t* in_words, uint32_t* out128, uint8_t bit_size) {
  opossum::SimdBp128Packing::unpack_block(reinterpret_cast<const opossum::uint128_t*>(in_words), out128, bit_size);
}
