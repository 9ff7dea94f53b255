// Stand-in for the reference's utils/assert.hpp (which needs Boost.Preprocessor, absent in this image), used ONLY when
// oracle/Makefile compiles reference sources in place into oracle/_ref/. Test infrastructure.
#pragma once
#include <stdexcept>
#include <string>
namespace opossum {
[[noreturn]] inline void Fail(const std::string& msg) { throw std::logic_error(msg); }
}  // namespace opossum
#define Assert(expr, msg) \
  if (!static_cast<bool>(expr)) opossum::Fail(msg)
#define DebugAssert(expr, msg)
