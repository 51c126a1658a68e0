// Stub standing in for the reference's platform/enforce.h (glog/gflags/boost closure,
// not buildable here).  Only what refer.h touches: the comparison enforce macros and
// platform::errors::InvalidArgument.  A failed check throws std::runtime_error, which
// the extern "C" wrappers in ref_shim.cc turn into a non-zero return code.
#pragma once
#include <cstdio>
#include <cstring>  // the real enforce.h closure provides std::memcpy to refer.h
#include <stdexcept>
#include <string>

namespace paddle {
namespace platform {
namespace errors {
template <typename... Args>
inline std::string InvalidArgument(const char* fmt, Args...) {
  return std::string("InvalidArgument: ") + fmt;
}
template <typename... Args>
inline std::string Unimplemented(const char* fmt, Args...) {
  return std::string("Unimplemented: ") + fmt;
}
}  // namespace errors
}  // namespace platform
}  // namespace paddle

#define PADDLE_THROW(err) throw std::runtime_error(err)
#define ORC_STUB_ENFORCE_CMP(a, b, op, err)                   \
  do {                                                        \
    if (!((a)op(b))) throw std::runtime_error(err);           \
  } while (0)
#define PADDLE_ENFORCE_EQ(a, b, err) ORC_STUB_ENFORCE_CMP(a, b, ==, err)
#define PADDLE_ENFORCE_NE(a, b, err) ORC_STUB_ENFORCE_CMP(a, b, !=, err)
#define PADDLE_ENFORCE_LT(a, b, err) ORC_STUB_ENFORCE_CMP(a, b, <, err)
#define PADDLE_ENFORCE_LE(a, b, err) ORC_STUB_ENFORCE_CMP(a, b, <=, err)
#define PADDLE_ENFORCE_GT(a, b, err) ORC_STUB_ENFORCE_CMP(a, b, >, err)
#define PADDLE_ENFORCE_GE(a, b, err) ORC_STUB_ENFORCE_CMP(a, b, >=, err)
