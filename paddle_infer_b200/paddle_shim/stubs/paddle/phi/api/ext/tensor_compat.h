// Syntax-check stand-in for paddle/phi/api/ext/tensor_compat.h: the real one re-exports the
// whole generated C++ API (absent here); this one re-exports the single function the shim uses.
#pragma once
#include "paddle/phi/api/include/api.h"
namespace paddle {
using experimental::empty;
}  // namespace paddle
