// Stand-in for the build-GENERATED paddle/phi/api/include/api.h, which is absent from the
// reference source tree.  Declares only what ctrfe_paddle_ops.cc calls, for a syntax check.
#pragma once
#include <cstdint>
#include <vector>
#include "paddle/phi/api/include/tensor.h"
#include "paddle/phi/common/data_type.h"
#include "paddle/phi/common/int_array.h"
#include "paddle/phi/common/place.h"
namespace paddle {
namespace experimental {
PADDLE_API Tensor empty(const IntArray& shape, DataType dtype = DataType::FLOAT32, const Place& place = CPUPlace());
}  // namespace experimental
}  // namespace paddle
