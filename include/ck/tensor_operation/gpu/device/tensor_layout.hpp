// layout tag types (include/ck/tensor_operation/gpu/device/tensor_layout.hpp)
#pragma once
namespace ck {
namespace tensor_layout {
struct BaseTensorLayout {};
namespace gemm {
struct RowMajor : public BaseTensorLayout { static constexpr const char* name = "RowMajor"; };
struct ColumnMajor : public BaseTensorLayout { static constexpr const char* name = "ColumnMajor"; };
} // namespace gemm
} // namespace tensor_layout
} // namespace ck
