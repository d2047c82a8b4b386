// GemmSpecialization (include/ck/tensor_operation/gpu/device/gemm_specialization.hpp).  The sm_100a kernels
// pad through TMA out-of-bounds fill, so every value behaves like MNKPadding; the enum is kept for source
// compatibility of instance declarations.
#pragma once
namespace ck {
namespace tensor_operation {
namespace device {
enum struct GemmSpecialization
{
    Default, MPadding, NPadding, KPadding, MNPadding, MKPadding, NKPadding, MNKPadding,
    OPadding, MOPadding, NOPadding, KOPadding, MNOPadding, MKOPadding, NKOPadding, MNKOPadding
};
} // namespace device
} // namespace tensor_operation
} // namespace ck
