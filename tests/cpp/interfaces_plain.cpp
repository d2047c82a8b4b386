// Acceptance test of the remaining device interfaces of the path, each verified against the CK-style host references:
//   DeviceGemm (int8 x int8 -> int32, exact)                                      vs ReferenceGemm
//   DeviceGemmSplitK (bf16 x int8 row-major B = mk_kn_mn, KBatch hint; workspace API, no allocation in MakeArgument)
//                                                                                 vs ReferenceGemmMultipleD with no D
//   DeviceGemmStreamK (bf16 x int8, decode-shaped M)                              vs ReferenceGemm
//   DeviceGemmMultiD_Xdl_CShuffle_V3 (reference argument order) int8 + MultiplyMultiply, exact   vs ReferenceGemmMultipleD
//   DeviceGemmMultipleD_ABScale ([EXT] W4A16 under the block-scale interface)     vs ReferenceGemm on the float operand
// plus ck::get_device_name() / is_sm100_supported() (host_utility/device_prop.hpp).
#include <cstring>
#include <iostream>

#include "ck/host_utility/device_prop.hpp"
#include "ck/library/reference_tensor_operation/cpu/reference_gemm.hpp"
#include "ck/library/reference_tensor_operation/cpu/reference_gemm_multiple_d.hpp"
#include "ck/library/utility/check_err.hpp"
#include "ck/library/utility/device_memory.hpp"
#include "ck/library/utility/host_tensor.hpp"
#include "ck/tensor_operation/gpu/device/impl/device_gemm_multiple_d_xdl_cshuffle_v3.hpp"
#include "ck/tensor_operation/gpu/device/impl/device_gemm_sm100_plain.hpp"

using BF16 = ck::bhalf_t; using F32 = float; using I8 = int8_t; using I32 = int32_t;
using Row = ck::tensor_layout::gemm::RowMajor; using Col = ck::tensor_layout::gemm::ColumnMajor;
using namespace ck::tensor_operation::element_wise;
namespace dev = ck::tensor_operation::device;
namespace host = ck::tensor_operation::host;

static int fails = 0;
static void report(const char* what, bool ok, const std::string& type)
{
    std::cout << (ok ? "PASSED  " : "FAILED  ") << what << "  [" << type << "]" << std::endl;
    if(!ok) ++fails;
}
template <typename T>
static void fill_int(Tensor<T>& t, int lo, int hi) { t.GenerateTensorValue([&](auto, auto) { return T(lo + std::rand() % (hi - lo + 1)); }); }
static void fill_bf16(Tensor<BF16>& t) { t.GenerateTensorValue([](auto, auto) { return ck::type_convert<BF16>(2.f * float(std::rand()) / float(RAND_MAX) - 1.f); }); }

int main()
{
    std::cout << "device: " << ck::get_device_name() << "  sm_100: " << ck::is_sm100_supported() << "  SMs: " << ck::get_sm_count() << std::endl;
    if(!ck::is_sm100_supported()) { std::cerr << "needs an sm_100 device" << std::endl; return 2; }
    {   // ---- DeviceGemm: int8 x int8 -> int32
        const ck::index_t M = 192, N = 256, K = 320;
        Tensor<I8> a({std::size_t(M), std::size_t(K)}), b_kn({std::size_t(K), std::size_t(N)}), bn({std::size_t(N), std::size_t(K)});
        Tensor<I32> c_host({std::size_t(M), std::size_t(N)}), c_dev(c_host.mDesc);
        fill_int(a, -128, 127); fill_int(bn, -128, 127);
        for(int k = 0; k < K; ++k) for(int n = 0; n < N; ++n) b_kn(k, n) = bn(n, k);
        DeviceMem a_buf(a.GetElementSpaceSize()), b_buf(bn.GetElementSpaceSize()), c_buf(4 * c_dev.GetElementSpaceSize());
        a_buf.ToDevice(a.mData.data()); b_buf.ToDevice(bn.mData.data());
        dev::DeviceGemm_Sm100<Row, Col, Row, I8, I8, I32, PassThrough, PassThrough, PassThrough> op;
        auto arg = op.MakeArgumentPointer(a_buf.GetDeviceBuffer(), b_buf.GetDeviceBuffer(), c_buf.GetDeviceBuffer(), M, N, K, K, K, N, {}, {}, {});
        bool ok = op.IsSupportedArgument(arg.get());
        if(ok)
        {
            op.MakeInvokerPointer()->Run(arg.get(), StreamConfig{nullptr, false});
            c_buf.FromDevice(c_dev.mData.data());
            using Ref = host::ReferenceGemm<I8, I8, I32, I32, PassThrough, PassThrough, PassThrough, I32, I32>;
            auto ref = Ref{}; auto inv = ref.MakeInvoker(); auto ra = ref.MakeArgument(a, b_kn, c_host, {}, {}, {});
            inv.Run(ra);
            ok = std::memcmp(c_dev.mData.data(), c_host.mData.data(), 4 * c_dev.mData.size()) == 0;
        }
        report("DeviceGemm int8 x int8 -> int32 (bit exact)", ok, op.GetTypeString());
    }
    for(int variant = 0; variant < 2; ++variant)
    {   // ---- DeviceGemmSplitK (row-major B, prefill M) / DeviceGemmStreamK (K-major B, decode M): bf16 x int8 -> bf16
        const ck::index_t M = variant == 0 ? 200 : 16, N = 512, K = 1024;
        Tensor<BF16> a({std::size_t(M), std::size_t(K)}), c_host({std::size_t(M), std::size_t(N)}), c_dev(c_host.mDesc);
        Tensor<I8> b_kn({std::size_t(K), std::size_t(N)}), bn({std::size_t(N), std::size_t(K)});
        fill_bf16(a); fill_int(b_kn, -8, 8);
        for(int k = 0; k < K; ++k) for(int n = 0; n < N; ++n) bn(n, k) = b_kn(k, n);
        DeviceMem a_buf(2 * a.GetElementSpaceSize()), b_buf(b_kn.GetElementSpaceSize()), c_buf(2 * c_dev.GetElementSpaceSize());
        a_buf.ToDevice(a.mData.data());
        b_buf.ToDevice(variant == 0 ? b_kn.mData.data() : bn.mData.data());
        std::unique_ptr<dev::BaseOperator> op;
        std::unique_ptr<dev::BaseArgument> arg;
        std::unique_ptr<dev::BaseInvoker> inv;
        if(variant == 0)
        {
            auto o = std::make_unique<dev::DeviceGemmSplitK_Sm100<Row, Row, Row, BF16, I8, BF16, PassThrough, PassThrough, PassThrough>>();
            arg = o->MakeArgumentPointer(a_buf.GetDeviceBuffer(), b_buf.GetDeviceBuffer(), c_buf.GetDeviceBuffer(), M, N, K, K, N, N, {}, {}, {}, 4);
            inv = o->MakeInvokerPointer(); op = std::move(o);
        }
        else
        {
            auto o = std::make_unique<dev::DeviceGemmStreamK_Sm100<Row, Col, Row, BF16, I8, BF16, PassThrough, PassThrough, PassThrough>>();
            arg = o->MakeArgumentPointer(a_buf.GetDeviceBuffer(), b_buf.GetDeviceBuffer(), c_buf.GetDeviceBuffer(), M, N, K, K, K, N, {}, {}, {}, 0);
            inv = o->MakeInvokerPointer(); op = std::move(o);
        }
        bool ok = op->IsSupportedArgument(arg.get());
        const size_t ws_bytes = op->GetWorkSpaceSize(arg.get());
        if(variant == 0) ok = ok && ws_bytes >= size_t(N) * K;   // the K-major copy of row-major B is reported, not hidden
        DeviceMem ws(ws_bytes);
        if(ok)
        {
            op->SetWorkSpacePointer(arg.get(), ws.GetDeviceBuffer());
            inv->Run(arg.get(), StreamConfig{nullptr, false});
            c_buf.FromDevice(c_dev.mData.data());
            if(variant == 0)
            {
                using Ref = host::ReferenceGemmMultipleD<BF16, I8, ck::Tuple<>, BF16, F32, PassThrough, PassThrough, PassThrough, BF16, BF16>;
                typename Ref::DsArray none{};
                auto ref = Ref{}; auto ri = ref.MakeInvoker(); auto ra = ref.MakeArgument(a, b_kn, none, c_host, {}, {}, {});
                ri.Run(ra);
            }
            else
            {
                using Ref = host::ReferenceGemm<BF16, I8, BF16, F32, PassThrough, PassThrough, PassThrough, BF16, BF16>;
                auto ref = Ref{}; auto ri = ref.MakeInvoker(); auto ra = ref.MakeArgument(a, b_kn, c_host, {}, {}, {});
                ri.Run(ra);
            }
            ok = ck::utils::check_err(c_dev, c_host, "Error: Incorrect results!", 1e-2, 1e-2);
        }
        report(variant == 0 ? "DeviceGemmSplitK bf16 x int8 mk_kn_mn, KBatch 4, workspace API" : "DeviceGemmStreamK bf16 x int8, M = 16", ok, op->GetTypeString());
    }
    {   // ---- DeviceGemmMultiD_Xdl_CShuffle_V3 (reference argument order): int8 x int8 with row / column fp32 scales -> bf16
        const ck::index_t M = 256, N = 384, K = 512;
        Tensor<I8> a({std::size_t(M), std::size_t(K)}), b_kn({std::size_t(K), std::size_t(N)}), bn({std::size_t(N), std::size_t(K)});
        Tensor<BF16> e_host({std::size_t(M), std::size_t(N)}), e_dev(e_host.mDesc);
        fill_int(a, -128, 127); fill_int(bn, -128, 127);
        for(int k = 0; k < K; ++k) for(int n = 0; n < N; ++n) b_kn(k, n) = bn(n, k);
        std::vector<float> d0(N), d1(M);
        for(auto& v : d0) v = 5e-4f + 1e-3f * float(std::rand()) / float(RAND_MAX);
        for(auto& v : d1) v = 5e-4f + 1e-3f * float(std::rand()) / float(RAND_MAX);
        DeviceMem a_buf(a.GetElementSpaceSize()), b_buf(bn.GetElementSpaceSize()), e_buf(2 * e_dev.GetElementSpaceSize()), d0_buf(4 * N), d1_buf(4 * M);
        a_buf.ToDevice(a.mData.data()); b_buf.ToDevice(bn.mData.data()); d0_buf.ToDevice(d0.data()); d1_buf.ToDevice(d1.data());
        using Ds = ck::Tuple<F32, F32>;
        dev::DeviceGemmMultiD_Xdl_CShuffle_V3<Row, Col, ck::Tuple<Row, Col>, Row, I8, I8, Ds, BF16, I32, I32, PassThrough, PassThrough, MultiplyMultiply> op;
        // D0 = per-column scale (Row layout, stride 0), D1 = per-row scale (Col layout, stride 0): example 65's broadcast strides
        auto arg = op.MakeArgumentPointer(a_buf.GetDeviceBuffer(), b_buf.GetDeviceBuffer(), {d0_buf.GetDeviceBuffer(), d1_buf.GetDeviceBuffer()},
                                          e_buf.GetDeviceBuffer(), M, N, K, K, K, {0, 0}, N, {}, {}, {});
        bool ok = op.IsSupportedArgument(arg.get());
        if(ok)
        {
            op.MakeInvokerPointer()->Run(arg.get(), StreamConfig{nullptr, false});
            e_buf.FromDevice(e_dev.mData.data());
            using Ref = host::ReferenceGemmMultipleD<I8, I8, Ds, BF16, I32, PassThrough, PassThrough, MultiplyMultiply, I32, I32>;
            typename Ref::DsArray ds{Tensor<F32>({std::size_t(M), std::size_t(N)}), Tensor<F32>({std::size_t(M), std::size_t(N)})};
            for(int m = 0; m < M; ++m) for(int n = 0; n < N; ++n) { ds[0](m, n) = d0[n]; ds[1](m, n) = d1[m]; }
            auto ref = Ref{}; auto ri = ref.MakeInvoker(); auto ra = ref.MakeArgument(a, b_kn, ds, e_host, {}, {}, {});
            ri.Run(ra);
            ok = std::memcmp(e_dev.mData.data(), e_host.mData.data(), 2 * e_dev.mData.size()) == 0;
        }
        report("DeviceGemmMultiD_Xdl_CShuffle_V3 int8 MultiplyMultiply (bit exact vs ReferenceGemmMultipleD)", ok, op.GetTypeString());
    }
    {   // ---- DeviceGemmMultipleD_ABScale: [EXT] W4A16, ScaleBlockK = 128
        const ck::index_t M = 8, N = 512, K = 1024, G = 128;
        Tensor<BF16> a({std::size_t(M), std::size_t(K)}), s({std::size_t(K / G), std::size_t(N)}), e_host({std::size_t(M), std::size_t(N)}), e_dev(e_host.mDesc);
        Tensor<uint8_t> q({std::size_t(K), std::size_t(N)});
        fill_bf16(a);
        q.GenerateTensorValue([](auto, auto) { return uint8_t(std::rand() & 15); });
        s.GenerateTensorValue([](auto, auto) { return ck::type_convert<BF16>(0.005f + 0.015f * float(std::rand()) / float(RAND_MAX)); });
        DeviceMem a_buf(2 * a.GetElementSpaceSize()), q_buf(q.GetElementSpaceSize()), s_buf(2 * s.GetElementSpaceSize()), e_buf(2 * e_dev.GetElementSpaceSize());
        DeviceMem packed(ckb_prepack_b_int4_bytes(N, K));
        a_buf.ToDevice(a.mData.data()); q_buf.ToDevice(q.mData.data()); s_buf.ToDevice(s.mData.data());
        bool ok = ckb_prepack_b_int4(q_buf.GetDeviceBuffer(), N, 1, N, K, packed.GetDeviceBuffer(), nullptr) == CKB_OK;
        dev::DeviceGemmMultipleD_ABScale_W4_Sm100<Row, ck::Tuple<>, Row, BF16, ck::Tuple<>, BF16, PassThrough, DequantizeGroup, PassThrough> op;
        auto arg = op.MakeArgumentPointer(a_buf.GetDeviceBuffer(), packed.GetDeviceBuffer(), {}, e_buf.GetDeviceBuffer(), M, N, K, K, 0, {}, N, nullptr,
                                          s_buf.GetDeviceBuffer(), {}, {}, {});
        ok = ok && op.IsSupportedArgument(arg.get());
        if(ok)
        {
            op.MakeInvokerPointer()->Run(arg.get(), StreamConfig{nullptr, false});
            e_buf.FromDevice(e_dev.mData.data());
            Tensor<F32> b({std::size_t(K), std::size_t(N)});
            for(int k = 0; k < K; ++k) for(int n = 0; n < N; ++n) DequantizeGroup{}(b(k, n), int(q(k, n)) - 8, s(k / G, n));
            using Ref = host::ReferenceGemm<BF16, F32, BF16, F32, PassThrough, PassThrough, PassThrough, F32, F32>;
            auto ref = Ref{}; auto ri = ref.MakeInvoker(); auto ra = ref.MakeArgument(a, b, e_host, {}, {}, {});
            ri.Run(ra);
            ok = ck::utils::check_err(e_dev, e_host, "Error: Incorrect results!", 1e-2, 1e-2);
        }
        report("DeviceGemmMultipleD_ABScale W4A16 (ScaleBlockK 128)", ok, op.GetTypeString());
    }
    return fails ? 1 : 0;
}
