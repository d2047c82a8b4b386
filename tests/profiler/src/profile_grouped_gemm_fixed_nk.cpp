// ckProfiler grouped_gemm_fixed_nk -- argument list of profiler/src/profile_grouped_gemm_fixed_nk.cpp:48-107 (comma lists of
// Ms, Ns, Ks and strides; all Ns and all Ks must be equal), data type 0 = bf16 @ int8.  The multi-ABD instances of
// client_example/31 are profiled: B = Multiply(int8 B0, bf16 per-channel B1), epilogue bias + FastGelu.
#include <iostream>
#include <sstream>

#include "ck/library/reference_tensor_operation/cpu/reference_gemm.hpp"
#include "ck/library/tensor_operation_instance/gpu/grouped_gemm_multi_abd_fixed_nk.hpp"
#include "ck/library/utility/check_err.hpp"
#include "ck/library/utility/device_memory.hpp"
#include "ck/library/utility/host_tensor.hpp"
#include "profiler/profile_util.hpp"
#include "profiler_operation_registry.hpp"

#define OP_NAME "grouped_gemm_fixed_nk"
#define OP_DESC "Grouped GEMM Fixed NK (bf16 A x int8 B multi-ABD, bias + FastGelu)"

namespace {
using BF16 = ck::bhalf_t; using I8 = int8_t; using F32 = float;
using Row = ck::tensor_layout::gemm::RowMajor; using Col = ck::tensor_layout::gemm::ColumnMajor;
using namespace ck::tensor_operation::element_wise;
using namespace ck::literals;

std::vector<int> int_list(const char* s)
{
    std::vector<int> out;
    std::istringstream in(s);
    for(std::string item; std::getline(in, item, ',');) out.push_back(std::stoi(item));
    return out;
}

template <typename BLayout>
bool profile(bool do_verification, int init_method, bool do_log, bool time_kernel, const std::vector<int>& Ms, int N, int K,
             std::vector<int> StrideAs, std::vector<int> StrideBs, std::vector<int> StrideCs, int kbatch, int n_warmup, int n_iter)
{
    constexpr bool b_row = std::is_same_v<BLayout, Row>;
    const int G = int(Ms.size());
    StrideAs.resize(G, -1); StrideBs.resize(G, -1); StrideCs.resize(G, -1);
    int sum_of_m = 0;
    for(int v : Ms) sum_of_m += v;
    std::vector<Tensor<BF16>> a, b1, d0, e_host, e_dev;
    std::vector<Tensor<I8>> b0;
    std::vector<std::unique_ptr<DeviceMem>> bufs;
    std::vector<ck::tensor_operation::device::GemmMultiABDDesc> descs;
    std::vector<ck::tensor_operation::device::GroupedGemmMultiABDKernelArgument<1, 2, 1>> kargs;
    auto dev = [&](const void* src, std::size_t bytes) {
        bufs.emplace_back(std::make_unique<DeviceMem>(bytes));
        if(src) bufs.back()->ToDevice(src);
        return bufs.back()->GetDeviceBuffer();
    };
    std::vector<void*> e_ptrs;
    const std::size_t n = N, k = K;
    for(int g = 0; g < G; ++g)
    {
        const std::size_t m = Ms[g];
        const int sa = StrideAs[g] < 0 ? K : StrideAs[g], sb = StrideBs[g] < 0 ? (b_row ? N : K) : StrideBs[g], sc = StrideCs[g] < 0 ? N : StrideCs[g];
        a.emplace_back(HostTensorDescriptor({m, k}, {std::size_t(sa), 1_uz}));
        b0.emplace_back(b_row ? HostTensorDescriptor({k, n}, {std::size_t(sb), 1_uz}) : HostTensorDescriptor({k, n}, {1_uz, std::size_t(sb)}));
        b1.emplace_back(HostTensorDescriptor({k, n}, {0_uz, 1_uz}));
        d0.emplace_back(HostTensorDescriptor({m, n}, {0_uz, 1_uz}));
        e_host.emplace_back(HostTensorDescriptor({m, n}, {std::size_t(sc), 1_uz}));
        e_dev.emplace_back(e_host.back().mDesc);
        const bool ints = init_method == 1;
        if(init_method)
        {
            a[g].GenerateTensorValue([&](auto, auto) { return ck::type_convert<BF16>(ints ? float(std::rand() % 5 - 2) : float(std::rand()) / float(RAND_MAX) - 0.5f); });
            b0[g].GenerateTensorValue([&](auto, auto) { return I8(ints ? std::rand() % 5 - 2 : std::rand() % 255 - 127); });
            b1[g].GenerateTensorValue([&](auto, auto) { return ck::type_convert<BF16>(ints ? float(1 + std::rand() % 2) : 0.001f + 0.01f * float(std::rand()) / float(RAND_MAX)); });
            d0[g].GenerateTensorValue([&](auto, auto) { return ck::type_convert<BF16>(ints ? float(std::rand() % 5 - 2) : float(std::rand()) / float(RAND_MAX) - 0.5f); });
        }
        descs.push_back({sum_of_m, N, K, {sa}, {sb, 0}, {0}, sc});
        const void* pa  = dev(a[g].mData.data(), 2 * a[g].GetElementSpaceSize());
        const void* pb0 = dev(b0[g].mData.data(), b0[g].GetElementSpaceSize());
        const void* pb1 = dev(b1[g].mData.data(), 2 * b1[g].GetElementSpaceSize());
        const void* pd0 = dev(d0[g].mData.data(), 2 * d0[g].GetElementSpaceSize());
        e_ptrs.push_back(dev(nullptr, 2 * e_dev[g].GetElementSpaceSize()));
        kargs.push_back({{pa}, {pb0, pb1}, {pd0}, e_ptrs.back(), Ms[g], N, K, {sa}, {sb, 0}, {0}, sc});
        if(do_verification)
        {
            Tensor<BF16> b(HostTensorDescriptor({k, n}, {n, 1_uz}));
            for(int j = 0; j < N; ++j) for(int i = 0; i < K; ++i) Multiply{}(b(i, j), b0[g](i, j), b1[g](i, j));
            Tensor<F32> c({m, n});
            using Ref = ck::tensor_operation::host::ReferenceGemm<BF16, BF16, F32, F32, PassThrough, PassThrough, PassThrough>;
            auto ref = Ref{}; auto inv = ref.MakeInvoker(); auto arg = ref.MakeArgument(a[g], b, c, PassThrough{}, PassThrough{}, PassThrough{});
            inv.Run(arg);
            for(int i = 0; i < Ms[g]; ++i) for(int j = 0; j < N; ++j) AddFastGelu{}(e_host[g](i, j), c(i, j), d0[g](i, j));
        }
    }
    using DeviceOp = ck::tensor_operation::device::DeviceGroupedGemmMultiABDFixedNK<ck::Tuple<Row>, ck::Tuple<BLayout, BLayout>, ck::Tuple<Row>, Row,
                                                                                    ck::Tuple<BF16>, ck::Tuple<I8, BF16>, ck::Tuple<BF16>, BF16,
                                                                                    PassThrough, Multiply, AddFastGelu>;
    auto op_ptrs = ck::tensor_operation::device::instance::DeviceOperationInstanceFactory<DeviceOp>::GetInstances();
    std::vector<std::array<const void*, 1>> p_As; std::vector<std::array<const void*, 2>> p_Bs; std::vector<std::array<const void*, 1>> p_Ds;
    std::vector<void*> p_Cs;
    const double flop = 2.0 * sum_of_m * N * K, bytes = 2.0 * sum_of_m * K + 1.0 * K * N * G + 2.0 * sum_of_m * N;
    return ck::profiler::run_instances(
        op_ptrs, [&](auto& op) { return op.MakeArgumentPointer(p_As, p_Bs, p_Ds, p_Cs, descs); },
        [&](auto& op, auto* arg) {
            auto kd = std::make_shared<DeviceMem>(op.GetDeviceKernelArgSize(arg));
            kd->ToDevice(kargs.data());
            op.SetDeviceKernelArgs(arg, kd->GetDeviceBuffer());
            op.SetElementwiseOps(arg, PassThrough{}, Multiply{}, AddFastGelu{});
            op.SetKBatch(arg, kbatch);
            return kd;
        },
        [&]() {
            bool ok = true;
            for(int g = 0; g < G; ++g)
            {
                if(Ms[g] == 0) continue;
                cudaMemcpy(e_dev[g].mData.data(), e_ptrs[g], 2 * e_dev[g].GetElementSpaceSize(), cudaMemcpyDeviceToHost);
                if(do_log) LogRangeAsType<float>(std::cout << "e_device[" << g << "]: ", e_dev[g].mData, ",") << std::endl;
                ok = ck::utils::check_err(e_dev[g], e_host[g], "Error: Incorrect results!", 1e-2, 1e-2) && ok;
            }
            return ok;
        },
        do_verification, time_kernel, n_warmup, n_iter, flop, bytes, 1.0);
}
} // namespace

int profile_grouped_gemm_fixed_nk(int argc, char* argv[])
{
    if(argc < 14)
    {
        std::cout << "arg1: tensor operation (" OP_NAME ": " OP_DESC ")\n"
                  << "arg2: data type (0: bf16@int8; others: not built for sm_100a)\n"
                  << "arg3: matrix layout (0: A[m, k] * B[k, n] = C[m, n]; 1: A[m, k] * B[n, k] = C[m, n])\n"
                  << "arg4: verification (0: no; 1: yes)\n"
                  << "arg5: initialization (0: no init; 1: integer value; 2: decimal value)\n"
                  << "arg6: print tensor value (0: no; 1: yes)\n"
                  << "arg7: time kernel (0=no, 1=yes)\n"
                  << "arg8 to 13: Ms, Ns, Ks, StrideAs, StrideBs, StrideCs (e.g., 256,256 128,128 64,64 -1 -1 -1)\n"
                  << "arg14: kbatch value (default 1; accepted and ignored like the reference's SetKBatch)\n"
                  << "optional: arg15: warm-up cycles (default 1), arg16: iterations (default 10)" << std::endl;
        return 1;
    }
    const int data_type = std::stoi(argv[2]), layout = std::stoi(argv[3]);
    const bool do_verification = std::stoi(argv[4]);
    const int init_method      = std::stoi(argv[5]);
    const bool do_log = std::stoi(argv[6]), time_kernel = std::stoi(argv[7]);
    const auto Ms = int_list(argv[8]), Ns = int_list(argv[9]), Ks = int_list(argv[10]);
    const auto StrideAs = int_list(argv[11]), StrideBs = int_list(argv[12]), StrideCs = int_list(argv[13]);
    const int kbatch = argc >= 15 ? std::stoi(argv[14]) : 1;
    int n_warmup = 1, n_iter = 10;
    if(argc == 17) { n_warmup = std::stoi(argv[15]); n_iter = std::stoi(argv[16]); }
    for(int v : Ns) if(v != Ns[0]) { std::cout << "fixed NK: all Ns must be equal" << std::endl; return 1; }
    for(int v : Ks) if(v != Ks[0]) { std::cout << "fixed NK: all Ks must be equal" << std::endl; return 1; }
    if(data_type != 0 || layout < 0 || layout > 1 || Ms.empty() || Ns.empty() || Ks.empty())
    {
        std::cout << "this data_type & layout is not implemented" << std::endl;
        return 1;
    }
    const bool pass = layout == 0 ? profile<Row>(do_verification, init_method, do_log, time_kernel, Ms, Ns[0], Ks[0], StrideAs, StrideBs, StrideCs, kbatch, n_warmup, n_iter)
                                  : profile<Col>(do_verification, init_method, do_log, time_kernel, Ms, Ns[0], Ks[0], StrideAs, StrideBs, StrideCs, kbatch, n_warmup, n_iter);
    return pass ? 0 : 1;
}

REGISTER_PROFILER_OPERATION(OP_NAME, OP_DESC, profile_grouped_gemm_fixed_nk);
