// Shared pieces of the profile_*_impl functions: the instance loop of
// profiler/include/profiler/profile_gemm_multiply_multiply_impl.hpp:181-300 (for every factory instance: supported? ->
// run once and verify against the host reference -> time -> "Perf:" line -> keep the best) plus the B200 roofline
// columns (fraction of the measured HBM copy bandwidth and of the dense tensor peak).
#pragma once
#include <algorithm>
#include <cstdlib>
#include <functional>
#include <iomanip>
#include <iostream>
#include <string>

#include "ck/stream_config.hpp"

namespace ck {
namespace profiler {

struct Peaks
{
    double hbm_gbs    = 6456.0; // MEASURED_PEAKS.json hbm_gbs of this pool's B200s; override with CKP_HBM_GBS
    double tensor_tfs = 1655.0; // dense bf16 burst; int8 is 2x; override with CKP_TENSOR_TFLOPS
    Peaks()
    {
        if(const char* e = std::getenv("CKP_HBM_GBS")) hbm_gbs = std::atof(e);
        if(const char* e = std::getenv("CKP_TENSOR_TFLOPS")) tensor_tfs = std::atof(e);
    }
};

// Cold-cache timing options (the reference's "memory for rotating buffer" argument of profile_gemm_multiply_multiply.cpp:31 and
// profile_gemm_universal*.cpp): rotating_mb > 0 -> StreamConfig::flush_cache with enough rotating A/B copies to cover that many
// MB, L2 evicted before every timed launch, median of the launches (include/ck/host_utility/flush_cache.hpp).
// Set from an op's own argument or, for every op, from CKP_ROTATING_MB.
struct Options
{
    int rotating_mb = 0;
    Options()
    {
        if(const char* e = std::getenv("CKP_ROTATING_MB")) rotating_mb = std::atoi(e);
    }
};
inline Options& options()
{
    static Options o;
    return o;
}

struct Best
{
    std::string name;
    float ms = 0, tflops = 0, gbs = 0;
    int instances = 0, supported = 0;
    bool pass = true;
};

inline void print_perf(const char* tag, float ms, double flop, double bytes, const std::string& name, double tensor_scale, Best* best)
{
    static const Peaks pk;
    const float tflops = float(flop / 1.E9 / ms), gbs = float(bytes / 1.E6 / ms);
    std::cout << tag << std::setw(10) << ms << " ms, " << tflops << " TFlops, " << gbs << " GB/s, " << name << "  [roofline: "
              << std::fixed << std::setprecision(1) << 100.0 * gbs / pk.hbm_gbs << "% HBM, "
              << 100.0 * tflops / (pk.tensor_tfs * tensor_scale) << "% tensor]" << std::defaultfloat << std::setprecision(6) << std::endl;
    // one machine-readable line per result (metrics / logging row)
    std::cout << "PerfJSON: {\"tag\": \"" << (best ? "perf" : "best") << "\", \"instance\": \"" << name << "\", \"ms\": " << ms << ", \"tflops\": " << tflops
              << ", \"gbs\": " << gbs << ", \"hbm_frac\": " << gbs / pk.hbm_gbs << ", \"tensor_frac\": " << tflops / (pk.tensor_tfs * tensor_scale)
              << ", \"rotating_mb\": " << options().rotating_mb << ", \"timing\": \"" << (options().rotating_mb > 0 ? "cold: rotating buffers + L2 flush, median" : "warm: back-to-back launches, mean")
              << "\"}" << std::endl;
    if(best && tflops > best->tflops) { best->tflops = tflops; best->gbs = gbs; best->ms = ms; best->name = name; }
}

// instance loop.  make_arg(op) -> argument pointer; prepare(op, arg) may attach workspaces / device arg arrays;
// verify() runs after the first launch and returns pass/fail (called only with do_verification).
template <typename OpPtrs, typename MakeArg, typename Prepare, typename Verify>
bool run_instances(OpPtrs& op_ptrs, MakeArg make_arg, Prepare prepare, Verify verify, bool do_verification, bool time_kernel, int n_warmup,
                   int n_iter, double flop, double bytes, double tensor_scale)
{
    Best best;
    std::cout << "found " << op_ptrs.size() << " instances" << std::endl;
    for(auto& op_ptr : op_ptrs)
    {
        ++best.instances;
        auto argument_ptr = make_arg(*op_ptr);
        auto invoker_ptr  = op_ptr->MakeInvokerPointer();
        if(!op_ptr->IsSupportedArgument(argument_ptr.get()))
        {
            std::cout << op_ptr->GetTypeString() << " does not support this problem" << std::endl;
            continue;
        }
        ++best.supported;
        auto keep = prepare(*op_ptr, argument_ptr.get()); // keeps device buffers alive for the runs below
        (void)keep;
        invoker_ptr->Run(argument_ptr.get(), StreamConfig{nullptr, false});
        if(do_verification) best.pass = verify() && best.pass;
        if(time_kernel)
        {
            StreamConfig sc{nullptr, true, 0, n_warmup, n_iter};
            if(options().rotating_mb > 0)
            {
                sc.flush_cache    = true;
                sc.rotating_count = std::max(2, int((double(options().rotating_mb) * 1.0e6 + bytes - 1) / bytes));
            }
            const float ms = invoker_ptr->Run(argument_ptr.get(), sc);
            print_perf("Perf: ", ms, flop, bytes, op_ptr->GetTypeString(), tensor_scale, &best);
        }
    }
    if(best.supported == 0) { std::cout << "no instance supports this problem" << std::endl; return false; }
    if(time_kernel) print_perf("Best Perf: ", best.ms, double(best.tflops) * 1.E9 * best.ms, double(best.gbs) * 1.E6 * best.ms, best.name, tensor_scale, nullptr);
    if(do_verification) std::cout << (best.pass ? "verification PASSED" : "verification FAILED") << std::endl;
    return best.pass;
}

} // namespace profiler
} // namespace ck
