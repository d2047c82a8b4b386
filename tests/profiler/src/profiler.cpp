// SPDX-License-Identifier: MIT
// Interface mirrored from Composable Kernel (Copyright (c) Advanced Micro Devices, Inc., MIT licence): declarations must match the
// reference for client code to compile unchanged; see the reference file cited below.
// ckProfiler <operation> <args...> -- dispatches to the registered operation (profiler/src/profiler.cpp:14-30).
#include <cstdlib>
#include <iostream>

#include "profiler_operation_registry.hpp"

static void print_help_message()
{
    std::cout << "arg1: tensor operation " << ProfilerOperationRegistry::GetInstance() << std::endl;
}

int main(int argc, char* argv[])
{
    if(argc == 1)
    {
        print_help_message();
        return EXIT_SUCCESS;
    }
    if(const auto operation = ProfilerOperationRegistry::GetInstance().Get(argv[1]); operation.has_value())
    {
        try { return (*operation)(argc, argv); }
        catch(const std::exception& e)
        {
            std::cerr << "ckProfiler: " << e.what() << std::endl;
            return EXIT_FAILURE;
        }
    }
    std::cerr << "cannot find operation: " << argv[1] << std::endl;
    return EXIT_FAILURE;
}
