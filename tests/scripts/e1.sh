python -m pytest tests/test_gpu_cpp_examples.py -m gpu -q -x 2>&1 | tail -12
tests/cpp/bin/client_gemm_bias_bf16_i8 4096 4096 4096 1 0 add | grep "Perf\|found"
tests/profiler/bin/ckProfiler gemm_quantization 0 0 2 0 1 4096 4096 4096 0.03 | grep "^Perf\|^Best\|found"
