python -m pytest tests/test_gpu_cpp_examples.py tests/test_gpu_parity.py -m gpu -q -x -k "grouped or cpp_example" 2>&1 | tail -8
