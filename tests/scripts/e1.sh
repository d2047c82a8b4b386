python -m pytest tests/test_gpu_guards.py tests/test_gpu_parity.py -m gpu -q -x -k "window or torch_custom" 2>&1 | tail -15
