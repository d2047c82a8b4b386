python -m pytest tests/test_gpu_parity.py -m gpu -q -x -k "prepack_equals or gptq or converters" 2>&1 | tail -25
