python -m pytest tests/test_gpu_parity.py -m gpu -q -x -k "passthrough_weight" 2>&1 | tail -8
