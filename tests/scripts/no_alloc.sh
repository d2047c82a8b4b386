# the allocation-free entry point: alone in a fresh process, then at the end of the whole GPU suite
timeout 600 python -m pytest tests/test_zz_no_alloc.py -x -q -m gpu --timeout 300 2>&1 | tail -15
timeout 1500 python -m pytest tests -x -q -m gpu --timeout 600 2>&1 | tail -4
