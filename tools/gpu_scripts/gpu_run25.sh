timeout 300 python -m pytest tests/test_zz_no_alloc.py -x -q -m gpu 2>&1 | tail -15
timeout 900 python -m pytest tests/test_gpu_parity.py tests/test_gpu_solver.py tests/test_zz_no_alloc.py -x -q -m gpu 2>&1 | tail -5
