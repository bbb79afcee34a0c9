timeout 900 python -m pytest tests/test_gpu_sharded.py tests/test_gpu_parity.py tests/test_gpu_solver.py -x -q -m gpu 2>&1 | tail -2
timeout 300 python tools/prof_build.py C3 1.0 exact 1 2 2>&1 | tail -1
