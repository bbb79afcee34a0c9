timeout 600 python -m pytest tests/test_gpu_parity.py tests/test_gpu_options.py -x -q -m gpu 2>&1 | tail -2
timeout 300 python tools/prof_build.py C3 1.0 exact 0 3 2>&1 | tail -2
timeout 300 python tools/prof_build.py C3 1.0 parallel 0 2 2>&1 | tail -1
