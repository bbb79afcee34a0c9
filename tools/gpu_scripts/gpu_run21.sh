timeout 900 python -m pytest tests/test_gpu_parity.py tests/test_gpu_sharded.py tests/test_gpu_options.py tests/test_gpu_builders.py -x -q -m gpu 2>&1 | tail -3
timeout 300 python tools/prof_build.py C2 1.0 exact 0 3 2>&1 | tail -1
timeout 300 python tools/prof_build.py C3 1.0 hashed 0 2 2>&1 | tail -1
timeout 300 python tools/prof_build.py C4 0.1 exact 0 2 2>&1 | tail -1
timeout 300 python tools/prof_build.py C4 0.1 hashed 0 2 2>&1 | tail -1
