timeout 90 python -m pytest tests/test_gpu_parity.py -x -q -m gpu 2>&1 | tail -2 || exit 1
timeout 60 python tools/prof_build.py C3 1.0 exact 0 2 2>&1 | tail -1
timeout 60 python tools/prof_build.py C3 1.0 parallel 0 2 2>&1 | tail -1
