timeout 100 python tools/prof_build.py C3 1.0 exact 1 3 2>&1 | tail -2
timeout 200 python -m pytest tests/test_gpu_parity_fullsize.py -x -q -m gpu -k "c3_full or c2_full" 2>&1 | tail -2
