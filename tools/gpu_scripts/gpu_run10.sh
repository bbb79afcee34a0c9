mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_parity.py tests/test_gpu_solver.py tests/test_gpu_sharded.py -x -q -m gpu 2>&1 | tail -8
GMG_TRACE=1 timeout 600 python tools/prof_build.py C3 1.0 exact 1 2 2>&1 | grep -E "galerkin:|^C3" | tail -8
timeout 600 python tools/prof_build.py C3 1.0 parallel 1 2 2>&1 | tail -1
timeout 600 python tools/prof_build.py C2 1.0 exact 1 2 2>&1 | tail -1
timeout 1500 python -m pytest tests/test_gpu_parity_fullsize.py -x -q -m gpu -s 2>&1 | grep -E "passed|failed|Error|error|count|assert" | tail -20
