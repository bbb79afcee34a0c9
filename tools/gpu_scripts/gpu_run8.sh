mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_gpu_parity.py -x -q -m gpu 2>&1 | tail -3
GMG_SAMPLER_TRACE=gpurun_out/trace_c3q.bin timeout 300 python tools/prof_build.py C3 0.25 exact 0 1 2>&1 | tail -1
timeout 300 python tools/prof_build.py C3 0.25 exact 0 2 2>&1 | tail -1
timeout 600 python tools/prof_build.py C3 1.0 exact 0 2 2>&1 | tail -1
timeout 600 python tools/prof_build.py C2 1.0 exact 0 2 2>&1 | tail -1
