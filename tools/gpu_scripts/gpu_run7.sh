mkdir -p gpurun_out
GMG_SAMPLER_DEBUG=1 timeout 300 python tools/prof_build.py C3 0.25 exact 0 2 2>&1 | grep -v "tile [0-9]* block" | tail -12
GMG_SAMPLER_DEBUG=1 timeout 300 python tools/prof_build.py C3 0.25 exact 0 1 2>&1 | grep "tile [0-9]* block" | head -30
timeout 600 python tools/prof_build.py C3 1.0 exact 0 2 2>&1 | tail -1
