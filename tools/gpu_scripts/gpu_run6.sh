mkdir -p gpurun_out
GMG_SAMPLER_DEBUG=1 timeout 600 python tools/prof_build.py C3 0.25 exact 0 2 2>&1 | tail -80
