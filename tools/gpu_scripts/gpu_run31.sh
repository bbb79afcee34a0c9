mkdir -p gpurun_out
cp gpurun_variants/lib_26_32_8.so gravo-mg_b200/lib/libgravomg_b200.so
GMG_SAMPLER_TRACE=gpurun_out/trace_c3q.bin timeout 300 python tools/prof_build.py C3 0.25 exact 0 1 2>&1 | tail -1
