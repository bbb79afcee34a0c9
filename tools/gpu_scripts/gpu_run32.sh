timeout 600 python -m pytest tests/test_gpu_parity.py tests/test_gpu_options.py -x -q -m gpu 2>&1 | tail -2
for v in 16_8_8 26_32_8 24_32_8 22_32_8 27_32_8 28_32_8; do
  cp gpurun_variants/lib_$v.so gravo-mg_b200/lib/libgravomg_b200.so
  echo "variant $v"
  timeout 300 python tools/prof_build.py C3 1.0 exact 0 2 2>&1 | tail -1
  timeout 300 python tools/prof_build.py C3 1.0 parallel 0 2 2>&1 | tail -1
done
cp gpurun_variants/lib_26_32_8.so gravo-mg_b200/lib/libgravomg_b200.so
GMG_SAMPLER_TRACE=gpurun_out/trace_c3q.bin timeout 300 python tools/prof_build.py C3 0.25 exact 0 1 2>&1 | tail -1
