mkdir -p gpurun_out
GMG_TRACE=1 timeout 600 python -m pytest tests/test_gpu_builders.py -x -q -m gpu 2>&1 | grep -v "^\[gmg\] \(galerkin\|transpose\|sampler\)" | tail -40
timeout 300 python tools/prof_build.py C3 0.25 exact 1 1 2>&1 | tail -1
timeout 900 ncu --set full --clock-control none --import-source on -k regex:'product1_kernel|product2_kernel|compact_columns|product._bound' -c 5 -f -o gpurun_out/r2_galerkin_q python tools/prof_build.py C3 0.25 exact 1 1 > gpurun_out/r2_galerkin_q.log 2>&1
tail -2 gpurun_out/r2_galerkin_q.log
ls -la gpurun_out
