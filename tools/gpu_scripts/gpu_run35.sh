mkdir -p gpurun_out
timeout 120 python tools/prof_build.py C3 1.0 exact 0 1 2>&1 | tail -1 || exit 1
timeout 240 ncu --set full --clock-control none --import-source on -k regex:'sampler_flow_kernel' -c 1 -f -o gpurun_out/r2_zz_flow python tools/prof_build.py C3 1.0 exact 0 1 > gpurun_out/r2_zz_flow.log 2>&1
tail -2 gpurun_out/r2_zz_flow.log
