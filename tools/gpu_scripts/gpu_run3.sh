set -x
mkdir -p gpurun_out
CMD="python tools/prof_build.py C3 0.25 exact 1"
$CMD > gpurun_out/r2c_plain.log 2>&1 && ncu --set full --clock-control none --import-source on -k regex:"galerkin_warp|sampler_build|sampler_flow|sampler_lane" -c 5 -o gpurun_out/r2c_prof $CMD > gpurun_out/r2c_ncu.log 2>&1
cat gpurun_out/r2c_plain.log; tail -5 gpurun_out/r2c_ncu.log
