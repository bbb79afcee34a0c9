#!/bin/bash
# ncu captures for profiles/ (run on the GPU box through gpurun; every ncu run follows a plain run of the same
# command that exited 0). Keep the reports small: gpurun copies back at most 64 MiB.
set -x
cd $GRAFT_REPO_ROOT
CMD="python bench.py --steps 1 --warmup 1 --no-cpu-baseline"
timeout 200 $CMD > gpurun_out/z_cap_plain.json 2> gpurun_out/z_cap_plain.err || exit 1
timeout 400 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/z_launches.csv $CMD > gpurun_out/z_ncu_launch.log 2>&1
timeout 600 ncu --set full --clock-control none --import-source on -k regex:'sampler_build|sampler_flow|parents_sweep|parents_pred|coarse_rows_warp|prolong_rows|edge_length|triangle_kernel' -c 12 -f -o gpurun_out/z_full $CMD > gpurun_out/z_ncu_full.log 2>&1
CMD3="python bench.py --workload C3 --sampler parallel --galerkin 1 --vcycle 1 --steps 1 --warmup 1 --no-cpu-baseline"
timeout 300 $CMD3 > gpurun_out/z_cap_c3_plain.json 2> gpurun_out/z_cap_c3_plain.err || exit 2
timeout 600 ncu --set full --clock-control none --import-source on -k regex:'galerkin_warp|sort_rows|transpose_fill|count_columns' -c 5 -f -o gpurun_out/z_full_c3 $CMD3 > gpurun_out/z_ncu_full_c3.log 2>&1
timeout 600 ncu --set full --clock-control none --import-source on -k regex:'spmv_kernel|gs_colour' -c 3 -f -o gpurun_out/z_full_c3_solver $CMD3 > gpurun_out/z_ncu_full_c3s.log 2>&1
ls -la gpurun_out/
du -sh gpurun_out
