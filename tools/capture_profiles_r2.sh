#!/bin/bash
# Round-2 evidence for profiles/ (run on the GPU box through gpurun). Every ncu run follows a plain run of the same
# command that exited 0; reports stay small (gpurun copies back at most 64 MiB).
cd $GRAFT_REPO_ROOT
mkdir -p gpurun_out
CMD="python bench.py --steps 3 --warmup 3 --no-cpu-baseline"
timeout 600 $CMD > gpurun_out/r2_cap_plain.json 2> gpurun_out/r2_cap_plain.err || { tail -5 gpurun_out/r2_cap_plain.err; exit 1; }
CMD1="python bench.py --steps 1 --warmup 1 --no-cpu-baseline"
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 1500 --csv --log-file gpurun_out/r2_launches.csv $CMD1 > gpurun_out/r2_ncu_launch.log 2>&1
timeout 1500 ncu --set full --clock-control none --import-source on -k regex:'sampler_build|sampler_flow|parents_sweep|parents_pred|coarse_rows_warp|prolong_rows|edge_length|triangle_kernel|product1_kernel|product2_kernel|compact_columns|transpose_fill|sort_rows|spmv_kernel|gs_colour' -c 16 -f -o gpurun_out/r2_full_c3 $CMD1 > gpurun_out/r2_ncu_full.log 2>&1
tail -3 gpurun_out/r2_ncu_full.log
ls -la gpurun_out/
