#!/bin/bash
# Round-2 evidence for profiles/ (run on the GPU box through gpurun). Every ncu run follows a plain run of the same
# command that exited 0; reports stay small (gpurun copies back at most 64 MiB).
cd $GRAFT_REPO_ROOT
mkdir -p gpurun_out
CMD="python bench.py --steps 5 --warmup 3"
timeout 900 $CMD > gpurun_out/r2_z_bench_c3_default.json 2> gpurun_out/r2_z_bench_c3_default.err || { tail -5 gpurun_out/r2_z_bench_c3_default.err; exit 1; }
CMD1="python bench.py --steps 1 --warmup 1 --no-cpu-baseline"
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 1500 --csv --log-file gpurun_out/r2_z_launches.csv $CMD1 > gpurun_out/r2_z_ncu_launch.log 2>&1
timeout 1500 ncu --set full --clock-control none --import-source on -k regex:'sampler_coherence|sampler_build|sampler_flow|parents_sweep|parents_pred|coarse_rows_warp|prolong_rows|prolong_fallback|edge_length|triangle_kernel|product1_kernel|product2_kernel|compact_columns|transpose_fill|sort_rows' -c 17 -f -o gpurun_out/r2_z_full_c3 $CMD1 > gpurun_out/r2_z_ncu_full.log 2>&1
tail -2 gpurun_out/r2_z_ncu_full.log
# the row-sharded kernels, world 1 on the 12.5 M-vertex torus (the N = 1 point of the scaling family)
CMD2="python bench.py --workload T12.5 --steps 1 --warmup 1 --no-cpu-baseline --vcycle 0"
timeout 600 $CMD2 > gpurun_out/r2_z_bench_T12.5_n1.json 2> gpurun_out/r2_z_bench_T12.5_n1.err || { tail -5 gpurun_out/r2_z_bench_T12.5_n1.err; exit 2; }
timeout 1500 ncu --set full --clock-control none --import-source on -k regex:'shard_' -c 14 -f -o gpurun_out/r2_z_full_shard $CMD2 > gpurun_out/r2_z_ncu_shard.log 2>&1
tail -2 gpurun_out/r2_z_ncu_shard.log
ls -la gpurun_out/ | head -30
