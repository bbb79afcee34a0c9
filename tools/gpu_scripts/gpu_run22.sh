cd $GRAFT_REPO_ROOT
mkdir -p gpurun_out
timeout 1800 python -m pytest tests -q -m gpu 2>&1 | tail -3 > gpurun_out/r2_z_pytest_gpu.log; cat gpurun_out/r2_z_pytest_gpu.log
timeout 1500 python bench.py --steps 5 --warmup 3 > gpurun_out/r2_z_bench_c3_default.json 2> gpurun_out/r2_z_bench_c3_default.err; tail -2 gpurun_out/r2_z_bench_c3_default.err
timeout 900 python bench.py --impl reference --steps 3 --warmup 1 > gpurun_out/r2_z_bench_reference_arm.json 2> gpurun_out/r2_z_bench_reference_arm.err; tail -2 gpurun_out/r2_z_bench_reference_arm.err
timeout 600 python bench.py --workload C2 --steps 5 --warmup 3 > gpurun_out/r2_z_bench_c2.json 2>/dev/null
timeout 900 python bench.py --workload C4 --reorder morton --steps 3 --warmup 3 --no-cpu-baseline > gpurun_out/r2_z_bench_c4_morton.json 2>/dev/null
timeout 900 python bench.py --workload C4 --steps 3 --warmup 3 --no-cpu-baseline > gpurun_out/r2_z_bench_c4.json 2>/dev/null
timeout 900 python bench.py --sampler parallel --steps 5 --warmup 3 --no-cpu-baseline > gpurun_out/r2_z_bench_c3_parallel.json 2>/dev/null
python tools/show_bench.py gpurun_out/r2_z_bench_*.json 2>&1 | grep -E "Mverts|cpu " 
