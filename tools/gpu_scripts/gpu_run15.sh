mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_parity.py tests/test_gpu_solver.py tests/test_gpu_sharded.py -x -q -m gpu 2>&1 | tail -3
GMG_TRACE=1 timeout 600 python tools/prof_build.py C3 1.0 exact 1 2 2>&1 | grep -E "large table|^C3" | tail -5
timeout 600 python tools/prof_build.py C2 1.0 exact 1 2 2>&1 | tail -1
timeout 1200 python -m pytest tests/test_gpu_parity_fullsize.py -x -q -m gpu -k "c2_full or c3_full" 2>&1 | tail -3
timeout 900 python bench.py --workload C4 --reorder morton --steps 3 --warmup 3 --no-cpu-baseline > gpurun_out/r2_bench_c4_morton.json 2> gpurun_out/r2_bench_c4_morton.err; python tools/show_bench.py gpurun_out/r2_bench_c4_morton.json 2>/dev/null | head -30; tail -3 gpurun_out/r2_bench_c4_morton.err
