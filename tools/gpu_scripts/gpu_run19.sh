mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_parity.py tests/test_capi_boundary.py -x -q 2>&1 | tail -4
timeout 900 python bench.py --steps 5 --warmup 3 --no-cpu-baseline > gpurun_out/r2_c_bench_c3_nocpu.json 2> gpurun_out/r2_c_bench_c3_nocpu.err; tail -3 gpurun_out/r2_c_bench_c3_nocpu.err; python tools/show_bench.py gpurun_out/r2_c_bench_c3_nocpu.json | head -4
timeout 900 python bench.py --workload C2 --steps 5 --warmup 3 --no-cpu-baseline > gpurun_out/r2_c_bench_c2_nocpu.json 2>/dev/null; python tools/show_bench.py gpurun_out/r2_c_bench_c2_nocpu.json | head -3
