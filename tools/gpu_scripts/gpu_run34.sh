mkdir -p gpurun_out
timeout 420 python -m pytest tests -x -q -m gpu 2>&1 | tail -3 | tee gpurun_out/r2_zz_pytest_gpu.log
timeout 300 python bench.py --steps 5 --warmup 3 > gpurun_out/r2_zz_bench_c3_default.json 2> gpurun_out/r2_zz_bench_c3_default.err; tail -2 gpurun_out/r2_zz_bench_c3_default.err
python tools/show_bench.py gpurun_out/r2_zz_bench_c3_default.json 2>/dev/null | head -2
