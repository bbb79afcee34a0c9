mkdir -p gpurun_out
timeout 200 python bench.py --sampler hashed --steps 5 --warmup 3 --no-cpu-baseline > gpurun_out/r2_zz_bench_c3_hashed.json 2>/dev/null
python tools/show_bench.py gpurun_out/r2_zz_bench_c3_hashed.json 2>/dev/null | head -1
