mkdir -p gpurun_out
timeout 120 python bench.py --workload C1 --steps 20 --warmup 5 > gpurun_out/r2_zz_bench_c1.json 2>/dev/null
timeout 200 python bench.py --workload C2 --steps 10 --warmup 3 > gpurun_out/r2_zz_bench_c2.json 2>/dev/null
python tools/show_bench.py gpurun_out/r2_zz_bench_c1.json gpurun_out/r2_zz_bench_c2.json 2>/dev/null | grep -E "Mverts|cpu "
