set -x
mkdir -p gpurun_out
timeout 300 python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/r2e_smoke.log 2>&1; echo smoke rc=$?; tail -2 gpurun_out/r2e_smoke.log
timeout 600 python tools/prof_build.py C3 0.25 exact 1 3; timeout 300 python tools/prof_build.py C3 0.25 parallel 1 2; timeout 300 python tools/prof_build.py C2 1 exact 1 3
timeout 900 python -m pytest tests -m gpu -q -x --deselect tests/test_gpu_parity_fullsize.py --deselect tests/test_gpu_fullsize.py 2>&1 | tail -15 > gpurun_out/r2e_pytest_old.log; tail -5 gpurun_out/r2e_pytest_old.log
timeout 600 python bench.py --workload C3 --sampler exact --no-cpu-baseline --steps 5 --warmup 3 > gpurun_out/r2e_bench_C3_exact.json 2> gpurun_out/r2e_bench_C3_exact.err; echo rc=$?; tail -c 300 gpurun_out/r2e_bench_C3_exact.err
timeout 1200 python -m pytest tests/test_gpu_parity_fullsize.py -q -x -k "c2_full or c3_full" 2>&1 | tail -8
