set -x
mkdir -p gpurun_out
timeout 300 python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/r2b_smoke.log 2>&1; echo smoke rc=$?; tail -2 gpurun_out/r2b_smoke.log
timeout 900 python -m pytest tests -m gpu -q -x --deselect tests/test_gpu_parity_fullsize.py --deselect tests/test_gpu_fullsize.py 2>&1 | tail -15 > gpurun_out/r2b_pytest_old.log; tail -5 gpurun_out/r2b_pytest_old.log
for cfg in "C3 exact" "C3 hashed" "C2 exact"; do set -- $cfg; timeout 600 python bench.py --workload $1 --sampler $2 --no-cpu-baseline --steps 5 --warmup 3 > gpurun_out/r2b_bench_$1_$2.json 2> gpurun_out/r2b_bench_$1_$2.err; echo bench $cfg rc=$?; tail -c 600 gpurun_out/r2b_bench_$1_$2.err; done
timeout 2400 python -m pytest tests/test_gpu_parity_fullsize.py -q -s -k "c2_full or c3_full or statistics" 2>&1 | tail -60 > gpurun_out/r2b_pytest_full.log; grep -E "count|passed|failed|Error" gpurun_out/r2b_pytest_full.log | tail -20
