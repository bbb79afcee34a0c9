set -x
mkdir -p gpurun_out
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm --format=csv
timeout 300 python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/r2a_smoke.log 2>&1; echo smoke rc=$?
timeout 1500 python -m pytest tests -m gpu -q -x --deselect tests/test_gpu_parity_fullsize.py 2>&1 | tail -15 > gpurun_out/r2a_pytest_old.log; cat gpurun_out/r2a_pytest_old.log | tail -5
timeout 2400 python -m pytest tests/test_gpu_parity_fullsize.py -q -s 2>&1 | tail -60 > gpurun_out/r2a_pytest_full.log; tail -30 gpurun_out/r2a_pytest_full.log
for cfg in "C3 exact" "C3 parallel" "C3 hashed" "C2 exact"; do set -- $cfg; timeout 600 python bench.py --workload $1 --sampler $2 --galerkin 1 --vcycle 1 --no-cpu-baseline --steps 5 --warmup 3 > gpurun_out/r2a_bench_$1_$2.json 2> gpurun_out/r2a_bench_$1_$2.err; echo bench $cfg rc=$?; done
