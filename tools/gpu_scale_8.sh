mkdir -p gpurun_out
timeout 1200 python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29518 bench.py --gpus 8 --steps 5 --warmup 3 > gpurun_out/r2_y_bench_sharded_n8.json 2> gpurun_out/r2_y_bench_sharded_n8.err; tail -c 400 gpurun_out/r2_y_bench_sharded_n8.err; python tools/show_bench.py gpurun_out/r2_y_bench_sharded_n8.json | head -3
