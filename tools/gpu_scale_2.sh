mkdir -p gpurun_out
timeout 1200 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29512 bench.py --gpus 2 --steps 5 --warmup 3 > gpurun_out/r2_y_bench_sharded_n2.json 2> gpurun_out/r2_y_bench_sharded_n2.err; tail -c 400 gpurun_out/r2_y_bench_sharded_n2.err; python tools/show_bench.py gpurun_out/r2_y_bench_sharded_n2.json | head -3
