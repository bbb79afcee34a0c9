mkdir -p gpurun_out
timeout 1200 python -m torch.distributed.run --nnodes=1 --nproc-per-node 4 --master-addr 127.0.0.1 --master-port 29514 bench.py --gpus 4 --steps 5 --warmup 3 > gpurun_out/r2_y_bench_sharded_n4.json 2> gpurun_out/r2_y_bench_sharded_n4.err; tail -c 400 gpurun_out/r2_y_bench_sharded_n4.err; python tools/show_bench.py gpurun_out/r2_y_bench_sharded_n4.json | head -3
