N=$1
mkdir -p gpurun_out
if [ "$N" = "1" ]; then
timeout 1200 python bench.py --workload C4 --sharded --steps 3 --warmup 2 > gpurun_out/r2_z_bench_c4_sharded_n1.json 2> gpurun_out/r2_z_bench_c4_sharded_n1.err
else
timeout 1200 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 2953$N bench.py --gpus $N --workload C4 --steps 3 --warmup 2 > gpurun_out/r2_z_bench_c4_sharded_n$N.json 2> gpurun_out/r2_z_bench_c4_sharded_n$N.err
fi
tail -c 600 gpurun_out/r2_z_bench_c4_sharded_n$N.err; python tools/show_bench.py gpurun_out/r2_z_bench_c4_sharded_n$N.json 2>/dev/null | head -2
