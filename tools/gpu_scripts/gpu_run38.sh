mkdir -p gpurun_out
timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29541 tools/dist_parity_check.py C3 0.05 2>&1 | grep -E "rank 0|DIST|Error|error" | tail -12 | tee gpurun_out/r2_zz_dist_parity_2gpu_c3.log
