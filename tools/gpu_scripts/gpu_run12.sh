mkdir -p gpurun_out
CUDA_LAUNCH_BLOCKING=1 timeout 600 python -m pytest tests/test_gpu_builders.py -x -q -m gpu 2>&1 | grep -E "Error|error|passed|failed|^tests|test_" | head -20
timeout 600 python -m pytest tests/test_gpu_builders.py -x -q -m gpu -k "bad or mesh" 2>&1 | grep -E "Error|error|passed|failed" | head
