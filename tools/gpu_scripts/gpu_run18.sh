mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_options.py tests/test_facade.py tests/test_gpu_builders.py -x -q -m gpu 2>&1 | tail -12
timeout 1500 python -m pytest tests -x -q -m gpu --deselect tests/test_gpu_parity_fullsize.py 2>&1 | tail -4
