mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_cli_gpu.py tests/test_fdlp_gpu.py -x -q -m gpu 2>&1 | tail -3
