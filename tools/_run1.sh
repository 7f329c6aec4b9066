python -m pytest tests/test_round2_gpu.py tests/test_model_gpu.py tests/test_kernels_gpu.py -m gpu -x -q 2>&1 | tail -3
