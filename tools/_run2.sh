python -m pytest tests/test_dp_gpu.py -m gpu -q -k "2 or trainer" 2>&1 | tail -4
