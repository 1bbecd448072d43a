timeout 1500 python -m pytest tests/test_gpu_primitives.py tests/test_gpu_scheme.py -m gpu -q 2>&1 | tail -60
