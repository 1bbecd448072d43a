set -x
nvidia-smi --query-gpu=name,memory.total --format=csv
timeout 900 python -m pytest tests/test_gpu_primitives.py -m gpu -x -q 2>&1 | tail -40
