(timeout 600 python -m pytest tests/test_b2frame.py tests/test_gpu_interleave.py -m gpu -x -q 2>&1 | tail -15)
