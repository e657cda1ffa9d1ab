timeout 300 python tools/bench_iterator.py 2>&1 | tail -3
(timeout 600 python -m pytest tests/test_gpu_api.py -m gpu -x -q 2>&1 | tail -3)
