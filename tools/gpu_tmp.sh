set -x
(timeout 900 python -m pytest tests -m gpu -x -q 2>&1 | tail -3)
B2G_TIMING=1 timeout 600 python bench.py --steps 4 --warmup 2 --kernels-only 2>&1 | grep -E "b2g timing|value" | cut -c1-200 | tail -5
