(timeout 900 python -m pytest tests -m gpu -x -q 2>&1 | tail -3)
(timeout 300 python -c "import __graft_entry__ as g; g.smoke()" 2>&1 | tail -1)
(timeout 300 python bench.py --steps 10 --warmup 3 2>/dev/null | tail -1 | cut -c1-200)
