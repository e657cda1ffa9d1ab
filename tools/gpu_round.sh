set -x
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm,power.draw --format=csv
(timeout 900 python -m pytest tests -m gpu -x -q 2>&1 | tail -25) > gpurun_out/r1_pytest.log 2>&1
(timeout 300 python -c "import __graft_entry__ as g; g.smoke()" ) > gpurun_out/r1_smoke.log 2>&1
(timeout 600 python bench.py ) > gpurun_out/r1_bench.log 2>&1
(timeout 600 python bench.py --impl reference --steps 3 --warmup 1) > gpurun_out/r1_bench_ref.log 2>&1
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/r1_launches.csv python bench.py --steps 2 --warmup 1 --kernels-only > gpurun_out/r1_ncu.log 2>&1
for f in gpurun_out/r1_pytest.log gpurun_out/r1_smoke.log; do tail -3 $f; done
