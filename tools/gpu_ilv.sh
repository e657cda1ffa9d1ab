(timeout 900 python -m pytest tests -m gpu -x -q 2>&1 | tail -4)
timeout 300 python tools/bench_interleave.py > gpurun_out/ilv_bench.jsonl 2> gpurun_out/ilv_bench.err; tail -3 gpurun_out/ilv_bench.err
cat gpurun_out/ilv_bench.jsonl
timeout 600 ncu --set full --clock-control none -k regex:k_deinterleave_vec -c 1 -s 3 -o gpurun_out/ilv_vec python tools/bench_interleave.py > gpurun_out/ilv_ncu.log 2>&1; tail -2 gpurun_out/ilv_ncu.log
