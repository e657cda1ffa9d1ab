(timeout 600 python -m pytest tests/test_gpu_interleave.py -x -q 2>&1 | tail -5)
timeout 300 python tools/bench_interleave.py > gpurun_out/ilv_bench.jsonl 2> gpurun_out/ilv_bench.err; tail -3 gpurun_out/ilv_bench.err
cat gpurun_out/ilv_bench.jsonl
cat MEASURED_PEAKS.json 2>/dev/null | head -20
