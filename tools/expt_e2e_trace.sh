B2G_TRACE=1 python bench.py --headline-only --steps 3 --warmup 3 > gpurun_out/e2e_trace.json 2> gpurun_out/e2e_trace.err
grep b2g-trace gpurun_out/e2e_trace.err | tail -120 > gpurun_out/e2e_trace.txt
wc -l gpurun_out/e2e_trace.txt
