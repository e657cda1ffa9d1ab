# round-2 baseline: GPU parity tests, the bench line, the other configs, launch list, ncu --set full of both codec kernels
set -x
nvidia-smi -L; nproc; free -g | head -2
(timeout 1500 python -m pytest tests -m gpu -x -q 2>&1 | tail -15) > gpurun_out/b_pytest.log 2>&1
tail -3 gpurun_out/b_pytest.log
(timeout 600 python bench.py) > gpurun_out/b_bench.json 2> gpurun_out/b_bench.err
tail -c 1500 gpurun_out/b_bench.json
(timeout 900 python tools/bench_configs.py 5) > gpurun_out/b_configs.jsonl 2> gpurun_out/b_configs.err
cat gpurun_out/b_configs.jsonl | cut -c1-400
timeout 900 ncu --set full --clock-control none --import-source on --kernel-name regex:"k_decode|k_compress" --launch-skip 2 --launch-count 2 -f -o gpurun_out/r2a_full python bench.py --steps 2 --warmup 1 --kernels-only > gpurun_out/ncu_full.log 2>&1
ls -la gpurun_out | tail
