# round-2 profile set: GPU tests, both bench arms, launch list + ncu --set full of the codec kernels on the bench workload,
# the zstd decode kernel on cfg 4 clevel 9 chunks, the blosclz kernels on cfg 3 (each ncu pass after a plain run of the
# same command has exited 0)
set -x
(timeout 1500 python -m pytest tests -m gpu -x -q 2>&1 | tail -4) > gpurun_out/p_pytest.log 2>&1; tail -2 gpurun_out/p_pytest.log
timeout 1200 python bench.py > gpurun_out/bench_full.json 2> gpurun_out/bench_full.err; tail -3 gpurun_out/bench_full.err
timeout 600 python bench.py --impl reference > gpurun_out/bench_ref.json 2> gpurun_out/bench_ref.err
timeout 300 python bench.py --steps 2 --warmup 1 --kernels-only > gpurun_out/p_plain.log 2>&1 || exit 1
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/r2_launches.csv python bench.py --steps 2 --warmup 1 --kernels-only > gpurun_out/p_launch.log 2>&1
timeout 900 ncu --set full --clock-control none --import-source on --kernel-name regex:"k_decode|k_compress" --launch-skip 2 --launch-count 2 -f -o gpurun_out/r2_full python bench.py --steps 2 --warmup 1 --kernels-only > gpurun_out/p_full.log 2>&1
timeout 300 python tools/expt_zstd9.py > gpurun_out/p_z.log 2>&1 && timeout 900 ncu --set full --clock-control none --import-source on --kernel-name k_decode --launch-skip 3 --launch-count 1 -f -o gpurun_out/r2_zstd9 python tools/expt_zstd9.py >> gpurun_out/p_z.log 2>&1
timeout 300 python tools/expt_cfg3.py > gpurun_out/p_c3.log 2>&1 && timeout 900 ncu --set full --clock-control none --import-source on --kernel-name regex:"k_decode|k_compress" --launch-skip 4 --launch-count 2 -f -o gpurun_out/r2_cfg3 python tools/expt_cfg3.py >> gpurun_out/p_c3.log 2>&1
ls -la gpurun_out | tail -14
