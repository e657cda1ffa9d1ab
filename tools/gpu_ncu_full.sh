# one ncu --set full capture of the decode and compress kernels of the bench workload (after a plain run exited 0)
set -x
timeout 300 python bench.py --steps 2 --warmup 1 --kernels-only > gpurun_out/ncu_plain.log 2>&1 || exit 1
timeout 900 ncu --set full --clock-control none --import-source on --kernel-name regex:"${1:-k_decode|k_compress}" --launch-skip ${2:-2} --launch-count ${3:-2} -f -o gpurun_out/${4:-prof} python bench.py --steps 2 --warmup 1 --kernels-only > gpurun_out/ncu_full.log 2>&1
ls -la gpurun_out
