# GPU tests, a short bench, and the DRAM traffic of one k_compress launch
(timeout 1500 python -m pytest tests -m gpu -x -q 2>&1 | tail -3)
timeout 600 python bench.py --steps 10 --warmup 3 --kernels-only > gpurun_out/q_bench.log 2>&1
python - <<'PY'
import json
for l in open('gpurun_out/q_bench.log'):
    if l.startswith('{'):
        d=json.loads(l); print('value',d['value'],'ratio',d['ratio'], {k:(v['ms_per_launch']) for k,v in d['roofline']['kernels'].items()})
PY
timeout 600 ncu --metrics gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum --clock-control none --kernel-name regex:"k_compress" --launch-skip 2 --launch-count 1 python bench.py --steps 2 --warmup 1 --kernels-only 2>&1 | grep "dram__\|gpu__time"
