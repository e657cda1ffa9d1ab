# quick GPU check: parity tests, then a short bench (kernel numbers only), two-phase decode on and off
set -x
(timeout 900 python -m pytest tests -m gpu -x -q 2>&1 | tail -15) > gpurun_out/q_pytest.log 2>&1
(timeout 600 python bench.py --steps 10 --warmup 3 --kernels-only) > gpurun_out/q_bench.log 2>&1
(B2G_TWO_PHASE=0 timeout 600 python bench.py --steps 10 --warmup 3 --kernels-only) > gpurun_out/q_bench_onepass.log 2>&1
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 60 --csv --log-file gpurun_out/q_launches.csv python bench.py --steps 2 --warmup 1 --kernels-only > gpurun_out/q_ncu.log 2>&1
cat gpurun_out/q_pytest.log | tail -5
python - <<'PY'
import json
for f in ('gpurun_out/q_bench.log','gpurun_out/q_bench_onepass.log'):
  for l in open(f):
    if l.startswith('{'):
        d=json.loads(l); print(f,'value',d['value'],'dec',d['decompress_gbs_kernel'],'comp',d['compress_gbs_kernel'],'ratio',d['ratio'])
        for k,v in d['roofline']['kernels'].items(): print(k, v['ms_per_launch'], v['frac'])
    else: print(l.rstrip()[-300:])
PY
grep -E "k_parse|k_decode|k_plan" gpurun_out/q_launches.csv | tail -8
