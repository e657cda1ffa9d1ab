# quick GPU check: parity tests, then a short bench (kernel numbers only)
set -x
(timeout 900 python -m pytest tests -m gpu -x -q 2>&1 | tail -15) > gpurun_out/q_pytest.log 2>&1
(timeout 600 python bench.py --steps 10 --warmup 3 --kernels-only) > gpurun_out/q_bench.log 2>&1
cat gpurun_out/q_pytest.log | tail -5
python - <<'PY'
import json
for l in open('gpurun_out/q_bench.log'):
    if l.startswith('{'):
        d=json.loads(l); print('value',d['value'],'dec',d['decompress_gbs_kernel'],'comp',d['compress_gbs_kernel'],'ratio',d['ratio'])
        for k,v in d['roofline']['kernels'].items(): print(k, v['ms_per_launch'], v['frac'])
    else: print(l.rstrip()[-300:])
PY
