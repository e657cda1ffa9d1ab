(timeout 900 python -m pytest tests -m gpu -x -q 2>&1 | tail -2)
B2G_TIMING=1 timeout 600 python bench.py --steps 6 --warmup 2 --kernels-only > gpurun_out/t1.log 2>&1
python - <<'PY'
import json
d=json.loads([l for l in open('gpurun_out/t1.log') if l.startswith('{')][0])
print(d['value'], d['ms_per_step']); 
for k,v in d['roofline']['kernels'].items(): print(k, v['ms_per_launch'])
PY
