(timeout 900 python -m pytest tests -m gpu -x -q 2>&1 | tail -2)
B2G_TIMING=1 timeout 600 python bench.py --steps 6 --warmup 2 --kernels-only > gpurun_out/t1.log 2>&1
grep "b2g timing" gpurun_out/t1.log | sed -n 3,4p
python - <<'PY'
import json
d=json.loads([l for l in open('gpurun_out/t1.log') if l.startswith('{')][0])
print(d['value'], d['ms_per_step']); 
for k,v in d['roofline']['kernels'].items(): print(k, v['ms_per_launch'])
PY
B2G_PAIR_PLANES=0 timeout 600 python bench.py --steps 6 --warmup 2 --kernels-only 2>&1 | python -c "
import json,sys
d=json.loads([l for l in sys.stdin if l.startswith('{')][0]); print('nopair', d['roofline']['kernels']['k_decode']['ms_per_launch'])"
