# quick GPU check: parity tests, a short bench (kernel numbers only); optional ncu --set full capture
#   usage: bash tools/gpu_quick.sh [ncu <kernel-regex> <launch-skip> <out-name>]
set -x
(timeout 1200 python -m pytest tests -m gpu -x -q 2>&1 | tail -15) > gpurun_out/q_pytest.log 2>&1
tail -3 gpurun_out/q_pytest.log
(timeout 600 python bench.py --steps 10 --warmup 3 --kernels-only) > gpurun_out/q_bench.log 2>&1
python - <<'PY'
import json
for f in ('gpurun_out/q_bench.log',):
  for l in open(f):
    if l.startswith('{'):
        d=json.loads(l); print(f,'value',d['value'],'dec',d['decompress_gbs_kernel'],'comp',d['compress_gbs_kernel'],'ratio',d['ratio'])
        for k,v in d['roofline']['kernels'].items(): print(k, v['ms_per_launch'], v['frac'])
    else: print(l.rstrip()[-300:])
PY
if [ "${1:-}" = "ncu" ]; then
timeout 900 ncu --set full --clock-control none --import-source on --kernel-name regex:"${2:-k_decode}" --launch-skip ${3:-2} --launch-count 1 -f -o gpurun_out/${4:-prof} python bench.py --steps 2 --warmup 1 --kernels-only > gpurun_out/ncu_full.log 2>&1
ls -la gpurun_out | tail -5
fi
