# encoder teams against one warp per stream: bench kernel numbers under the experiment switches
set -x
run() {
  echo "== $*"
  env "$@" timeout 600 python bench.py --steps 10 --warmup 3 --kernels-only 2>&1 | python -c "
import json,sys
for l in sys.stdin:
    if l.startswith('{'):
        d=json.loads(l); print('value',d['value'],'comp',d['compress_gbs_kernel'],'ratio',d['ratio'], 'k_compress ms', d['roofline']['kernels']['k_compress']['ms_per_launch'], 'k_decode ms', d['roofline']['kernels']['k_decode']['ms_per_launch'])
    else: print(l.rstrip()[-300:])
"
}
run B2G_TEAM=1
run B2G_TEAM=4 B2G_TEAM_PRIME=0
run B2G_TEAM=4 B2G_TEAM_PRIME=2048
run B2G_TEAM=4 B2G_TEAM_PRIME=4096
run B2G_TEAM=4 B2G_TEAM_PRIME=16384
run B2G_TEAM=2 B2G_TEAM_PRIME=8192
(timeout 900 python -m pytest tests/test_gpu_parity.py tests/test_gpu_configs.py -m gpu -x -q 2>&1 | tail -8)
