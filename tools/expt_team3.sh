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
run B2G_TEAM=2 B2G_TEAM_PRIME=8192 B2G_TEAM_SPARSE=1
run B2G_TEAM=2 B2G_TEAM_PRIME=4096 B2G_TEAM_SPARSE=1
run B2G_TEAM=4 B2G_TEAM_PRIME=12288 B2G_TEAM_SPARSE=1
run B2G_TEAM=4 B2G_TEAM_PRIME=4096 B2G_TEAM_SPARSE=1
fmt='
import json,sys
for l in sys.stdin:
    if l.startswith("{"):
        d=json.loads(l); print(d["case"][:40], "ratio", d["ratio"], "cpu", d.get("cpu_ratio"), "comp", d.get("compress_gbs"), d.get("compress_ms"), "dec", d["decompress_gbs"])
    else: print(l.rstrip()[-200:])
'
for cs in "cfg1" "bitshuffle"; do
for sp in 0 1; do
  echo "== $cs default team SPARSE=$sp"; B2G_CASE=$cs B2G_TEAM_SPARSE=$sp timeout 300 python tools/bench_configs.py 3 2>&1 | python -c "$fmt"
done; done
