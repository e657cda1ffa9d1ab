fmt='
import json,sys
for l in sys.stdin:
    if l.startswith("{"):
        d=json.loads(l); print(d["case"][:40], "ratio", d["ratio"], "cpu", d.get("cpu_ratio"), "comp", d.get("compress_gbs"), d.get("compress_ms"), "dec", d["decompress_gbs"])
    else: print(l.rstrip()[-200:])
'
for cs in "cfg1" "bitshuffle"; do
for t in 1 2 4 8; do for pr in 0 2048 8192; do
  if [ $t = 1 ] && [ $pr != 0 ]; then continue; fi
  echo "== $cs TEAM=$t PRIME=$pr"; B2G_CASE=$cs B2G_TEAM=$t B2G_TEAM_PRIME=$pr timeout 300 python tools/bench_configs.py 3 2>&1 | python -c "$fmt"
done; done; done
