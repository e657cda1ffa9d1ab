# kernel-scope numbers of every config (tools/bench_configs.py) in one table
timeout 900 python tools/bench_configs.py 3 2>&1 | python -c '
import json,sys
for l in sys.stdin:
    if l.startswith("{"):
        d=json.loads(l); print(d["case"][:44], "ratio", d["ratio"], "cpu", d.get("cpu_ratio"), "comp", d.get("compress_gbs"), d.get("compress_ms"), "dec", d["decompress_gbs"], d["decompress_ms"])
    else: print(l.rstrip()[-200:])
'
