# developer tooling: pageable e2e of the headline step against lanes (worker threads) and staged sub-batch size
for cfg in "6 12" "8 12" "8 6" "8 24" "6 24" "6 6"; do set -- $cfg
B2G_LANES=$1 B2G_STAGED_SUBBATCH_MB=$2 python bench.py --headline-only --steps 10 --warmup 3 > gpurun_out/e2e_pg.json 2> gpurun_out/e2e_pg.err
python - "$1" "$2" <<'PY'
import sys, json
for l in open('gpurun_out/e2e_pg.json'):
    if l.startswith('{'):
        d = json.loads(l); e = d['e2e']; p = d['e2e_pageable']
        print('lanes %s staged sub-batch %s MiB: pinned %.1f GB/s, pageable %.1f GB/s (%.2f ms), host memcpy %.1f' % (sys.argv[1], sys.argv[2], e['value'], p['value'], p['ms_per_step'], p['host_memcpy_gbs']))
PY
done
