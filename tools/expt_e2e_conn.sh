# diagnosis: pinned e2e against the number of hardware work queues (CUDA_DEVICE_MAX_CONNECTIONS) and lanes
for cfg in "8 6" "32 6" "32 4" "32 8" "16 6"; do set -- $cfg
CUDA_DEVICE_MAX_CONNECTIONS=$1 B2G_LANES=$2 python bench.py --headline-only --steps 10 --warmup 3 > gpurun_out/e2e_conn.json 2> gpurun_out/e2e_conn.err
python - "$1" "$2" <<'PY'
import sys, json
for l in open('gpurun_out/e2e_conn.json'):
    if l.startswith('{'):
        d = json.loads(l); e = d['e2e']
        print('connections %s lanes %s: e2e %.1f GB/s, %.2f ms, link %.1f, frac %.3f, pageable %.1f' % (sys.argv[1], sys.argv[2], e['value'], e['ms_per_step'], e['link_gbs'], e['frac_of_link'], d['e2e_pageable']['value']))
PY
done
