# developer tooling: pinned e2e of the headline step against sub-batch size / lanes
for cfg in "32 6" "16 6" "8 6" "16 8" "64 6" "32 3" "24 8"; do set -- $cfg
B2G_SUBBATCH_MB=$1 B2G_LANES=$2 python bench.py --headline-only --steps 10 --warmup 3 | python -c "
import sys,json; d=json.loads(sys.stdin.read()); print('sub $1 MiB lanes $2: e2e', round(d['e2e']['value'],1), 'ms', round(d['e2e']['ms_per_step'],2), 'link', round(d['e2e']['link_gbs'],1), 'frac', round(d['e2e']['frac_of_link'],3), 'pageable', round(d['e2e_pageable']['value'],1))"
done
