# diagnosis: the e2e step's two directions alone (B2G_E2E_ONLY), next to the combined step
for o in "" dec comp; do
B2G_E2E_ONLY=$o python bench.py --headline-only --steps 10 --warmup 3 > gpurun_out/e2e_only_$o.json 2> gpurun_out/e2e_only_$o.err
python - "$o" <<'PY'
import sys, json
o = sys.argv[1]
for l in open('gpurun_out/e2e_only_%s.json' % o):
    if l.startswith('{'):
        d = json.loads(l); e = d['e2e']
        print('only=%s e2e ms %.2f link %.1f pageable ms %.2f' % (o, e['ms_per_step'], e['link_gbs'], d['e2e_pageable']['ms_per_step']))
PY
done
