# pinned e2e of the headline step, a few runs in a row (run-to-run spread)
for i in 1 2 3; do
python bench.py --headline-only --steps 10 --warmup 3 > gpurun_out/e2e_rep.json 2> gpurun_out/e2e_rep.err
python - <<'PY'
import json
for l in open('gpurun_out/e2e_rep.json'):
    if l.startswith('{'):
        d = json.loads(l); e = d['e2e']
        print('e2e %.1f GB/s %.2f ms link %.1f frac %.3f pageable %.1f value %.1f' % (e['value'], e['ms_per_step'], e['link_gbs'], e['frac_of_link'], d['e2e_pageable']['value'], d['value']))
PY
done
