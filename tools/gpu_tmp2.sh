for L in 3 6; do for S in 16 32 64; do
B2G_LANES=$L B2G_SUBBATCH_MB=$S timeout 600 python bench.py --steps 5 --warmup 3 2>/dev/null | python -c "
import json,sys
d=json.loads([l for l in sys.stdin if l.startswith('{')][0]); print('lanes $L sub $S e2e', round(d['e2e']['value'],1), round(d['e2e']['ms_per_step'],2), 'value', round(d['value'],1))"
done; done
