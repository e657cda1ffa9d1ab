# bench.py under torchrun on N GPUs of one box: usage bash tools/gpu_ngpu.sh N
N=${1:-2}
timeout 1200 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29533 bench.py --gpus $N --steps 10 --warmup 3 > gpurun_out/bench_n$N.out 2> gpurun_out/bench_n$N.err
echo rc=$?
grep -h '^{' gpurun_out/bench_n$N.out > gpurun_out/bench_n$N.json
python - <<PY
import json
d=json.loads(open('gpurun_out/bench_n$N.json').readline())
print('N=$N value', d['value'], 'ms', d['ms_per_step'], 'e2e', d['e2e']['value'], 'frac_of_link', d['e2e'].get('frac_of_link'), 'strong', d.get('strong',{}).get('value'), 'gather ms', d.get('gather',{}).get('ms_max_over_ranks'))
PY
