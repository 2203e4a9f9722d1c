n=${NGPU:-8}
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port 29541 bench.py --gpus $n --steps 5 --warmup 3 > gpurun_out/bench_n$n.json 2> gpurun_out/bench_n$n.err; tail -c 300 gpurun_out/bench_n$n.err; python3 -c "
import json
d=json.loads(open('gpurun_out/bench_n$n.json').read().strip().splitlines()[-1])
print({k:d[k] for k in ('value','ms_per_step','scaling','n_gpus')}); print('e2e', d['e2e']['value'] if d['e2e'] else None, (d['e2e'].get('side_format') or {}).get('value') if d['e2e'] else None); print(d['weak_replicas']); print(d['strong_check']); print(d['ksharded']); print(d['breakdown_ms'])"
