nvidia-smi topo -m > gpurun_out/topo.txt 2>&1
lscpu | grep -E "NUMA|Model name|Socket|^CPU\(s\)" > gpurun_out/lscpu.txt 2>&1; free -g >> gpurun_out/lscpu.txt
for n in 1 2 4 8; do timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port 29531 tools/h2d_scaling_probe.py 2>&1 | grep -E "^N=|^rank 0" ; done > gpurun_out/h2d_scaling.log 2>&1
cat gpurun_out/h2d_scaling.log
for n in 8 4; do timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port 29541 bench.py --gpus $n --steps 5 --warmup 3 > gpurun_out/bench_n$n.json 2> gpurun_out/bench_n$n.err; tail -c 600 gpurun_out/bench_n$n.err; python3 -c "
import json
d=json.loads(open('gpurun_out/bench_n$n.json').read().strip().splitlines()[-1])
print({k:d[k] for k in ('value','ms_per_step','scaling','n_gpus')}); print('e2e', d['e2e']['value'] if d['e2e'] else None, d['e2e'].get('ms_per_step') if d['e2e'] else None, (d['e2e'].get('side_format') or {}).get('value') if d['e2e'] else None); print(d['weak_replicas']); print(d['strong_check']); print(d['ksharded']); print(d['breakdown_ms'])"; done
